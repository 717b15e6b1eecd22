"""The C-ABI library loads (no GPU needed) and exports every symbol include/*.h declares;
without a usable B200 every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    syms = set()
    for h in ("spartan_b200.h", "spartan_b200_kernels.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        syms |= set(re.findall(r"\b(sb200_\w+)\s*\(", text))
    return sorted(syms)


def test_library_exports_every_declared_symbol():
    import otti_b200
    assert os.path.exists(otti_b200.LIB_PATH), "run __graft_entry__.build() first"
    lib = C.CDLL(otti_b200.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 30
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_headers_cite_the_reference():
    text = open(os.path.join(ROOT, "include", "spartan_b200.h")).read()
    for needle in ("lib.rs:121-225", "lib.rs:493-535", "lib.rs:538-574", "lib.rs:466-477", "lib.rs:228-257"):
        assert needle in text


def test_no_cpu_fallback_without_gpu():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    import otti_b200 as sb
    with pytest.raises(sb.BackendError):
        sb.NIZKGens.new(16, 16, 3)
    with pytest.raises(sb.BackendError):
        sb.Instance.produce_synthetic_r1cs(16, 16, 3)
    # host-only pieces still work: the transcript object and the scalar canonicity check
    t = sb.Transcript.new(b"test protocol")
    t.append_message(b"some label", b"some data")
    assert t.challenge_bytes(b"challenge", 32).hex() == "d5a21972d0d5fe320c0d263fac7fffb8145aa640af6e9bca177c03c7efcf0615"
    with pytest.raises(sb.R1CSError):
        sb.Assignment.new([b"\xff" * 32])


def test_product_does_not_touch_the_oracle():
    """Nothing under otti_b200/ or include/ may include, link or import oracle/ code."""
    bad = []
    for base in ("otti_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "_obj" in dp or "__pycache__" in dp:
                continue
            for f in files:
                if f.endswith((".so", ".o", ".pyc")):
                    continue
                text = open(os.path.join(dp, f), errors="replace").read()
                if re.search(r'#include\s+"[^"]*oracle|import\s+oracle|from\s+oracle|oracle_lib|liboracle', text):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad
    out = os.popen("ldd %s" % os.path.join(ROOT, "otti_b200", "libspartan_b200.so")).read()
    assert "oracle" not in out


def test_bench_reference_arm_prints_one_json_line():
    """bench.py's stdout is the one JSON line of the contract, whatever libraries print (the reference arm runs on CPU)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--log-size", "14"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.splitlines()
    assert len(lines) == 1 and lines[0].startswith("{")
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "constraints/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["e2e"]["h2d_bytes_per_step"] == 0
    assert line["config"]["workload"] == "synthetic Spartan R1CS 2^14 constraints = vars, 10 inputs, NIZK::prove"  # == the GPU arm's string


def test_keccak_variants_agree():
    """host/hmerlin.hpp picks the portable Keccak-f by default and the AVX-512 one with SB200_KECCAK=avx512 (when the CPU has
    it): both must give the Merlin known-answer vector and the same challenges over a long transcript."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); import otti_b200 as sb; t = sb.Transcript.new(b'test protocol'); "
            "t.append_message(b'some label', b'some data'); out = t.challenge_bytes(b'challenge', 32).hex()\n"
            "for i in range(300):\n"
            "    t.append_message(b'l%%d' %% i, bytes(range(i %% 200)))\n"
            "    out += t.challenge_bytes(b'c', 64 + i %% 7).hex()\n"
            "import hashlib; print(out[:64], hashlib.sha256(out.encode()).hexdigest())") % ROOT
    outs = []
    for mode in ("scalar", "avx512"):
        env = dict(os.environ, SB200_KECCAK=mode)
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        outs.append(r.stdout.split())
    assert outs[0][0] == "d5a21972d0d5fe320c0d263fac7fffb8145aa640af6e9bca177c03c7efcf0615"
    assert outs[0] == outs[1]


def test_host_inversion_matches_fermat_and_python():
    """host/hfl.hpp fl_invert (binary extended Euclid with batched halvings) against a^(l-2) and Python's pow(x, -1, l): the
    bullet rounds invert their challenge with it (nizk/bullet.rs:100-101, ristretto255.rs:548-602)."""
    import ctypes as C
    import random
    import otti_b200
    L = otti_b200.lib()
    ell = 2 ** 252 + 27742317777372353535851937790883648493
    rng = random.Random(99)
    xs = [0, 1, 2, 3, ell - 1, ell - 2, (ell + 1) // 2, 2 ** 64, 2 ** 128, 2 ** 192, 2 ** 252, 2 ** 252 - 1, 2 ** 63, 2 ** 126 + 1]
    xs += [1 << k for k in range(0, 252, 7)] + [(1 << k) - 1 for k in range(2, 252, 11)]
    xs += [rng.randrange(ell) for _ in range(3000)] + [rng.randrange(1 << rng.randrange(1, 252)) for _ in range(500)]
    out = (C.c_uint8 * 32)()
    for x in xs:
        b = (C.c_uint8 * 32).from_buffer_copy(x.to_bytes(32, "little"))
        want = pow(x, -1, ell) if x else 0
        for fermat in (0, 1):
            assert L.sb200_k_host_invert(b, out, fermat) == 0
            assert int.from_bytes(bytes(out), "little") == want, (x, fermat)
    bad = (C.c_uint8 * 32).from_buffer_copy(ell.to_bytes(32, "little"))
    assert L.sb200_k_host_invert(bad, out, 0) == -5
