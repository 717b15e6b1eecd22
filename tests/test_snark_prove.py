"""SNARK::prove / SNARK::verify (lib.rs:335-457) -- SURVEY.md section 8(f) N3, second half.

CPU part: the C++ oracle (oracle/snark.hpp) against the proof lengths the reference publishes (README.md:305-345),
prove -> verify round trips and rejection of tampered proofs.
GPU part: the CUDA path through the C ABI, byte for byte against the oracle's proof on the same seeded inputs; each side's
verifier on the other side's proof.  Bar: bit-exact.
"""
import pytest

import oracle_lib as O
from common import TAPE_SEED, ragged_r1cs, otti_shaped_r1cs, coo_of

LABEL = b"snark_example"  # Spartan/examples + README.md SNARK example


def ilog2(n):
    return n.bit_length() - 1


def next_pow2(n):
    p = 1
    while p < n:
        p *= 2
    return p


# ---------------------------------------------------------------------------------- closed-form bincode lengths
def len_dotlog(lg_n):
    return 2 * (8 + 32 * lg_n) + 4 * 32


def len_zk_sumcheck(rounds, degree):
    return 2 * (8 + 32 * rounds) + 8 + rounds * (32 + 32 + 8 + 32 * (degree + 1) + 32 + 32)


def len_r1cs_sat_proof(num_cons, num_vars):
    s_x, s_y, ell = ilog2(num_cons), ilog2(2 * num_vars), ilog2(num_vars)
    return (8 + 32 * (1 << (ell // 2)) + len_zk_sumcheck(s_x, 3) + 4 * 32 + 96 + (96 + 160) + 64 + len_zk_sumcheck(s_y, 2) + 32 +
            len_dotlog(ell - ell // 2) + 64)


def len_batched(ninst, length, ndotp):
    layers = ilog2(length)
    rounds = layers * (layers - 1) // 2
    return 8 + layers * (8 + 2 * (8 + 32 * ninst)) + rounds * (8 + 96) + 3 * (8 + 32 * ndotp)


def len_product_layer_proof(N, cells):
    return 2 * (32 + 2 * (8 + 96) + 32) + 2 * (8 + 96) + len_batched(4, cells, 0) + len_batched(12, N, 6)


def len_r1cs_eval_proof(num_cons, num_vars, nz):
    N = next_pow2(nz)
    cells = 1 << max(ilog2(num_cons), ilog2(2 * num_vars))
    nv_ops, nv_mem, nv_derefs = ilog2(N) + 4, ilog2(cells) + 1, ilog2(N) + 3
    derefs = 8 + 32 * (1 << (nv_derefs // 2))
    hash_layer = 4 * (8 + 96) + 2 * 32 + 3 * (8 + 96) + sum(len_dotlog(nv - nv // 2) for nv in (nv_ops, nv_mem, nv_derefs))
    return derefs + len_product_layer_proof(N, cells) + hash_layer


def test_closed_form_lengths_match_the_reference_profile():
    """README.md:305-345 (2^20 constraints, variables and non-zero entries): len_r1cs_sat_proof 47024,
    len_product_layer_proof 64712, len_r1cs_eval_proof 133720."""
    n = 1 << 20
    assert len_r1cs_sat_proof(n, n) == 47024
    assert len_product_layer_proof(n, 2 * n) == 64712
    assert len_r1cs_eval_proof(n, n, n) == 133720


def _golden():
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "snark_prove_golden.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("case", _golden()["cases"], ids=lambda c: "%dx%d" % (c["num_cons"], c["num_vars"]))
def test_oracle_snark_matches_committed_digests(case):
    """tests/golden/snark_prove_golden.json (made by tests/golden/make_snark_prove_golden.py): the oracle's proof bytes for
    fixed seeds, so that the checker itself cannot drift unnoticed."""
    import hashlib
    inst, v, i = O.Instance.synthetic(case["num_cons"], case["num_vars"], case["num_inputs"], case["seed"])
    assert max(inst.nnz) == case["num_nz_entries"]
    pf = O.snark_prove(inst, case["num_nz_entries"], v, i, LABEL, TAPE_SEED)
    assert len(pf) == case["proof_len"] and pf[:64].hex() == case["proof_head"] and pf[-64:].hex() == case["proof_tail"]
    assert hashlib.sha256(pf).hexdigest() == case["proof_sha256"]
    assert hashlib.sha256(O.snark_encode(inst, case["num_nz_entries"])).hexdigest() == case["commitment_sha256"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", _golden()["cases"], ids=lambda c: "%dx%d" % (c["num_cons"], c["num_vars"]))
def test_gpu_snark_matches_committed_digests(case):
    """The CUDA path against the committed digests directly (no oracle call in the loop)."""
    import hashlib
    import otti_b200 as sb
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(case["num_cons"], case["num_vars"], case["num_inputs"], case["seed"])
    gens = sb.SNARKGens.new(case["num_cons"], case["num_vars"], case["num_inputs"], case["num_nz_entries"])
    comm = sb.SNARK.encode(inst, gens)
    assert hashlib.sha256(comm.to_bytes()).hexdigest() == case["commitment_sha256"]
    pf = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED).to_bytes()
    assert len(pf) == case["proof_len"] and hashlib.sha256(pf).hexdigest() == case["proof_sha256"]


def _py_golden():
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "snark_prove_py_golden.json")) as f:
        return json.load(f)["cases"]


def _py_case_id(c):
    return "%s%dx%d" % (c["kind"][:3], c["num_cons"], c["num_vars"])


def _py_case_inputs(case, make_instance, synthetic):
    """(instance, vars bytes, inputs bytes) of a golden case for the oracle or the product"""
    if case["kind"] == "synthetic":
        return synthetic(case["num_cons"], case["num_vars"], case["num_inputs"], case["seed"])
    mats = [[(r, c, bytes.fromhex(v)) for r, c, v in M] for M in case["coo"]]
    inst = make_instance(case["num_cons"], case["num_vars"], case["num_inputs"], mats[0], mats[1], mats[2])
    return inst, bytes.fromhex(case["vars"]), bytes.fromhex(case["inputs"])


@pytest.mark.parametrize("case", _py_golden(), ids=_py_case_id)
def test_oracle_snark_matches_python_restatement(case):
    """tests/golden/snark_prove_py_golden.json holds whole proofs made by oracle/pyspec/snark_prove_spec.py, a second
    restatement written from the Rust sources (big-int Python, small sizes): the C++ oracle must give the same bytes --
    synthetic instances with num_cons != num_vars, and shuffled COO instances with unequal matrices."""
    inst, v, i = _py_case_inputs(case, O.Instance.new, O.Instance.synthetic)
    pf = O.snark_prove(inst, case["num_nz_entries"], v, i, LABEL, TAPE_SEED)
    assert pf.hex() == case["proof"]
    assert O.snark_verify(O.snark_encode(inst, case["num_nz_entries"]), case["num_nz_entries"], i, LABEL, pf) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("case", _py_golden(), ids=_py_case_id)
def test_gpu_snark_matches_python_restatement(case):
    import otti_b200 as sb

    def synthetic(nc, nv, ni, seed):
        inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(nc, nv, ni, seed)
        return inst, bytes(vars_.data)[:32 * nv], bytes(inputs.data)[:32 * ni]

    inst, v, i = _py_case_inputs(case, sb.Instance.new, synthetic)
    gens = sb.SNARKGens.new(case["num_cons"], case["num_vars"], case["num_inputs"], case["num_nz_entries"])
    comm = sb.SNARK.encode(inst, gens)
    pf = sb.SNARK.prove(inst, comm, None, sb.Assignment.new(v), sb.Assignment.new(i), gens, sb.Transcript.new(LABEL), TAPE_SEED)
    assert pf.to_bytes().hex() == case["proof"]
    pf.verify(comm, sb.Assignment.new(i), sb.Transcript.new(LABEL), gens)


SIZES = [(8, 8, 2), (16, 8, 3), (32, 64, 10), (64, 64, 10), (256, 256, 10)]


@pytest.mark.parametrize("nc,nv,ni", SIZES)
def test_oracle_snark_roundtrip_and_lengths(nc, nv, ni):
    inst, v, i = O.Instance.synthetic(nc, nv, ni, 0x11)
    nz = max(inst.nnz)
    comm = O.snark_encode(inst, nz)
    pf = O.snark_prove(inst, nz, v, i, LABEL, TAPE_SEED)
    assert len(pf) == len_r1cs_sat_proof(inst.num_cons, inst.num_vars) + 96 + len_r1cs_eval_proof(inst.num_cons, inst.num_vars, nz)
    assert O.snark_verify(comm, nz, i, LABEL, pf) == 0
    assert O.snark_verify(comm, nz, i, b"other label", pf) == -1
    # a flipped bit anywhere must be caught (malformed scalar / point, or a failed check)
    for pos in (40, len(pf) // 3, len(pf) // 2, 2 * len(pf) // 3, len(pf) - 40):
        bad = bytearray(pf)
        bad[pos] ^= 1
        assert O.snark_verify(comm, nz, i, LABEL, bytes(bad)) != 0, pos
    if ni:
        other = bytearray(i)
        other[0] ^= 1
        assert O.snark_verify(comm, nz, bytes(other), LABEL, pf) != 0


def test_oracle_snark_ragged_instance():
    A, B, Cm, vars_, inputs = ragged_r1cs(37, 29, 3)
    from common import canon
    inst = O.Instance.new(37, 29, 3, A, B, Cm)
    v, i = canon(vars_), canon(inputs)
    assert inst.is_sat(v, i) == 1
    nz = max(inst.nnz)
    comm = O.snark_encode(inst, nz)
    pf = O.snark_prove(inst, nz, v, i, LABEL, TAPE_SEED, threads=2)
    assert O.snark_verify(comm, nz, i, LABEL, pf) == 0


def test_oracle_snark_rejects_wrong_commitment():
    """A proof for one instance does not verify against the commitment of another of the same shape."""
    a, va, ia = O.Instance.synthetic(64, 64, 4, 1)
    b, vb, ib = O.Instance.synthetic(64, 64, 4, 2)
    nz = max(a.nnz)
    pf = O.snark_prove(a, nz, va, ia, LABEL, TAPE_SEED)
    assert O.snark_verify(O.snark_encode(a, nz), nz, ia, LABEL, pf) == 0
    assert O.snark_verify(O.snark_encode(b, nz), nz, ia, LABEL, pf) != 0


# ------------------------------------------------------------------------------------------------ GPU
def gpu_case(nc, nv, ni, seed):
    import otti_b200 as sb
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(nc, nv, ni, seed)
    oi, ov, oin = O.Instance.synthetic(nc, nv, ni, seed)
    assert bytes(vars_.data)[:len(ov)] == ov
    return sb, inst, vars_, inputs, oi, ov, oin


@pytest.mark.gpu
@pytest.mark.parametrize("logn", [3, 4, 6, 8, 10, 12])
def test_gpu_snark_prove_matches_oracle(logn):
    n = 1 << logn
    sb, inst, vars_, inputs, oi, ov, oin = gpu_case(n, n, min(10, n - 1), 0x0771 + logn)
    nz = max(oi.nnz)
    gens = sb.SNARKGens.new(n, n, inputs.n, nz)
    comm = sb.SNARK.encode(inst, gens)
    assert comm.to_bytes() == O.snark_encode(oi, nz, threads=4)
    proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    want = O.snark_prove(oi, nz, ov, oin, LABEL, TAPE_SEED, threads=8)
    got = proof.to_bytes()
    assert len(got) == len(want)
    if got != want:
        first = next(k for k in range(len(got)) if got[k] != want[k])
        raise AssertionError("GPU SNARK proof differs from the oracle at byte %d of %d" % (first, len(got)))
    # each verifier on the other side's proof
    assert O.snark_verify(comm.to_bytes(), nz, oin, LABEL, got) == 0
    proof.verify(comm, inputs, sb.Transcript.new(LABEL), gens)
    sb.SNARK(want).verify(comm, inputs, sb.Transcript.new(LABEL), gens)


@pytest.mark.gpu
def test_gpu_snark_unequal_sizes_and_ragged():
    """num_cons != num_vars (equalize pads the shorter point), several entries per row, duplicate entries."""
    import otti_b200 as sb
    from common import canon
    for (nc, nv, ni) in [(37, 29, 3), (16, 128, 5), (256, 16, 2)]:
        A, B, Cm, vars_, inputs = ragged_r1cs(nc, nv, ni)
        oi = O.Instance.new(nc, nv, ni, A, B, Cm)
        inst = sb.Instance.new(nc, nv, ni, A, B, Cm)
        v, i = canon(vars_), canon(inputs)
        nz = max(oi.nnz)
        gens = sb.SNARKGens.new(nc, nv, ni, nz)
        comm = sb.SNARK.encode(inst, gens)
        proof = sb.SNARK.prove(inst, comm, None, sb.Assignment.new(v), sb.Assignment.new(i), gens, sb.Transcript.new(LABEL), TAPE_SEED)
        want = O.snark_prove(oi, nz, v, i, LABEL, TAPE_SEED, threads=4)
        assert proof.to_bytes() == want, (nc, nv, ni)
        proof.verify(comm, sb.Assignment.new(i), sb.Transcript.new(LABEL), gens)


@pytest.mark.gpu
def test_gpu_snark_verify_rejects():
    n = 1 << 8
    sb, inst, vars_, inputs, oi, ov, oin = gpu_case(n, n, 10, 0x51)
    nz = max(oi.nnz)
    gens = sb.SNARKGens.new(n, n, inputs.n, nz)
    comm = sb.SNARK.encode(inst, gens)
    proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    got = proof.to_bytes()
    proof.verify(comm, inputs, sb.Transcript.new(LABEL), gens)
    with pytest.raises(sb.ProofVerifyError):
        proof.verify(comm, inputs, sb.Transcript.new(b"other label"), gens)
    for pos in (40, len(got) // 3, len(got) // 2, 2 * len(got) // 3, len(got) - 40):
        bad = bytearray(got)
        bad[pos] ^= 1
        with pytest.raises(sb.ProofVerifyError):
            sb.SNARK(bytes(bad)).verify(comm, inputs, sb.Transcript.new(LABEL), gens)
        assert O.snark_verify(comm.to_bytes(), nz, oin, LABEL, bytes(bad)) != 0
    with pytest.raises(sb.ProofVerifyError):
        sb.SNARK(got[:-1]).verify(comm, inputs, sb.Transcript.new(LABEL), gens)
    # the commitment of another instance of the same shape
    inst2, _, _ = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x52)
    with pytest.raises(sb.ProofVerifyError):
        proof.verify(sb.SNARK.encode(inst2, gens), inputs, sb.Transcript.new(LABEL), gens)


@pytest.mark.gpu
def test_gpu_snark_full_size_properties():
    """2^16: bytes against the oracle (a few seconds of CPU); 2^20: the proof has the reference's published length
    (README.md:305-345), the device verifier and the oracle verifier accept it, a second run reproduces it."""
    n = 1 << 16
    sb, inst, vars_, inputs, oi, ov, oin = gpu_case(n, n, 10, 0x0771)
    gens = sb.SNARKGens.new(n, n, 10, n)
    comm = sb.SNARK.encode(inst, gens)
    proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == O.snark_prove(oi, n, ov, oin, LABEL, TAPE_SEED, threads=16)
    proof.verify(comm, inputs, sb.Transcript.new(LABEL), gens)
    del gens, inst, proof, comm
    n = 1 << 20
    sb, inst, vars_, inputs, oi, ov, oin = gpu_case(n, n, 10, 0x0771)
    gens = sb.SNARKGens.new(n, n, 10, n, proofs_expected=2)
    comm = sb.SNARK.encode(inst, gens)
    proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    got = proof.to_bytes()
    assert len(got) == 47024 + 96 + 133720
    proof.verify(comm, inputs, sb.Transcript.new(LABEL), gens)
    assert O.snark_verify(comm.to_bytes(), n, oin, LABEL, got) == 0
    assert sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED).to_bytes() == got


@pytest.mark.gpu
def test_gpu_snark_doorbell_path_gives_the_same_proof():
    """SB200_BELL=1 (round kernels launched ahead of their challenge, released through a doorbell in mapped memory) is off
    by default; it must stay byte-identical.  The variable is read once per process, hence the subprocess."""
    import hashlib
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    case = _golden()["cases"][-1]
    code = ("import sys; sys.path.insert(0, %r); import hashlib, otti_b200 as sb\n"
            "inst, v, i = sb.Instance.produce_synthetic_r1cs(%d, %d, %d, %d)\n"
            "g = sb.SNARKGens.new(%d, %d, %d, %d)\n"
            "c = sb.SNARK.encode(inst, g)\n"
            "p = sb.SNARK.prove(inst, c, None, v, i, g, sb.Transcript.new(b'snark_example'), bytes(range(64)))\n"
            "p.verify(c, i, sb.Transcript.new(b'snark_example'), g)\n"
            "print(hashlib.sha256(p.to_bytes()).hexdigest())") % (
                root, case["num_cons"], case["num_vars"], case["num_inputs"], case["seed"], case["num_cons"], case["num_vars"],
                case["num_inputs"], case["num_nz_entries"])
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, SB200_BELL="1"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-800:]
    assert r.stdout.split()[-1] == case["proof_sha256"]
