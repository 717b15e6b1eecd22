"""Generates tests/golden/snark_encode_golden.json from the independent pure-Python restatement of SNARK::encode
(oracle/pyspec/snark_spec.py).  Run here (CPU container); the JSON is committed so that the GPU box, which has no
/root/reference, can pin both the C++ oracle and the CUDA path.

    python tests/golden/make_snark_golden.py
"""
import hashlib
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle", "pyspec"))
sys.path.insert(0, os.path.join(HERE, ".."))
import snark_spec as K  # noqa: E402
import spartan_spec as S  # noqa: E402
from common import ragged_r1cs  # noqa: E402


def digest(xs):
    return hashlib.sha256(b"".join(S.fl_to_bytes(x) for x in xs)).hexdigest()


def finish(case, inst, nz):
    b, ops, mem = K.snark_encode(inst, nz)
    case.update({"num_nz_entries": nz, "commitment": b.hex(), "comb_ops_sha256": digest(ops), "comb_mem_sha256": digest(mem),
                 "comb_ops_len": len(ops), "comb_mem_len": len(mem)})
    # the point the sat proof would end at: rx over the constraint variables, ry over the 2 num_vars columns
    rng = random.Random(len(b))
    rx = [rng.randrange(S.ELL) for _ in range(S.log2(inst.num_cons))]
    ry = [rng.randrange(S.ELL) for _ in range(S.log2(2 * inst.num_vars))]
    pts, comb = K.commit_derefs(inst, nz, rx, ry)
    case.update({"rx": b"".join(S.fl_to_bytes(x) for x in rx).hex(), "ry": b"".join(S.fl_to_bytes(x) for x in ry).hex(),
                 "derefs_commitment": b"".join(pts).hex(), "derefs_comb_sha256": digest(comb), "derefs_comb_len": len(comb)})
    rh, rm = rng.randrange(S.ELL), rng.randrange(S.ELL)
    prods = K.network_products(inst, rx, ry, rh, rm)
    case.update({"r_hash": S.fl_to_bytes(rh).hex(), "r_multiset_check": S.fl_to_bytes(rm).hex(),
                 "network_products": [S.fl_to_bytes(x).hex() for x in prods]})
    return case


def synthetic_case(s, seed):
    n = 1 << s
    ni = min(10, n - 1)
    inst, _, _ = S.synthetic_r1cs(n, n, ni, seed)
    return finish({"kind": "synthetic", "log_size": s, "num_inputs": ni, "seed": seed}, inst, n)


def ragged_case(nc, nv, ni, shuffle_seed):
    """Entries in shuffled order: the read timestamps depend on the order of the ops, not on the matrix."""
    A, B, Cm, _, _ = ragged_r1cs(nc, nv, ni)
    rng = random.Random(shuffle_seed)
    for M in (A, B, Cm):
        rng.shuffle(M)
    inst = S.Instance(nc, nv, ni, A, B, Cm)
    coo = [[[r, c, v.hex()] for r, c, v in M] for M in (A, B, Cm)]
    return finish({"kind": "coo", "num_cons": nc, "num_vars": nv, "num_inputs": ni, "coo": coo}, inst, max(len(A), len(B), len(Cm)))


if __name__ == "__main__":
    cases = [synthetic_case(3, 0x0771), synthetic_case(5, 0x0771), ragged_case(11, 6, 2, 1), ragged_case(20, 40, 3, 2),
             ragged_case(33, 9, 5, 3)]
    with open(os.path.join(HERE, "snark_encode_golden.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_snark_golden.py (oracle/pyspec/snark_spec.py)", "cases": cases}, f, indent=1)
    print("wrote", len(cases), "cases")
