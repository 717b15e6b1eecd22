#!/usr/bin/env python
"""Generates tests/golden/snark_prove_py_golden.json from the independent pure-Python restatement of SNARK::prove
(oracle/pyspec/snark_prove_spec.py, written from the Rust sources).  The JSON pins the C++ oracle (tests/test_snark_prove.py)
and, on the GPU box, the CUDA path.  Slow (big-int Python): small instances only.

    python tests/golden/make_snark_prove_py_golden.py
"""
import hashlib
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle", "pyspec"))
sys.path.insert(0, os.path.join(HERE, ".."))
import snark_prove_spec as P  # noqa: E402
import spartan_spec as S  # noqa: E402
from common import canon, ragged_r1cs  # noqa: E402

LABEL = b"snark_example"
TAPE = bytes(range(64))
cases = []


def add(case, inst, vars_, inputs, nz):
    pf = P.snark_prove(inst, vars_, inputs, nz, S.Transcript(LABEL), S.fl_from_bytes_wide(TAPE))
    case.update({"num_nz_entries": nz, "proof_len": len(pf), "proof_sha256": hashlib.sha256(pf).hexdigest(), "proof": pf.hex()})
    cases.append(case)


for (nc, nv, ni, seed) in [(4, 4, 1, 1), (8, 8, 2, 2), (16, 8, 3, 3), (8, 32, 4, 4)]:
    inst, vars_, inputs = S.synthetic_r1cs(nc, nv, ni, seed)
    add({"kind": "synthetic", "num_cons": nc, "num_vars": nv, "num_inputs": ni, "seed": seed}, inst, vars_, inputs,
        max(len(inst.A), len(inst.B), len(inst.C)))
for (nc, nv, ni, sh) in [(11, 9, 2, 5), (6, 20, 3, 6)]:
    A, B, Cm, vars_, inputs = ragged_r1cs(nc, nv, ni)
    rng = random.Random(sh)
    for M in (A, B, Cm):
        rng.shuffle(M)
    inst = S.Instance(nc, nv, ni, A, B, Cm)
    assert inst.is_sat(vars_, inputs)
    coo = [[[r, c, v.hex()] for r, c, v in M] for M in (A, B, Cm)]
    add({"kind": "coo", "num_cons": nc, "num_vars": nv, "num_inputs": ni, "coo": coo, "vars": canon(vars_).hex(),
         "inputs": canon(inputs).hex()}, inst, vars_, inputs, max(len(A), len(B), len(Cm)))
with open(os.path.join(HERE, "snark_prove_py_golden.json"), "w") as f:
    json.dump({"label": LABEL.decode(), "tape_seed": TAPE.hex(), "cases": cases}, f)
print("wrote %d cases" % len(cases))
