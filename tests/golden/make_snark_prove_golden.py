#!/usr/bin/env python
"""Golden digests of SNARK::prove on the CPU oracle (oracle/snark.hpp): regression pins for the oracle itself, so that the
checker cannot drift unnoticed between rounds.  python tests/golden/make_snark_prove_golden.py > tests/golden/snark_prove_golden.json"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402

LABEL = b"snark_example"
TAPE = bytes(range(64))
cases = []
for (nc, nv, ni, seed) in [(8, 8, 2, 1), (16, 8, 3, 2), (32, 64, 10, 3), (64, 64, 10, 4), (256, 256, 10, 5), (1024, 1024, 10, 6)]:
    inst, v, i = O.Instance.synthetic(nc, nv, ni, seed)
    nz = max(inst.nnz)
    comm = O.snark_encode(inst, nz)
    pf = O.snark_prove(inst, nz, v, i, LABEL, TAPE)
    assert O.snark_verify(comm, nz, i, LABEL, pf) == 0
    cases.append({"num_cons": nc, "num_vars": nv, "num_inputs": ni, "seed": seed, "num_nz_entries": nz, "proof_len": len(pf),
                  "proof_sha256": hashlib.sha256(pf).hexdigest(), "commitment_sha256": hashlib.sha256(comm).hexdigest(),
                  "proof_head": pf[:64].hex(), "proof_tail": pf[-64:].hex()})
json.dump({"label": LABEL.decode(), "tape_seed": TAPE.hex(), "cases": cases}, sys.stdout, indent=1)
