"""Generates tests/golden/nizk_golden.json from the independent pure-Python restatement
(oracle/pyspec/spartan_spec.py).  Run here (CPU container); the JSON is committed so the GPU
box, which has no /root/reference, can still pin both the oracle and the CUDA path.

    python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle", "pyspec"))
import spartan_spec as S  # noqa: E402

LABEL = b"nizk_example"
TAPE = bytes(range(64))


def synthetic_case(s, seed):
    n = 1 << s
    ni = min(10, n - 1)
    inst, vars_, inputs = S.synthetic_r1cs(n, n, ni, seed)
    gens = S.NIZKGens(n, n, ni)
    pf = S.nizk_prove(inst, vars_, inputs, gens, S.Transcript(LABEL), S.fl_from_bytes_wide(TAPE))
    assert S.nizk_verify(pf, inst, inputs, gens, S.Transcript(LABEL))
    b = S.serialize_nizk(pf)
    return {"kind": "synthetic", "log_size": s, "num_inputs": ni, "seed": seed, "label": LABEL.decode(), "tape_seed64": TAPE.hex(),
            "vars_sha256": hashlib.sha256(b"".join(S.fl_to_bytes(v) for v in vars_)).hexdigest(), "proof_len": len(b),
            "r1cs_proof_len": len(S.serialize_r1cs_proof(pf)), "proof": b.hex(),
            "comm_vars0": pf["comm_vars"][0].hex(), "rx0": S.fl_to_bytes(pf["rx"][0]).hex()}


def padded_case():
    """lib.rs:653-731 test_padded_constraints (NIZK part): 1 constraint, 0 vars, 3 inputs."""
    one = S.fl_to_bytes(1)
    A = [(0, 2, one)]
    B = [(0, 2, one)]
    C = [(0, 1, one), (0, 0, S.fl_to_bytes(S.ELL - 13)), (0, 3, S.fl_to_bytes(S.ELL - 1))]
    inputs = [16, 1, 2]
    inst = S.Instance(1, 0, 3, A, B, C)
    assert inst.is_sat([], inputs)
    gens = S.NIZKGens(1, 0, 3)
    pf = S.nizk_prove(inst, [], inputs, gens, S.Transcript(LABEL), S.fl_from_bytes_wide(TAPE))
    assert S.nizk_verify(pf, inst, inputs, gens, S.Transcript(LABEL))
    b = S.serialize_nizk(pf)
    return {"kind": "padded_constraints", "label": LABEL.decode(), "tape_seed64": TAPE.hex(), "proof": b.hex(), "proof_len": len(b)}


def primitives():
    g = S.MultiCommitGens.new(1, b"test-knowledgeproof")
    gens = S.NIZKGens(16, 16, 3)
    t = S.Transcript(b"test protocol")
    t.append_message(b"some label", b"some data")
    return {
        "merlin_test_protocol": t.challenge_bytes(b"challenge", 32).hex(),
        "ristretto_basepoint": S.BASEPOINT.compress().hex(),
        "ristretto_2B": (2 * S.BASEPOINT).compress().hex(),
        "rfc9496_hash_to_group_1": S.from_uniform_bytes(hashlib.sha512(b"Ristretto is traditionally a short shot of espresso coffee").digest()).compress().hex(),
        "gens_test_knowledgeproof": [g.G[0].compress().hex(), g.h.compress().hex()],
        "gens_r1cs_sat_16": {"gens_n_G": [p.compress().hex() for p in gens.gens_n.G], "h": gens.gens_n.h.compress().hex(),
                             "gens_1_G0": gens.gens_1.G[0].compress().hex(), "gens_4_h": gens.gens_4.h.compress().hex()},
        "eq_evals_r_4_3": [hex(x) for x in S.eq_evals([4, 3])],
        "proof_len_closed_form": {str(s): S.r1cs_proof_len(s, s) for s in (10, 16, 20, 22, 24)},
    }


if __name__ == "__main__":
    out = {"generator": "tests/golden/make_golden.py over oracle/pyspec/spartan_spec.py",
           "primitives": primitives(), "cases": [synthetic_case(s, 0x0771) for s in (1, 2, 3, 4, 5, 6)] + [padded_case()]}
    with open(os.path.join(HERE, "nizk_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(out["cases"]), "cases")
