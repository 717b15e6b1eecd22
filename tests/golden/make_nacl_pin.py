"""Generates tests/golden/nacl_pin.json: an INDEPENDENT pin for the oracle's F_l, Edwards and multiscalar arithmetic.

The reference's group arithmetic lives in curve25519-dalek (un-vendored, Rust: cannot run here).  libsodium (through
PyNaCl 1.6.2, present in this image) is a third implementation of the same curve and scalar field written by other
people; nothing in it comes from this repository.  It has no ristretto255 API in PyNaCl, so points are exchanged as
Edwards points (the affine representative oracle/ristretto.hpp's decompression yields, encoded the ed25519 way) and
compared modulo the 4-torsion coset a Ristretto point stands for (RFC 9496 section 4.3.3:
x1*y2 == y1*x2  or  y1*y2 == x1*x2).

What libsodium computes here (SURVEY.md section 8(c) names these entry points):
  scalars   crypto_core_ed25519_scalar_{mul,add,sub,invert,negate}, scalar_reduce of 64 bytes  (Scalar ops, from_bytes_wide)
  base      crypto_scalarmult_ed25519_base_noclamp(k)                                           (k * B)
  add       crypto_core_ed25519_add on the representatives of two Spartan generators            (Edwards addition)
  msm       sum_i crypto_scalarmult_ed25519_noclamp(k_i, 4*G_i) = 4 * Com(k)                    (the commitment MSM)
            (4*G_i by two doublings through crypto_core_ed25519_add: a decoded Ristretto representative lies in 2E and
             may carry 4-torsion, which scalarmult_noclamp refuses; times four it is in the prime-order subgroup)

The only oracle outputs that enter the file are the generator points themselves (inputs of `add` and `msm`); the test
recomputes them, so a change in the oracle's generator derivation shows up as well.

    python tests/golden/make_nacl_pin.py
"""
import json
import os
import random
import sys

import nacl.bindings as nb

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import oracle_lib as O  # noqa: E402

P = 2**255 - 19
ELL = 2**252 + 27742317777372353535851937790883648493


def ed_encode(xy64):
    """affine (x, y) as 2 x 32 LE bytes -> ed25519 encoding (y with the sign of x in bit 255)"""
    x = int.from_bytes(xy64[:32], "little")
    y = int.from_bytes(xy64[32:], "little")
    return (y | ((x & 1) << 255)).to_bytes(32, "little")


def quadruple(ed):
    d = nb.crypto_core_ed25519_add(ed, ed)
    return nb.crypto_core_ed25519_add(d, d)


def main():
    rng = random.Random(0xED25519)
    sc = lambda: rng.randrange(ELL).to_bytes(32, "little")
    out = {"about": "libsodium (PyNaCl %s) outputs; see make_nacl_pin.py" % __import__("nacl").__version__,
           "scalars": [], "wide": [], "base": [], "add": [], "msm": []}
    edge = [0, 1, 2, ELL - 1, ELL - 2, (ELL - 1) // 2, 2**252, 2**128 - 1]
    pairs = [(a.to_bytes(32, "little"), b.to_bytes(32, "little")) for a in edge for b in (1, ELL - 1, 2**200 + 12345)]
    pairs += [(sc(), sc()) for _ in range(40)]
    for a, b in pairs:
        e = {"a": a.hex(), "b": b.hex(), "mul": nb.crypto_core_ed25519_scalar_mul(a, b).hex(),
             "add": nb.crypto_core_ed25519_scalar_add(a, b).hex(), "sub": nb.crypto_core_ed25519_scalar_sub(a, b).hex(),
             "neg": nb.crypto_core_ed25519_scalar_negate(a).hex()}
        if int.from_bytes(a, "little") != 0:
            e["inv"] = nb.crypto_core_ed25519_scalar_invert(a).hex()
        out["scalars"].append(e)
    for w in [bytes(64), b"\xff" * 64, bytes(range(64))] + [rng.randbytes(64) for _ in range(20)]:
        out["wide"].append({"in": w.hex(), "reduced": nb.crypto_core_ed25519_scalar_reduce(w).hex()})
    for k in [1, 2, 3, 8, ELL - 1, 2**252, 2**16, 2**16 - 1] + [rng.randrange(1, ELL) for _ in range(24)]:
        kb = k.to_bytes(32, "little")
        out["base"].append({"k": kb.hex(), "ed": nb.crypto_scalarmult_ed25519_base_noclamp(kb).hex()})
    # Spartan generators of a 2^8-variable NIZKGens: gens_n (R = 16 points) and its h
    gens = O.Gens(1 << 8, 1 << 8, 10)
    G = [gens.point(0, i) for i in range(gens.n)] + [gens.point(1)]  # which = 0: gens_n.G[i], 1: gens_n.h
    reps = [ed_encode(O.pt_affine(g)) for g in G]
    out["gens"] = {"num_vars": 1 << 8, "points": [g.hex() for g in G]}
    for i in range(len(G) - 1):
        out["add"].append({"i": i, "j": i + 1, "ed_sum_of_reps": nb.crypto_core_ed25519_add(reps[i], reps[i + 1]).hex()})
    q4 = [quadruple(r) for r in reps]
    for case in range(6):
        n = [1, 2, 5, gens.n, gens.n, gens.n][case]
        ks = [sc() for _ in range(n)]
        if case == 4:
            ks = [(rng.randrange(2)).to_bytes(32, "little") for _ in range(n)]  # bits, as an Otti witness row
            ks[0] = (1).to_bytes(32, "little")
        blind = sc()
        terms = [(k, q4[i]) for i, k in enumerate(ks)] + [(blind, q4[-1])]
        acc = None
        for k, pt in terms:
            if int.from_bytes(k, "little") == 0:
                continue
            t = nb.crypto_scalarmult_ed25519_noclamp(k, pt)
            acc = t if acc is None else nb.crypto_core_ed25519_add(acc, t)
        out["msm"].append({"scalars": [k.hex() for k in ks], "blind": blind.hex(), "ed_4_times_commitment": acc.hex()})
    with open(os.path.join(HERE, "nacl_pin.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("wrote nacl_pin.json: %d scalar cases, %d base multiples, %d additions, %d commitments" % (
        len(out["scalars"]), len(out["base"]), len(out["add"]), len(out["msm"])))


if __name__ == "__main__":
    main()
