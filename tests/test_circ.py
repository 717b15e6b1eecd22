"""rust-circ adapter: libspartan_b200's `sb200_circ_r1cs` / `sb200_circ_int_to_scalar`
against the pure-Python restatement of rust-circ/src/target/r1cs/spartan.rs:31-186 (oracle/pyspec/circ_spec.py).
GPU: prove what was converted and compare the proof bytes with the oracle's."""
import random

import pytest

import circ_spec
import oracle_lib as O
from common import ELL, LABEL, TAPE_SEED, otti_shaped_r1cs


def _circ_of(cons, vars_, inputs, nv, signed=True, seed=3):
    """A CirC-style R1cs of an "Otti-shaped" constraint list over z = vars || 1 || inputs.  CirC ids are a
    permutation; values are listed in shuffled order; coefficients use signed / oversized integers where that
    does not change |i| mod l (the adapter reads magnitudes only, spartan.rs:141-160)."""
    from otti_b200.circ import CircR1cs, Lc
    rng = random.Random(seed)
    ni = len(inputs)
    n = nv + ni
    ids = list(range(n))
    # the reference panics on a monomial id >= num_vars (spartan.rs:168): keep used ids small
    wid = ids[:nv]
    iid = ids[nv:]
    names, values = {}, {}
    order = [("w", j) for j in range(nv)] + [("i", j) for j in range(ni)]
    rng.shuffle(order)
    for kind, j in order:
        if kind == "w":
            k = wid[j]
            names[k] = ("main_f0_lex0_w%d_v0" % j) if j % 3 else ("f1_ret_%d" % j)
            v = vars_[j]
            values[k] = v + ELL * rng.randrange(3) if j % 5 == 0 else v  # oversized: reduced by int_to_scalar
        else:
            k = iid[j]
            names[k] = "main_f0_lex0_x%d_v0" % j
            values[k] = inputs[j]

    def lc(terms):
        out = Lc()
        for c, v in terms:
            if c == nv:
                out.constant = (out.constant + v) % ELL
            elif c < nv:
                out.monomials[wid[c]] = (out.monomials.get(wid[c], 0) + v) % ELL
            else:
                raise AssertionError("inputs cannot appear in constraints (spartan.rs:168)")
        return out
    constraints = [(lc(a), lc(b), lc(c)) for a, b, c in cons]
    return CircR1cs(idxs_signals=names, values=values, constraints=constraints, next_idx=n)


def _spec_of(r):
    return circ_spec.r1cs_to_spartan(r.idxs_signals, r.values, [tuple((l.constant, l.monomials) for l in c) for c in r.constraints],
                                     r.next_idx, r.public_idxs)


def _no_inputs(cons, nv):
    """drop the terms on input columns (they cannot be expressed, see above) -- C is recomputed by the caller"""
    return [tuple([(c, v) for c, v in lc if c <= nv] for lc in con) for con in cons]


def _satisfiable(nc, nv, ni, seed):
    cons, v, i = otti_shaped_r1cs(nc, nv, ni, seed=seed)
    cons = _no_inputs(cons, nv)
    z = v + [1]
    fixed = []
    for a, b, c in cons:
        p = (sum(x * z[k] for k, x in a) % ELL) * (sum(x * z[k] for k, x in b) % ELL) % ELL
        fixed.append((a, b, [(nv, p)]))  # C = the product on the constant column
    return fixed, v, i


def test_int_to_scalar():
    from otti_b200.circ import int_to_scalar
    rng = random.Random(1)
    cases = [0, 1, 2**63, 2**64 - 1, 2**64, ELL - 1, ELL, ELL + 1, 2**255, 2**256 - 1, 2**300 + 5, -1, -ELL - 7, 2**1000 + 12345]
    cases += [rng.randrange(2**(8 * rng.randrange(1, 70))) for _ in range(50)]
    for i in cases:
        want = abs(i) % ELL
        assert circ_spec.int_to_scalar(i) == want
        assert int.from_bytes(int_to_scalar(i), "little") == want


@pytest.mark.parametrize("dims", [(7, 5, 0), (100, 64, 3), (300, 1000, 0), (513, 200, 17)])
def test_conversion_matches_spec(dims):
    from otti_b200.circ import circ_to_r1cs
    nc, nv, ni = dims
    cons, v, i = _satisfiable(nc, nv, ni, seed=nc)
    r = _circ_of(cons, v, i, nv)
    spec = _spec_of(r)
    got = circ_to_r1cs(r)
    assert (got.num_cons, got.num_vars, got.num_inputs) == (spec["num_cons"], spec["num_vars"], spec["num_inputs"]) == (nc, nv, ni)
    for m, key in enumerate("ABC"):
        assert got.matrix(m) == spec[key]
    assert got.vars_assignment().data == spec["vars"] and got.inputs_assignment().data == spec["inputs"]
    # the instance the reference would build is satisfied (spartan.rs:135-136), on the CPU oracle
    oi = O.Instance.new(nc, nv, ni, spec["A"], spec["B"], spec["C"])
    assert oi.is_sat(spec["vars"], spec["inputs"]) == 1


def test_signs_and_zero_terms():
    """Magnitudes only: a coefficient of -3 becomes 3, a constant of -5 becomes 5; a zero coefficient is still an
    entry, a zero constant is not (spartan.rs:170-184)."""
    from otti_b200.circ import CircR1cs, Lc, circ_to_r1cs
    names = {0: "main_f0_lex0_w0_v0", 1: "main_f0_lex0_w1_v0", 2: "main_f0_lex0_in_v0"}
    values = {2: 9, 1: -4, 0: 2**260 + 1}
    cons = [(Lc(-5, {0: -3, 1: 0}), Lc(0, {1: 1}), Lc(7, {}))]
    r = CircR1cs(names, values, cons, next_idx=3)
    spec = _spec_of(r)
    got = circ_to_r1cs(r)
    b = lambda x: int(x).to_bytes(32, "little")
    # values are listed as ids 2, 1, 0: id 1 is witness 0, id 0 is witness 1, id 2 the input
    assert spec["A"] == [(0, 1, b(3)), (0, 0, b(0)), (0, 2, b(5))] and spec["B"] == [(0, 0, b(1))] and spec["C"] == [(0, 2, b(7))]
    assert spec["vars"] == b(4) + b((2**260 + 1) % ELL) and spec["inputs"] == b(9)
    for m, key in enumerate("ABC"):
        assert got.matrix(m) == spec[key]
    assert got.vars_assignment().data == spec["vars"] and got.inputs_assignment().data == spec["inputs"]


def test_panics_become_errors():
    from otti_b200.circ import CircR1cs, Lc, circ_to_r1cs
    names = {0: "main_f0_lex0_w0_v0", 1: "main_f0_lex0_w1_v0", 5: "main_f0_lex0_in_v0"}
    values = {0: 1, 1: 2, 5: 3}
    ok = [(Lc(0, {0: 1}), Lc(0, {1: 1}), Lc(2, {}))]
    circ_to_r1cs(CircR1cs(names, values, ok, next_idx=3))
    bad = [
        CircR1cs(names, values, ok, next_idx=3, public_idxs=frozenset([5])),          # :42
        CircR1cs(names, values, ok, next_idx=4),                                       # :80
        CircR1cs(names, values, [(Lc(0, {5: 1}), Lc(), Lc())], next_idx=3),            # :168 id >= num_vars
        CircR1cs({**names, 9: "x"}, values, [(Lc(0, {9: 1}), Lc(), Lc())], next_idx=3),
    ]
    for r in bad:
        with pytest.raises(circ_spec.Panic):
            _spec_of(r)
        with pytest.raises(AssertionError):
            circ_to_r1cs(r)
    # id 1 < num_vars but never assigned: trans.get(k).unwrap()
    r = CircR1cs({0: "main_f0_lex0_w0_v0", 7: "main_f0_lex0_w1_v0"}, {0: 1, 7: 2}, [(Lc(0, {1: 1}), Lc(), Lc())], next_idx=2)
    with pytest.raises(circ_spec.Panic):
        _spec_of(r)
    with pytest.raises(AssertionError):
        circ_to_r1cs(r)


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(100, 64, 3), (2000, 1500, 0)])
def test_prove_from_circ(dims):
    """circ.rs:246-273: r1cs_to_spartan -> NIZKGens -> prove -> verify; proof bytes equal the oracle's on the
    instance the restatement of the adapter builds."""
    import otti_b200 as sb
    from otti_b200.circ import r1cs_to_spartan
    nc, nv, ni = dims
    cons, v, i = _satisfiable(nc, nv, ni, seed=nc + 1)
    r = _circ_of(cons, v, i, nv)
    spec = _spec_of(r)
    inst, wit, inp, num_cons, num_vars, num_inputs = r1cs_to_spartan(r)
    assert (num_cons, num_vars, num_inputs) == (nc, nv, ni)
    gens = sb.NIZKGens.new(num_cons, num_vars, num_inputs)
    proof = sb.NIZK.prove(inst, wit, inp, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    oi = O.Instance.new(nc, nv, ni, spec["A"], spec["B"], spec["C"])
    og = O.Gens(nc, nv, ni)
    assert proof.to_bytes() == O.nizk_prove(oi, spec["vars"], spec["inputs"], og, LABEL, TAPE_SEED)
    assert proof.verify(inst, inp, sb.Transcript.new(LABEL), gens) is None
