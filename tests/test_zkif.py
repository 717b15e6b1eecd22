"""zkinterface ingestion: libspartan_b200's reader against the pure-Python restatement of
spartan-zkinterface/src/lib.rs:75-260 (oracle/pyspec/zkif_spec.py), on the reference's own fixtures
(tests/golden/zkif/, which the reference requires to parse and be satisfiable, lib.rs:318-332), on generated
files with the id-translation corner cases, and on malformed input.  GPU: prove what was ingested."""
import os
import random

import pytest

import oracle_lib as O
import zkif_spec
import zkif_writer as W
from common import ELL, LABEL, TAPE_SEED, canon, coo_of, otti_shaped_r1cs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "zkif")


def _fixture(name):
    return [open(os.path.join(GOLD, name + ext), "rb").read() for ext in (".zkif", ".inp.zkif", ".wit.zkif")]


def _same(r, spec):
    assert (r.num_cons, r.num_vars, r.num_inputs) == (spec["num_cons"], spec["num_vars"], spec["num_inputs"])
    assert r.non_zero_entries == spec["non_zero_entries"]
    for m, key in enumerate("ABC"):
        assert r.matrix(m) == spec[key]
    assert r.vars_assignment().data == spec["vars"] and r.inputs_assignment().data == spec["inputs"]


@pytest.mark.parametrize("name", ["add", "foo", "inv"])
def test_reference_fixtures(name):
    from otti_b200 import zkif
    c, h, w = _fixture(name)
    spec = zkif_spec.parse(c, h, w)
    # the reference's test: the instance built from the fixture is satisfied by its witness (lib.rs:283-290)
    oi = O.Instance.new(spec["num_cons"], spec["num_vars"], spec["num_inputs"], spec["A"], spec["B"], spec["C"])
    assert oi.is_sat(spec["vars"], spec["inputs"]) == 1
    _same(zkif.R1cs.from_buffers(c, h, w), spec)
    r = zkif.R1cs.from_files(*[os.path.join(GOLD, name + ext) for ext in (".zkif", ".inp.zkif", ".wit.zkif")])
    assert r.num_cons == spec["num_cons"]


def _files_of(cons, vars_, inputs, nv, width=32, wid=lambda i: 1000 + 3 * i, iid=lambda i: 1 + i):
    """zkif bytes of a constraint list over z = vars || 1 || inputs (constant: id 0, in neither file)."""
    def vid(c):
        return wid(c) if c < nv else (0 if c == nv else iid(c - nv - 1))
    cs = W.constraint_system([tuple([(vid(c), v) for c, v in lc] for lc in con) for con in cons], width)
    return cs, W.header([iid(i) for i in range(len(inputs))], inputs, ELL, width), W.witness([wid(i) for i in range(nv)], vars_, width)


@pytest.mark.parametrize("dims", [(7, 5, 2), (100, 64, 3), (300, 1000, 0), (513, 200, 17)])
def test_generated_files(dims):
    from otti_b200 import zkif
    nc, nv, ni = dims
    cons, v, i = otti_shaped_r1cs(nc, nv, ni, seed=nc)
    c, h, w = _files_of(cons, v, i, nv)
    spec = zkif_spec.parse(c, h, w)
    A, B, Cm = coo_of(cons)
    assert (spec["A"], spec["B"], spec["C"]) == (A, B, Cm)  # writer + spec agree with the generator
    assert spec["vars"] == canon(v) and spec["inputs"] == canon(i)
    _same(zkif.R1cs.from_buffers(c, h, w), spec)


def test_translation_corner_cases():
    """duplicate ids (first occurrence wins), ids in neither file (-> constant column), an id in both files
    (witness wins), short values (zero-extended), empty linear combinations"""
    from otti_b200 import zkif
    wit_ids, wit_vals = [5, 9, 5, 11], [3, 4, 99, 6]
    in_ids, in_vals = [9, 2, 2], [7, 8, 55]
    cons = [([(5, 1), (9, 2)], [(2, 1)], [(777, 3), (11, 1)]), ([], [(0, 1)], []), ([(2, 200), (5, 1)], [(11, 255)], [(9, 1)])]
    c = W.constraint_system(cons, 1)
    h = W.header(in_ids, in_vals, ELL, 1)
    w = W.witness(wit_ids, wit_vals, 1)
    spec = zkif_spec.parse(c, h, w)
    assert [(r, col) for r, col, _ in spec["A"]] == [(0, 0), (0, 1), (2, 6), (2, 0)]
    assert [(r, col) for r, col, _ in spec["C"]] == [(0, 4), (0, 3), (2, 1)]
    assert spec["B"][1][1] == 4 and spec["C"][0][2] == bytes([3]) + bytes(31)
    _same(zkif.R1cs.from_buffers(c, h, w), spec)


def test_malformed_input_is_refused_not_crashed():
    from otti_b200 import zkif
    c, h, w = _fixture("add")
    with pytest.raises(zkif.ZkifError):
        zkif.R1cs.from_buffers(h, c, w)  # wrong message types ("Input file is not a flatbuffer ...", lib.rs:166-185)
    with pytest.raises(zkif.ZkifError):
        zkif.R1cs.from_buffers(W.constraint_system([]), h, w)  # "No constraints given!" (lib.rs:235-237)
    with pytest.raises(zkif.ZkifError):
        zkif.R1cs.from_buffers(c, h, W.witness([1, 2], [1, 2], 33))  # value wider than 32 bytes
    rng = random.Random(1)
    outcomes = set()
    for trial in range(400):
        bufs = [bytearray(c), bytearray(h), bytearray(w)]
        b = bufs[trial % 3]
        if trial % 2:
            del b[rng.randrange(len(b)):]
        else:
            for _ in range(rng.randrange(1, 4)):
                b[rng.randrange(len(b))] = rng.randrange(256)
        try:
            r = zkif.R1cs.from_buffers(*bufs)
            r.matrix(0), r.matrix(1), r.matrix(2)
            outcomes.add("parsed")
        except zkif.ZkifError:
            outcomes.add("refused")
    assert "refused" in outcomes


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["add", "foo", "inv"])
def test_prove_reference_fixtures(name):
    """spzk prove/verify --nizk on the reference's fixtures (main.rs:57-106): is_sat, prove == oracle bytes, verify."""
    import otti_b200 as sb
    from otti_b200 import zkif
    assert sb.lib().sb200_init(0) == 0
    c, h, w = _fixture(name)
    r = zkif.R1cs.from_buffers(c, h, w)
    inst, vars_, inputs, gens = r.instance(), r.vars_assignment(), r.inputs_assignment(), r.nizk_public_params()
    assert inst.is_sat(vars_, inputs)
    spec = zkif_spec.parse(c, h, w)
    oi = O.Instance.new(spec["num_cons"], spec["num_vars"], spec["num_inputs"], spec["A"], spec["B"], spec["C"])
    og = O.Gens(spec["num_cons"], spec["num_vars"], spec["num_inputs"])
    want = O.nizk_prove(oi, spec["vars"], spec["inputs"], og, LABEL, TAPE_SEED)
    proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
    assert O.nizk_verify(oi, spec["inputs"], og, LABEL, proof.to_bytes()) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(100, 64, 3), (3000, 1500, 10), (5000, 20000, 4)])
def test_prove_otti_shaped_through_zkif(dims):
    import otti_b200 as sb
    from otti_b200 import zkif
    assert sb.lib().sb200_init(0) == 0
    nc, nv, ni = dims
    cons, v, i = otti_shaped_r1cs(nc, nv, ni, seed=nv)
    c, h, w = _files_of(cons, v, i, nv)
    r = zkif.R1cs.from_buffers(c, h, w)
    inst, vars_, inputs, gens = r.instance(), r.vars_assignment(), r.inputs_assignment(), r.nizk_public_params()
    assert inst.is_sat(vars_, inputs)
    A, B, Cm = coo_of(cons)
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    og = O.Gens(nc, nv, ni)
    want = O.nizk_prove(oi, canon(v), canon(i), og, LABEL, TAPE_SEED)
    proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)


def test_generated_heavy_constant_column_files():
    """tools/zkifgen.cpp (the C++ writer behind the 2^18 / 2^20 Otti-shaped cases): its files parse to its own COO
    triples under both the Python restatement of the reference reader and the product's reader; ~30 % of all entries
    sit in the constant column."""
    import zkifgen
    from otti_b200 import zkif
    g = zkifgen.OttiShaped(700, 450, 9, seed=5)
    assert g.const_entries >= 0.25 * sum(g.nnz)
    c, h, w = (g.file_bytes(k) for k in range(3))
    spec = zkif_spec.parse(c, h, w)
    assert [spec[k] for k in "ABC"] == [g.coo_list(m) for m in range(3)]
    assert spec["vars"] == g.vars_bytes() and spec["inputs"] == g.inputs_bytes()
    oi = O.Instance.new_arrays(700, 450, 9, [g.coo_arrays(m) + (g.nnz[m],) for m in range(3)])
    assert oi.is_sat(g.vars_bytes(), g.inputs_bytes()) == 1
    _same(zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2)), spec)


def test_threaded_reader_keeps_file_order():
    """Large constraint systems are walked by several threads (host/zkif.hpp): the triples must come out in file order,
    exactly the generator's COO lists."""
    import zkifgen
    from otti_b200 import zkif
    g = zkifgen.OttiShaped(30000, 17000, 5, seed=3)
    r = zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2))
    assert (r.num_cons, r.num_vars, r.num_inputs) == (30000, 17000, 5) and r.nnz == g.nnz
    for m in range(3):
        rows, cols, vals = g.coo_arrays(m)
        assert r.matrix_arrays(m) == (bytes(rows)[:8 * g.nnz[m]], bytes(cols)[:8 * g.nnz[m]], bytes(vals)[:32 * g.nnz[m]])
    assert bytes(r.vars_assignment().data) == g.vars_bytes()


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(1 << 18, 1 << 18, 10), (200000, 1 << 17, 3)])
def test_prove_otti_shaped_heavy_column_2p18(dims):
    """An Otti-shaped instance at 2^18 constraints whose constant column holds ~30 % of all non-zero entries (one column of
    ~2 * 10^5 entries per matrix), through zkinterface: reader == generator's COO, is_sat, proof bytes == oracle,
    device verify and oracle verify accept."""
    import otti_b200 as sb
    import zkifgen
    from otti_b200 import zkif
    assert sb.lib().sb200_init(0) == 0
    nc, nv, ni = dims
    g = zkifgen.OttiShaped(nc, nv, ni, seed=11)
    assert g.const_entries >= 0.25 * sum(g.nnz)
    r = zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2))
    assert (r.num_cons, r.num_vars, r.num_inputs) == dims and r.nnz == g.nnz
    for m in range(3):
        rows, cols, vals = g.coo_arrays(m)
        assert r.matrix_arrays(m) == (bytes(rows)[:8 * g.nnz[m]], bytes(cols)[:8 * g.nnz[m]], bytes(vals)[:32 * g.nnz[m]])
    inst, vars_, inputs, gens = r.instance(), r.vars_assignment(), r.inputs_assignment(), r.nizk_public_params()
    assert bytes(vars_.data) == g.vars_bytes()
    assert inst.is_sat(vars_, inputs)
    oi = O.Instance.new_arrays(nc, nv, ni, [g.coo_arrays(m) + (g.nnz[m],) for m in range(3)])
    og = O.Gens(nc, nv, ni)
    want = O.nizk_prove(oi, g.vars_bytes(), g.inputs_bytes(), og, LABEL, TAPE_SEED, os.cpu_count() or 8)
    proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
    assert O.nizk_verify(oi, g.inputs_bytes(), og, LABEL, proof.to_bytes()) == 0
