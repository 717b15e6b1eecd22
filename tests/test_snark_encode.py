"""SNARK::encode -- the computation commitment (lib.rs:321-333, sparse_mlpoly.rs:371-506), SURVEY.md section 8(f) N3.

CPU part: the C++ oracle against the golden vectors of the independent Python restatement (tests/golden/
snark_encode_golden.json, made by tests/golden/make_snark_golden.py) and the properties of the timestamps.
GPU part: the CUDA path (dense representation kernels + fixed-base MSM) through the C ABI against the oracle and the
golden vectors.  Bar: bit-exact.
"""
import ctypes as C
import hashlib
import json
import os
import random
import struct

import pytest

import oracle_lib as O
from common import ELL, otti_shaped_r1cs, coo_of, ragged_r1cs

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "snark_encode_golden.json")) as f:
    GOLDEN = json.load(f)["cases"]


def case_coo(case):
    """(num_cons, num_vars, num_inputs, A, B, C) of a golden case, or None for a synthetic one."""
    if case["kind"] != "coo":
        return None
    mats = [[(r, c, bytes.fromhex(v)) for r, c, v in M] for M in case["coo"]]
    return case["num_cons"], case["num_vars"], case["num_inputs"], mats[0], mats[1], mats[2]


def oracle_instance(case):
    if case["kind"] == "synthetic":
        n = 1 << case["log_size"]
        return O.Instance.synthetic(n, n, case["num_inputs"], case["seed"])[0]
    nc, nv, ni, A, B, Cm = case_coo(case)
    return O.Instance.new(nc, nv, ni, A, B, Cm)


def case_id(case):
    return "syn%d" % case["log_size"] if case["kind"] == "synthetic" else "coo%dx%d" % (case["num_cons"], case["num_vars"])


# ------------------------------------------------------------------------------------------------ CPU: the oracle
@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_oracle_matches_python_restatement(case):
    inst = oracle_instance(case)
    N, cells, ops, mem = O.snark_dense(inst)
    assert len(ops) == 32 * case["comb_ops_len"] and len(mem) == 32 * case["comb_mem_len"]
    assert hashlib.sha256(ops).hexdigest() == case["comb_ops_sha256"]
    assert hashlib.sha256(mem).hexdigest() == case["comb_mem_sha256"]
    assert O.snark_encode(inst, case["num_nz_entries"], threads=2).hex() == case["commitment"]


@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_oracle_derefs_match_python_restatement(case):
    """commit_nondet_witness of SparseMatPolyEvalProof::prove (sparse_mlpoly.rs:1485-1503), incl. equalize when rx and ry
    differ in length."""
    inst = oracle_instance(case)
    pts, comb = O.snark_commit_derefs(inst, case["num_nz_entries"], bytes.fromhex(case["rx"]), bytes.fromhex(case["ry"]),
                                      threads=2, want_comb=True)
    assert len(comb) == 32 * case["derefs_comb_len"] and hashlib.sha256(comb).hexdigest() == case["derefs_comb_sha256"]
    assert pts.hex() == case["derefs_commitment"]


def multiset_identity_holds(prods):
    """hashed_write_set == hashed_read_set (sparse_mlpoly.rs:631-643): init * prod(write) == prod(read) * audit."""
    v = [int.from_bytes(p, "little") for p in prods]
    for s0 in (0, 8):
        lhs = v[s0] * v[s0 + 4] * v[s0 + 5] * v[s0 + 6] % ELL
        rhs = v[s0 + 1] * v[s0 + 2] * v[s0 + 3] * v[s0 + 7] % ELL
        if lhs != rhs:
            return False
    return True


@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_oracle_network_products_match_python_restatement(case):
    """PolyEvalNetwork::new (sparse_mlpoly.rs:652-668): hash layers + product circuits, as ProductCircuit::evaluate values."""
    inst = oracle_instance(case)
    got = O.snark_network_products(inst, bytes.fromhex(case["rx"]), bytes.fromhex(case["ry"]), bytes.fromhex(case["r_hash"]),
                                   bytes.fromhex(case["r_multiset_check"]))
    assert [g.hex() for g in got] == case["network_products"]
    assert multiset_identity_holds(got)


def test_oracle_timestamps_are_a_memory_trace():
    """AddrTimestamps::new (sparse_mlpoly.rs:233-266): per cell the read timestamps, taken in op order over the three
    matrices, count 0, 1, 2, ... and the audit timestamp is the number of accesses; padding ops read cell 0."""
    A, B, Cm, _, _ = ragged_r1cs(13, 7, 2)
    rng = random.Random(4)
    for M in (A, B, Cm):
        rng.shuffle(M)
    inst = O.Instance.new(13, 7, 2, A, B, Cm)
    N, cells, ops, mem = O.snark_dense(inst)
    val = lambda b, i: int.from_bytes(b[32 * i:32 * i + 32], "little")
    for base_addr, base_ts, audit0 in ((0, 3 * N, 0), (6 * N, 9 * N, cells)):
        seen = [0] * cells
        for i in range(3 * N):  # matrix-major, op order
            a, ts = val(ops, base_addr + i), val(ops, base_ts + i)
            assert ts == seen[a]
            seen[a] += 1
        assert [val(mem, audit0 + a) for a in range(cells)] == seen
        assert sum(seen) == 3 * N
    # the row addresses of the padding ops are zero, the last block of comb_ops is zero padding
    for m, M in enumerate((A, B, Cm)):
        assert all(val(ops, m * N + i) == 0 for i in range(len(M), N))
    assert ops[32 * 15 * N:] == bytes(32 * N)
    # val block: the entries in the caller's order
    assert [val(ops, 12 * N + i) for i in range(len(A))] == [int.from_bytes(e[2], "little") for e in A]


def test_oracle_asserts_like_the_reference():
    """[Scalar]::commit asserts gens_n.n == row length (commitments.rs:81): generators for another num_nz_entries panic."""
    inst = O.Instance.synthetic(16, 16, 3, 1)[0]
    O.snark_encode(inst, 16)
    with pytest.raises(RuntimeError):
        O.snark_encode(inst, 64)


def test_commitment_wire_form_parses():
    import otti_b200
    case = GOLDEN[0]
    cc = otti_b200.ComputationCommitment(bytes.fromhex(case["commitment"]))
    n = 1 << case["log_size"]
    assert (cc.num_cons, cc.num_vars, cc.num_inputs, cc.batch_size) == (n, n, case["num_inputs"], 3)
    assert cc.num_ops == n and cc.num_mem_cells == 2 * n
    assert len(cc.comm_comb_ops) * len(cc.comm_comb_ops) in (16 * n, 8 * n) and len(cc.comm_comb_mem) ** 2 in (4 * n, 2 * n)
    with pytest.raises(ValueError):
        otti_b200.ComputationCommitment(cc.bytes + b"\0")


# ------------------------------------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def sb():
    import otti_b200
    L = otti_b200.lib()
    assert L.sb200_init(0) == 0, L.sb200_last_error()
    return otti_b200


def product_instance(sb, case):
    if case["kind"] == "synthetic":
        n = 1 << case["log_size"]
        return sb.Instance.produce_synthetic_r1cs(n, n, case["num_inputs"], seed=case["seed"])[0], (n, n, case["num_inputs"])
    nc, nv, ni, A, B, Cm = case_coo(case)
    return sb.Instance.new(nc, nv, ni, A, B, Cm), (nc, nv, ni)


def product_dense(sb, inst):
    d = (C.c_uint64 * 2)()
    assert sb.lib().sb200_k_snark_dense(inst.h, d, None, None) == 0, sb.lib().sb200_last_error()
    ops, mem = (C.c_uint8 * (32 * 16 * d[0]))(), (C.c_uint8 * (32 * 2 * d[1]))()
    assert sb.lib().sb200_k_snark_dense(inst.h, d, ops, mem) == 0, sb.lib().sb200_last_error()
    return d[0], d[1], bytes(ops), bytes(mem)


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_encode_matches_golden(sb, case):
    inst, (nc, nv, ni) = product_instance(sb, case)
    N, cells, ops, mem = product_dense(sb, inst)
    assert hashlib.sha256(ops).hexdigest() == case["comb_ops_sha256"]
    assert hashlib.sha256(mem).hexdigest() == case["comb_mem_sha256"]
    # SNARKGens::new takes log2 of the num_cons it is given (r1csinstance.rs:37) while the instance pads num_cons to a
    # power of two (lib.rs:143-156): as in the reference, generators made from an unpadded count whose log2 decides
    # num_vars_mem do not fit the instance (the assert of commitments.rs:81) -- the golden vectors use the padded sizes
    if (1 << (nc.bit_length() - 1)) != nc and inst.num_cons > 2 * inst.num_vars:
        with pytest.raises(AssertionError):
            sb.SNARK.encode(inst, sb.SNARKGens.new(nc, nv, ni, case["num_nz_entries"]))
    gens = sb.SNARKGens.new(inst.num_cons, nv, ni, case["num_nz_entries"])
    cc = sb.SNARK.encode(inst, gens)
    assert cc.bytes.hex() == case["commitment"]
    assert sb.SNARK.encode(inst, gens).bytes == cc.bytes  # deterministic (no tape, no transcript)


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_commit_derefs_matches_golden(sb, case):
    inst, (nc, nv, ni) = product_instance(sb, case)
    gens = sb.SNARKGens.new(inst.num_cons, nv, ni, case["num_nz_entries"])
    pts, comb = sb.SNARK.commit_derefs(inst, gens, bytes.fromhex(case["rx"]), bytes.fromhex(case["ry"]), want_comb=True)
    assert hashlib.sha256(comb).hexdigest() == case["derefs_comb_sha256"]
    assert b"".join(pts).hex() == case["derefs_commitment"]


@pytest.mark.gpu
@pytest.mark.parametrize("s", [10, 16])
def test_commit_derefs_matches_oracle(sb, s):
    """Full-size scalars in every term (eq-table values): the commitment kernel's dense case.  2^16 = configs[1]."""
    from common import canon, rand_scalars
    n = 1 << s
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
    oinst = O.Instance.synthetic(n, n, 10, 0x0771)[0]
    rx, ry = canon(rand_scalars(s, 40 + s)), canon(rand_scalars(s + 1, 50 + s))
    gens = sb.SNARKGens.new(n, n, 10, n)
    pts, comb = sb.SNARK.commit_derefs(inst, gens, rx, ry, want_comb=True)
    opts, ocomb = O.snark_commit_derefs(oinst, n, rx, ry, threads=os.cpu_count() or 1, want_comb=True)
    assert comb == ocomb
    assert b"".join(pts) == opts
    with pytest.raises(AssertionError):  # a point shorter than the memory: mem_val[a] out of bounds in the reference
        sb.SNARK.commit_derefs(inst, gens, rx[:32 * (s - 1)], ry[:32 * (s - 1)])
    with pytest.raises(sb.R1CSError):
        sb.SNARK.commit_derefs(inst, gens, (ELL).to_bytes(32, "little") * s, ry)


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLDEN, ids=case_id)
def test_network_products_match_golden(sb, case):
    inst, _ = product_instance(sb, case)
    got = sb.SNARK.network_products(inst, bytes.fromhex(case["rx"]), bytes.fromhex(case["ry"]), bytes.fromhex(case["r_hash"]),
                                    bytes.fromhex(case["r_multiset_check"]))
    assert [g.hex() for g in got] == case["network_products"]


@pytest.mark.gpu
@pytest.mark.parametrize("s", [10, 13, 16])
def test_network_products_match_oracle(sb, s):
    """Sizes on both sides of the single-block tail of the product trees; 2^16 = configs[1]."""
    from common import canon, rand_scalars
    n = 1 << s
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
    oinst = O.Instance.synthetic(n, n, 10, 0x0771)[0]
    rx, ry = canon(rand_scalars(s, 60 + s)), canon(rand_scalars(s + 1, 70 + s))
    rh, rm = canon(rand_scalars(1, 80 + s)), canon(rand_scalars(1, 90 + s))
    got = sb.SNARK.network_products(inst, rx, ry, rh, rm)
    assert got == O.snark_network_products(oinst, rx, ry, rh, rm)
    assert multiset_identity_holds(got)
    with pytest.raises(AssertionError):  # eval_table.len() != audit_ts.len()
        sb.SNARK.network_products(inst, rx, ry[:32 * s], rh, rm)


@pytest.mark.gpu
def test_network_products_full_size_property(sb):
    """2^20 (the README's SNARK profile size), Otti-like unequal matrices are covered above: here only the size-independent
    property -- the memory-checking multiset identity init * prod(write) == prod(read) * audit on both sides."""
    from common import canon, rand_scalars
    s = 20
    n = 1 << s
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
    got = sb.SNARK.network_products(inst, canon(rand_scalars(s, 1)), canon(rand_scalars(s + 1, 2)), canon(rand_scalars(1, 3)),
                                    canon(rand_scalars(1, 4)))
    assert multiset_identity_holds(got)
    # in the synthetic instance op i of A, B and C all touch row i, so on the row side write[A] == read[B], write[B] == read[C]
    assert got[4] == got[2] and got[5] == got[3] and len(set(got)) == 14


@pytest.mark.gpu
@pytest.mark.parametrize("s", [8, 12])
def test_encode_synthetic_matches_oracle(sb, s):
    n = 1 << s
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
    oinst = O.Instance.synthetic(n, n, 10, 0x0771)[0]
    assert product_dense(sb, inst) == O.snark_dense(oinst)
    cc = sb.SNARK.encode(inst, sb.SNARKGens.new(n, n, 10, n))
    assert cc.bytes == O.snark_encode(oinst, n, threads=os.cpu_count() or 1)
    assert (cc.num_ops, cc.num_mem_cells) == (n, 2 * n)


@pytest.mark.gpu
@pytest.mark.parametrize("nc,nv,ni,seed", [(700, 300, 4, 1), (257, 1500, 10, 2), (1024, 1024, 0, 3)])
def test_encode_otti_shaped_matches_oracle(sb, nc, nv, ni, seed):
    """Rows of 1..64 entries, a heavy constant column, num_cons != num_vars, entries in shuffled order, different
    numbers of entries per matrix (padding ops read cell 0)."""
    cons, _, _ = otti_shaped_r1cs(nc, nv, ni, seed=seed)
    A, B, Cm = coo_of(cons)
    rng = random.Random(seed)
    for M in (A, B, Cm):
        rng.shuffle(M)
    nz = max(len(A), len(B), len(Cm))
    inst = sb.Instance.new(nc, nv, ni, A, B, Cm)
    oinst = O.Instance.new(nc, nv, ni, A, B, Cm)
    assert product_dense(sb, inst) == O.snark_dense(oinst)
    cc = sb.SNARK.encode(inst, sb.SNARKGens.new(inst.num_cons, nv, ni, nz))
    assert cc.bytes == O.snark_encode(oinst, nz, threads=os.cpu_count() or 1)


@pytest.mark.gpu
def test_encode_error_behaviour(sb):
    n = 64
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 3, seed=5)[0]
    with pytest.raises(AssertionError):  # commitments.rs:81
        sb.SNARK.encode(inst, sb.SNARKGens.new(n, n, 3, 8 * n))
    with pytest.raises(AssertionError):  # r1csinstance.rs:36
        sb.SNARKGens.new(0, 4, 3, 4)
    one = (1).to_bytes(32, "little")
    tiny = sb.Instance.new(1, 2, 1, [(0, 0, one)], [(0, 1, one)], [(0, 2, one)])  # lib.rs:191-197 pads such an instance
    with pytest.raises(sb.BackendError):
        sb.SNARK.encode(tiny, sb.SNARKGens.new(2, 2, 1, 2))
    shard = sb.Instance.produce_synthetic_r1cs(n, n, 3, seed=5, rank=0, world=2)[0]
    with pytest.raises(sb.BackendError):
        sb.SNARK.encode(shard, sb.SNARKGens.new(n, n, 3, n))


@pytest.mark.gpu
def test_encode_full_size_dense_and_memory_rows(sb):
    """configs[1] size (2^16): the dense representation against the oracle's, the whole commitment against the oracle
    with every host core (the small scalars make its MSM quick), and the profile keys."""
    n = 1 << 16
    inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
    oinst = O.Instance.synthetic(n, n, 10, 0x0771)[0]
    got, want = product_dense(sb, inst), O.snark_dense(oinst)
    assert got[:2] == want[:2] and hashlib.sha256(got[2]).digest() == hashlib.sha256(want[2]).digest()
    assert got[3] == want[3]
    cc = sb.SNARK.encode(inst, sb.SNARKGens.new(n, n, 10, n))
    assert cc.bytes == O.snark_encode(oinst, n, threads=os.cpu_count() or 1)
    prof = sb.SNARK.last_encode_profile()
    assert prof["commit_ops_ms"] > 0 and prof["total_ms"] >= prof["commit_ops_ms"]
    head = struct.unpack_from("<7Q", cc.bytes, 0)
    assert head == (n, n, 10, 3, n, 2 * n, 1024)
