"""Parity of every sm_100a kernel and of the whole NIZK::prove against the CPU oracle.

All calls go through the C ABI (include/spartan_b200.h, spartan_b200_kernels.h).
Bar: bit-exact (integer / byte work).
"""
import ctypes as C
import os
import random

import pytest

import oracle_lib as O
from common import ELL, LABEL, TAPE_SEED, canon, mont, ragged_r1cs, rand_scalars

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sb():
    import otti_b200
    L = otti_b200.lib()
    assert L.sb200_init(0) == 0, L.sb200_last_error()
    return otti_b200


def _buf(b):
    return (C.c_uint8 * max(len(b), 1)).from_buffer_copy(bytes(b)) if len(b) else (C.c_uint8 * 1)()


def _out(n):
    return (C.c_uint8 * max(n, 1))()


EDGE = [0, 1, 2, ELL - 1, ELL - 2, 2**252, 2**252 - 1, 2**128, 2**64 - 1, (ELL - 1) // 2]


def test_fl_binops(sb):
    """Scalar add/sub/mul on the device == ristretto255.rs:702-755 (via the oracle), incl. edge values."""
    n = 4096
    xs = EDGE + rand_scalars(n - len(EDGE), 1)
    ys = list(reversed(EDGE)) + rand_scalars(n - len(EDGE), 2)
    a, b = mont(xs), mont(ys)
    for op, name in ((0, "add"), (1, "sub"), (2, "mul")):
        o = _out(32 * n)
        assert sb.lib().sb200_k_fl_binop(op, _buf(a), _buf(b), o, C.c_size_t(n)) == 0
        got = bytes(o)
        for i in range(0, n, 97):
            assert got[32 * i:32 * i + 32] == O.fl_binop(name, a[32 * i:32 * i + 32], b[32 * i:32 * i + 32]), (name, i)
        f = {"add": lambda x, y: (x + y) % ELL, "sub": lambda x, y: (x - y) % ELL, "mul": lambda x, y: x * y % ELL}[name]
        assert got == mont([f(x, y) for x, y in zip(xs, ys)])


def test_fp_binops(sb):
    """Coordinate field 2^255-19 on the device (PTX carry chains) == Python integers, on loose 256-bit inputs."""
    import random
    P = 2**255 - 19
    rnd = random.Random(5)
    edge = [0, 1, 2, 18, 19, 37, 38, 39, P - 1, P, P + 1, 2 * P, 2 * P + 37, 2**256 - 1, 2**256 - 38, 2**256 - 39, 2**255,
            2**255 - 1, 2**224, 2**32 - 1, 2**64 - 1, (2**256 - 1) ^ (2**128 - 1), 2**128 - 1]
    n = 4096
    xs = edge + [rnd.getrandbits(256) for _ in range(n - len(edge))]
    ys = list(reversed(edge)) + [rnd.getrandbits(256) for _ in range(n - len(edge))]
    for i in range(len(edge)):  # every edge pair at least once
        for j in range(len(edge)):
            xs.append(edge[i]); ys.append(edge[j])
    n = len(xs)
    a = b"".join(x.to_bytes(32, "little") for x in xs)
    b = b"".join(y.to_bytes(32, "little") for y in ys)
    ops = {0: lambda x, y: (x + y) % P, 1: lambda x, y: (x - y) % P, 2: lambda x, y: x * y % P, 3: lambda x, y: x * x % P,
           4: lambda x, y: pow(x % P, P - 2, P)}
    for op, f in ops.items():
        o = _out(32 * n)
        assert sb.lib().sb200_k_fp_binop(op, _buf(a), _buf(b), o, C.c_size_t(n)) == 0
        got = bytes(o)
        for i, (x, y) in enumerate(zip(xs, ys)):
            assert int.from_bytes(got[32 * i:32 * i + 32], "little") == f(x, y), (op, i, hex(x), hex(y))


def test_to_from_mont_and_canonicity(sb):
    """Scalar::from_bytes / to_bytes (ristretto255.rs:398-438) incl. rejection of values >= l."""
    xs = EDGE + rand_scalars(1000, 3)
    o = _out(32 * len(xs))
    bad = C.c_int(7)
    assert sb.lib().sb200_k_to_mont(_buf(canon(xs)), o, C.c_size_t(len(xs)), C.byref(bad)) == 0
    assert bad.value == 0
    want, _ = O.fl_batch_from_bytes(canon(xs))
    assert bytes(o) == want
    back = _out(32 * len(xs))
    assert sb.lib().sb200_k_from_mont(o, back, C.c_size_t(len(xs))) == 0
    assert bytes(back) == canon(xs)
    for v in (ELL, ELL + 1, 2**256 - 1):
        assert sb.lib().sb200_k_to_mont(_buf(canon([5, v, 6])), _out(96), C.c_size_t(3), C.byref(bad)) == 0
        assert bad.value == 1


@pytest.mark.parametrize("ell", [0, 1, 2, 3, 7, 8, 9, 10, 13, 16])
def test_eq_evals(sb, ell):
    """EqPolynomial::evals (dense_mlpoly.rs:67-83)."""
    r = mont(rand_scalars(ell, 10 + ell))
    o = _out(32 << ell)
    assert sb.lib().sb200_k_eq_evals(_buf(r), ell, o) == 0
    assert bytes(o) == O.eq_evals(r)


def _random_instance(num_cons, num_vars, num_inputs, seed, max_row=6, heavy=True):
    """Ragged R1CS: empty rows, multi-entry rows, duplicate (row,col) entries, a heavy constant column."""
    rng = random.Random(seed)
    ncols = num_vars + 1 + num_inputs
    mats = []
    for m in range(3):
        M = []
        for row in range(num_cons):
            k = rng.choice([0, 1, 1, 1, 2, 3, max_row])
            for _ in range(k):
                col = rng.randrange(ncols)
                if heavy and rng.random() < 0.3:
                    col = num_vars  # the constant-1 column
                val = rng.choice([1, ELL - 1, 2, rng.randrange(ELL), rng.randrange(2**32)])
                M.append((row, col, int(val).to_bytes(32, "little")))
        if M:
            M.append(M[0])  # duplicate entry: contributions must add up
        mats.append(M)
    return mats


@pytest.mark.parametrize("dims", [(8, 8, 3), (37, 20, 5), (256, 256, 10), (1000, 700, 17), (1, 0, 3), (2, 1, 0)])
def test_multiply_vec_and_eval_table(sb, dims):
    """R1CSInstance::multiply_vec and compute_eval_table_sparse+combine on ragged instances
    (r1csinstance.rs:248-278, sparse_mlpoly.rs:452-483, r1csproof.rs:268-279)."""
    nc, nv, ni = dims
    A, B, Cm = _random_instance(nc, nv, ni, 100 + nc)
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    pi = sb.Instance.new(nc, nv, ni, A, B, Cm)
    assert (pi.num_cons, pi.num_vars, pi.num_inputs) == (oi.num_cons, oi.num_vars, oi.num_inputs)
    z = mont(rand_scalars(2 * oi.num_vars, 5))
    oa, ob, oc = oi.multiply_vec(z)
    ga, gb, gc = _out(len(oa)), _out(len(oa)), _out(len(oa))
    assert sb.lib().sb200_k_multiply_vec(pi.h, _buf(z), ga, gb, gc) == 0
    assert (bytes(ga), bytes(gb), bytes(gc)) == (oa, ob, oc)
    e = mont(rand_scalars(oi.num_cons, 6))
    rabc = mont(rand_scalars(3, 7))
    want = oi.eval_table_abc(e, rabc)
    got = _out(len(want))
    assert sb.lib().sb200_k_eval_table_abc(pi.h, _buf(e), _buf(rabc), got) == 0
    assert bytes(got) == want


def test_sparse_kernels_long_segments(sb):
    """Rows / columns longer than one thread's share (dev.hpp kLongSeg = 16, kSegChunk = 512, kSumInline = 8): segments of
    exactly 16, 17, 512, 513, 4096, 4097 and 9000 entries, as rows AND as columns, coefficients +1 / -1 / general
    (the +-1 fast path), duplicates included.  Same bytes as the oracle's insertion-order loops."""
    rng = random.Random(99)
    nc, nv, ni = 9500, 9400, 6
    ncols = nv + 1 + ni
    lens = [16, 17, 512, 513, 4096, 4097, 9000]
    val = lambda: rng.choice([1, 1, ELL - 1, 2, rng.randrange(ELL), rng.randrange(2**32)]).to_bytes(32, "little")
    mats = []
    for m in range(3):
        M = []
        for k, ln in enumerate(lens):
            row = 3 * k + m  # a long row: ln entries in distinct columns (plus one duplicate)
            cols = rng.sample(range(ncols), min(ln, ncols))
            M += [(row, c, val()) for c in cols] + [(row, cols[0], val())]
            col = 40 + 5 * k + m  # a long column: ln entries in distinct rows
            rows = rng.sample(range(nc), min(ln, nc))
            M += [(r, col if col != nv else col + 1, val()) for r in rows]
        M += [(rng.randrange(nc), nv, val()) for _ in range(7000)]  # the constant column on top
        M += [(rng.randrange(nc), rng.randrange(ncols), val()) for _ in range(5000)]
        rng.shuffle(M)
        mats.append(M)
    A, B, Cm = mats
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    pi = sb.Instance.new(nc, nv, ni, A, B, Cm)
    z = mont(rand_scalars(2 * oi.num_vars, 5))
    oa, ob, oc = oi.multiply_vec(z)
    ga, gb, gc = _out(len(oa)), _out(len(oa)), _out(len(oa))
    assert sb.lib().sb200_k_multiply_vec(pi.h, _buf(z), ga, gb, gc) == 0
    assert (bytes(ga), bytes(gb), bytes(gc)) == (oa, ob, oc)
    e = mont(rand_scalars(oi.num_cons, 6))
    rabc = mont(rand_scalars(3, 7))
    want = oi.eval_table_abc(e, rabc)
    got = _out(len(want))
    assert sb.lib().sb200_k_eval_table_abc(pi.h, _buf(e), _buf(rabc), got) == 0
    assert bytes(got) == want


@pytest.mark.parametrize("c", [2, 5, 8, 11, 13, 16])
def test_window_sweep_short_scalars(sb, c):
    """The commitment kernel stops a short scalar after (bitlen-1)/c + 2 windows (kernels.cu msm_fixed_kernel): for every
    window width, scalars 2^k - 1, 2^k, 2^k + 1 around the multiples of c (where the recoding carry crosses a window
    boundary), all-ones words, bits and full-size scalars give the oracle's commitment rows."""
    nv = 256
    og = O.Gens(nv, nv, 1)
    os.environ["SB200_MSM_WINDOW"] = str(c)
    try:
        pg = sb.NIZKGens.new(nv, nv, 1)
    finally:
        os.environ.pop("SB200_MSM_WINDOW", None)
    assert pg.window_bits == c
    vals = [0, 1, 2**32 - 1, 2**64 - 1, ELL - 1, 2**252, 2**252 - 1]
    for m in range(1, 253 // c + 1):
        k = m * c
        for kk in (k - 1, k, k + 1):
            if 0 < kk < 252:
                vals += [2**kk - 1, 2**kk, 2**kk + 1, 2**kk - 2**(kk - 1) if kk > 1 else 1]
    vals += [2**(c - 1), 2**(c - 1) + 1, 2**(c - 1) - 1, (2**c - 1) * (2**c + 1)]
    rng = random.Random(c)
    while len(vals) % 16:
        vals.append(rng.randrange(2))
    vals = vals[:(len(vals) // 16) * 16]
    L = len(vals) // 16
    blinds = rand_scalars(L, 50 + c)
    o = _out(32 * L)
    assert sb.lib().sb200_k_commit_rows(pg.h, _buf(canon(vals)), C.c_size_t(len(vals)), _buf(canon(blinds)), C.c_size_t(L), o) == 0
    want = b"".join(og.commit_rows(mont(vals[16 * i:16 * i + 16]), mont([blinds[i]])) for i in range(L))
    assert bytes(o) == want


@pytest.mark.parametrize("s,expected", [(10, 1), (16, 1), (16, 50), (18, 1)])
def test_one_shot_generators_same_proof(sb, s, expected):
    """sb200_gens_new_for: the hint changes the window width of the tables (and where the host tables of the tail
    generators come from), never a proof byte."""
    n = 1 << s
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x0771)
    g1 = sb.NIZKGens.new(n, n, 10, proofs_expected=expected)
    g0 = sb.NIZKGens.new(n, n, 10)
    assert g1.window_bits <= g0.window_bits
    if expected == 1:
        assert g1.window_bits < 13
    oi, ov, oin = O.Instance.synthetic(n, n, 10, 0x0771)
    want = O.nizk_prove(oi, ov, oin, O.Gens(n, n, 10), LABEL, TAPE_SEED, 8)
    for g in (g1, g0):
        proof = sb.NIZK.prove(pi, pv, pin, g, sb.Transcript.new(LABEL), TAPE_SEED)
        assert proof.to_bytes() == want
        assert proof.verify(pi, pin, sb.Transcript.new(LABEL), g) is None


@pytest.mark.parametrize("ntab", [4, 2])
@pytest.mark.parametrize("logn", [1, 2, 5, 9, 12, 15])
def test_sumcheck_rounds(sb, ntab, logn):
    """A full chain of fused bind+evaluate rounds == bound_poly_var_top + round evaluations of the
    reference (sumcheck.rs:453-484, 618-674; dense_mlpoly.rs:214-221)."""
    n = 1 << logn
    tabs = [mont(rand_scalars(n, 20 + k)) for k in range(ntab)]
    dev = [_buf(t) for t in tabs]
    ptrs = (C.c_void_p * ntab)(*[C.cast(d, C.c_void_p) for d in dev])
    ev = _out(96)
    assert sb.lib().sb200_k_sumcheck_round(ptrs, ntab, C.c_size_t(n), None, ev) == 0
    e = O.sumcheck_round(tabs)
    assert bytes(ev)[:64] == e[0] + e[1]
    if ntab == 4:
        assert bytes(ev)[64:] == e[2]
    length = n
    rng = random.Random(logn)
    while length > 1:
        r = mont([rng.randrange(ELL)])
        assert sb.lib().sb200_k_sumcheck_round(ptrs, ntab, C.c_size_t(length), _buf(r), ev) == 0
        tabs = [O.bind_top(t, r) for t in tabs]
        length //= 2
        for d, t in zip(dev, tabs):
            assert bytes(d)[:32 * length] == t
        if length > 1:
            e = O.sumcheck_round(tabs)
            assert bytes(ev)[:64] == e[0] + e[1]
            if ntab == 4:
                assert bytes(ev)[64:] == e[2]


@pytest.mark.parametrize("shape", [(1, 2), (2, 2), (4, 8), (16, 16), (64, 128), (256, 256)])
def test_poly_bound_and_dot(sb, shape):
    """DensePolynomial::bound and compute_dotproduct (dense_mlpoly.rs:205-212, nizk/mod.rs:430-433)."""
    Ls, Rs = shape
    Z = mont(rand_scalars(Ls * Rs, 31))
    Lv = mont(rand_scalars(Ls, 32))
    o = _out(32 * Rs)
    assert sb.lib().sb200_k_poly_bound(_buf(Z), C.c_size_t(Ls * Rs), _buf(Lv), C.c_size_t(Ls), o) == 0
    assert bytes(o) == O.poly_bound(Z, Lv)
    b = mont(rand_scalars(Rs, 33))
    d = _out(32)
    assert sb.lib().sb200_k_dot(o, _buf(b), C.c_size_t(Rs), d) == 0
    assert bytes(d) == O.dot(bytes(o), b)


@pytest.mark.parametrize("nv", [2, 16, 64, 1024])
def test_generators_and_commit_rows(sb, nv):
    """MultiCommitGens::new derivation (commitments.rs:15-58) and commit_inner rows
    (dense_mlpoly.rs:163-176) incl. zero, small and maximal scalars."""
    og = O.Gens(nv, nv, 1)
    pg = sb.NIZKGens.new(nv, nv, 1)
    n = sb.lib().sb200_k_gens_n(pg.h)
    assert n == og.n
    for which, cnt in ((0, min(n, 8)), (1, 1), (2, 1), (3, 3), (4, 1), (5, 4), (6, 1)):
        for i in range(cnt):
            o = _out(32)
            assert sb.lib().sb200_k_gens_point(pg.h, which, C.c_size_t(i), o) == 0
            assert bytes(o) == og.point(which, i), (which, i)
    ell = nv.bit_length() - 1
    Ls = 1 << (ell // 2)
    xs = rand_scalars(nv, 41)
    for i, v in enumerate([0, 1, ELL - 1, 2**252, 255, 256, 2**64]):
        if i < nv:
            xs[i] = v
    blinds = rand_scalars(Ls, 42)
    blinds[0] = 0
    o = _out(32 * Ls)
    assert sb.lib().sb200_k_commit_rows(pg.h, _buf(canon(xs)), C.c_size_t(nv), _buf(canon(blinds)), C.c_size_t(Ls), o) == 0
    assert bytes(o) == og.commit_rows(mont(xs), mont(blinds))
    c = _out(32)
    assert sb.lib().sb200_k_commit_scalar(pg.h, _buf(canon([xs[-1]])), _buf(canon([blinds[-1]])), c) == 0
    assert bytes(c) == O.pt_msm(mont([xs[-1], blinds[-1]]), og.point(2) + og.point(1))


@pytest.mark.parametrize("c", [2, 5, 8, 11, 13, 16])
def test_commit_rows_short_scalars_every_window_width(sb, c):
    """The commitment kernel stops a short scalar after (bitlen - 1) / c + 2 windows (the recoding carry can reach one
    window above the top bit).  Scalars 2^k - 1, 2^k, 2^k + 1 for k around the multiples of c, all-ones words and
    bits, under window widths where (bitlen - 1) / c sits on a window boundary: a dropped carry would change a row."""
    nv = 256
    old = os.environ.get("SB200_MSM_WINDOW")
    os.environ["SB200_MSM_WINDOW"] = str(c)
    try:
        pg = sb.NIZKGens.new(nv, nv, 1)
    finally:
        if old is None:
            os.environ.pop("SB200_MSM_WINDOW", None)
        else:
            os.environ["SB200_MSM_WINDOW"] = old
    assert pg.window_bits == c
    og = O.Gens(nv, nv, 1)
    xs = []
    for m in range(1, 10):
        k = m * c
        for kk in (k - 1, k, k + 1):
            for dv in (-1, 0, 1):
                v = (1 << kk) + dv
                if 0 <= v < ELL:
                    xs.append(v)
    xs += [0, 1, 2**32 - 1, 2**64 - 1, 2**16 - 1, 2**15, 2**15 + 1, 2**251, 2**252, ELL - 1, ELL - 2]
    rng = rand_scalars(nv, 97)
    xs = (xs + rng)[:nv]
    Ls = 1 << ((nv.bit_length() - 1) // 2)
    blinds = rand_scalars(Ls, 98)
    o = _out(32 * Ls)
    assert sb.lib().sb200_k_commit_rows(pg.h, _buf(canon(xs)), C.c_size_t(nv), _buf(canon(blinds)), C.c_size_t(Ls), o) == 0
    assert bytes(o) == og.commit_rows(mont(xs), mont(blinds))


def test_merlin(sb):
    o = _out(32)
    sb.lib().sb200_k_merlin(_buf(b"test protocol"), C.c_size_t(13), b"some label", _buf(b"some data"), C.c_size_t(9),
                            b"challenge", o, C.c_size_t(32))
    assert bytes(o).hex() == "d5a21972d0d5fe320c0d263fac7fffb8145aa640af6e9bca177c03c7efcf0615"


@pytest.mark.parametrize("s", [1, 2, 3, 4, 5, 6, 8, 10, 12])
def test_nizk_prove_byte_identical_synthetic(sb, s):
    """NIZK::prove on produce_synthetic_r1cs: proof bytes == oracle's, oracle verify accepts."""
    n = 1 << s
    ni = min(10, n - 1)
    oi, ov, oin = O.Instance.synthetic(n, n, ni, 0x0771)
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, ni, 0x0771)
    assert (pv.data, pin.data) == (ov, oin)
    assert pi.is_sat(pv, pin)
    og = O.Gens(n, n, ni)
    pg = sb.NIZKGens.new(n, n, ni)
    want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    assert O.nizk_verify(oi, oin, og, LABEL, proof.to_bytes()) == 0


@pytest.mark.parametrize("dims", [(37, 20, 5), (1, 0, 3), (300, 100, 64), (64, 600, 2), (5, 3, 8)])
def test_nizk_prove_ragged(sb, dims):
    """Non-power-of-two sizes (padding of Instance::new / Assignment::pad), num_vars < num_inputs+1,
    heavy columns, duplicate entries: find a satisfying assignment by construction, prove, compare."""
    nc, nv, ni = dims
    rng = random.Random(nc * 1000 + nv)
    vars_ = [rng.randrange(ELL) for _ in range(nv)]
    inputs = [rng.randrange(ELL) for _ in range(ni)]
    z = vars_ + [1] + inputs
    ncols = len(z)
    A, B, Cm = [], [], []
    for row in range(nc):
        ra = [(rng.randrange(ncols), rng.choice([1, 2, ELL - 1, rng.randrange(ELL)])) for _ in range(rng.choice([1, 2, 3]))]
        rb = [(rng.randrange(ncols), rng.choice([1, rng.randrange(2**20)])) for _ in range(rng.choice([1, 2]))]
        if row % 7 == 0:
            ra.append((nv, 5))  # the constant column
        az = sum(v * z[c] for c, v in ra) % ELL
        bz = sum(v * z[c] for c, v in rb) % ELL
        # C row: one entry on the constant column carrying az*bz
        A += [(row, c, v.to_bytes(32, "little")) for c, v in ra]
        B += [(row, c, v.to_bytes(32, "little")) for c, v in rb]
        Cm.append((row, nv, (az * bz % ELL).to_bytes(32, "little")))
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    pi = sb.Instance.new(nc, nv, ni, A, B, Cm)
    vb, ib = canon(vars_), canon(inputs)
    assert oi.is_sat(vb, ib) == 1
    pv, pin = sb.VarsAssignment.new(vb), sb.InputsAssignment.new(ib)
    assert pi.is_sat(pv, pin)
    og = O.Gens(nc, nv, ni)
    pg = sb.NIZKGens.new(nc, nv, ni)
    want = O.nizk_prove(oi, vb, ib, og, LABEL, TAPE_SEED)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    assert O.nizk_verify(oi, ib, og, LABEL, proof.to_bytes()) == 0
    # an unsatisfying assignment is detected
    bad = canon([(vars_[0] + 1) % ELL] + vars_[1:]) if nv else None
    if bad is not None and any(c == 0 for _, c, _ in A + B):
        assert not pi.is_sat(sb.VarsAssignment.new(bad), pin)


def test_first_error_wins_like_the_reference(sb):
    """Instance::new reports the failure it meets first: matrices in order, entries in order, row/column before the scalar
    (lib.rs:163-187).  The library checks indices on the host and scalars on the device; the oracle is the judge."""
    zero, one = bytes(32), bytes([1]) + bytes(31)
    larger = bytes([0xEE] * 32)
    ok = [(k % 4, k % 8, one) for k in range(12)]

    def with_(base, at, entry):
        L = list(base)
        L[at] = entry
        return L

    bad_s, bad_r, bad_c = (1, 1, larger), (4, 1, zero), (1, 10, zero)
    cases = [
        (with_(ok, 5, bad_s)[:9] + [bad_r] + ok[10:], ok, ok),          # A: scalar at 5, row at 9
        (with_(ok, 3, bad_c)[:7] + [bad_s] + ok[8:], ok, ok),           # A: column at 3, scalar at 7
        (ok, with_(ok, 11, bad_s), with_(ok, 0, bad_r)),                # B scalar before C row
        (with_(ok, 11, bad_s), with_(ok, 0, bad_r), ok),                # A scalar (late) before B row (early)
        (ok, with_(ok, 2, bad_r), with_(ok, 1, bad_s)),                 # B row before C scalar
        (ok, ok, with_(ok, 6, (4, 1, larger))),                         # one entry with both: the row is checked first
        (ok, ok, with_(ok, 6, bad_s)),
    ]
    for A, B, Cm in cases:
        with pytest.raises(ValueError) as oe:
            O.Instance.new(4, 8, 1, A, B, Cm)
        with pytest.raises(sb.R1CSError) as e:
            sb.Instance.new(4, 8, 1, A, B, Cm)
        assert e.value.kind in str(oe.value), (e.value.kind, str(oe.value))


def test_error_behaviour(sb):
    """check_r1cs_invalid_index / check_r1cs_invalid_scalar (lib.rs:607-650) and prove's asserts."""
    zero = bytes(32)
    with pytest.raises(sb.R1CSError) as e:
        sb.Instance.new(4, 8, 1, [(0, 0, zero)], [(100, 1, zero)], [(1, 1, zero)])
    assert e.value.kind == "InvalidIndex"
    larger = bytes([3, 0, 0, 0, 255, 255, 255, 255, 254, 91, 254, 255, 2, 164, 189, 83, 5, 216, 161, 9, 8, 216, 57, 51, 72,
                    125, 157, 41, 83, 167, 237, 115])
    with pytest.raises(sb.R1CSError) as e:
        sb.Instance.new(4, 8, 1, [(0, 0, zero)], [(1, 1, larger)], [(1, 1, zero)])
    assert e.value.kind == "InvalidScalar"
    with pytest.raises(sb.R1CSError) as e:
        sb.VarsAssignment.new([larger])
    assert e.value.kind == "InvalidScalar"
    inst, v, i = sb.Instance.produce_synthetic_r1cs(16, 16, 3, 1)
    gens = sb.NIZKGens.new(16, 16, 3)
    with pytest.raises(AssertionError):  # wrong number of inputs: the reference asserts z.len() == num_cols
        sb.NIZK.prove(inst, v, sb.InputsAssignment.new(i.data[:64]), gens, sb.Transcript.new(LABEL), TAPE_SEED)
    with pytest.raises(sb.R1CSError) as e:
        inst.is_sat(v, sb.InputsAssignment.new(i.data[:64]))
    assert e.value.kind == "InvalidNumberOfInputs"


def test_test_padded_constraints(sb):
    """lib.rs:653-731 test_padded_constraints: 1 constraint, 0 vars, 3 inputs (NIZK part)."""
    one = (1).to_bytes(32, "little")
    A = [(0, 2, one)]
    B = [(0, 2, one)]
    Cm = [(0, 1, one), (0, 0, (ELL - 13).to_bytes(32, "little")), (0, 3, (ELL - 1).to_bytes(32, "little"))]
    inputs = canon([16, 1, 2])
    pi = sb.Instance.new(1, 0, 3, A, B, Cm)
    oi = O.Instance.new(1, 0, 3, A, B, Cm)
    pv, pin = sb.VarsAssignment.new(b""), sb.InputsAssignment.new(inputs)
    assert pi.is_sat(pv, pin)
    pg, og = sb.NIZKGens.new(1, 0, 3), O.Gens(1, 0, 3)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == O.nizk_prove(oi, b"", inputs, og, LABEL, TAPE_SEED)
    assert O.nizk_verify(oi, inputs, og, LABEL, proof.to_bytes()) == 0


def test_transcript_state_advances_and_tape_randomness(sb):
    """The caller's transcript is mutated in place (lib.rs:497-504); with tape_seed=None two proofs differ
    (random.rs:11-19) yet both verify."""
    inst, v, i = sb.Instance.produce_synthetic_r1cs(64, 64, 5, 9)
    oi, ov, oin = O.Instance.synthetic(64, 64, 5, 9)
    gens, og = sb.NIZKGens.new(64, 64, 5), O.Gens(64, 64, 5)
    t = sb.Transcript.new(LABEL)
    p1 = sb.NIZK.prove(inst, v, i, gens, t, TAPE_SEED)
    fresh = sb.Transcript.new(LABEL)
    assert t.challenge_bytes(b"x", 16) != fresh.challenge_bytes(b"x", 16)
    pa = sb.NIZK.prove(inst, v, i, gens, sb.Transcript.new(LABEL))
    pb = sb.NIZK.prove(inst, v, i, gens, sb.Transcript.new(LABEL))
    assert pa.to_bytes() != pb.to_bytes()
    for p in (p1, pa, pb):
        assert O.nizk_verify(oi, oin, og, LABEL, p.to_bytes()) == 0
    assert p1.r1cs_sat_proof_len == len(p1.to_bytes()) - (8 + 32 * 6) - (8 + 32 * 7)


@pytest.mark.parametrize("s", [1, 4, 6, 10])
def test_nizk_verify_accepts_and_rejects(sb, s):
    """NIZK::verify (lib.rs:538-574) on the device path: accepts the GPU proof AND the oracle's proof,
    agrees with the oracle verifier on tampered proofs, wrong inputs, wrong label."""
    n = 1 << s
    ni = min(10, n - 1)
    oi, ov, oin = O.Instance.synthetic(n, n, ni, 77)
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, ni, 77)
    og, pg = O.Gens(n, n, ni), sb.NIZKGens.new(n, n, ni)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.verify(pi, pin, sb.Transcript.new(LABEL), pg) is None
    oracle_proof = O.nizk_prove(oi, ov, oin, og, LABEL, bytes(64))  # different tape seed -> different valid proof
    assert sb.NIZK(oracle_proof).verify(pi, pin, sb.Transcript.new(LABEL), pg) is None
    b = proof.to_bytes()
    rng = random.Random(s)
    for off in [40, len(b) - 1, len(b) // 2] + [rng.randrange(8, len(b)) for _ in range(12)]:
        bad = bytearray(b)
        bad[off] ^= 1 << rng.randrange(8)
        want = O.nizk_verify(oi, oin, og, LABEL, bytes(bad))
        if want == 0:
            continue  # (a flipped bit inside a length prefix may still parse to the same proof -- never happens in practice)
        with pytest.raises(sb.ProofVerifyError):
            sb.NIZK(bytes(bad)).verify(pi, pin, sb.Transcript.new(LABEL), pg)
    with pytest.raises(sb.ProofVerifyError):
        proof.verify(pi, pin, sb.Transcript.new(b"other"), pg)
    if ni:
        wrong = canon([(int.from_bytes(oin[:32], "little") + 1) % ELL]) + oin[32:]
        with pytest.raises(sb.ProofVerifyError):
            proof.verify(pi, sb.InputsAssignment.new(wrong), sb.Transcript.new(LABEL), pg)
    with pytest.raises(sb.ProofVerifyError):
        sb.NIZK(b[:-1]).verify(pi, pin, sb.Transcript.new(LABEL), pg)


def test_large_round_trip_properties(sb):
    """BASELINE size (2^16, configs[1]) and 2^18: proof length equals the closed form (App. A.4), prove -> verify
    round trip on the device, commitment rows are linear in the witness rows (homomorphism check on 2 rows)."""
    for s in (16, 18):
        n = 1 << s
        pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, 10, 5)
        pg = sb.NIZKGens.new(n, n, 10)
        proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
        lL = s // 2
        dpp = lambda k: 32 + 32 + 8 + 32 * k + 32 + 32
        sc = lambda rounds, k: (8 + 32 * rounds) * 2 + 8 + rounds * dpp(k)
        want_len = (8 + 32 * (1 << lL)) + sc(s, 4) + 128 + 352 + 64 + sc(s + 1, 3) + 32 + (2 * (8 + 32 * (s - lL)) + 128) + 64
        assert proof.r1cs_sat_proof_len == want_len
        assert proof.verify(pi, pin, sb.Transcript.new(LABEL), pg) is None
    # 2^16 byte-identity against the oracle (the oracle needs ~2 s here)
    oi, ov, oin = O.Instance.synthetic(1 << 16, 1 << 16, 10, 5)
    og = O.Gens(1 << 16, 1 << 16, 10)
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(1 << 16, 1 << 16, 10, 5)
    pg = sb.NIZKGens.new(1 << 16, 1 << 16, 10)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED, 8)
    # Pedersen homomorphism: Com(x; r) + Com(y; s) == Com(x + y; r + s), row-wise on the device MSM
    R = 256
    xs, ys = rand_scalars(R, 1), rand_scalars(R, 2)
    zs = [(a + b) % ELL for a, b in zip(xs, ys)]
    bl = rand_scalars(2, 3)
    o = _out(96)
    Z = canon(xs + ys + zs + [0] * R)
    blinds = canon([bl[0], bl[1], (bl[0] + bl[1]) % ELL, 0])
    assert sb.lib().sb200_k_commit_rows(pg.h, _buf(Z), C.c_size_t(4 * R), _buf(blinds), C.c_size_t(4), _out(128)) == 0
    out = _out(128)
    sb.lib().sb200_k_commit_rows(pg.h, _buf(Z), C.c_size_t(4 * R), _buf(blinds), C.c_size_t(4), out)
    out = bytes(out)
    assert O.pt_add(out[:32], out[32:64]) == out[64:96]
    assert out[96:] == bytes(32)  # the all-zero row commits to the identity


def test_pipelined_witness_upload(sb):
    """A witness of >= 8 MiB coming from host memory is copied in pieces and committed piece by piece
    (prover.inc): same bytes as the single-copy path and as the oracle, also when the padded tail of a
    non-power-of-two witness cuts through the pieces."""
    s = 18
    n = 1 << s
    # (a) synthetic 2^18, pieces 1 / default / maximum
    oi, ov, oin = O.Instance.synthetic(n, n, 10, 11)
    og = O.Gens(n, n, 10)
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, 10, 11)
    pg = sb.NIZKGens.new(n, n, 10)
    want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED, 8)
    got = {}
    try:
        for pieces in ("1", None, "8", "3"):
            if pieces is None:
                os.environ.pop("SB200_H2D_PIECES", None)
            else:
                os.environ["SB200_H2D_PIECES"] = pieces
            got[pieces] = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED).to_bytes()
    finally:
        os.environ.pop("SB200_H2D_PIECES", None)
    for k, v in got.items():
        assert v == want, "pieces=%s" % k
    # (b) few constraints, 2^18 - 4133 variables: the zero padding starts inside the last piece
    nc, nv, ni = 48, n - 4133, 3
    A, B, Cm, vars_, inputs = ragged_r1cs(nc, nv, ni)
    oi = O.Instance.new(nc, nv, ni, A, B, Cm)
    pi = sb.Instance.new(nc, nv, ni, A, B, Cm)
    vb, ib = canon(vars_), canon(inputs)
    pv, pin = sb.VarsAssignment.new(vb), sb.InputsAssignment.new(ib)
    assert pi.is_sat(pv, pin)
    og = O.Gens(nc, nv, ni)
    pg = sb.NIZKGens.new(nc, nv, ni)
    want = O.nizk_prove(oi, vb, ib, og, LABEL, TAPE_SEED, 8)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    assert proof.to_bytes() == want
    assert proof.verify(pi, pin, sb.Transcript.new(LABEL), pg) is None


@pytest.mark.parametrize("s", [20, 22])
def test_full_size_properties(sb, s):
    """BASELINE.json's full sizes (2^20 of the README comparison, 2^22 of configs[4]), where the oracle is too slow
    to be the checker: size-independent properties.  Proof length = closed form (App. A.4); prove -> verify round trip
    on the device; determinism (same tape seed -> same bytes, resident witness == witness from host memory, i.e. the
    pipelined upload); a different tape seed gives a different proof that also verifies; a flipped proof byte, a wrong
    input and an unsatisfying witness are rejected."""
    n = 1 << s
    pi, pv, pin = sb.Instance.produce_synthetic_r1cs(n, n, 10, 0x51 + s)
    pg = sb.NIZKGens.new(n, n, 10)
    assert pi.is_sat(pv, pin)
    proof = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED)
    b = proof.to_bytes()
    lL = s // 2
    dpp = lambda k: 32 + 32 + 8 + 32 * k + 32 + 32
    sc = lambda rounds, k: (8 + 32 * rounds) * 2 + 8 + rounds * dpp(k)
    want_len = (8 + 32 * (1 << lL)) + sc(s, 4) + 128 + 352 + 64 + sc(s + 1, 3) + 32 + (2 * (8 + 32 * (s - lL)) + 128) + 64
    assert proof.r1cs_sat_proof_len == want_len
    if s == 20:
        assert want_len == 47024  # Spartan/README.md:376, the only byte-level figure the reference publishes
    assert proof.verify(pi, pin, sb.Transcript.new(LABEL), pg) is None
    # the oracle sees the full-size proof: its own verifier accepts the GPU's bytes, and its own prover (all host
    # threads; ~4 s at 2^20, ~15-25 s at 2^22) produces the same bytes
    oi, ov, oin = O.Instance.synthetic(n, n, 10, 0x51 + s)
    assert ov == bytes(pv.data) and oin == bytes(pin.data)
    og = O.Gens(n, n, 10)
    assert O.nizk_verify(oi, oin, og, LABEL, b) == 0
    assert O.nizk_prove(oi, ov, oin, og, LABEL, TAPE_SEED, os.cpu_count() or 8) == b
    del oi, og
    os.environ["SB200_H2D_PIECES"] = "1"
    try:
        assert sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), TAPE_SEED).to_bytes() == b
    finally:
        os.environ.pop("SB200_H2D_PIECES", None)
    other = sb.NIZK.prove(pi, pv, pin, pg, sb.Transcript.new(LABEL), bytes(64))
    assert other.to_bytes() != b and len(other.to_bytes()) == len(b)
    assert other.verify(pi, pin, sb.Transcript.new(LABEL), pg) is None
    oi, ov, oin = O.Instance.synthetic(n, n, 10, 0x51 + s)
    assert O.nizk_verify(oi, oin, O.Gens(n, n, 10), LABEL, other.to_bytes()) == 0
    del oi
    rng = random.Random(s)
    for off in (40, len(b) // 3, len(b) - 5):
        bad = bytearray(b)
        bad[off] ^= 1 << rng.randrange(8)
        with pytest.raises(sb.ProofVerifyError):
            sb.NIZK(bytes(bad)).verify(pi, pin, sb.Transcript.new(LABEL), pg)
    wrong = canon([(int.from_bytes(bytes(pin.data[:32]), "little") + 1) % ELL]) + bytes(pin.data[32:])
    with pytest.raises(sb.ProofVerifyError):
        proof.verify(pi, sb.InputsAssignment.new(wrong), sb.Transcript.new(LABEL), pg)
    vb = bytearray(bytes(pv.data))
    vb[0] ^= 1
    assert not pi.is_sat(sb.VarsAssignment.new(bytes(vb)), pin)


@pytest.mark.parametrize("env", [{"SB200_SC_LAT_PAIRS": "0", "SB200_SC_HALF_ZERO": "2"},
                                 {"SB200_SC_LAT_PAIRS": "0", "SB200_SC_HALF_ZERO": "2", "SB200_HOST_TAIL_LOG": "0"},
                                 {"SB200_SC_EQ_FACTOR": "0", "SB200_SC_HALF_ZERO": "1"}, {"SB200_SC_HALF_ZERO": "0"},
                                 {"SB200_SC_LAT_PAIRS": "0", "SB200_SC_HALF_ZERO": "2", "SB200_SC_HALF_ZERO_FALLBACK": "1"}])
def test_factored_first_sumcheck_gives_the_same_proof(env):
    """Phase one's long rounds run with eq(tau) factored out of the round polynomial (kernels.cu sc_cubic_eq_kernel,
    prover.inc EqFactor).  By default that covers the rounds above 8192 pairs (instances from 2^15 on, which the full-size
    tests compare with the oracle); here SB200_SC_LAT_PAIRS=0 pushes every device round of small instances through it, and
    SB200_SC_EQ_FACTOR=0 switches it off: the proof bytes stay the oracle's (sumcheck.rs:618-674).  The same runs cover phase
    two's first rounds, which skip the (almost empty) upper half of z and keep the factor 1 - r_0 outside the tables
    (sc_quad_half0_kernel, prover.inc HalfZero; SB200_SC_HALF_ZERO = least pairs per rank, default 2^17, 0 = off) -- with 3, 10 and 100 inputs, and with the final
    claims read from the device (SB200_HOST_TAIL_LOG=0) or from the host's tail rounds, and through the path kept for the
    challenge r_0 = 1 (SB200_SC_HALF_ZERO_FALLBACK=1: the ordinary bind after clearing what was left unwritten).  The variables are read once per
    process, hence the subprocess."""
    import hashlib
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    shapes = [(1 << 12, 1 << 12, 10), (1 << 13, 1 << 12, 3), (1 << 14, 1 << 15, 100), (1 << 16, 1 << 16, 10)]
    want = []
    for nc, nv, ni in shapes:
        oi, ov, oin = O.Instance.synthetic(nc, nv, ni, 0x5eed)
        want.append(hashlib.sha256(O.nizk_prove(oi, ov, oin, O.Gens(nc, nv, ni), LABEL, TAPE_SEED, 8)).hexdigest())
    code = ("import sys; sys.path.insert(0, %r); import hashlib, otti_b200 as sb\n"
            "for nc, nv, ni in %r:\n"
            "    inst, v, i = sb.Instance.produce_synthetic_r1cs(nc, nv, ni, 0x5eed)\n"
            "    g = sb.NIZKGens.new(nc, nv, ni)\n"
            "    p = sb.NIZK.prove(inst, v, i, g, sb.Transcript.new(%r), bytes(%r))\n"
            "    print(hashlib.sha256(p.to_bytes()).hexdigest())\n") % (root, shapes, LABEL, list(TAPE_SEED))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-800:]
    assert r.stdout.split()[-len(shapes):] == want
