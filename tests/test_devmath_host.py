"""The __host__ __device__ limb arithmetic the sm_100a kernels run (otti_b200/csrc/{fl,fp,ge}.cuh),
compiled for the host and checked against Python big integers and the oracle: F_l Montgomery ops,
loose-form F_p ops at every wrap-around edge, the signed-window MSM recoding for every window width,
extended-coordinate addition and ristretto compress.  No GPU needed."""
import ctypes as C
import os
import random
import subprocess

import pytest

import oracle_lib as O
from common import ELL, MONT_R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = 2**255 - 19


@pytest.fixture(scope="module")
def dm(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("devmath") / "devmath.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", so,
                           os.path.join(ROOT, "tests", "csrc", "devmath_host.cpp")])
    return C.CDLL(so)


def buf(b):
    return (C.c_uint8 * len(b)).from_buffer_copy(b)


def call(f, *a, n=32):
    o = (C.c_uint8 * n)()
    f(*a, o)
    return bytes(o)


def b32(x):
    return int(x).to_bytes(32, "little")


def test_fl(dm):
    rng = random.Random(1)
    edge = [0, 1, ELL - 1, ELL - 2, 2**252, 2**128 - 1]
    for it in range(1500):
        a, b = rng.randrange(ELL), rng.randrange(ELL)
        if it < len(edge) ** 2:
            a, b = edge[it % len(edge)], edge[it // len(edge)]
        am, bm = b32(a * MONT_R % ELL), b32(b * MONT_R % ELL)
        for op, ref in ((0, (a + b) % ELL), (1, (a - b) % ELL), (2, a * b % ELL)):
            assert int.from_bytes(call(dm.dm_fl_binop, op, buf(am), buf(bm)), "little") == ref * MONT_R % ELL, (op, a, b)
        assert call(dm.dm_fl_to_mont, buf(b32(a))) == am
        assert call(dm.dm_fl_from_mont, buf(am)) == b32(a)
        assert call(dm.dm_fl_neg, buf(am)) == b32((-a) % ELL * MONT_R % ELL)
    assert dm.dm_fl_is_canonical(buf(b32(ELL - 1))) == 1
    assert dm.dm_fl_is_canonical(buf(b32(ELL))) == 0
    assert dm.dm_fl_is_canonical(buf(b32(2**256 - 1))) == 0


def test_fp_loose_form_edges(dm):
    rng = random.Random(2)
    edge = [0, 1, P - 1, P, P + 1, 2 * P, 2 * P + 1, 2**256 - 1, 2**256 - 38, 2**256 - 39, 37, 38, 39, 2**255, 2**255 - 1, 19, 18]
    for it in range(2500):
        a, b = rng.randrange(2**256), rng.randrange(2**256)
        if it < len(edge) ** 2:
            a, b = edge[it % len(edge)], edge[it // len(edge)]
        for op, ref in ((0, (a + b) % P), (1, (a - b) % P), (2, a * b % P)):
            assert int.from_bytes(call(dm.dm_fp_binop, op, buf(b32(a)), buf(b32(b))), "little") == ref, (op, hex(a), hex(b))
        assert int.from_bytes(call(dm.dm_fp_canon, buf(b32(a))), "little") == a % P
    for it in range(30):
        a = rng.randrange(1, 2**256)
        if a % P:
            assert int.from_bytes(call(dm.dm_fp_invert, buf(b32(a))), "little") == pow(a, P - 2, P)


@pytest.mark.parametrize("c", [2, 3, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16])
def test_msm_recoding_and_fixed_base_mul(dm, c):
    """sum_w digit_w * 2^(c*w) * P == k * P for every window width, incl. the scalars whose top window is
    the tight case (l - 1, 2^252, 2^252 + 2^124), and no carry ever leaves the last window."""
    rng = random.Random(c)
    g = O.Gens(16, 16, 3)
    pts = [g.point(0, i) for i in range(4)] + [g.point(1), g.point(2), O.basepoint()]
    special = [ELL - 1, 2**252, 2**252 + 2**124, 1, 0, 2**252 - 1, (1 << (c - 1)), (1 << (c - 1)) + 1, (1 << c) - 1]
    for it, k in enumerate(special + [rng.randrange(ELL) for _ in range(4)]):
        assert dm.dm_final_carry(buf(b32(k)), c) == 0
        pc = pts[it % len(pts)]
        assert call(dm.dm_fixed_base_mul, buf(O.pt_affine(pc)), buf(b32(k)), c) == O.pt_mul(pc, b32(k)), (c, k)
    for it in range(1500):
        k = rng.randrange(ELL) if it > 60 else ELL - 1 - it
        assert dm.dm_final_carry(buf(b32(k)), c) == 0


def test_add_and_compress(dm):
    g = O.Gens(64, 64, 3)
    pts = [g.point(0, i) for i in range(8)]
    for a, b in zip(pts, pts[1:]):
        assert call(dm.dm_add_compress, buf(O.pt_affine(a)), buf(O.pt_affine(b))) == O.pt_add(a, b)
    # doubling through the unified addition law
    assert call(dm.dm_add_compress, buf(O.pt_affine(pts[0])), buf(O.pt_affine(pts[0]))) == O.pt_add(pts[0], pts[0])
