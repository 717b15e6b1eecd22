"""sb200_poly_bound_noalloc: the allocation-free variant of sb200_k_poly_bound (otti_b200/csrc/noalloc.cu).  No counterpart
in the reference, so the checker here is the existing entry point --
byte-exact, as for every F_l kernel -- with the oracle's DensePolynomial::bound beside it.  The file sorts last on purpose."""
import ctypes as C

import pytest

import oracle_lib as O
from common import mont, rand_scalars

pytestmark = pytest.mark.gpu


def _buf(b):
    return (C.c_uint8 * len(b)).from_buffer_copy(bytes(b))


def _noalloc(L, torch, Z, Lv, Ls):
    """One call through the C ABI on device buffers held by torch (plumbing); returns the output bytes."""
    n = len(Z) // 32
    dZ = torch.frombuffer(bytearray(Z), dtype=torch.uint8).cuda()
    dL = torch.frombuffer(bytearray(Lv), dtype=torch.uint8).cuda()
    dout = torch.full((32 * (n // Ls),), 0xAB, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = L.sb200_poly_bound_noalloc(C.c_void_p(dZ.data_ptr()), C.c_size_t(n), C.c_void_p(dL.data_ptr()), C.c_size_t(Ls),
                                    C.c_void_p(dout.data_ptr()), C.c_void_p(0))
    assert rc == 0, L.sb200_last_error()
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return bytes(dout.cpu().numpy().tobytes()), free0 - free1


def test_poly_bound_noalloc_matches_the_allocating_entry_point():
    import torch
    import otti_b200
    L = otti_b200.lib()
    L.sb200_last_error.restype = C.c_char_p
    assert L.sb200_init(0) == 0, L.sb200_last_error()
    torch.cuda.set_device(0)
    # two calls in a row on different shapes and data (the second sees the first's leftovers in the shared scratch), then
    # the first again; shapes cover one chunk, many chunks, a ragged last chunk and the largest column count
    shapes = [(64, 128), (1000, 96), (3, 4096), (256, 256), (64, 128)]
    grown = []
    for k, (Ls, Rs) in enumerate(shapes):
        Z = mont(rand_scalars(Ls * Rs, 700 + k))
        Lv = mont(rand_scalars(Ls, 800 + k))
        want = (C.c_uint8 * (32 * Rs))()
        assert L.sb200_k_poly_bound(_buf(Z), C.c_size_t(Ls * Rs), _buf(Lv), C.c_size_t(Ls), want) == 0, L.sb200_last_error()
        got, used = _noalloc(L, torch, Z, Lv, Ls)
        assert got == bytes(want)            # bit-exact: integer arithmetic in F_l
        assert got == O.poly_bound(Z, Lv)    # and the oracle's DensePolynomial::bound (dense_mlpoly.rs:205-212)
        grown.append(used)
    # after the first call on the device (module load + symbol look-up) a call takes no device memory
    assert all(g <= 0 for g in grown[1:]), grown


def test_poly_bound_noalloc_rejects_what_the_scratch_cannot_hold():
    import torch
    import otti_b200
    L = otti_b200.lib()
    L.sb200_last_error.restype = C.c_char_p
    assert L.sb200_init(0) == 0
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    p = C.c_void_p(d.data_ptr())
    # 8192 columns: beyond the fixed array.  Refused before anything is launched (the buffers are never touched).
    assert L.sb200_poly_bound_noalloc(p, C.c_size_t(2 * 8192), p, C.c_size_t(2), p, C.c_void_p(0)) == -102
    assert b"4096" in L.sb200_last_error()
    assert L.sb200_poly_bound_noalloc(p, C.c_size_t(10), p, C.c_size_t(3), p, C.c_void_p(0)) == -102   # n % Ls != 0
    assert L.sb200_poly_bound_noalloc(None, C.c_size_t(4), p, C.c_size_t(2), p, C.c_void_p(0)) == -102
    torch.cuda.synchronize()
