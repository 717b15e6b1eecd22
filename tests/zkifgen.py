"""ctypes binding of tools/zkifgen.cpp: Otti-shaped R1CS instances with a heavy constant column, as COO arrays and as the
three zkinterface files spzk reads.  Input infrastructure for tests/ and bench.py (not product, not oracle)."""
import ctypes as C
import os
import subprocess

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SRC = os.path.join(_ROOT, "tools", "zkifgen.cpp")
_SO = os.path.join(_ROOT, "tools", "_build", "libzkifgen.so")
_lib = None


def build():
    if not os.path.exists(_SO) or (os.path.exists(_SRC) and os.path.getmtime(_SRC) > os.path.getmtime(_SO)):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-o", _SO, _SRC])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.zg_new.restype = C.c_void_p
        _lib.zg_vars.restype = C.c_void_p
        _lib.zg_inputs.restype = C.c_void_p
        _lib.zg_file.restype = C.c_size_t
    return _lib


class OttiShaped:
    """nc constraints over nv variables and ni inputs (tools/zkifgen.cpp describes the row mix)."""

    def __init__(self, nc, nv, ni, seed=7, files=True):
        self.h = C.c_void_p(lib().zg_new(C.c_uint64(nc), C.c_uint64(nv), C.c_uint64(ni), C.c_uint64(seed), C.c_int(1 if files else 0)))
        if not self.h:
            raise ValueError("zkifgen: bad sizes")
        d = (C.c_uint64 * 7)()
        lib().zg_dims(self.h, d)
        self.num_cons, self.num_vars, self.num_inputs = int(d[0]), int(d[1]), int(d[2])
        self.nnz = (int(d[3]), int(d[4]), int(d[5]))
        self.const_entries = int(d[6])
        self.has_files = files

    def coo_arrays(self, m):
        """(rows u64[nnz], cols u64[nnz], vals u8[32 nnz]) as ctypes arrays"""
        n = self.nnz[m]
        r, c, v = (C.c_uint64 * max(n, 1))(), (C.c_uint64 * max(n, 1))(), (C.c_uint8 * max(32 * n, 1))()
        lib().zg_coo(self.h, C.c_int(m), r, c, v)
        return r, c, v

    def coo_list(self, m):
        r, c, v = self.coo_arrays(m)
        vb = bytes(v)
        return [(int(r[k]), int(c[k]), vb[32 * k:32 * k + 32]) for k in range(self.nnz[m])]

    def vars_bytes(self):
        return C.string_at(lib().zg_vars(self.h), 32 * self.num_vars)

    def inputs_bytes(self):
        return C.string_at(lib().zg_inputs(self.h), 32 * self.num_inputs) if self.num_inputs else b""

    def file_ptr(self, which):
        """(pointer, length) of the circuit (0), header / inputs (1) or witness (2) flatbuffer; owned by this object"""
        p = C.POINTER(C.c_uint8)()
        n = lib().zg_file(self.h, C.c_int(which), C.byref(p))
        return p, int(n)

    def file_bytes(self, which):
        p, n = self.file_ptr(which)
        return C.string_at(p, n)

    def __del__(self):
        if getattr(self, "h", None):
            lib().zg_free(self.h)
            self.h = None
