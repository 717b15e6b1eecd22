"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).

Test infrastructure: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "_build", "liboracle.so")
_lib = None

u8p = C.POINTER(C.c_uint8)


def build(force=False):
    srcs = [os.path.join(_ROOT, "oracle", f) for f in
            ("oracle_capi.cpp", "spartan.hpp", "snark.hpp", "transcript.hpp", "ristretto.hpp", "fl.hpp")]
    if force or not os.path.exists(_SO) or any(os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", os.path.join(_ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO) or os.path.exists(os.path.join(_ROOT, "oracle", "oracle_capi.cpp")):
            try:
                build()
            except Exception:
                if not os.path.exists(_SO):
                    raise
        _lib = C.CDLL(_SO)
        _lib.orc_gens_new.restype = C.c_void_p
        _lib.orc_instance_new.restype = C.c_void_p
        _lib.orc_synthetic_new.restype = C.c_void_p
        _lib.orc_gens_n.restype = C.c_size_t
        _lib.orc_fl_batch_from_bytes.restype = C.c_long
    return _lib


def _buf(b):
    return (C.c_uint8 * len(b)).from_buffer_copy(bytes(b)) if len(b) else (C.c_uint8 * 1)()


def _out(n):
    return (C.c_uint8 * max(n, 1))()


# ---- F_l (raw Montgomery limbs as 32 bytes)
def fl_from_bytes(b):
    o = _out(32)
    ok = lib().orc_fl_from_bytes(_buf(b), o)
    return bytes(o) if ok else None


def fl_to_bytes(m):
    o = _out(32)
    lib().orc_fl_to_bytes(_buf(m), o)
    return bytes(o)


def fl_from_bytes_wide(b):
    o = _out(32)
    lib().orc_fl_from_bytes_wide(_buf(b), o)
    return bytes(o)


def fl_binop(op, a, b):
    o = _out(32)
    lib().orc_fl_binop({"add": 0, "sub": 1, "mul": 2}[op], _buf(a), _buf(b), o)
    return bytes(o)


def fl_unop(op, a):
    o = _out(32)
    lib().orc_fl_unop({"neg": 0, "square": 1, "invert": 2}[op], _buf(a), o)
    return bytes(o)


def fl_batch_from_bytes(b):
    n = len(b) // 32
    o = _out(32 * n)
    bad = lib().orc_fl_batch_from_bytes(_buf(b), C.c_size_t(n), o)
    return bytes(o)[:32 * n], bad


def fl_batch_to_bytes(m):
    n = len(m) // 32
    o = _out(32 * n)
    lib().orc_fl_batch_to_bytes(_buf(m), C.c_size_t(n), o)
    return bytes(o)[:32 * n]


def mont(x):
    """python int -> Montgomery limb bytes"""
    return fl_from_bytes(int(x).to_bytes(32, "little"))


def unmont(m):
    return int.from_bytes(fl_to_bytes(m), "little")


# ---- group
def basepoint():
    o = _out(32)
    lib().orc_basepoint(o)
    return bytes(o)


def pt_mul(pc, k_canon):
    o = _out(32)
    ok = lib().orc_pt_mul(_buf(pc), _buf(k_canon), o)
    return bytes(o) if ok else None


def pt_add(a, b):
    o = _out(32)
    ok = lib().orc_pt_add(_buf(a), _buf(b), o)
    return bytes(o) if ok else None


def pt_affine(a):
    o = _out(64)
    ok = lib().orc_pt_affine(_buf(a), o)
    return bytes(o) if ok else None


def pt_decompress_ok(a):
    return bool(lib().orc_pt_decompress_ok(_buf(a)))


def from_uniform_bytes(b):
    o = _out(32)
    lib().orc_from_uniform_bytes(_buf(b), o)
    return bytes(o)


def pt_msm(k_mont, pts):
    n = len(pts) // 32
    o = _out(32)
    ok = lib().orc_pt_msm(_buf(k_mont), _buf(pts), C.c_size_t(n), o)
    return bytes(o) if ok else None


def merlin_test(label, mlabel, msg, clabel, n):
    o = _out(n)
    lib().orc_merlin_test(_buf(label), C.c_size_t(len(label)), mlabel + b"\0", _buf(msg), C.c_size_t(len(msg)),
                          clabel + b"\0", o, C.c_size_t(n))
    return bytes(o)[:n]


def shake256(data, n):
    o = _out(n)
    lib().orc_shake256(_buf(data), C.c_size_t(len(data)), o, C.c_size_t(n))
    return bytes(o)[:n]


# ---- hot loops
def eq_evals(r_mont):
    ell = len(r_mont) // 32
    o = _out(32 << ell)
    lib().orc_eq_evals(_buf(r_mont), C.c_size_t(ell), o)
    return bytes(o)[:32 << ell]


def bind_top(Z, r):
    n = len(Z) // 32
    b = _buf(Z)
    lib().orc_bind_top(b, C.c_size_t(n), _buf(r))
    return bytes(b)[:16 * n]


def sumcheck_round(tables):
    n = len(tables[0]) // 32
    bufs = [_buf(t) for t in tables]
    arr = (C.c_void_p * len(tables))(*[C.cast(b, C.c_void_p) for b in bufs])
    o = _out(96)
    lib().orc_sumcheck_round(arr, len(tables), C.c_size_t(n), o)
    o = bytes(o)
    return o[:32], o[32:64], o[64:96]


def dot(a, b):
    o = _out(32)
    lib().orc_dot(_buf(a), _buf(b), C.c_size_t(len(a) // 32), o)
    return bytes(o)


def poly_bound(Z, Lv):
    n, Ls = len(Z) // 32, len(Lv) // 32
    o = _out(32 * (n // Ls))
    lib().orc_poly_bound(_buf(Z), C.c_size_t(n), _buf(Lv), C.c_size_t(Ls), o)
    return bytes(o)[:32 * (n // Ls)]


def unipoly_from_evals(evals):
    n = len(evals) // 32
    o = _out(32 * n)
    lib().orc_unipoly_from_evals(_buf(evals), C.c_size_t(n), o)
    return bytes(o)[:32 * n]


def unipoly_eval(coeffs, r):
    o = _out(32)
    lib().orc_unipoly_eval(_buf(coeffs), C.c_size_t(len(coeffs) // 32), _buf(r), o)
    return bytes(o)


class Gens:
    def __init__(self, num_cons, num_vars, num_inputs, label=b"gens_r1cs_sat"):
        self.h = C.c_void_p(lib().orc_gens_new(C.c_size_t(num_cons), C.c_size_t(num_vars), C.c_size_t(num_inputs),
                                               _buf(label), C.c_size_t(len(label))))
        self.n = lib().orc_gens_n(self.h)

    def point(self, which, i=0):
        o = _out(32)
        lib().orc_gens_point(self.h, which, C.c_size_t(i), o)
        return bytes(o)

    def commit_rows(self, Z_mont, blinds_mont, threads=1):
        n, L = len(Z_mont) // 32, len(blinds_mont) // 32
        o = _out(32 * L)
        lib().orc_commit_rows(self.h, _buf(Z_mont), C.c_size_t(n), _buf(blinds_mont), C.c_size_t(L), threads, o)
        return bytes(o)[:32 * L]

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_gens_free(self.h)
            self.h = None


class Instance:
    """Oracle-side Instance (lib.rs:121-225).  A/B/C: lists of (row, col, 32-byte canonical)."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)
        d = (C.c_uint64 * 6)()
        lib().orc_instance_dims(self.h, d)
        self.num_cons, self.num_vars, self.num_inputs = d[0], d[1], d[2]
        self.nnz = (d[3], d[4], d[5])

    @staticmethod
    def new(num_cons, num_vars, num_inputs, A, B, Cm):
        rows, cols, vals, keep = [], [], [], []
        for M in (A, B, Cm):
            r = (C.c_uint64 * max(len(M), 1))(*[e[0] for e in M])
            c = (C.c_uint64 * max(len(M), 1))(*[e[1] for e in M])
            v = _buf(b"".join(e[2] for e in M))
            keep += [r, c, v]
            rows.append(C.cast(r, C.c_void_p))
            cols.append(C.cast(c, C.c_void_p))
            vals.append(C.cast(v, C.c_void_p))
        nnz = (C.c_uint64 * 3)(len(A), len(B), len(Cm))
        err = C.c_int(0)
        h = lib().orc_instance_new(C.c_size_t(num_cons), C.c_size_t(num_vars), C.c_size_t(num_inputs),
                                   (C.c_void_p * 3)(*rows), (C.c_void_p * 3)(*cols), (C.c_void_p * 3)(*vals), nnz,
                                   C.byref(err))
        if not h:
            raise ValueError({-5: "InvalidScalar", -6: "InvalidIndex"}.get(err.value, str(err.value)))
        return Instance(h)

    @staticmethod
    def new_arrays(num_cons, num_vars, num_inputs, mats):
        """mats: three (rows u64 ctypes array, cols u64 ctypes array, vals u8 ctypes array, nnz)"""
        rows = (C.c_void_p * 3)(*[C.cast(m[0], C.c_void_p) for m in mats])
        cols = (C.c_void_p * 3)(*[C.cast(m[1], C.c_void_p) for m in mats])
        vals = (C.c_void_p * 3)(*[C.cast(m[2], C.c_void_p) for m in mats])
        nnz = (C.c_uint64 * 3)(*[m[3] for m in mats])
        err = C.c_int(0)
        h = lib().orc_instance_new(C.c_size_t(num_cons), C.c_size_t(num_vars), C.c_size_t(num_inputs), rows, cols, vals, nnz,
                                   C.byref(err))
        if not h:
            raise ValueError({-5: "InvalidScalar", -6: "InvalidIndex"}.get(err.value, str(err.value)))
        return Instance(h)

    @staticmethod
    def synthetic(num_cons, num_vars, num_inputs, seed):
        v = _out(32 * num_vars)
        i = _out(32 * num_inputs)
        h = lib().orc_synthetic_new(C.c_size_t(num_cons), C.c_size_t(num_vars), C.c_size_t(num_inputs),
                                    C.c_uint64(seed), v, i)
        return Instance(h), bytes(v)[:32 * num_vars], bytes(i)[:32 * num_inputs]

    def coo(self, m):
        n = self.nnz[m]
        r = (C.c_uint64 * max(n, 1))()
        c = (C.c_uint64 * max(n, 1))()
        v = _out(32 * n)
        lib().orc_instance_coo(self.h, m, r, c, v)
        vb = bytes(v)
        return [(r[k], c[k], vb[32 * k:32 * k + 32]) for k in range(n)]

    def is_sat(self, vars_b, inputs_b):
        return lib().orc_instance_is_sat(self.h, _buf(vars_b), C.c_size_t(len(vars_b) // 32), _buf(inputs_b),
                                         C.c_size_t(len(inputs_b) // 32))

    def multiply_vec(self, z_mont):
        o = _out(96 * self.num_cons)
        lib().orc_multiply_vec(self.h, _buf(z_mont), o)
        o = bytes(o)
        k = 32 * self.num_cons
        return o[:k], o[k:2 * k], o[2 * k:3 * k]

    def eval_table_abc(self, rx_evals, rabc):
        o = _out(64 * self.num_vars)
        lib().orc_eval_table_abc(self.h, _buf(rx_evals), _buf(rabc), o)
        return bytes(o)[:64 * self.num_vars]

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_instance_free(self.h)
            self.h = None


def nizk_prove(inst, vars_b, inputs_b, gens, label, tape_seed64, threads=1, want_times=False):
    out = u8p()
    n = C.c_size_t(0)
    t = (C.c_double * 5)()
    rc = lib().orc_nizk_prove(inst.h, _buf(vars_b), C.c_size_t(len(vars_b) // 32), _buf(inputs_b),
                              C.c_size_t(len(inputs_b) // 32), gens.h, _buf(label), C.c_size_t(len(label)),
                              _buf(tape_seed64), threads, C.byref(out), C.byref(n), t)
    if rc != 0:
        raise RuntimeError("oracle prove failed: %d" % rc)
    b = C.string_at(out, n.value)
    lib().orc_free(out)
    return (b, list(t)) if want_times else b


def nizk_verify(inst, inputs_b, gens, label, proof):
    return lib().orc_nizk_verify(inst.h, _buf(inputs_b), C.c_size_t(len(inputs_b) // 32), gens.h, _buf(label),
                                 C.c_size_t(len(label)), _buf(proof), C.c_size_t(len(proof)))


def snark_dense(inst):
    """(N, num_mem_cells, comb_ops, comb_mem): the dense representation SNARK::encode commits to, canonical bytes."""
    d = (C.c_uint64 * 2)()
    if lib().orc_snark_dense(inst.h, d, None, None) != 0:
        raise RuntimeError("oracle sparse_to_dense panicked")
    ops, mem = _out(32 * 16 * d[0]), _out(32 * 2 * d[1])
    lib().orc_snark_dense(inst.h, d, ops, mem)
    return d[0], d[1], bytes(ops)[:32 * 16 * d[0]], bytes(mem)[:32 * 2 * d[1]]


def snark_encode(inst, num_nz_entries, threads=1):
    """SNARK::encode (lib.rs:321-333): num_cons, num_vars, num_inputs + bincode(SparseMatPolyCommitment)."""
    out = u8p()
    n = C.c_size_t(0)
    rc = lib().orc_snark_encode(inst.h, C.c_size_t(num_nz_entries), threads, C.byref(out), C.byref(n), None, None)
    if rc != 0:
        raise RuntimeError("oracle SNARK::encode panicked (an assert of the reference)")
    b = C.string_at(out, n.value)
    lib().orc_free(out)
    return b


def snark_commit_derefs(inst, num_nz_entries, rx_canon, ry_canon, threads=1, want_comb=False):
    """commit_nondet_witness of SparseMatPolyEvalProof::prove (sparse_mlpoly.rs:1485-1503): the compressed points of
    DerefsCommitment.comm_ops_val (joined), optionally Derefs.comb as canonical bytes."""
    N = snark_dense(inst)[0]
    cap = 8 * N
    out = _out(32 * cap)
    comb = _out(32 * 8 * N) if want_comb else None
    lib().orc_snark_commit_derefs.restype = C.c_long
    k = lib().orc_snark_commit_derefs(inst.h, C.c_size_t(num_nz_entries), _buf(rx_canon), C.c_size_t(len(rx_canon) // 32),
                                      _buf(ry_canon), C.c_size_t(len(ry_canon) // 32), threads, out, C.c_size_t(cap), comb)
    if k < 0:
        raise RuntimeError("oracle commit_derefs panicked (%d)" % k)
    pts = bytes(out)[:32 * k]
    return (pts, bytes(comb)) if want_comb else pts


def snark_network_products(inst, rx_canon, ry_canon, r_hash_canon, r_multiset_canon):
    """PolyEvalNetwork::new (sparse_mlpoly.rs:652-668) reduced to the 16 ProductCircuit::evaluate values (canonical):
    row side init, read x3, write x3, audit, then the column side."""
    out = _out(512)
    rc = lib().orc_snark_network_products(inst.h, _buf(rx_canon), C.c_size_t(len(rx_canon) // 32), _buf(ry_canon),
                                          C.c_size_t(len(ry_canon) // 32), _buf(r_hash_canon), _buf(r_multiset_canon), out)
    if rc != 0:
        raise RuntimeError("oracle network products failed: %d" % rc)
    b = bytes(out)
    return [b[32 * i:32 * i + 32] for i in range(16)]


def snark_prove(inst, num_nz_entries, vars_b, inputs_b, label, tape_seed64, threads=1, want_times=False):
    """SNARK::prove (lib.rs:335-403): bincode of SNARK { r1cs_sat_proof, inst_evals, r1cs_eval_proof }."""
    out = u8p()
    n = C.c_size_t(0)
    t = (C.c_double * 3)()
    rc = lib().orc_snark_prove(inst.h, C.c_size_t(num_nz_entries), _buf(vars_b), C.c_size_t(len(vars_b) // 32), _buf(inputs_b),
                               C.c_size_t(len(inputs_b) // 32), _buf(label), C.c_size_t(len(label)), _buf(tape_seed64), threads,
                               C.byref(out), C.byref(n), t)
    if rc != 0:
        raise RuntimeError("oracle SNARK::prove failed: %d" % rc)
    b = C.string_at(out, n.value)
    lib().orc_free(out)
    return (b, list(t)) if want_times else b


def snark_verify(comm, num_nz_entries, inputs_b, label, proof):
    """SNARK::verify (lib.rs:405-457) against the bytes snark_encode returned.  0 accepted, -1 rejected, -2 malformed."""
    return lib().orc_snark_verify(_buf(comm), C.c_size_t(len(comm)), C.c_size_t(num_nz_entries), _buf(inputs_b),
                                  C.c_size_t(len(inputs_b) // 32), _buf(label), C.c_size_t(len(label)), _buf(proof),
                                  C.c_size_t(len(proof)))
