"""zkinterface ingestion, mirroring spartan-zkinterface's `R1csReader` / `R1cs`
(spartan-zkinterface/src/lib.rs:60-155): three flatbuffers -> Instance + assignments.

    r1cs = R1cs.from_files("c.zkif", "c.inp.zkif", "c.wit.zkif")      # main.rs:36-52
    inst = r1cs.instance()                                            # lib.rs:109-135
    vars_, inputs = r1cs.vars_assignment(), r1cs.inputs_assignment()  # lib.rs:79-93
    gens = r1cs.nizk_public_params()                                  # lib.rs:146-152

Parsing happens in libspartan_b200 (host code, no GPU needed); this module only marshals."""
import ctypes as C

from . import Assignment, Instance, NIZKGens, BackendError, lib, _check


class ZkifError(ValueError):
    """Input on which the reference `unwrap()`s or panics."""


class R1cs:
    def __init__(self, handle):
        self.h = handle
        d = (C.c_uint64 * 7)()
        _check(lib().sb200_r1cs_dims(self.h, d))
        self.num_cons, self.num_vars, self.num_inputs = int(d[0]), int(d[1]), int(d[2])
        self.nnz = (int(d[3]), int(d[4]), int(d[5]))
        self.non_zero_entries = int(d[6])

    @staticmethod
    def from_buffers(circuit, header, witness):
        """`R1cs::from(R1csReader::new(header, circuit, witness))`"""
        L = lib()
        h = C.c_void_p()
        bufs = [bytes(circuit), bytes(header), bytes(witness)]
        rc = L.sb200_zkif_parse(bufs[0], C.c_size_t(len(bufs[0])), bufs[1], C.c_size_t(len(bufs[1])), bufs[2],
                                C.c_size_t(len(bufs[2])), C.byref(h))
        if rc == -104:
            raise ZkifError((L.sb200_last_error() or b"").decode())
        _check(rc)
        return R1cs(h)

    @staticmethod
    def from_pointers(circuit, header, witness):
        """the same from three (pointer, length) pairs: no copy of a multi-hundred-megabyte circuit file"""
        L = lib()
        h = C.c_void_p()
        rc = L.sb200_zkif_parse(circuit[0], C.c_size_t(circuit[1]), header[0], C.c_size_t(header[1]), witness[0],
                                C.c_size_t(witness[1]), C.byref(h))
        if rc == -104:
            raise ZkifError((L.sb200_last_error() or b"").decode())
        _check(rc)
        return R1cs(h)

    @staticmethod
    def from_files(circuit_fn, inputs_fn, witness_fn):
        """Argument order of `spzk prove --nizk <circuit.zkif> <inputs.zkif> <witness.zkif>` (main.rs:18)."""
        with open(circuit_fn, "rb") as f:
            c = f.read()
        with open(inputs_fn, "rb") as f:
            i = f.read()
        with open(witness_fn, "rb") as f:
            w = f.read()
        return R1cs.from_buffers(c, i, w)

    def matrix_arrays(self, m):
        """(rows, cols, vals) of matrix m as bytes (u64 LE, u64 LE, 32 bytes each)"""
        rows, cols, vals = C.POINTER(C.c_uint64)(), C.POINTER(C.c_uint64)(), C.POINTER(C.c_uint8)()
        _check(lib().sb200_r1cs_matrix(self.h, C.c_int(m), C.byref(rows), C.byref(cols), C.byref(vals)))
        n = self.nnz[m]
        return (C.string_at(rows, 8 * n), C.string_at(cols, 8 * n), C.string_at(vals, 32 * n)) if n else (b"", b"", b"")

    def matrix(self, m):
        """COO triples of A (m=0), B (1) or C (2): [(row, col, 32 bytes)] in the reference's order."""
        lib().sb200_r1cs_vars.restype = C.c_void_p
        rows, cols, vals = C.POINTER(C.c_uint64)(), C.POINTER(C.c_uint64)(), C.POINTER(C.c_uint8)()
        _check(lib().sb200_r1cs_matrix(self.h, C.c_int(m), C.byref(rows), C.byref(cols), C.byref(vals)))
        n = self.nnz[m]
        vb = C.string_at(vals, 32 * n) if n else b""
        return [(int(rows[k]), int(cols[k]), vb[32 * k:32 * k + 32]) for k in range(n)]

    def _bytes(self, fn, n):
        fn.restype = C.c_void_p
        p = fn(self.h)
        return C.string_at(p, 32 * n) if n else b""

    def vars_assignment(self):
        return Assignment.new(self._bytes(lib().sb200_r1cs_vars, self.num_vars))

    def inputs_assignment(self):
        return Assignment.new(self._bytes(lib().sb200_r1cs_inputs, self.num_inputs))

    def instance(self, rank=0, world=1):
        """`R1cs::instance` -> `Instance::new(constraints.len(), witness.len(), inputs.len(), A, B, C)`"""
        h = C.c_void_p()
        _check(lib().sb200_r1cs_instance(self.h, C.c_int(rank), C.c_int(world), C.byref(h)))
        return Instance(h)

    def nizk_public_params(self, proofs_expected=None):
        """`R1cs::nizk_public_params` (lib.rs:146-152).  spzk makes one proof per process and times the generators
        with it (main.rs:71-85): pass proofs_expected=1 for tables sized for that."""
        return NIZKGens.new(self.num_cons, self.num_vars, self.num_inputs, proofs_expected)

    def __del__(self):
        if getattr(self, "h", None):
            try:
                lib().sb200_r1cs_free(self.h)
            except Exception:
                pass
            self.h = None
