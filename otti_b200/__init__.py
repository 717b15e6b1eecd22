"""otti_b200 -- host-side mirror of libspartan's public API over libspartan_b200.so.

The reference is Rust (`/root/reference/Spartan/src/lib.rs`); no Rust toolchain
exists in this image, so the host side above the C ABI (include/spartan_b200.h)
is mirrored here in Python with the reference's names, argument meaning and
error behaviour:

    Assignment / VarsAssignment / InputsAssignment      lib.rs:58-113
    Instance.new / is_sat / produce_synthetic_r1cs      lib.rs:115-272
    NIZKGens.new                                        lib.rs:459-477
    NIZK.prove / NIZK.verify                            lib.rs:479-575
    SNARKGens.new / SNARK.encode                        lib.rs:273-333 (computation commitment only)
    Transcript.new (merlin)                             spartan-zkinterface/src/main.rs:78

All arithmetic happens in the CUDA library; this module only marshals
pointers.  There is no CPU fallback: importing works anywhere (so the symbol
table can be checked without a GPU), every compute call raises
`BackendError` when no B200 is usable.
"""
import ctypes as C
import os

__all__ = ["Assignment", "VarsAssignment", "InputsAssignment", "Instance", "NIZKGens", "NIZK", "SNARKGens", "SNARK", "ComputationCommitment", "Transcript", "R1CSError",
           "ProofVerifyError", "BackendError", "lib", "LIB_PATH"]

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libspartan_b200.so")


class R1CSError(Exception):
    """errors.rs:19-32"""

    def __init__(self, kind, text=""):
        super().__init__("%s%s" % (kind, (": " + text) if text else ""))
        self.kind = kind


class ProofVerifyError(Exception):
    """errors.rs:4-10 (and the reference's verify-side panics)"""


class BackendError(RuntimeError):
    pass


_R1CS_KINDS = {-1: "NonPowerOfTwoCons", -2: "NonPowerOfTwoVars", -3: "InvalidNumberOfInputs", -4: "InvalidNumberOfVars",
               -5: "InvalidScalar", -6: "InvalidIndex"}

_lib = None


def lib():
    """The loaded libspartan_b200.so (built in-tree by __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BackendError("libspartan_b200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.sb200_last_error.restype = C.c_char_p
        L.sb200_k_gens_n.restype = C.c_size_t
        L.sb200_k_gens_window.restype = C.c_uint32
        L.sb200_k_launch_count.restype = C.c_uint64
        L.sb200_instance_stream.restype = C.c_uint64
        _lib = L
    return _lib


def _check(rc):
    if rc == 0:
        return
    text = (lib().sb200_last_error() or b"").decode("utf-8", "replace")
    if rc in _R1CS_KINDS:
        raise R1CSError(_R1CS_KINDS[rc], text)
    if rc in (-20, -21):
        raise ProofVerifyError(text or ("malformed proof" if rc == -21 else "Proof verification failed"))
    if rc == -30:
        raise AssertionError(text)  # the reference panics here
    raise BackendError("libspartan_b200 error %d: %s" % (rc, text))


def _ptr(obj):
    """(address, keepalive) for bytes / bytearray / numpy array / torch tensor."""
    if obj is None:
        return None, None
    if hasattr(obj, "data_ptr"):  # torch tensor (possibly pinned)
        return C.c_void_p(obj.data_ptr()), obj
    if hasattr(obj, "ctypes"):  # numpy
        return C.c_void_p(obj.ctypes.data), obj
    if isinstance(obj, bytes):  # no copy: the bytes object itself is the keep-alive
        return C.cast(C.c_char_p(obj), C.c_void_p), obj
    if isinstance(obj, C.Array):  # no copy either
        return C.cast(obj, C.c_void_p), obj
    b = (C.c_uint8 * max(len(obj), 1)).from_buffer_copy(bytes(obj)) if len(obj) else (C.c_uint8 * 1)()
    return C.cast(b, C.c_void_p), b


def _nbytes(obj):
    if hasattr(obj, "data_ptr"):
        return obj.numel() * obj.element_size()
    if hasattr(obj, "nbytes"):
        return obj.nbytes
    return len(obj)


class Assignment:
    """`Assignment::new(&[[u8; 32]])` (lib.rs:65-89): values are canonical little-endian scalars."""

    def __init__(self, data, n):
        self.data, self.n = data, n

    @staticmethod
    def new(assignment):
        if isinstance(assignment, (list, tuple)):
            for v in assignment:
                if len(v) != 32:
                    raise ValueError("each value must be 32 bytes")
            assignment = b"".join(bytes(v) for v in assignment)
        nb = _nbytes(assignment)
        if nb % 32:
            raise ValueError("assignment must be n x 32 bytes")
        p, keep = _ptr(assignment)
        _check(lib().sb200_assignment_check(p, C.c_uint64(nb // 32)))  # InvalidScalar
        return Assignment(assignment, nb // 32)

    def __len__(self):
        return self.n


VarsAssignment = Assignment
InputsAssignment = Assignment


class Transcript:
    """merlin::Transcript (the caller-owned `&mut Transcript` of NIZK::prove / verify)."""

    def __init__(self, label):
        self.h = C.c_void_p()
        lb, _ = _ptr(label)
        _check(lib().sb200_transcript_new(lb, C.c_size_t(len(label)), C.byref(self.h)))

    @staticmethod
    def new(label):
        return Transcript(label)

    def append_message(self, label, msg):
        lp, k1 = _ptr(label)
        mp, k2 = _ptr(msg)
        lib().sb200_transcript_append_message(self.h, lp, C.c_size_t(len(label)), mp, C.c_size_t(len(msg)))

    def challenge_bytes(self, label, n):
        out = (C.c_uint8 * n)()
        lp, k1 = _ptr(label)
        lib().sb200_transcript_challenge_bytes(self.h, lp, C.c_size_t(len(label)), out, C.c_size_t(n))
        return bytes(out)

    def _vt(self):
        vt = _TranscriptVT()
        lib().sb200_transcript_vtable(self.h, C.byref(vt))
        return vt

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.sb200_transcript_free(self.h)
            self.h = None


class _TranscriptVT(C.Structure):
    _fields_ = [("ctx", C.c_void_p), ("append_message", C.c_void_p), ("challenge_bytes", C.c_void_p)]


def _coo(M):
    """list of (row, col, 32 bytes) or (rows u64 array, cols u64 array, vals n x 32 uint8) -> pointers"""
    if isinstance(M, (list,)) or (isinstance(M, tuple) and (len(M) == 0 or isinstance(M[0], (tuple, list)))):
        n = len(M)
        rows = (C.c_uint64 * max(n, 1))(*[e[0] for e in M])
        cols = (C.c_uint64 * max(n, 1))(*[e[1] for e in M])
        vals = b"".join(bytes(e[2]) for e in M)
        vp, vk = _ptr(vals)
        return n, C.cast(rows, C.c_void_p), C.cast(cols, C.c_void_p), vp, (rows, cols, vk)
    r, c, v = M
    n = len(r)
    rp, rk = _ptr(r)
    cp, ck = _ptr(c)
    vp, vk = _ptr(v)
    return n, rp, cp, vp, (rk, ck, vk)


class Instance:
    """`Instance` (lib.rs:115-272): the R1CS matrices, resident in HBM as CSR + CSC."""

    def __init__(self, handle):
        self.h = handle
        d = (C.c_uint64 * 3)()
        _check(lib().sb200_instance_dims(self.h, d))
        self.num_cons, self.num_vars, self.num_inputs = int(d[0]), int(d[1]), int(d[2])

    @staticmethod
    def new(num_cons, num_vars, num_inputs, A, B, Cm, rank=0, world=1):
        """`Instance::new` (lib.rs:121-225); raises R1CSError("InvalidIndex" | "InvalidScalar").
        world > 1: this process holds shard `rank` of one proof sharded over `world` GPUs."""
        parts = [_coo(M) for M in (A, B, Cm)]
        rows = (C.c_void_p * 3)(*[p[1] for p in parts])
        cols = (C.c_void_p * 3)(*[p[2] for p in parts])
        vals = (C.c_void_p * 3)(*[p[3] for p in parts])
        nnz = (C.c_uint64 * 3)(*[p[0] for p in parts])
        h = C.c_void_p()
        _check(lib().sb200_instance_new_sharded(C.c_uint64(num_cons), C.c_uint64(num_vars), C.c_uint64(num_inputs), rows, cols,
                                                vals, nnz, C.c_int(rank), C.c_int(world), C.byref(h)))
        return Instance(h)

    @staticmethod
    def produce_synthetic_r1cs(num_cons, num_vars, num_inputs, seed=0x0771, pinned=False, rank=0, world=1):
        """`Instance::produce_synthetic_r1cs` (lib.rs:260-271) with a seeded PRNG in place of OsRng.
        Returns (Instance, VarsAssignment, InputsAssignment)."""
        if pinned:
            import torch
            vbuf = torch.empty(max(num_vars, 1) * 32, dtype=torch.uint8).pin_memory()
            ibuf = torch.empty(max(num_inputs, 1) * 32, dtype=torch.uint8).pin_memory()
        else:
            vbuf = bytearray(max(num_vars, 1) * 32)
            ibuf = bytearray(max(num_inputs, 1) * 32)
            vbuf = (C.c_uint8 * len(vbuf)).from_buffer(vbuf)
            ibuf = (C.c_uint8 * len(ibuf)).from_buffer(ibuf)
        vp = C.c_void_p(vbuf.data_ptr()) if pinned else C.cast(vbuf, C.c_void_p)
        ip = C.c_void_p(ibuf.data_ptr()) if pinned else C.cast(ibuf, C.c_void_p)
        h = C.c_void_p()
        _check(lib().sb200_instance_synthetic_sharded(C.c_uint64(num_cons), C.c_uint64(num_vars), C.c_uint64(num_inputs),
                                                      C.c_uint64(seed), C.c_int(rank), C.c_int(world), C.byref(h), vp, ip))
        if pinned:
            v, i = vbuf[:num_vars * 32], ibuf[:num_inputs * 32]
        else:
            v, i = bytes(vbuf)[:num_vars * 32], bytes(ibuf)[:num_inputs * 32]
        return Instance(h), Assignment(v, num_vars), Assignment(i, num_inputs)

    def upload_witness(self, gens, vars_):
        """Place the witness in HBM (bench.py's "inputs already resident" measurement)."""
        vp, vk = _ptr(vars_.data)
        _check(lib().sb200_witness_upload(self.h, gens.h, vp, C.c_uint64(vars_.n)))

    @property
    def stream(self):
        return int(lib().sb200_instance_stream(self.h))

    def is_sat(self, vars_, inputs):
        """`Instance::is_sat` (lib.rs:228-257) -> bool; raises R1CSError("InvalidNumberOfInputs")."""
        vp, vk = _ptr(vars_.data)
        ip, ik = _ptr(inputs.data)
        sat = C.c_int(0)
        _check(lib().sb200_instance_is_sat(self.h, vp, C.c_uint64(vars_.n), ip, C.c_uint64(inputs.n), C.byref(sat)))
        return bool(sat.value)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.sb200_instance_free(self.h)
            self.h = None


class NIZKGens:
    """`NIZKGens::new(num_cons, num_vars, num_inputs)` (lib.rs:466-477)."""

    def __init__(self, num_cons, num_vars, num_inputs, label=None, proofs_expected=None):
        self.h = C.c_void_p()
        lp, keep = _ptr(label)
        if proofs_expected is None:
            _check(lib().sb200_gens_new(lp, C.c_size_t(len(label) if label else 0), C.c_uint64(num_cons), C.c_uint64(num_vars),
                                        C.c_uint64(num_inputs), C.byref(self.h)))
        else:
            _check(lib().sb200_gens_new_for(lp, C.c_size_t(len(label) if label else 0), C.c_uint64(num_cons),
                                            C.c_uint64(num_vars), C.c_uint64(num_inputs), C.c_uint64(proofs_expected),
                                            C.byref(self.h)))

    @staticmethod
    def new(num_cons, num_vars, num_inputs, proofs_expected=None):
        """proofs_expected: None / 0 = many proofs (widest window tables that fit the HBM, built once: 0.2-1.2 s);
        k > 0 = tables sized for k proofs (1: a one-shot caller such as spzk, whose timer covers gens + prove)."""
        return NIZKGens(num_cons, num_vars, num_inputs, proofs_expected=proofs_expected)

    @property
    def window_bits(self):
        return int(lib().sb200_k_gens_window(self.h))

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.sb200_gens_free(self.h)
            self.h = None


class NIZK:
    """`NIZK` (lib.rs:479-575).  Holds bincode(NIZK) exactly as the reference would serialise it."""

    def __init__(self, proof_bytes):
        self.bytes = proof_bytes

    @staticmethod
    def prove(inst, vars_, inputs, gens, transcript, tape_seed64=None, resident=False, comm=None):
        """`NIZK::prove(&inst, vars, &inputs, &gens, &mut transcript)` (lib.rs:493-535).
        tape_seed64 (64 bytes) pins the RandomTape seed for reproducible proofs; None = OS randomness.
        comm: exchange hook (otti_b200.comm) when `inst` is one shard of a proof spread over several GPUs."""
        vp, vk = (None, None) if resident else _ptr(vars_.data)
        ip, ik = _ptr(inputs.data)
        sp, sk = _ptr(tape_seed64)
        if tape_seed64 is not None and len(tape_seed64) != 64:
            raise ValueError("tape_seed64 must be 64 bytes")
        vt = transcript._vt()
        out = C.POINTER(C.c_uint8)()
        n = C.c_size_t(0)
        _check(lib().sb200_nizk_prove(inst.h, vp, C.c_uint64(vars_.n), ip, C.c_uint64(inputs.n), gens.h, C.byref(vt), None,
                                      C.c_size_t(0), sp, C.byref(comm.vt) if comm is not None else None, C.byref(out),
                                      C.byref(n)))
        b = C.string_at(out, n.value)
        lib().sb200_free(out)
        return NIZK(b)

    def verify(self, inst, inputs, transcript, gens):
        """`NIZK::verify(&self, &inst, &inputs, &mut transcript, &gens)` (lib.rs:538-574).
        Returns None on success, raises ProofVerifyError otherwise."""
        ip, ik = _ptr(inputs.data)
        pp, pk = _ptr(self.bytes)
        vt = transcript._vt()
        _check(lib().sb200_nizk_verify(inst.h, ip, C.c_uint64(inputs.n), gens.h, C.byref(vt), None, C.c_size_t(0), pp,
                                       C.c_size_t(len(self.bytes))))

    def to_bytes(self):
        return self.bytes

    @property
    def r1cs_sat_proof_len(self):
        """len(bincode(proof.r1cs_sat_proof)), what spzk reports (main.rs:87-89)."""
        import struct
        b = self.bytes
        off = 8 + 32 * struct.unpack_from("<Q", b, 0)[0]              # comm_vars
        s_x = struct.unpack_from("<Q", b, off)[0]                      # sc_proof_phase1.comm_polys.len()
        off += 2 * (8 + 32 * s_x) + 8 + s_x * (32 + 32 + 8 + 32 * 4 + 32 + 32)
        off += 128 + (96 + 256) + 64                                   # claims, pok + product proof, eq proof
        s_y = struct.unpack_from("<Q", b, off)[0]
        return len(b) - (8 + 32 * s_x) - (8 + 32 * s_y)

    @staticmethod
    def last_profile():
        """Timer scopes of the last prove (timer.rs; r1csproof.rs:155-307)."""
        d = (C.c_double * 16)()
        lib().sb200_last_prove_profile(d)
        keys = ["polycommit", "prove_sc_phase_one", "prove_sc_phase_two", "polyeval", "total", "h2d", "host_fiat_shamir",
                "kernel_launches", "h2d_bytes", "d2h_bytes", "msm_commit_kernel", "sc_cubic_round0_kernel", "exchange",
                "exchange_calls", "world", "reserved"]
        return dict(zip(keys, list(d)))


class SNARKGens:
    """`SNARKGens::new(num_cons, num_vars, num_inputs, num_nz_entries)` (lib.rs:279-303) -- the generators
    SNARK::encode reads (`gens_r1cs_eval`) and their window tables in HBM."""

    def __init__(self, num_cons, num_vars, num_inputs, num_nz_entries, proofs_expected=None):
        self.h = C.c_void_p()
        if proofs_expected is None:
            _check(lib().sb200_snark_gens_new(C.c_uint64(num_cons), C.c_uint64(num_vars), C.c_uint64(num_inputs),
                                              C.c_uint64(num_nz_entries), C.byref(self.h)))
        else:
            _check(lib().sb200_snark_gens_new_for(C.c_uint64(num_cons), C.c_uint64(num_vars), C.c_uint64(num_inputs),
                                                  C.c_uint64(num_nz_entries), C.c_uint64(proofs_expected), C.byref(self.h)))

    @staticmethod
    def new(num_cons, num_vars, num_inputs, num_nz_entries, proofs_expected=None):
        return SNARKGens(num_cons, num_vars, num_inputs, num_nz_entries, proofs_expected)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.sb200_snark_gens_free(self.h)
            self.h = None


class ComputationCommitment:
    """`ComputationCommitment` (lib.rs:47-50): R1CSCommitment's sizes and the two polynomial commitments of
    SparseMatPolyCommitment (sparse_mlpoly.rs:329-337).  `bytes` is the library's wire form (include/spartan_b200.h)."""

    def __init__(self, b):
        import struct
        self.bytes = b
        (self.num_cons, self.num_vars, self.num_inputs, self.batch_size, self.num_ops, self.num_mem_cells,
         n_ops) = struct.unpack_from("<7Q", b, 0)
        off = 56
        self.comm_comb_ops = [b[off + 32 * i:off + 32 * i + 32] for i in range(n_ops)]
        off += 32 * n_ops
        n_mem = struct.unpack_from("<Q", b, off)[0]
        off += 8
        self.comm_comb_mem = [b[off + 32 * i:off + 32 * i + 32] for i in range(n_mem)]
        if off + 32 * n_mem != len(b):
            raise ValueError("trailing bytes after the computation commitment")

    def to_bytes(self):
        return self.bytes


class SNARK:
    """`SNARK` (lib.rs:305-457).  An object holds bincode(SNARK) exactly as the reference would serialise it; the static
    methods below `verify` are the steps of `prove` exposed one by one (tests, profiling)."""

    def __init__(self, proof_bytes=b""):
        self.bytes = proof_bytes

    @staticmethod
    def prove(inst, comm, decomm, vars_, inputs, gens, transcript, tape_seed64=None, resident=False):
        """`SNARK::prove(&inst, &comm, &decomm, vars, &inputs, &gens, &mut transcript)` (lib.rs:335-403).  `comm` and
        `decomm` are accepted for signature parity and unused: the commitment does not enter the prover's transcript and
        the decommitment is rebuilt on the device from `inst`."""
        vp, vk = (None, None) if resident else _ptr(vars_.data)
        ip, ik = _ptr(inputs.data)
        sp, sk = _ptr(tape_seed64)
        if tape_seed64 is not None and len(tape_seed64) != 64:
            raise ValueError("tape_seed64 must be 64 bytes")
        vt = transcript._vt()
        out = C.POINTER(C.c_uint8)()
        n = C.c_size_t(0)
        _check(lib().sb200_snark_prove(inst.h, gens.h, vp, C.c_uint64(vars_.n), ip, C.c_uint64(inputs.n), C.byref(vt), None,
                                       C.c_size_t(0), sp, C.byref(out), C.byref(n)))
        b = C.string_at(out, n.value)
        lib().sb200_free(out)
        return SNARK(b)

    def verify(self, comm, inputs, transcript, gens):
        """`SNARK::verify(&self, &comm, &inputs, &mut transcript, &gens)` (lib.rs:405-457).  Returns None on success,
        raises ProofVerifyError otherwise."""
        cb = comm.to_bytes() if hasattr(comm, "to_bytes") else comm
        cp, ck = _ptr(cb)
        ip, ik = _ptr(inputs.data)
        pp, pk = _ptr(self.bytes)
        vt = transcript._vt()
        _check(lib().sb200_snark_verify(cp, C.c_size_t(len(cb)), gens.h, ip, C.c_uint64(inputs.n), C.byref(vt), None, C.c_size_t(0),
                                        pp, C.c_size_t(len(self.bytes))))

    def to_bytes(self):
        return self.bytes

    @staticmethod
    def last_prove_profile():
        """Timer scopes of the last `prove`, seconds (the reference's profiler lines, README.md:305-345)."""
        d = (C.c_double * 8)()
        lib().sb200_last_snark_profile(d)
        return dict(zip(["r1cs_sat_proof", "eval_sparse_polys", "commit_nondet_witness", "build_layered_network",
                         "product_layer_proof", "hash_layer_proof", "total", "kernel_launches"], list(d)))

    @staticmethod
    def encode(inst, gens):
        """`SNARK::encode(&inst, &gens)` (lib.rs:321-333).  Returns the ComputationCommitment; the decommitment (the dense
        representation of the matrices) is rebuilt on the device from `inst` whenever it is needed."""
        out = C.POINTER(C.c_uint8)()
        n = C.c_size_t(0)
        _check(lib().sb200_snark_encode(inst.h, gens.h, C.byref(out), C.byref(n)))
        b = C.string_at(out, n.value)
        lib().sb200_free(out)
        return ComputationCommitment(b)

    @staticmethod
    def commit_derefs(inst, gens, rx, ry, want_comb=False):
        """The `commit_nondet_witness` step of `SparseMatPolyEvalProof::prove` (sparse_mlpoly.rs:1485-1503): `rx`, `ry` are
        canonical 32-byte scalars concatenated.  Returns the compressed points of `DerefsCommitment.comm_ops_val`
        (and, with want_comb, `Derefs.comb` as canonical bytes)."""
        out = C.POINTER(C.c_uint8)()
        n = C.c_size_t(0)
        comb = None
        if want_comb:
            N = 1
            d = (C.c_uint64 * 2)()
            _check(lib().sb200_k_snark_dense(inst.h, d, None, None))
            N = int(d[0])
            comb = (C.c_uint8 * (32 * 8 * N))()
        xp, xk = _ptr(rx)
        yp, yk = _ptr(ry)
        _check(lib().sb200_snark_commit_derefs(inst.h, gens.h, xp, C.c_uint64(len(rx) // 32), yp, C.c_uint64(len(ry) // 32),
                                               C.byref(out), C.byref(n), comb))
        b = C.string_at(out, n.value)
        lib().sb200_free(out)
        pts = [b[i:i + 32] for i in range(0, len(b), 32)]
        return (pts, bytes(comb)) if want_comb else pts

    @staticmethod
    def network_products(inst, rx, ry, r_hash, r_multiset):
        """`PolyEvalNetwork::new` (sparse_mlpoly.rs:652-668) reduced to the 16 `ProductCircuit::evaluate` values: row side
        init, read x3, write x3, audit, then the column side (canonical 32-byte scalars)."""
        out = (C.c_uint8 * 512)()
        xp, xk = _ptr(rx)
        yp, yk = _ptr(ry)
        hp, hk = _ptr(r_hash)
        mp, mk = _ptr(r_multiset)
        _check(lib().sb200_snark_network_products(inst.h, xp, C.c_uint64(len(rx) // 32), yp, C.c_uint64(len(ry) // 32), hp, mp, out))
        b = bytes(out)
        return [b[32 * i:32 * i + 32] for i in range(16)]

    @staticmethod
    def last_encode_profile():
        d = (C.c_double * 4)()
        lib().sb200_last_encode_profile(d)
        return dict(zip(["dense_ms", "commit_ops_ms", "commit_mem_ms", "total_ms"], list(d)))
