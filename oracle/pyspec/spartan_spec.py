"""Pure-Python (big-int) restatement of the Spartan NIZK prove/verify path.

TEST INFRASTRUCTURE ONLY. Nothing in the product (`otti_b200/`) may import this
module; it exists to pin the C++ oracle (`oracle/*.cpp`) against an independent
implementation and to generate the small golden vectors under `tests/golden/`.
Pure-Python loops => small cases only (<= 2^8 constraints).

PARITY UNPINNED against a real `cargo` build of the reference (no Rust toolchain
in this image). What pins it instead: the reference's own F_l known answers
(Spartan/src/scalar/ristretto255.rs:784-1214), merlin's published test vector,
RFC 9496 ristretto255 vectors, hashlib SHAKE256, and the README proof length
47,024 B @ 2^20 (closed form, Spartan/README.md:376).

Each function cites the reference file:line (relative to /root/reference) it
follows.
"""
import hashlib
import struct

# ----------------------------------------------------------------------------
# F_l  (Spartan/src/scalar/ristretto255.rs)
# ----------------------------------------------------------------------------
ELL = 2**252 + 27742317777372353535851937790883648493  # ristretto255.rs:239-244
MONT_R = pow(2, 256, ELL)  # ristretto255.rs:306-311


def fl_from_bytes(b):
    """Scalar::from_bytes, ristretto255.rs:398-422: canonical LE, None if >= l."""
    v = int.from_bytes(b, "little")
    return v if v < ELL else None


def fl_to_bytes(a):
    """Scalar::to_bytes, ristretto255.rs:426-438 (canonical LE)."""
    return int(a % ELL).to_bytes(32, "little")


def fl_from_bytes_wide(b):
    """Scalar::from_bytes_wide, ristretto255.rs:442-453."""
    return int.from_bytes(b, "little") % ELL


def fl_inv(a):
    return pow(a, ELL - 2, ELL)


def fl_ser(a):
    """serde of Scalar([u64;4]) = raw Montgomery limbs (ristretto255.rs:199-200)."""
    return ((a * MONT_R) % ELL).to_bytes(32, "little")


def fl_deser(b):
    return (int.from_bytes(b, "little") * fl_inv(MONT_R)) % ELL


# ----------------------------------------------------------------------------
# F_p, Edwards25519 extended coordinates, ristretto255 (curve25519-dalek 3.2,
# restated from RFC 9496 / SURVEY App. A.6)
# ----------------------------------------------------------------------------
P = 2**255 - 19
D = (-121665 * pow(121666, P - 2, P)) % P
SQRT_M1 = pow(2, (P - 1) // 4, P)
INVSQRT_A_MINUS_D = 54469307008909316920995813868745141605393597292927456921205312896311721017578
ONE_MINUS_D_SQ = (1 - D * D) % P
D_MINUS_ONE_SQ = ((D - 1) * (D - 1)) % P
SQRT_AD_MINUS_ONE = 25063068953384623474111414158702152701244531502492656460079210482610430750235
BASE_X = 15112221349535400772501151409588531511454012693041857206046113283949847762202
BASE_Y = 46316835694926478169428394003475163141307993866256225615783033603165251855960


def fp_neg(x):
    return (x % P) & 1 == 1


def fp_abs(x):
    x %= P
    return (P - x) % P if x & 1 else x


def sqrt_ratio_m1(u, v):
    u %= P
    v %= P
    v3 = v * v % P * v % P
    v7 = v3 * v3 % P * v % P
    r = (u * v3) % P * pow(u * v7 % P, (P - 5) // 8, P) % P
    c = v * r % P * r % P
    ok = c == u
    f = c == (-u) % P
    fi = c == (-u * SQRT_M1) % P
    if f or fi:
        r = r * SQRT_M1 % P
    return (ok or f), fp_abs(r)


class Pt:
    """Edwards extended (X:Y:Z:T), a = -1."""
    __slots__ = ("X", "Y", "Z", "T")

    def __init__(self, X, Y, Z, T):
        self.X, self.Y, self.Z, self.T = X % P, Y % P, Z % P, T % P

    def __add__(self, o):
        A = (self.Y - self.X) * (o.Y - o.X) % P
        B = (self.Y + self.X) * (o.Y + o.X) % P
        C = 2 * D * self.T % P * o.T % P
        Dd = 2 * self.Z * o.Z % P
        E, F, G, H = B - A, Dd - C, Dd + C, B + A
        return Pt(E * F, G * H, F * G, E * H)

    def __neg__(self):
        return Pt(-self.X, self.Y, self.Z, -self.T)

    def __sub__(self, o):
        return self + (-o)

    def __rmul__(self, k):
        k %= ELL
        acc = IDENT
        base = self
        while k:
            if k & 1:
                acc = acc + base
            base = base + base
            k >>= 1
        return acc

    def __eq__(self, o):
        # ristretto equality (group.rs via dalek RistrettoPoint::ct_eq)
        return (self.X * o.Y - self.Y * o.X) % P == 0 or (self.X * o.X - self.Y * o.Y) % P == 0

    def compress(self):
        """RistrettoPoint::compress (RFC 9496 4.3.2)."""
        X, Y, Z, T = self.X, self.Y, self.Z, self.T
        u1 = (Z + Y) * (Z - Y) % P
        u2 = X * Y % P
        _, inv = sqrt_ratio_m1(1, u1 * u2 % P * u2 % P)
        d1 = inv * u1 % P
        d2 = inv * u2 % P
        zinv = d1 * d2 % P * T % P
        if fp_neg(T * zinv):
            x, y, dn = Y * SQRT_M1 % P, X * SQRT_M1 % P, d1 * INVSQRT_A_MINUS_D % P
        else:
            x, y, dn = X, Y, d2
        if fp_neg(x * zinv):
            y = (-y) % P
        s = fp_abs(dn * (Z - y))
        return s.to_bytes(32, "little")


IDENT = Pt(0, 1, 1, 0)
BASEPOINT = Pt(BASE_X, BASE_Y, 1, BASE_X * BASE_Y)


def decompress(b):
    """CompressedRistretto::decompress (RFC 9496 4.3.1); None on failure."""
    s = int.from_bytes(b, "little")
    if s >= P or s & 1:
        return None
    ss = s * s % P
    u1 = (1 - ss) % P
    u2 = (1 + ss) % P
    u2s = u2 * u2 % P
    v = (-D * u1 % P * u1 - u2s) % P
    ok, inv = sqrt_ratio_m1(1, v * u2s % P)
    dx = inv * u2 % P
    dy = inv * dx % P * v % P
    x = fp_abs(2 * s * dx)
    y = u1 * dy % P
    t = x * y % P
    if (not ok) or fp_neg(t) or y == 0:
        return None
    return Pt(x, y, 1, t)


def elligator(t):
    """RistrettoPoint::elligator_ristretto_flavor (RFC 9496 MAP)."""
    r = SQRT_M1 * t % P * t % P
    u = (r + 1) * ONE_MINUS_D_SQ % P
    v = (-1 - r * D) * (r + D) % P
    sq, s = sqrt_ratio_m1(u, v)
    sp = (-fp_abs(s * t)) % P
    if not sq:
        s = sp
    c = (P - 1) if sq else r
    N = (c * (r - 1) % P * D_MINUS_ONE_SQ - v) % P
    w0 = 2 * s * v % P
    w1 = N * SQRT_AD_MINUS_ONE % P
    w2 = (1 - s * s) % P
    w3 = (1 + s * s) % P
    return Pt(w0 * w3, w2 * w1, w1 * w3, w0 * w2)


def from_uniform_bytes(b):
    """RistrettoPoint::from_uniform_bytes (commitments.rs:25)."""
    lo = int.from_bytes(b[:32], "little") & (2**255 - 1)
    hi = int.from_bytes(b[32:], "little") & (2**255 - 1)
    return elligator(lo % P) + elligator(hi % P)


def msm(scalars, points):
    """GroupElement::vartime_multiscalar_mul (group.rs:98-117); value only."""
    acc = IDENT
    for s, p in zip(scalars, points):
        acc = acc + (s * p)
    return acc


# ----------------------------------------------------------------------------
# Keccak-f[1600], STROBE-128, Merlin (merlin 3.0.0)
# ----------------------------------------------------------------------------
_RC = [
    0x0000000000000001, 0x0000000000008082, 0x800000000000808A, 0x8000000080008000,
    0x000000000000808B, 0x0000000080000001, 0x8000000080008081, 0x8000000000008009,
    0x000000000000008A, 0x0000000000000088, 0x0000000080008009, 0x000000008000000A,
    0x000000008000808B, 0x800000000000008B, 0x8000000000008089, 0x8000000000008003,
    0x8000000000008002, 0x8000000000000080, 0x000000000000800A, 0x800000008000000A,
    0x8000000080008081, 0x8000000000008080, 0x0000000080000001, 0x8000000080008008,
]
_ROT = [[0, 36, 3, 41, 18], [1, 44, 10, 45, 2], [62, 6, 43, 15, 61], [28, 55, 25, 21, 56], [27, 20, 39, 8, 14]]
_M64 = (1 << 64) - 1


def _rol(x, n):
    n %= 64
    return ((x << n) | (x >> (64 - n))) & _M64 if n else x


def keccak_f1600(state):
    lanes = list(struct.unpack("<25Q", bytes(state)))
    A = [[lanes[x + 5 * y] for y in range(5)] for x in range(5)]
    for rnd in range(24):
        C = [A[x][0] ^ A[x][1] ^ A[x][2] ^ A[x][3] ^ A[x][4] for x in range(5)]
        Dd = [C[(x - 1) % 5] ^ _rol(C[(x + 1) % 5], 1) for x in range(5)]
        A = [[A[x][y] ^ Dd[x] for y in range(5)] for x in range(5)]
        B = [[0] * 5 for _ in range(5)]
        for x in range(5):
            for y in range(5):
                B[y][(2 * x + 3 * y) % 5] = _rol(A[x][y], _ROT[x][y])
        A = [[B[x][y] ^ ((~B[(x + 1) % 5][y]) & B[(x + 2) % 5][y] & _M64) for y in range(5)] for x in range(5)]
        A[0][0] ^= _RC[rnd]
    out = [0] * 25
    for x in range(5):
        for y in range(5):
            out[x + 5 * y] = A[x][y]
    return bytearray(struct.pack("<25Q", *out))


class Strobe128:
    R = 166
    FLAG_I, FLAG_A, FLAG_C, FLAG_T, FLAG_M, FLAG_K = 1, 2, 4, 8, 16, 32

    def __init__(self, label):
        st = bytearray(200)
        st[0:6] = bytes([1, self.R + 2, 1, 0, 1, 96])
        st[6:18] = b"STROBEv1.0.2"
        self.st = keccak_f1600(st)
        self.pos = 0
        self.pos_begin = 0
        self.cur_flags = 0
        self.meta_ad(label, False)

    def _run_f(self):
        self.st[self.pos] ^= self.pos_begin
        self.st[self.pos + 1] ^= 0x04
        self.st[self.R + 1] ^= 0x80
        self.st = keccak_f1600(self.st)
        self.pos = 0
        self.pos_begin = 0

    def _absorb(self, data):
        for b in data:
            self.st[self.pos] ^= b
            self.pos += 1
            if self.pos == self.R:
                self._run_f()

    def _squeeze(self, n):
        out = bytearray(n)
        for i in range(n):
            out[i] = self.st[self.pos]
            self.st[self.pos] = 0
            self.pos += 1
            if self.pos == self.R:
                self._run_f()
        return bytes(out)

    def _begin_op(self, flags, more):
        if more:
            assert self.cur_flags == flags
            return
        assert flags & self.FLAG_T == 0
        old_begin = self.pos_begin
        self.pos_begin = self.pos + 1
        self.cur_flags = flags
        self._absorb(bytes([old_begin, flags]))
        if flags & (self.FLAG_C | self.FLAG_K) and self.pos != 0:
            self._run_f()

    def meta_ad(self, data, more):
        self._begin_op(self.FLAG_M | self.FLAG_A, more)
        self._absorb(data)

    def ad(self, data, more):
        self._begin_op(self.FLAG_A, more)
        self._absorb(data)

    def prf(self, n, more):
        self._begin_op(self.FLAG_I | self.FLAG_A | self.FLAG_C, more)
        return self._squeeze(n)


class Transcript:
    """merlin::Transcript + Spartan's ProofTranscript (transcript.rs:13-37)."""

    def __init__(self, label):
        self.s = Strobe128(b"Merlin v1.0")
        self.append_message(b"dom-sep", label)

    def append_message(self, label, msg):
        self.s.meta_ad(label, False)
        self.s.meta_ad(struct.pack("<I", len(msg)), True)
        self.s.ad(msg, False)

    def challenge_bytes(self, label, n):
        self.s.meta_ad(label, False)
        self.s.meta_ad(struct.pack("<I", n), True)
        return self.s.prf(n, False)

    # ProofTranscript
    def append_protocol_name(self, name):
        self.append_message(b"protocol-name", name)

    def append_scalar(self, label, s):
        self.append_message(label, fl_to_bytes(s))

    def append_point(self, label, cbytes):
        self.append_message(label, cbytes)

    def challenge_scalar(self, label):
        return fl_from_bytes_wide(self.challenge_bytes(label, 64))

    def challenge_vector(self, label, n):
        return [self.challenge_scalar(label) for _ in range(n)]


class RandomTape:
    """random.rs:11-27 with the OsRng seed scalar injected."""

    def __init__(self, name, seed_scalar):
        self.t = Transcript(name)
        self.t.append_scalar(b"init_randomness", seed_scalar)

    def random_scalar(self, label):
        return self.t.challenge_scalar(label)

    def random_vector(self, label, n):
        return self.t.challenge_vector(label, n)


# ----------------------------------------------------------------------------
# Generators (commitments.rs:15-58, nizk/mod.rs:409-414, dense_mlpoly.rs:28-35,
# r1csproof.rs:45-70, lib.rs:466-477)
# ----------------------------------------------------------------------------
BASEPOINT_COMPRESSED = bytes.fromhex("e2f2ae0a6abc4e71a884a961c500515f58e30b6aa582dd8db6a65945e08d2d76")


class MultiCommitGens:
    def __init__(self, n, G, h):
        self.n, self.G, self.h = n, G, h

    @staticmethod
    def new(n, label):
        stream = hashlib.shake_256(label + BASEPOINT_COMPRESSED).digest(64 * (n + 1))
        pts = [from_uniform_bytes(stream[64 * i:64 * (i + 1)]) for i in range(n + 1)]
        return MultiCommitGens(n, pts[:n], pts[n])

    def split_at(self, mid):
        return (MultiCommitGens(mid, self.G[:mid], self.h),
                MultiCommitGens(self.n - mid, self.G[mid:], self.h))


def commit_scalar(x, blind, g1):
    """impl Commitments for Scalar, commitments.rs:65-70."""
    assert g1.n == 1
    return msm([x, blind], [g1.G[0], g1.h])


def commit_vec(xs, blind, gn):
    """impl Commitments for [Scalar], commitments.rs:79-84."""
    assert gn.n == len(xs)
    return msm(xs, gn.G) + (blind * gn.h)


def log2(n):
    return n.bit_length() - 1


class NIZKGens:
    def __init__(self, num_cons, num_vars, num_inputs, label=b"gens_r1cs_sat"):
        nv = max(num_vars, num_inputs + 1)
        if nv & (nv - 1):
            nv = 1 << nv.bit_length()
        self.num_vars_padded = nv
        ell = log2(nv)
        right = ell - ell // 2
        n = 1 << right
        self.gens_n, self.gens_1 = MultiCommitGens.new(n + 1, label).split_at(n)
        self.gens_3 = MultiCommitGens.new(3, label)
        self.gens_4 = MultiCommitGens.new(4, label)


# ----------------------------------------------------------------------------
# Polynomials
# ----------------------------------------------------------------------------
def eq_evals(r):
    """EqPolynomial::evals, dense_mlpoly.rs:67-83."""
    ev = [1] * (1 << len(r))
    size = 1
    for j in range(len(r)):
        size *= 2
        for i in range(size - 1, -1, -2):
            s = ev[i // 2]
            ev[i] = s * r[j] % ELL
            ev[i - 1] = (s - ev[i]) % ELL
    return ev


def factored_lens(ell):
    return ell // 2, ell - ell // 2


def unipoly_from_evals(e):
    """UniPoly::from_evals, unipoly.rs:23-54."""
    two_inv = fl_inv(2)
    if len(e) == 3:
        c = e[0]
        a = two_inv * (e[2] - e[1] - e[1] + c) % ELL
        b = (e[1] - c - a) % ELL
        return [c, b, a]
    six_inv = fl_inv(6)
    d = e[0]
    a = six_inv * (e[3] - 3 * e[2] + 3 * e[1] - e[0]) % ELL
    b = two_inv * (2 * e[0] - 5 * e[1] + 4 * e[2] - e[3]) % ELL
    c = (e[1] - d - a - b) % ELL
    return [d, c, b, a]


def unipoly_eval(co, r):
    """UniPoly::evaluate, unipoly.rs:72-80."""
    ev, power = co[0], r
    for i in range(1, len(co)):
        ev = (ev + power * co[i]) % ELL
        power = power * r % ELL
    return ev


# ----------------------------------------------------------------------------
# R1CS instance (lib.rs:121-225, r1csinstance.rs)
# ----------------------------------------------------------------------------
class Instance:
    """Instance::new: entries are (row, col, canonical 32-byte value)."""

    def __init__(self, num_cons, num_vars, num_inputs, A, B, C):
        nv = max(num_vars, num_inputs + 1)
        if nv & (nv - 1):
            nv = 1 << nv.bit_length()
        # lib.rs:143-156 (note: tests the *unpadded* num_cons in the second if; 0 -> 1, 1 -> 2)
        nc = num_cons
        if nc in (0, 1):
            nc = 2
        npt = 1 if num_cons == 0 else 1 << (num_cons - 1).bit_length()
        if npt != num_cons:
            nc = npt
        self.num_cons, self.num_vars, self.num_inputs = nc, nv, num_inputs

        def conv(tups):
            mat = []
            for row, col, vb in tups:
                if row >= num_cons or col >= num_vars + 1 + num_inputs:
                    raise ValueError("InvalidIndex")
                v = fl_from_bytes(vb)
                if v is None:
                    raise ValueError("InvalidScalar")
                mat.append((row, col + nv - num_vars if col >= num_vars else col, v))
            if num_cons in (0, 1):
                for i in range(len(tups), nc):
                    mat.append((i, num_vars, 0))
            return mat

        self.A, self.B, self.C = conv(A), conv(B), conv(C)

    def multiply_vec(self, z):
        out = []
        for M in (self.A, self.B, self.C):
            v = [0] * self.num_cons
            for r, c, val in M:
                v[r] = (v[r] + val * z[c]) % ELL
            out.append(v)
        return out

    def eval_table(self, rx_evals, num_cols):
        out = []
        for M in (self.A, self.B, self.C):
            v = [0] * num_cols
            for r, c, val in M:
                v[c] = (v[c] + rx_evals[r] * val) % ELL
            out.append(v)
        return out

    def evaluate(self, rx, ry):
        ex, ey = eq_evals(rx), eq_evals(ry)
        return [sum(ex[r] * ey[c] % ELL * v for r, c, v in M) % ELL for M in (self.A, self.B, self.C)]

    def is_sat(self, vars_, inputs):
        vars_ = list(vars_) + [0] * (self.num_vars - len(vars_))
        z = vars_ + [1] + list(inputs)
        z = z + [0] * (2 * self.num_vars - len(z))
        Az, Bz, Cz = self.multiply_vec(z)
        return all((a * b - c) % ELL == 0 for a, b, c in zip(Az, Bz, Cz))


def splitmix64(state):
    state = (state + 0x9E3779B97F4A7C15) & _M64
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return state, z ^ (z >> 31)


def synthetic_r1cs(num_cons, num_vars, num_inputs, seed):
    """R1CSInstance::produce_synthetic_r1cs (r1csinstance.rs:136-214), OsRng
    replaced by SplitMix64(seed): each Scalar::random = from_u512(8 x next_u64)."""
    size_z = num_vars + num_inputs + 1
    st = seed
    Z = []
    for _ in range(size_z):
        limbs = []
        for _ in range(8):
            st, v = splitmix64(st)
            limbs.append(v)
        Z.append(int.from_bytes(struct.pack("<8Q", *limbs), "little") % ELL)
    Z[num_vars] = 1
    A, B, C = [], [], []
    one = fl_to_bytes(1)
    for i in range(num_cons):
        ai, bi, ci = i % size_z, (i + 2) % size_z, (i + 3) % size_z
        A.append((i, ai, one))
        B.append((i, bi, one))
        ab = Z[ai] * Z[bi] % ELL
        if Z[ci] == 0:
            C.append((i, num_vars, fl_to_bytes(ab)))
        else:
            C.append((i, ci, fl_to_bytes(ab * fl_inv(Z[ci]) % ELL)))
    inst = Instance(num_cons, num_vars, num_inputs, A, B, C)
    return inst, Z[:num_vars], Z[num_vars + 1:]


# ----------------------------------------------------------------------------
# Sigma protocols (nizk/mod.rs, nizk/bullet.rs)
# ----------------------------------------------------------------------------
def knowledge_prove(g1, tr, tape, x, r):
    """KnowledgeProof::prove, nizk/mod.rs:26-51."""
    tr.append_protocol_name(b"knowledge proof")
    t1, t2 = tape.random_scalar(b"t1"), tape.random_scalar(b"t2")
    C = commit_scalar(x, r, g1).compress()
    tr.append_point(b"C", C)
    alpha = commit_scalar(t1, t2, g1).compress()
    tr.append_point(b"alpha", alpha)
    c = tr.challenge_scalar(b"c")
    return {"alpha": alpha, "z1": (x * c + t1) % ELL, "z2": (r * c + t2) % ELL}, C


def knowledge_verify(pf, g1, tr, C):
    tr.append_protocol_name(b"knowledge proof")
    tr.append_point(b"C", C)
    tr.append_point(b"alpha", pf["alpha"])
    c = tr.challenge_scalar(b"c")
    lhs = commit_scalar(pf["z1"], pf["z2"], g1).compress()
    rhs = ((c * decompress(C)) + decompress(pf["alpha"])).compress()
    return lhs == rhs


def equality_prove(g1, tr, tape, v1, s1, v2, s2):
    """EqualityProof::prove, nizk/mod.rs:87-115."""
    tr.append_protocol_name(b"equality proof")
    r = tape.random_scalar(b"r")
    C1 = commit_scalar(v1, s1, g1).compress()
    tr.append_point(b"C1", C1)
    C2 = commit_scalar(v2, s2, g1).compress()
    tr.append_point(b"C2", C2)
    alpha = (r * g1.h).compress()
    tr.append_point(b"alpha", alpha)
    c = tr.challenge_scalar(b"c")
    return {"alpha": alpha, "z": (c * (s1 - s2) + r) % ELL}, C1, C2


def equality_verify(pf, g1, tr, C1, C2):
    tr.append_protocol_name(b"equality proof")
    tr.append_point(b"C1", C1)
    tr.append_point(b"C2", C2)
    tr.append_point(b"alpha", pf["alpha"])
    c = tr.challenge_scalar(b"c")
    Cd = decompress(C1) - decompress(C2)
    rhs = ((c * Cd) + decompress(pf["alpha"])).compress()
    lhs = (pf["z"] * g1.h).compress()
    return lhs == rhs


def product_prove(g1, tr, tape, x, rX, y, rY, z, rZ):
    """ProductProof::prove, nizk/mod.rs:158-228."""
    tr.append_protocol_name(b"product proof")
    b1, b2, b3, b4, b5 = [tape.random_scalar(l) for l in (b"b1", b"b2", b"b3", b"b4", b"b5")]
    X = commit_scalar(x, rX, g1).compress()
    tr.append_point(b"X", X)
    Y = commit_scalar(y, rY, g1).compress()
    tr.append_point(b"Y", Y)
    Z = commit_scalar(z, rZ, g1).compress()
    tr.append_point(b"Z", Z)
    alpha = commit_scalar(b1, b2, g1).compress()
    tr.append_point(b"alpha", alpha)
    beta = commit_scalar(b3, b4, g1).compress()
    tr.append_point(b"beta", beta)
    gX = MultiCommitGens(1, [decompress(X)], g1.h)
    delta = commit_scalar(b3, b5, gX).compress()
    tr.append_point(b"delta", delta)
    c = tr.challenge_scalar(b"c")
    zs = [(b1 + c * x) % ELL, (b2 + c * rX) % ELL, (b3 + c * y) % ELL, (b4 + c * rY) % ELL,
          (b5 + c * (rZ - rX * y)) % ELL]
    return {"alpha": alpha, "beta": beta, "delta": delta, "z": zs}, X, Y, Z


def product_verify(pf, g1, tr, X, Y, Z):
    tr.append_protocol_name(b"product proof")
    for l, v in ((b"X", X), (b"Y", Y), (b"Z", Z), (b"alpha", pf["alpha"]), (b"beta", pf["beta"]),
                 (b"delta", pf["delta"])):
        tr.append_point(l, v)
    z1, z2, z3, z4, z5 = pf["z"]
    c = tr.challenge_scalar(b"c")

    def chk(Pc, Xc, g, za, zb):
        lhs = (decompress(Pc) + (c * decompress(Xc))).compress()
        return lhs == commit_scalar(za, zb, g).compress()

    gX = MultiCommitGens(1, [decompress(X)], g1.h)
    return chk(pf["alpha"], X, g1, z1, z2) and chk(pf["beta"], Y, g1, z3, z4) and chk(pf["delta"], Z, gX, z3, z5)


def dotproduct_prove(g1, gn, tr, tape, x_vec, blind_x, a_vec, y, blind_y):
    """DotProductProof::prove, nizk/mod.rs:310-367."""
    tr.append_protocol_name(b"dot product proof")
    n = len(x_vec)
    d_vec = tape.random_vector(b"d_vec", n)
    r_delta = tape.random_scalar(b"r_delta")
    r_beta = tape.random_scalar(b"r_beta")
    Cx = commit_vec(x_vec, blind_x, gn).compress()
    tr.append_point(b"Cx", Cx)
    Cy = commit_scalar(y, blind_y, g1).compress()
    tr.append_point(b"Cy", Cy)
    delta = commit_vec(d_vec, r_delta, gn).compress()
    tr.append_point(b"delta", delta)
    dp = sum(a * d for a, d in zip(a_vec, d_vec)) % ELL
    beta = commit_scalar(dp, r_beta, g1).compress()
    tr.append_point(b"beta", beta)
    c = tr.challenge_scalar(b"c")
    z = [(c * x + d) % ELL for x, d in zip(x_vec, d_vec)]
    return {"delta": delta, "beta": beta, "z": z, "z_delta": (c * blind_x + r_delta) % ELL,
            "z_beta": (c * blind_y + r_beta) % ELL}, Cx, Cy


def dotproduct_verify(pf, g1, gn, tr, a, Cx, Cy):
    tr.append_protocol_name(b"dot product proof")
    tr.append_point(b"Cx", Cx)
    tr.append_point(b"Cy", Cy)
    tr.append_point(b"delta", pf["delta"])
    tr.append_point(b"beta", pf["beta"])
    c = tr.challenge_scalar(b"c")
    ok = ((c * decompress(Cx)) + decompress(pf["delta"])) == commit_vec(pf["z"], pf["z_delta"], gn)
    dz = sum(x * y for x, y in zip(pf["z"], a)) % ELL
    ok &= ((c * decompress(Cy)) + decompress(pf["beta"])) == commit_scalar(dz, pf["z_beta"], g1)
    return ok


def bullet_prove(tr, Q, G_vec, H, a_vec, b_vec, blind, blinds_vec):
    """BulletReductionProof::prove, nizk/bullet.rs:31-131."""
    G, a, b = list(G_vec), list(a_vec), list(b_vec)
    n = len(G)
    L_vec, R_vec = [], []
    blind_fin = blind
    it = iter(blinds_vec)
    while n != 1:
        n //= 2
        aL, aR, bL, bR, GL, GR = a[:n], a[n:2 * n], b[:n], b[n:2 * n], G[:n], G[n:2 * n]
        cL = sum(x * y for x, y in zip(aL, bR)) % ELL
        cR = sum(x * y for x, y in zip(aR, bL)) % ELL
        blL, blR = next(it)
        Lp = msm(aL + [cL, blL], GR + [Q, H])
        Rp = msm(aR + [cR, blR], GL + [Q, H])
        tr.append_point(b"L", Lp.compress())
        tr.append_point(b"R", Rp.compress())
        u = tr.challenge_scalar(b"u")
        ui = fl_inv(u)
        a = [(aL[i] * u + ui * aR[i]) % ELL for i in range(n)]
        b = [(bL[i] * ui + u * bR[i]) % ELL for i in range(n)]
        G = [(ui * GL[i]) + (u * GR[i]) for i in range(n)]
        blind_fin = (blind_fin + blL * u * u + blR * ui * ui) % ELL
        L_vec.append(Lp.compress())
        R_vec.append(Rp.compress())
    return {"L_vec": L_vec, "R_vec": R_vec}, a[0], b[0], G[0], blind_fin


def bullet_verify(pf, n, a, tr, Gamma, G):
    """BulletReductionProof::verify, nizk/bullet.rs:136-224."""
    lg_n = len(pf["L_vec"])
    if n != (1 << lg_n):
        return None
    ch = []
    for Lc, Rc in zip(pf["L_vec"], pf["R_vec"]):
        tr.append_point(b"L", Lc)
        tr.append_point(b"R", Rc)
        ch.append(tr.challenge_scalar(b"u"))
    ch_inv = [fl_inv(c) for c in ch]
    allinv = 1
    for c in ch_inv:
        allinv = allinv * c % ELL
    u_sq = [c * c % ELL for c in ch]
    u_inv_sq = [c * c % ELL for c in ch_inv]
    s = [allinv]
    for i in range(1, n):
        lg_i = i.bit_length() - 1
        k = 1 << lg_i
        s.append(s[i - k] * u_sq[(lg_n - 1) - lg_i] % ELL)
    Ls = [decompress(p) for p in pf["L_vec"]]
    Rs = [decompress(p) for p in pf["R_vec"]]
    if any(p is None for p in Ls + Rs):
        return None
    G_hat = msm(s, G)
    a_hat = sum(x * y for x, y in zip(a, s)) % ELL
    Gamma_hat = msm(u_sq + u_inv_sq + [1], Ls + Rs + [Gamma])
    return G_hat, Gamma_hat, a_hat


def dotproductlog_prove(gens_n, gens_1, tr, tape, x_vec, blind_x, a_vec, y, blind_y):
    """DotProductProofLog::prove, nizk/mod.rs:435-512."""
    tr.append_protocol_name(b"dot product proof (log)")
    n = len(x_vec)
    d = tape.random_scalar(b"d")
    r_delta = tape.random_scalar(b"r_delta")
    r_beta = tape.random_scalar(b"r_delta")  # sic: nizk/mod.rs:454
    v1 = tape.random_vector(b"blinds_vec_1", 2 * log2(n))
    v2 = tape.random_vector(b"blinds_vec_2", 2 * log2(n))
    blinds_vec = list(zip(v1, v2))
    Cx = commit_vec(x_vec, blind_x, gens_n).compress()
    tr.append_point(b"Cx", Cx)
    Cy = commit_scalar(y, blind_y, gens_1).compress()
    tr.append_point(b"Cy", Cy)
    blind_Gamma = (blind_x + blind_y) % ELL
    bp, x_hat, a_hat, g_hat, rhat = bullet_prove(tr, gens_1.G[0], gens_n.G, gens_n.h, x_vec, a_vec,
                                                 blind_Gamma, blinds_vec)
    y_hat = x_hat * a_hat % ELL
    delta = commit_scalar(d, r_delta, MultiCommitGens(1, [g_hat], gens_1.h)).compress()
    tr.append_point(b"delta", delta)
    beta = commit_scalar(d, r_beta, gens_1).compress()
    tr.append_point(b"beta", beta)
    c = tr.challenge_scalar(b"c")
    z1 = (d + c * y_hat) % ELL
    z2 = (a_hat * (c * rhat + r_beta) + r_delta) % ELL
    return {"bullet": bp, "delta": delta, "beta": beta, "z1": z1, "z2": z2}, Cx, Cy


def dotproductlog_verify(pf, n, gens_n, gens_1, tr, a, Cx, Cy):
    tr.append_protocol_name(b"dot product proof (log)")
    tr.append_point(b"Cx", Cx)
    tr.append_point(b"Cy", Cy)
    Gamma = decompress(Cx) + decompress(Cy)
    res = bullet_verify(pf["bullet"], n, a, tr, Gamma, gens_n.G)
    if res is None:
        return False
    g_hat, Gamma_hat, a_hat = res
    tr.append_point(b"delta", pf["delta"])
    tr.append_point(b"beta", pf["beta"])
    c = tr.challenge_scalar(b"c")
    lhs = ((a_hat * ((c * Gamma_hat) + decompress(pf["beta"]))) + decompress(pf["delta"])).compress()
    rhs = ((pf["z1"] * (g_hat + (a_hat * gens_1.G[0]))) + (pf["z2"] * gens_1.h)).compress()
    return lhs == rhs


# ----------------------------------------------------------------------------
# ZK sum-check (sumcheck.rs:426-774)
# ----------------------------------------------------------------------------
def zk_sumcheck_prove(claim, blind_claim, num_rounds, tables, degree, comb, g1, gn, tr, tape):
    """prove_quad (degree 2, 2 tables) / prove_cubic_with_additive_term (degree 3,
    4 tables).  tables are mutated (bound) in place."""
    blinds_poly = tape.random_vector(b"blinds_poly", num_rounds)
    blinds_evals = tape.random_vector(b"blinds_evals", num_rounds)
    claim_per_round = claim
    comm_claim_per_round = commit_scalar(claim, blind_claim, g1).compress()
    r, comm_polys, comm_evals, proofs = [], [], [], []
    for j in range(num_rounds):
        h = len(tables[0]) // 2
        e0 = e2 = e3 = 0
        for i in range(h):
            lo = [t[i] for t in tables]
            hi = [t[h + i] for t in tables]
            e0 += comb(*lo)
            p2 = [(2 * b - a) % ELL for a, b in zip(lo, hi)]
            e2 += comb(*p2)
            if degree == 3:
                p3 = [(3 * b - 2 * a) % ELL for a, b in zip(lo, hi)]
                e3 += comb(*p3)
        e0, e2, e3 = e0 % ELL, e2 % ELL, e3 % ELL
        evals = [e0, (claim_per_round - e0) % ELL, e2] + ([e3] if degree == 3 else [])
        poly = unipoly_from_evals(evals)
        comm_poly = commit_vec(poly, blinds_poly[j], gn).compress()
        tr.append_point(b"comm_poly", comm_poly)
        comm_polys.append(comm_poly)
        r_j = tr.challenge_scalar(b"challenge_nextround")
        for t in tables:  # bound_poly_var_top, dense_mlpoly.rs:214-221
            for i in range(h):
                t[i] = (t[i] + r_j * (t[h + i] - t[i])) % ELL
            del t[h:]
        ev = unipoly_eval(poly, r_j)
        comm_eval = commit_scalar(ev, blinds_evals[j], g1).compress()
        tr.append_point(b"comm_claim_per_round", comm_claim_per_round)
        tr.append_point(b"comm_eval", comm_eval)
        w = tr.challenge_vector(b"combine_two_claims_to_one", 2)
        target = (w[0] * claim_per_round + w[1] * ev) % ELL
        comm_target = msm(w, [decompress(comm_claim_per_round), decompress(comm_eval)]).compress()
        blind_sc = blind_claim if j == 0 else blinds_evals[j - 1]
        blind = (w[0] * blind_sc + w[1] * blinds_evals[j]) % ELL
        assert commit_scalar(target, blind, g1).compress() == comm_target
        a_sc = [2] + [1] * degree
        a_ev = [1]
        for _ in range(degree):
            a_ev.append(a_ev[-1] * r_j % ELL)
        a = [(w[0] * x + w[1] * y) % ELL for x, y in zip(a_sc, a_ev)]
        pf, _, _ = dotproduct_prove(g1, gn, tr, tape, poly, blinds_poly[j], a, target, blind)
        proofs.append(pf)
        claim_per_round = ev
        comm_claim_per_round = comm_eval
        r.append(r_j)
        comm_evals.append(comm_eval)
    return ({"comm_polys": comm_polys, "comm_evals": comm_evals, "proofs": proofs}, r,
            [t[0] for t in tables], blinds_evals[num_rounds - 1])


def zk_sumcheck_verify(pf, comm_claim, num_rounds, degree, g1, gn, tr):
    """ZKSumcheckInstanceProof::verify, sumcheck.rs:84-177."""
    assert gn.n == degree + 1
    if len(pf["comm_polys"]) != num_rounds or len(pf["comm_evals"]) != num_rounds:
        return None
    r = []
    for i in range(num_rounds):
        comm_poly = pf["comm_polys"][i]
        tr.append_point(b"comm_poly", comm_poly)
        r_i = tr.challenge_scalar(b"challenge_nextround")
        ccpr = comm_claim if i == 0 else pf["comm_evals"][i - 1]
        comm_eval = pf["comm_evals"][i]
        tr.append_point(b"comm_claim_per_round", ccpr)
        tr.append_point(b"comm_eval", comm_eval)
        w = tr.challenge_vector(b"combine_two_claims_to_one", 2)
        comm_target = msm(w, [decompress(ccpr), decompress(comm_eval)]).compress()
        a_sc = [2] + [1] * degree
        a_ev = [1]
        for _ in range(degree):
            a_ev.append(a_ev[-1] * r_i % ELL)
        a = [(w[0] * x + w[1] * y) % ELL for x, y in zip(a_sc, a_ev)]
        if not dotproduct_verify(pf["proofs"][i], g1, gn, tr, a, comm_poly, comm_target):
            return None
        r.append(r_i)
    return pf["comm_evals"][-1], r


# ----------------------------------------------------------------------------
# NIZK::prove / verify (lib.rs:493-574, r1csproof.rs:141-498)
# ----------------------------------------------------------------------------
def nizk_prove(inst, vars_, inputs, gens, tr, tape_seed):
    """NIZK::prove, lib.rs:493-535."""
    tape = RandomTape(b"proof", tape_seed)
    tr.append_protocol_name(b"Spartan NIZK proof")
    vars_ = list(vars_) + [0] * (inst.num_vars - len(vars_))
    return r1cs_prove(inst, vars_, inputs, gens, tr, tape)


def r1cs_prove(inst, vars_, inputs, gens, tr, tape):
    """R1CSProof::prove, r1csproof.rs:141-345 (shared by NIZK::prove and SNARK::prove; vars_ already padded)."""
    tr.append_protocol_name(b"R1CS proof")
    assert len(inputs) < len(vars_)
    nv = len(vars_)
    # polycommit (dense_mlpoly.rs:178-203)
    ell = log2(nv)
    lnv, rnv = factored_lens(ell)
    Ls, Rs = 1 << lnv, 1 << rnv
    blinds_vars = tape.random_vector(b"poly_blinds", Ls)
    comm_vars = [commit_vec(vars_[Rs * i:Rs * (i + 1)], blinds_vars[i], gens.gens_n).compress() for i in range(Ls)]
    tr.append_message(b"poly_commitment", b"poly_commitment_begin")
    for c in comm_vars:
        tr.append_point(b"poly_commitment_share", c)
    tr.append_message(b"poly_commitment", b"poly_commitment_end")
    # phase 1
    z = vars_ + [1] + list(inputs) + [0] * (nv - len(inputs) - 1)
    nrx, nry = log2(inst.num_cons), log2(len(z))
    tau = tr.challenge_vector(b"challenge_tau", nrx)
    poly_tau = eq_evals(tau)
    Az, Bz, Cz = inst.multiply_vec(z)
    comb3 = lambda a, b, c, d: a * (b * c - d) % ELL
    sc1, rx, claims1, blind_post1 = zk_sumcheck_prove(0, 0, nrx, [poly_tau, Az, Bz, Cz], 3, comb3,
                                                      gens.gens_1, gens.gens_4, tr, tape)
    tau_claim, Az_claim, Bz_claim, Cz_claim = claims1
    Az_blind, Bz_blind, Cz_blind, prod_blind = [tape.random_scalar(l) for l in
                                                (b"Az_blind", b"Bz_blind", b"Cz_blind", b"prod_Az_Bz_blind")]
    pok_Cz, comm_Cz = knowledge_prove(gens.gens_1, tr, tape, Cz_claim, Cz_blind)
    prod = Az_claim * Bz_claim % ELL
    proof_prod, comm_Az, comm_Bz, comm_prod = product_prove(gens.gens_1, tr, tape, Az_claim, Az_blind, Bz_claim,
                                                            Bz_blind, prod, prod_blind)
    tr.append_point(b"comm_Az_claim", comm_Az)
    tr.append_point(b"comm_Bz_claim", comm_Bz)
    tr.append_point(b"comm_Cz_claim", comm_Cz)
    tr.append_point(b"comm_prod_Az_Bz_claims", comm_prod)
    blind_exp1 = tau_claim * (prod_blind - Cz_blind) % ELL
    claim_post1 = (Az_claim * Bz_claim - Cz_claim) * tau_claim % ELL
    eq1, _, _ = equality_prove(gens.gens_1, tr, tape, claim_post1, blind_exp1, claim_post1, blind_post1)
    # phase 2
    r_A = tr.challenge_scalar(b"challenege_Az")
    r_B = tr.challenge_scalar(b"challenege_Bz")
    r_C = tr.challenge_scalar(b"challenege_Cz")
    claim2 = (r_A * Az_claim + r_B * Bz_claim + r_C * Cz_claim) % ELL
    blind2 = (r_A * Az_blind + r_B * Bz_blind + r_C * Cz_blind) % ELL
    evals_rx = eq_evals(rx)
    eA, eB, eC = inst.eval_table(evals_rx, len(z))
    evals_ABC = [(r_A * a + r_B * b + r_C * c) % ELL for a, b, c in zip(eA, eB, eC)]
    comb2 = lambda a, b: a * b % ELL
    sc2, ry, claims2, blind_post2 = zk_sumcheck_prove(claim2, blind2, nry, [list(z), evals_ABC], 2, comb2,
                                                      gens.gens_1, gens.gens_3, tr, tape)
    # polyeval (dense_mlpoly.rs:309-362)
    chis = eq_evals(ry[1:])
    eval_vars = sum(a * b for a, b in zip(vars_, chis)) % ELL
    blind_eval = tape.random_scalar(b"blind_eval")
    tr.append_protocol_name(b"polynomial evaluation proof")
    r = ry[1:]
    Lv, Rv = eq_evals(r[:lnv]), eq_evals(r[lnv:])
    LZ = [sum(Lv[j] * vars_[j * Rs + i] for j in range(Ls)) % ELL for i in range(Rs)]
    LZ_blind = sum(b * l for b, l in zip(blinds_vars, Lv)) % ELL
    dpl, _, comm_vars_at_ry = dotproductlog_prove(gens.gens_n, gens.gens_1, tr, tape, LZ, LZ_blind, Rv, eval_vars,
                                                  blind_eval)
    blind_eval_Z = (1 - ry[0]) * blind_eval % ELL
    blind_exp2 = claims2[1] * blind_eval_Z % ELL
    claim_post2 = claims2[0] * claims2[1] % ELL
    eq2, _, _ = equality_prove(gens.gens_1, tr, tape, claim_post2, blind_exp2, claim_post2, blind_post2)
    return {
        "comm_vars": comm_vars, "sc1": sc1, "claims_phase2": (comm_Az, comm_Bz, comm_Cz, comm_prod),
        "pok": pok_Cz, "prod": proof_prod, "eq1": eq1, "sc2": sc2, "comm_vars_at_ry": comm_vars_at_ry,
        "dpl": dpl, "eq2": eq2, "rx": rx, "ry": ry,
    }


def sparse_poly_evaluate(num_vars, entries, r):
    """SparsePolynomial::evaluate, sparse_mlpoly.rs:1619-1628."""
    tot = 0
    for idx, val in entries:
        chi = 1
        for j in range(num_vars):
            bit = (idx >> (num_vars - j - 1)) & 1
            chi = chi * (r[j] if bit else (1 - r[j])) % ELL
        tot += chi * val
    return tot % ELL


def nizk_verify(pf, inst, inputs, gens, tr):
    """NIZK::verify lib.rs:538-574 + R1CSProof::verify r1csproof.rs:347-498.
    Returns True/False (panics of the reference become False)."""
    try:
        return _nizk_verify(pf, inst, inputs, gens, tr)
    except (TypeError, AttributeError, AssertionError):
        return False  # decompress() returned None somewhere


def _nizk_verify(pf, inst, inputs, gens, tr):
    tr.append_protocol_name(b"Spartan NIZK proof")
    evA, evB, evC = inst.evaluate(pf["rx"], pf["ry"])
    assert len(inputs) == inst.num_inputs
    tr.append_protocol_name(b"R1CS proof")
    n = inst.num_vars
    tr.append_message(b"poly_commitment", b"poly_commitment_begin")
    for c in pf["comm_vars"]:
        tr.append_point(b"poly_commitment_share", c)
    tr.append_message(b"poly_commitment", b"poly_commitment_end")
    nrx, nry = log2(inst.num_cons), log2(2 * n)
    tau = tr.challenge_vector(b"challenge_tau", nrx)
    claim1 = commit_scalar(0, 0, gens.gens_1).compress()
    res = zk_sumcheck_verify(pf["sc1"], claim1, nrx, 3, gens.gens_1, gens.gens_4, tr)
    if res is None:
        return False
    comm_post1, rx = res
    comm_Az, comm_Bz, comm_Cz, comm_prod = pf["claims_phase2"]
    if not knowledge_verify(pf["pok"], gens.gens_1, tr, comm_Cz):
        return False
    if not product_verify(pf["prod"], gens.gens_1, tr, comm_Az, comm_Bz, comm_prod):
        return False
    tr.append_point(b"comm_Az_claim", comm_Az)
    tr.append_point(b"comm_Bz_claim", comm_Bz)
    tr.append_point(b"comm_Cz_claim", comm_Cz)
    tr.append_point(b"comm_prod_Az_Bz_claims", comm_prod)
    taus = 1
    for a, b in zip(rx, tau):
        taus = taus * (a * b + (1 - a) * (1 - b)) % ELL
    exp1 = (taus * (decompress(comm_prod) - decompress(comm_Cz))).compress()
    if not equality_verify(pf["eq1"], gens.gens_1, tr, exp1, comm_post1):
        return False
    r_A = tr.challenge_scalar(b"challenege_Az")
    r_B = tr.challenge_scalar(b"challenege_Bz")
    r_C = tr.challenge_scalar(b"challenege_Cz")
    comm_claim2 = msm([r_A, r_B, r_C], [decompress(comm_Az), decompress(comm_Bz), decompress(comm_Cz)]).compress()
    res = zk_sumcheck_verify(pf["sc2"], comm_claim2, nry, 2, gens.gens_1, gens.gens_3, tr)
    if res is None:
        return False
    comm_post2, ry = res
    # PolyEvalProof::verify, dense_mlpoly.rs:364-386
    tr.append_protocol_name(b"polynomial evaluation proof")
    r = ry[1:]
    lnv, _ = factored_lens(len(r))
    Lv, Rv = eq_evals(r[:lnv]), eq_evals(r[lnv:])
    C_LZ = msm(Lv, [decompress(c) for c in pf["comm_vars"]]).compress()
    if not dotproductlog_verify(pf["dpl"], len(Rv), gens.gens_n, gens.gens_1, tr, Rv, C_LZ, pf["comm_vars_at_ry"]):
        return False
    entries = [(0, 1)] + [(i + 1, v) for i, v in enumerate(inputs)]
    poly_input_eval = sparse_poly_evaluate(log2(n), entries, ry[1:])
    comm_eval_Z = msm([(1 - ry[0]) % ELL, ry[0]],
                      [decompress(pf["comm_vars_at_ry"]), commit_scalar(poly_input_eval, 0, gens.gens_1)])
    exp2 = (((r_A * evA + r_B * evB + r_C * evC) % ELL) * comm_eval_Z).compress()
    if not equality_verify(pf["eq2"], gens.gens_1, tr, exp2, comm_post2):
        return False
    return rx == pf["rx"] and ry == pf["ry"]


# ----------------------------------------------------------------------------
# bincode 1.3 layout of NIZK
# ----------------------------------------------------------------------------
def _vec_pts(v):
    return struct.pack("<Q", len(v)) + b"".join(v)


def _vec_sc(v):
    return struct.pack("<Q", len(v)) + b"".join(fl_ser(x) for x in v)


def _ser_dpp(p):
    return p["delta"] + p["beta"] + _vec_sc(p["z"]) + fl_ser(p["z_delta"]) + fl_ser(p["z_beta"])


def _ser_sc(sc):
    return (_vec_pts(sc["comm_polys"]) + _vec_pts(sc["comm_evals"]) + struct.pack("<Q", len(sc["proofs"])) +
            b"".join(_ser_dpp(p) for p in sc["proofs"]))


def serialize_r1cs_proof(pf):
    out = _vec_pts(pf["comm_vars"])
    out += _ser_sc(pf["sc1"])
    out += b"".join(pf["claims_phase2"])
    out += pf["pok"]["alpha"] + fl_ser(pf["pok"]["z1"]) + fl_ser(pf["pok"]["z2"])
    out += pf["prod"]["alpha"] + pf["prod"]["beta"] + pf["prod"]["delta"] + b"".join(fl_ser(x) for x in pf["prod"]["z"])
    out += pf["eq1"]["alpha"] + fl_ser(pf["eq1"]["z"])
    out += _ser_sc(pf["sc2"])
    out += pf["comm_vars_at_ry"]
    d = pf["dpl"]
    out += _vec_pts(d["bullet"]["L_vec"]) + _vec_pts(d["bullet"]["R_vec"]) + d["delta"] + d["beta"] + fl_ser(d["z1"]) + fl_ser(d["z2"])
    out += pf["eq2"]["alpha"] + fl_ser(pf["eq2"]["z"])
    return out


def serialize_nizk(pf):
    return serialize_r1cs_proof(pf) + _vec_sc(pf["rx"]) + _vec_sc(pf["ry"])


def r1cs_proof_len(s_x, s_v):
    """Closed form of App. A.4."""
    s_y = s_v + 1
    lL = s_v // 2
    lR = s_v - lL
    dpp = lambda n: 32 + 32 + 8 + 32 * n + 32 + 32
    sc = lambda rounds, n: (8 + 32 * rounds) * 2 + 8 + rounds * dpp(n)
    return ((8 + 32 * (1 << lL)) + sc(s_x, 4) + 128 + (96 + 256) + 64 + sc(s_y, 3) + 32 +
            (2 * (8 + 32 * lR) + 128) + 64)
