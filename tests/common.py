"""Shared helpers for the parity tests: seeded scalars and the oracle <-> product glue."""
import hashlib
import random

ELL = 2**252 + 27742317777372353535851937790883648493
MONT_R = pow(2, 256, ELL)
TAPE_SEED = bytes(range(64))
LABEL = b"nizk_example"  # spartan-zkinterface/src/main.rs:78


def rand_scalars(n, seed):
    """n uniform scalars as python ints (deterministic)."""
    rng = random.Random(seed)
    return [rng.randrange(ELL) for _ in range(n)]


def canon(xs):
    return b"".join(int(x).to_bytes(32, "little") for x in xs)


def mont(xs):
    return b"".join((int(x) * MONT_R % ELL).to_bytes(32, "little") for x in xs)


def unmont(b):
    rinv = pow(MONT_R, ELL - 2, ELL)
    return [int.from_bytes(b[i:i + 32], "little") * rinv % ELL for i in range(0, len(b), 32)]


def sha(b):
    return hashlib.sha256(b).hexdigest()


def ragged_r1cs(nc, nv, ni):
    """A satisfiable R1CS with non-power-of-two sizes, several entries per row, duplicate entries and the
    constant column in use; returns (A, B, C, vars, inputs) with COO triples (row, col, 32 bytes)."""
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
        A += [(row, c, v.to_bytes(32, "little")) for c, v in ra]
        B += [(row, c, v.to_bytes(32, "little")) for c, v in rb]
        Cm.append((row, nv, (az * bz % ELL).to_bytes(32, "little")))
    return A, B, Cm, vars_, inputs


def otti_shaped_r1cs(nc, nv, ni, seed=7):
    """"Otti-shaped" satisfiable R1CS: what CirC emits for the LP/SDP/SGD checkers is dominated by
    bit decompositions of fixed-point / float values (compiler/src/IR/SMT/ToPf.hs:558-560,644-646), so ~80 % of the
    witness is boolean, the rest are 32-bit words; coefficients are +-1, powers of two and small integers; rows have
    1..64 entries (booleanity rows, 32-term decomposition rows, short product rows); num_cons != num_vars.
    Shape-substituted: the real Otti circuits cannot be generated without the CirC toolchain.
    Returns (constraints, vars, inputs) with constraints = [(a, b, c)], each a list of (z index, coefficient int),
    z = vars || 1 || inputs (the constant is index nv)."""
    rng = random.Random(seed)
    nb = max(1, (nv * 4) // 5)
    nw = nv - nb
    bits = [rng.randrange(2) for _ in range(nb)]
    words = [sum(bits[(j * 32 + k) % nb] << k for k in range(32)) for j in range(nw)]
    vars_ = bits + words
    inputs = [rng.randrange(1 << 32) for _ in range(ni)]
    z = vars_ + [1] + inputs
    ONE = nv

    def combo():
        n = 1
        while n < 64 and rng.random() < 0.6:
            n += 1
        return [(rng.randrange(len(z)), rng.choice([1, 1, ELL - 1, 2, 3, 1 << rng.randrange(1, 31)])) for _ in range(n)]

    cons = []
    for r in range(nc):
        kind = r % 3
        if kind == 0 or (kind == 1 and nw == 0):
            b = r % nb
            cons.append(([(b, 1)], [(b, 1), (ONE, ELL - 1)], []))  # b * (b - 1) = 0
        elif kind == 1:
            j = r % nw
            cons.append(([((j * 32 + k) % nb, 1 << k) for k in range(32)], [(ONE, 1)], [(nb + j, 1)]))
        else:
            a, b = combo(), combo()
            p = (sum(v * z[c] for c, v in a) % ELL) * (sum(v * z[c] for c, v in b) % ELL) % ELL
            if nw:
                t = nb + rng.randrange(nw)
                cons.append((a, b, [(t, 1), (ONE, (p - z[t]) % ELL)]))
            else:
                cons.append((a, b, [(ONE, p)]))
    return cons, vars_, inputs


def coo_of(cons):
    """constraints -> three COO lists (row, col, 32 bytes) in the order R1cs::instance pushes them."""
    M = ([], [], [])
    for r, con in enumerate(cons):
        for m in range(3):
            M[m].extend((r, c, int(v).to_bytes(32, "little")) for c, v in con[m])
    return M
