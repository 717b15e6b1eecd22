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
