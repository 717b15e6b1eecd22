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
