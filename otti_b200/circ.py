"""rust-circ adapter, mirroring `r1cs_to_spartan` / `int_to_scalar`
(rust-circ/src/target/r1cs/spartan.rs:31-160): a CirC `R1cs` -> Instance + assignments, in memory.

    inst, assn_witness, assn_inputs, num_cons, num_vars, num_inputs = r1cs_to_spartan(r1cs)   # spartan.rs:31
    gens = NIZKGens.new(num_cons, num_vars, num_inputs)                                        # circ.rs:257
    proof = NIZK.prove(inst, assn_witness, assn_inputs, gens, Transcript.new(b"nizk_example")) # circ.rs:262-263

`r1cs` is a `CircR1cs` (the fields of rust-circ's `R1cs<S>` this path reads, src/target/r1cs/mod.rs).  The
reference iterates `values` and `lc.monomials` as Rust HashMaps, i.e. in no defined order; here the iteration
order of the Python dicts is the order.  The conversion itself runs in libspartan_b200 (host code)."""
import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

from . import NIZKGens, lib, _check
from .zkif import R1cs


@dataclass
class Lc:
    """`Lc { constant: Integer, monomials: HashMap<usize, Integer> }`"""
    constant: int = 0
    monomials: Dict[int, int] = field(default_factory=dict)


@dataclass
class CircR1cs:
    """`R1cs<S>`: idxs_signals (id -> name), values (id -> Integer), public_idxs, next_idx, constraints."""
    idxs_signals: Dict[int, str]
    values: Dict[int, int]
    constraints: List[Tuple[Lc, Lc, Lc]]
    next_idx: int
    public_idxs: frozenset = frozenset()


def _mag(i):
    i = abs(int(i))
    return i.to_bytes((i.bit_length() + 7) // 8, "little")


class _BigInts:
    def __init__(self, ints):
        parts = [_mag(i) for i in ints]
        off = [0]
        for p in parts:
            off.append(off[-1] + len(p))
        blob = b"".join(parts)
        self.bytes = (C.c_uint8 * max(len(blob), 1)).from_buffer_copy(blob or b"\0")
        self.off = (C.c_uint64 * len(off))(*off)


def int_to_scalar(i):
    """`int_to_scalar` (spartan.rs:141-160): |i| mod l, as 32 canonical little-endian bytes."""
    m = _mag(i)
    out = (C.c_uint8 * 32)()
    _check(lib().sb200_circ_int_to_scalar(m, C.c_size_t(len(m)), out))
    return bytes(out)


def circ_to_r1cs(r1cs):
    """The conversion alone: an `otti_b200.zkif.R1cs` (COO triples + assignments) from a `CircR1cs`."""
    if r1cs.values is None:
        raise AssertionError("Tried to run Spartan without inputs/witness")  # spartan.rs:77
    ids = list(r1cs.values.keys())
    names = [r1cs.idxs_signals[k].encode() for k in ids]  # idxs_signals.get(k).unwrap()
    noff = [0]
    for nm in names:
        noff.append(noff[-1] + len(nm))
    nblob = b"".join(names)
    vals = _BigInts([r1cs.values[k] for k in ids])
    u64 = C.c_uint64
    sides = []
    for m in range(3):
        lc_off, mono, coeff, const = [0], [], [], []
        for cons in r1cs.constraints:
            lc = cons[m]
            for k, v in lc.monomials.items():
                mono.append(k)
                coeff.append(v)
            lc_off.append(len(mono))
            const.append(lc.constant)
        sides.append(((u64 * len(lc_off))(*lc_off), (u64 * max(len(mono), 1))(*mono), _BigInts(coeff), _BigInts(const)))
    P64, P8 = C.POINTER(u64), C.POINTER(C.c_uint8)
    arr64 = lambda xs: (P64 * 3)(*[C.cast(x, P64) for x in xs])
    arr8 = lambda xs: (P8 * 3)(*[C.cast(x, P8) for x in xs])
    h = C.c_void_p()
    rc = lib().sb200_circ_r1cs(
        u64(len(ids)), (u64 * max(len(ids), 1))(*ids), nblob or b"\0", (u64 * len(noff))(*noff), vals.bytes, vals.off,
        u64(len(r1cs.public_idxs)), u64(r1cs.next_idx), u64(len(r1cs.constraints)),
        arr64([s[0] for s in sides]), arr64([s[1] for s in sides]), arr8([s[2].bytes for s in sides]),
        arr64([s[2].off for s in sides]), arr8([s[3].bytes for s in sides]), arr64([s[3].off for s in sides]), C.byref(h))
    _check(rc)
    return R1cs(h)


def r1cs_to_spartan(r1cs):
    """`r1cs_to_spartan` (spartan.rs:31-139): (inst, assn_witness, assn_inputs, num_cons, num_vars, num_inputs).
    Like the reference it asserts that the instance it built is satisfied by the assignment (needs the GPU)."""
    r = circ_to_r1cs(r1cs)
    inst = r.instance()
    assn_witness, assn_inputs = r.vars_assignment(), r.inputs_assignment()
    if not inst.is_sat(assn_witness, assn_inputs):  # spartan.rs:135-136
        raise AssertionError("assert_eq!(inst.is_sat(..).unwrap(), true)")
    return inst, assn_witness, assn_inputs, r.num_cons, r.num_vars, r.num_inputs


def nizk_public_params(num_cons, num_vars, num_inputs):
    return NIZKGens.new(num_cons, num_vars, num_inputs)
