"""ORACLE (test infrastructure only): pure-Python restatement of rust-circ's Spartan adapter,
rust-circ/src/target/r1cs/spartan.rs:31-186 (`r1cs_to_spartan`, `int_to_scalar`, `lc_to_v`).

Parity unpinned against a real cargo build (no Rust toolchain here): the reference holds no test or
fixture for this file; the restatement follows the source statement by statement.  Where the
reference iterates a HashMap the order of the Python dict is used (the product does the same)."""
ELL = 2**252 + 27742317777372353535851937790883648493


def int_to_scalar(i):
    """spartan.rs:141-160: Horner over the 64-bit limbs of rug's magnitude (`as_ref()` carries no
    sign), most significant first, in the scalar field."""
    i = abs(int(i))
    limbs = []
    while i:
        limbs.append(i & (2**64 - 1))
        i >>= 64
    m = (2**63 % ELL) * 2 % ELL
    acc = 0
    for digit in reversed(limbs):
        acc = (acc * m + digit) % ELL
    return acc


def _b(x):
    return int(x).to_bytes(32, "little")


class Panic(Exception):
    """the reference assert!s, unwrap()s a None or panic!s here"""


def r1cs_to_spartan(idxs_signals, values, constraints, next_idx, public_idxs=()):
    """-> dict(num_cons, num_vars, num_inputs, A, B, C (COO triples (row, col, 32 bytes)), vars, inputs).
    constraints: [(lc_a, lc_b, lc_c)], lc = (constant, {id: coeff}) with dict order = iteration order."""
    wit, inp, trans, itrans = [], [], {}, {}
    if len(public_idxs) != 0:
        raise Panic("assert!(r1cs.public_idxs.len() == 0)")  # :42
    if values is None:
        raise Panic("Tried to run Spartan without inputs/witness")  # :77
    for k, v in values.items():  # :47-75
        name = idxs_signals[k]
        scalar = int_to_scalar(v)
        if "main_f0_lex0_w" in name:
            trans[k] = len(wit)
            wit.append(_b(scalar))
        elif "main_f0_lex0_" in name:
            itrans[k] = len(inp)
            inp.append(_b(scalar))
        else:
            trans[k] = len(wit)
            wit.append(_b(scalar))
    if len(wit) + len(inp) != next_idx:
        raise Panic("assert_eq!(wit.len() + inp.len(), r1cs.next_idx)")  # :80
    num_vars = const_id = len(wit)
    num_inputs = len(inp)
    for cid, sid in itrans.items():  # :96-99
        trans[cid] = sid + const_id + 1

    def lc_to_v(lc):  # :164-186
        constant, monomials = lc
        v = []
        for k, m in monomials.items():
            if k >= const_id:
                raise Panic("Error: variable number off")
            if k not in trans:
                raise Panic("trans.get(k).unwrap()")
            v.append((trans[k], _b(int_to_scalar(m))))
        if constant != 0:
            v.append((const_id, _b(int_to_scalar(constant))))
        return v

    A, B, C = [], [], []
    i = 0
    for lc_a, lc_b, lc_c in constraints:  # :110-130
        a, b, c = lc_to_v(lc_a), lc_to_v(lc_b), lc_to_v(lc_c)
        A += [(i, sid, val) for sid, val in a]
        B += [(i, sid, val) for sid, val in b]
        C += [(i, sid, val) for sid, val in c]
        i += 1
    return {"num_cons": i, "num_vars": num_vars, "num_inputs": num_inputs, "A": A, "B": B, "C": C,
            "vars": b"".join(wit), "inputs": b"".join(inp)}
