"""ORACLE (test infrastructure only): pure-Python restatement of spartan-zkinterface's ingestion,
spartan-zkinterface/src/lib.rs:75-260 -- flatbuffer walk, get_variables (lib.rs:194-219), translate with the
reference's first-match `position` semantics (lib.rs:96-107), matrix order of R1cs::instance (lib.rs:109-135).
Pinned by the reference's own fixtures test/{add,foo,inv}*.zkif, which its tests require to parse and be
satisfiable (lib.rs:318-332)."""
import struct


def _u32(b, o):
    return struct.unpack_from("<I", b, o)[0]


def _field(b, t, fid):
    vt = t - struct.unpack_from("<i", b, t)[0]
    vsize = struct.unpack_from("<H", b, vt)[0]
    slot = 4 + 2 * fid
    if slot + 2 > vsize:
        return 0
    off = struct.unpack_from("<H", b, vt + slot)[0]
    return t + off if off else 0


def _ind(b, at):
    return at + _u32(b, at)


def _message(b, want):
    root = _ind(b, 0)
    ftype, fmsg = _field(b, root, 0), _field(b, root, 1)
    if not ftype or not fmsg or b[ftype] != want:
        raise ValueError("wrong message type")
    return _ind(b, fmsg)


def _variables(b, fpos):
    if not fpos:
        raise ValueError("missing Variables")  # .unwrap()
    t = _ind(b, fpos)
    fid, fval = _field(b, t, 0), _field(b, t, 1)
    if not fid or not fval:
        raise ValueError("missing ids/values")
    v = _ind(b, fid)
    n = _u32(b, v)
    ids = list(struct.unpack_from("<%dQ" % n, b, v + 4))
    w = _ind(b, fval)
    nv = _u32(b, w)
    raw = bytes(b[w + 4:w + 4 + nv])
    if n == 0:
        return []
    ba = nv // n
    if ba > 32:
        raise ValueError("value too wide")
    return [(ids[i], raw[i * ba:(i + 1) * ba] + bytes(32 - ba)) for i in range(n)]


def parse(circuit, header, witness):
    """-> dict(num_cons, num_vars, num_inputs, A, B, C, vars, inputs, non_zero_entries)"""
    th, tc, tw = _message(header, 1), _message(circuit, 2), _message(witness, 3)
    inputs = _variables(header, _field(header, th, 0))
    wit = _variables(witness, _field(witness, tw, 0))
    num_vars = len(wit)
    wit_ids, in_ids = [i for i, _ in wit], [i for i, _ in inputs]

    def translate(i):  # lib.rs:96-107
        if i in wit_ids:
            return wit_ids.index(i)
        if i in in_ids:
            return in_ids.index(i) + num_vars + 1
        return num_vars

    fc = _field(circuit, tc, 0)
    v = _ind(circuit, fc)
    ncons = _u32(circuit, v)
    if ncons == 0:
        raise ValueError("No constraints given!")
    M, nz = ([], [], []), [0, 0, 0]
    for r in range(ncons):
        t = _ind(circuit, v + 4 + 4 * r)
        for m in range(3):
            for i, val in _variables(circuit, _field(circuit, t, m)):
                M[m].append((r, translate(i), val))
                nz[m] += any(val)
    return dict(num_cons=ncons, num_vars=num_vars, num_inputs=len(inputs), A=M[0], B=M[1], C=M[2],
                vars=b"".join(x for _, x in wit), inputs=b"".join(x for _, x in inputs), non_zero_entries=max(nz))
