"""Minimal zkinterface flatbuffer WRITER for tests (schema: spartan-zkinterface/schema/zkinterface.fbs).
Forward layout: a table is written before the objects it points to, so every uoffset is positive;
each table is preceded by its own vtable.  Only what the prover's ingestion reads is emitted."""
import struct

MSG_HEADER, MSG_CS, MSG_WITNESS = 1, 2, 3


class _B:
    def __init__(self):
        self.b = bytearray()

    def align(self, n):
        while len(self.b) % n:
            self.b.append(0)

    def table(self, fields):
        """fields: list of None | ("u8", v) | ("u64", v) | ("ref", writer) in field-id order.
        Returns the table position; children are written right after it."""
        # layout of the inline part: soffset(4) then each present field, naturally aligned
        off, offs = 4, []
        for f in fields:
            if f is None:
                offs.append(0)
                continue
            size = {"u8": 1, "u64": 8, "ref": 4}[f[0]]
            off = (off + size - 1) // size * size
            offs.append(off)
            off += size
        tsize = off
        vt = struct.pack("<HH", 4 + 2 * len(fields), tsize) + b"".join(struct.pack("<H", o) for o in offs)
        # place the vtable so that the table that follows is 8-aligned
        self.align(8)
        pad = (-(len(self.b) + len(vt))) % 8
        self.b += b"\0" * pad
        vpos = len(self.b)
        self.b += vt
        tpos = len(self.b)
        self.b += b"\0" * tsize
        struct.pack_into("<i", self.b, tpos, tpos - vpos)
        refs = []
        for f, o in zip(fields, offs):
            if f is None:
                continue
            if f[0] == "u8":
                self.b[tpos + o] = f[1]
            elif f[0] == "u64":
                struct.pack_into("<Q", self.b, tpos + o, f[1])
            else:
                refs.append((tpos + o, f[1]))
        for at, writer in refs:
            child = writer(self)
            struct.pack_into("<I", self.b, at, child - at)
        return tpos

    def vector(self, elem_size, payload, count):
        self.align(4)
        if elem_size == 8:
            while (len(self.b) + 4) % 8:
                self.b.append(0)
        pos = len(self.b)
        self.b += struct.pack("<I", count) + payload
        return pos

    def offset_vector(self, writers):
        self.align(4)
        pos = len(self.b)
        self.b += struct.pack("<I", len(writers)) + b"\0" * (4 * len(writers))
        for i, w in enumerate(writers):
            at = pos + 4 + 4 * i
            child = w(self)
            struct.pack_into("<I", self.b, at, child - at)
        return pos


def _variables(ids, values, width):
    """Variables{variable_ids, values}: values are `width`-byte little-endian each."""
    payload = b"".join(int(v).to_bytes(width, "little") for v in values)
    return lambda b: b.table([("ref", lambda b2: b2.vector(8, struct.pack("<%dQ" % len(ids), *ids), len(ids))),
                              ("ref", lambda b2: b2.vector(1, payload, len(payload)))])


def _root(msg_type, msg_writer):
    b = _B()
    b.b += b"\0\0\0\0zkif"
    root = b.table([("u8", msg_type), ("ref", msg_writer)])
    struct.pack_into("<I", b.b, 0, root)
    return bytes(b.b)


def header(input_ids, input_values, field_max, width=32, free_variable_id=0):
    fm = int(field_max).to_bytes(32, "little")
    return _root(MSG_HEADER, lambda b: b.table([("ref", _variables(input_ids, input_values, width)), ("u64", free_variable_id),
                                               ("ref", lambda b2: b2.vector(1, fm, 32))]))


def witness(ids, values, width=32):
    return _root(MSG_WITNESS, lambda b: b.table([("ref", _variables(ids, values, width))]))


def constraint_system(constraints, width=32):
    """constraints: list of (a, b, c), each a list of (variable id, coefficient)."""
    def one(con):
        return lambda b: b.table([("ref", _variables([i for i, _ in lc], [v for _, v in lc], width)) for lc in con])
    return _root(MSG_CS, lambda b: b.table([("ref", lambda b2: b2.offset_vector([one(c) for c in constraints]))]))
