#!/usr/bin/env python3
"""Print a RAPTOR ``.rap`` flowchart document as plain text.

Purpose: make the finding in SURVEY.md section 0 / section 2 ("the reference is three
scalar teaching flowcharts and nothing else") reproducible with one command,
without Windows or the RAPTOR application:

    python tools/rap_dump.py "/root/reference/RAPTOR/BMI.rap"

This is host-side VERIFICATION TOOLING. It is not a hot path, not an oracle
and not an interpreter: it never evaluates an expression, it only lists the
nodes a document holds, in control-flow order, with the byte offset of each
node's serialization record (the ``file@offset`` citations SURVEY.md uses).

Format notes: a ``.rap`` file is a run of back-to-back .NET
BinaryFormatter (MS-NRBF) streams. Stream 5 carries the chart object graph:
``raptor.Oval`` / ``Parallelogram`` / ``Rectangle`` / ``IF_Control`` / ``Loop``
objects linked by ``_Successor``, ``_left_Child`` / ``_right_Child`` and
``_before_Child`` / ``_after_Child``.
"""
from __future__ import annotations

import struct
import sys

# MS-NRBF PrimitiveTypeEnumeration -> struct format (fixed-width ones).
_PRIM_FMT = {
    1: "<?", 2: "<B", 6: "<d", 7: "<h", 8: "<i", 9: "<q", 10: "<b",
    11: "<f", 12: "<q", 13: "<Q", 14: "<H", 15: "<I", 16: "<Q",
}


class Ref:
    """Unresolved pointer to another object of the same stream."""

    __slots__ = ("id",)

    def __init__(self, id_):
        self.id = id_


class Obj:
    """One deserialized class instance."""

    __slots__ = ("cls", "offset", "fields")

    def __init__(self, cls, offset):
        self.cls = cls
        self.offset = offset
        self.fields = {}

    def __repr__(self):
        return f"<{self.cls}@{self.offset}>"


class _Stream:
    """Parser for ONE BinaryFormatter stream starting at ``pos``."""

    def __init__(self, data: bytes, pos: int):
        self.d = data
        self.p = pos
        self.objects = {}     # objectId -> Obj | str | list
        self.classes = {}     # objectId -> (name, member names, member types)
        self.libraries = {}   # libraryId -> assembly name
        self.root_id = None
        self.top = []         # top-level records, in order

    # -- low level ---------------------------------------------------------
    def _u8(self):
        v = self.d[self.p]
        self.p += 1
        return v

    def _i32(self):
        v = struct.unpack_from("<i", self.d, self.p)[0]
        self.p += 4
        return v

    def _lps(self):
        n = shift = 0
        while True:
            b = self._u8()
            n |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
        s = self.d[self.p:self.p + n].decode("utf-8")
        self.p += n
        return s

    def _prim(self, t):
        if t in _PRIM_FMT:
            fmt = _PRIM_FMT[t]
            v = struct.unpack_from(fmt, self.d, self.p)[0]
            self.p += struct.calcsize(fmt)
            return v
        if t in (5, 18):          # Decimal / String: length-prefixed text
            return self._lps()
        if t == 3:                # Char: one UTF-8 code point
            b0 = self.d[self.p]
            n = 1 if b0 < 0x80 else 2 if b0 < 0xE0 else 3 if b0 < 0xF0 else 4
            s = self.d[self.p:self.p + n].decode("utf-8")
            self.p += n
            return s
        raise ValueError(f"primitive type {t} at byte {self.p}")

    # -- class metadata ----------------------------------------------------
    def _class_info(self):
        oid = self._i32()
        name = self._lps()
        names = [self._lps() for _ in range(self._i32())]
        return oid, name, names

    def _member_types(self, n):
        kinds = [self._u8() for _ in range(n)]
        out = []
        for k in kinds:
            if k in (0, 7):       # Primitive / PrimitiveArray
                out.append((k, self._u8()))
            elif k == 3:          # SystemClass
                out.append((k, self._lps()))
            elif k == 4:          # Class (+ library id)
                out.append((k, (self._lps(), self._i32())))
            else:                 # String, Object, ObjectArray, StringArray
                out.append((k, None))
        return out

    def _members(self, obj, names, types):
        for name, (kind, extra) in zip(names, types):
            obj.fields[name] = self._prim(extra) if kind == 0 else self.record()

    # -- records -----------------------------------------------------------
    def record(self):
        at = self.p
        rt = self._u8()
        if rt == 0:               # SerializedStreamHeader
            self.root_id = self._i32()
            self._i32(); self._i32(); self._i32()
            return None
        if rt == 1:               # ClassWithId
            oid, mid = self._i32(), self._i32()
            name, names, types = self.classes[mid]
            obj = Obj(name, at)
            self.objects[oid] = obj
            self._members(obj, names, types)
            return obj
        if rt in (4, 5):          # (System)ClassWithMembersAndTypes
            oid, name, names = self._class_info()
            types = self._member_types(len(names))
            if rt == 5:
                self._i32()       # library id
            self.classes[oid] = (name, names, types)
            obj = Obj(name, at)
            self.objects[oid] = obj
            self._members(obj, names, types)
            return obj
        if rt == 6:               # BinaryObjectString
            oid = self._i32()
            s = self._lps()
            self.objects[oid] = s
            return s
        if rt == 8:               # MemberPrimitiveTyped
            return self._prim(self._u8())
        if rt == 9:               # MemberReference
            return Ref(self._i32())
        if rt == 10:              # ObjectNull
            return None
        if rt == 11:              # MessageEnd
            return StopIteration
        if rt == 12:              # BinaryLibrary
            lid = self._i32()
            self.libraries[lid] = self._lps()
            return self.record()  # a library record always precedes a real one
        if rt == 13:              # ObjectNullMultiple256
            return [None] * self._u8()
        if rt == 14:              # ObjectNullMultiple
            return [None] * self._i32()
        if rt == 15:              # ArraySinglePrimitive
            oid, n, t = self._i32(), self._i32(), self._u8()
            arr = [self._prim(t) for _ in range(n)]
            self.objects[oid] = arr
            return arr
        if rt in (16, 17):        # ArraySingleObject / ArraySingleString
            oid, n = self._i32(), self._i32()
            arr = []
            while len(arr) < n:
                v = self.record()
                arr.extend(v) if isinstance(v, list) and v and v[0] is None else arr.append(v)
            self.objects[oid] = arr
            return arr
        raise ValueError(f"NRBF record type {rt} at byte {at} not handled")

    def parse(self):
        while True:
            r = self.record()
            if r is StopIteration:
                return self
            if r is not None:
                self.top.append(r)

    def deref(self, v):
        while isinstance(v, Ref):
            v = self.objects.get(v.id)
        return v


def read_streams(data: bytes):
    """Split ``data`` into its BinaryFormatter streams; returns [(start, end, _Stream)]."""
    out, pos = [], 0
    while pos < len(data):
        s = _Stream(data, pos).parse()
        out.append((pos, s.p, s))
        pos = s.p
    return out


def _short(cls):
    return cls.rsplit(".", 1)[-1]


def _describe(st, node):
    f = node.fields
    kind = _short(node.cls)
    text = st.deref(f.get("_text_str"))
    if kind == "Parallelogram":
        if f.get("_is_input"):
            return f'input  prompt={st.deref(f.get("_prompt"))!r} -> {text}'
        return f"output {text}"
    if kind == "Rectangle":
        k = st.deref(f.get("_kind"))
        k = k.fields.get("value__") if isinstance(k, Obj) else k
        return f"{'assign' if not k else f'call(kind={k})'} {text}"
    if kind == "IF_Control":
        return f"if {text}"
    if kind == "Loop":
        return f"loop, exit when {text}"
    return text or ""


def walk(st, node, depth=0, out=None):
    """List the chart from ``node`` on, as (depth, offset, class, description) rows."""
    out = [] if out is None else out
    node = st.deref(node)
    while isinstance(node, Obj):
        kind = _short(node.cls)
        out.append((depth, node.offset, kind, _describe(st, node)))
        f = node.fields
        if kind == "IF_Control":
            for label, key in (("yes", "_left_Child"), ("no", "_right_Child")):
                child = st.deref(f.get(key))
                if isinstance(child, Obj):
                    out.append((depth + 1, None, label, ""))
                    walk(st, child, depth + 2, out)
        elif kind == "Loop":
            for label, key in (("before test", "_before_Child"), ("after test", "_after_Child")):
                child = st.deref(f.get(key))
                if isinstance(child, Obj):
                    out.append((depth + 1, None, label, ""))
                    walk(st, child, depth + 2, out)
        node = st.deref(f.get("_Successor"))
    return out


def chart_rows(data: bytes):
    """Rows of the first chart (the stream whose root object is a ``raptor.Oval``)."""
    for _, _, st in read_streams(data):
        root = st.deref(Ref(st.root_id)) if st.root_id is not None else None
        if isinstance(root, Obj) and root.cls == "raptor.Oval":
            return walk(st, root)
    raise ValueError("no raptor.Oval-rooted stream found: not a RAPTOR chart document")


def dump(path: str) -> str:
    with open(path, "rb") as fh:
        data = fh.read()
    lines = [f"{path.rsplit('/', 1)[-1]}  ({len(data)} bytes, {len(read_streams(data))} streams)"]
    rows = chart_rows(data)
    for depth, off, kind, desc in rows:
        where = f"@{off:<6}" if off is not None else " " * 7
        lines.append(f"  {where} {'  ' * depth}{kind}{': ' + desc if desc else ''}")
    lines.append(f"  nodes: {sum(1 for r in rows if r[1] is not None)}")
    return "\n".join(lines)


if __name__ == "__main__":
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    print("\n\n".join(dump(p) for p in sys.argv[1:]))
