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

import os
import struct
import sys

# MS-NRBF PrimitiveTypeEnumeration -> struct format (fixed-width ones).
# Deepest record nesting accepted. BinaryFormatter writes object members as references
# and the objects themselves later at top level, so real documents stay shallow (the
# three surveyed files reach depth 2).
MAX_DEPTH = 64

_PRIM_FMT = {
    1: "<?", 2: "<B", 6: "<d", 7: "<h", 8: "<i", 9: "<q", 10: "<b",
    11: "<f", 12: "<q", 13: "<Q", 14: "<H", 15: "<I", 16: "<Q",
}


class Ref:
    """Unresolved pointer to another object of the same stream."""

    __slots__ = ("id",)

    def __init__(self, id_):
        self.id = id_


class NullRun:
    """``n`` consecutive null array elements (record types 13 and 14).

    A type of its own, so the array reader can tell a run of nulls from an inline
    nested array whose first element happens to be null. ``n`` may be 0.
    """

    __slots__ = ("n",)

    def __init__(self, n):
        self.n = n


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
        self.depth = 0        # records currently open (object members, array elements)

    # -- low level ---------------------------------------------------------
    def _take(self, n):
        """The next ``n`` bytes. Every read goes through here, so a short file always
        fails the same way: ``ValueError`` naming the byte where data ran out."""
        if n < 0 or self.p + n > len(self.d):
            raise ValueError(f"truncated NRBF stream at byte {len(self.d)} "
                             f"(needed {n} bytes from byte {self.p})")
        b = self.d[self.p:self.p + n]
        self.p += n
        return b

    def _u8(self):
        return self._take(1)[0]

    def _i32(self):
        return struct.unpack("<i", self._take(4))[0]

    def _array_len(self, at):
        """A declared array length. An element costs at least one byte except inside a
        null run, so a length beyond the bytes left is treated as corruption. This is a
        limit of this reader, not of the format: it keeps a 31-byte file from asking for
        gigabytes."""
        n = self._count()
        left = len(self.d) - self.p
        if n > left:
            raise ValueError(f"array at byte {at} declares {n} elements with only "
                             f"{left} bytes left")
        return n

    def _count(self):
        """A non-negative Int32 length field."""
        at, n = self.p, self._i32()
        if n < 0:
            raise ValueError(f"negative length {n} at byte {at}")
        return n

    def _text(self, n):
        at = self.p
        try:
            return self._take(n).decode("utf-8")
        except UnicodeDecodeError as e:
            raise ValueError(f"bad UTF-8 in string at byte {at}: {e.reason}") from None

    def _lps(self):
        n = shift = 0
        while True:
            b = self._u8()
            n |= (b & 0x7F) << shift
            shift += 7
            if not b & 0x80:
                break
            if shift > 28:
                raise ValueError(f"string length prefix too long at byte {self.p}")
        return self._text(n)

    def _prim(self, t):
        if t in _PRIM_FMT:
            fmt = _PRIM_FMT[t]
            return struct.unpack(fmt, self._take(struct.calcsize(fmt)))[0]
        if t in (5, 18):          # Decimal / String: length-prefixed text
            return self._lps()
        if t == 3:                # Char: one UTF-8 code point
            b0 = self._take(1)[0]
            self.p -= 1
            return self._text(1 if b0 < 0x80 else 2 if b0 < 0xE0 else 3 if b0 < 0xF0 else 4)
        raise ValueError(f"primitive type {t} at byte {self.p}")

    # -- class metadata ----------------------------------------------------
    def _class_info(self):
        oid = self._i32()
        name = self._lps()
        names = [self._lps() for _ in range(self._count())]
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
        nulls = 0                 # members still covered by an earlier null run
        for name, (kind, extra) in zip(names, types):
            if nulls:
                if kind == 0:
                    raise ValueError(f"null run covers primitive member {name!r} "
                                     f"of the object at byte {obj.offset}")
                nulls -= 1
                obj.fields[name] = None
                continue
            v = self._prim(extra) if kind == 0 else self.record()
            if v is StopIteration:
                raise ValueError(f"MessageEnd as member {name!r} of the object at byte "
                                 f"{obj.offset}")
            if isinstance(v, NullRun):
                if v.n < 1:
                    raise ValueError(f"empty null run as member {name!r} "
                                     f"of the object at byte {obj.offset}")
                nulls, v = v.n - 1, None
            obj.fields[name] = v
        if nulls:
            raise ValueError(f"null run overflows the members of the object at byte "
                             f"{obj.offset} by {nulls}")

    # -- records -----------------------------------------------------------
    def record(self):
        """One record, with whatever it nests. Nesting deeper than ``MAX_DEPTH`` is
        refused, so a crafted file ends in ``ValueError`` and never ``RecursionError``."""
        if self.depth >= MAX_DEPTH:
            raise ValueError(f"records nested deeper than {MAX_DEPTH} at byte {self.p}")
        self.depth += 1
        try:
            return self._record()
        finally:
            self.depth -= 1

    def _record(self):
        at = self.p
        rt = self._u8()
        while rt == 12:           # BinaryLibrary: precedes a real record; may repeat
            lid = self._i32()
            self.libraries[lid] = self._lps()
            at = self.p
            rt = self._u8()
        if rt == 0:               # SerializedStreamHeader
            self.root_id = self._i32()
            self._i32(); self._i32(); self._i32()
            return None
        if rt == 1:               # ClassWithId
            oid, mid = self._i32(), self._i32()
            if mid not in self.classes:
                raise ValueError(f"ClassWithId at byte {at} names unknown class id {mid}")
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
        if rt == 13:              # ObjectNullMultiple256
            return NullRun(self._u8())
        if rt == 14:              # ObjectNullMultiple
            return NullRun(self._count())
        if rt == 15:              # ArraySinglePrimitive
            oid, n, t = self._i32(), self._array_len(at), self._u8()
            arr = [self._prim(t) for _ in range(n)]
            self.objects[oid] = arr
            return arr
        if rt in (16, 17):        # ArraySingleObject / ArraySingleString
            oid, n = self._i32(), self._array_len(at)
            arr = []
            while len(arr) < n:
                v = self.record()
                if isinstance(v, NullRun):
                    if v.n > n - len(arr):
                        raise ValueError(f"null run of {v.n} overflows the {n}-element "
                                         f"array at byte {at}")
                    arr.extend([None] * v.n)
                elif v is StopIteration:
                    raise ValueError(f"MessageEnd inside the array that starts at byte {at}")
                else:
                    arr.append(v)
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


_CHILDREN = {
    "IF_Control": (("yes", "_left_Child"), ("no", "_right_Child")),
    "Loop": (("before test", "_before_Child"), ("after test", "_after_Child")),
}


def walk(st, node, depth=0, out=None, seen=None):
    """List the chart from ``node`` on, as (depth, offset, class, description) rows.

    Label rows ("yes" / "no" / "before test" / "after test") carry offset ``None``.
    A node reached twice means a successor or child cycle (or two parents sharing a
    node); neither occurs in a well-formed chart, so it raises instead of looping.
    """
    out = [] if out is None else out
    seen = set() if seen is None else seen
    node = st.deref(node)
    while isinstance(node, Obj):
        if node.offset in seen:
            raise ValueError(f"chart revisits the node at byte {node.offset}: "
                             "successor/child cycle or shared node")
        seen.add(node.offset)
        kind = _short(node.cls)
        out.append((depth, node.offset, kind, _describe(st, node)))
        f = node.fields
        for label, key in _CHILDREN.get(kind, ()):
            child = st.deref(f.get(key))
            if isinstance(child, Obj):
                out.append((depth + 1, None, label, ""))
                walk(st, child, depth + 2, out, seen)
        node = st.deref(f.get("_Successor"))
    return out


def _root(st):
    return st.deref(Ref(st.root_id)) if st.root_id is not None else None


def declared_chart_count(streams):
    """The tab count a ``.rap`` document declares, or ``None`` if it declares none.

    SURVEY.md section 1: stream 0 is the format version (boxed Int32), stream 1 a
    Boolean, stream 2 the number of chart tabs (boxed Int32), all before the first chart.
    """
    if len(streams) < 3:
        return None
    roots = [_root(st) for _, _, st in streams[:3]]
    if any(isinstance(r, Obj) and r.cls == "raptor.Oval" for r in roots):
        return None
    r = roots[2]
    if isinstance(r, Obj) and r.cls == "System.Int32":
        return r.fields.get("m_value")
    return None


def charts(data: bytes):
    """Every chart in the document: ``[(tab name, rows)]``, in file order.

    A chart is a stream whose root object is a ``raptor.Oval`` (its Start node). The tab
    name is the nearest earlier stream whose root is a string. Raises if the document
    declares a different number of tabs than were found, so a multi-tab document can
    never be listed as if its first tab were the whole program.
    """
    streams = read_streams(data)
    found, name = [], None
    for _, _, st in streams:
        root = _root(st)
        if isinstance(root, str):
            name = root
        elif isinstance(root, Obj) and root.cls == "raptor.Oval":
            found.append((name, walk(st, root)))
            name = None
    if not found:
        raise ValueError("no raptor.Oval-rooted stream found: not a RAPTOR chart document")
    declared = declared_chart_count(streams)
    if declared is not None and declared != len(found):
        raise ValueError(f"document declares {declared} chart tab(s) but {len(found)} "
                         "raptor.Oval-rooted stream(s) were found: refusing to list a partial program")
    return found


def chart_rows(data: bytes):
    """Rows of the first chart (``main``), after the tab-count check in ``charts``."""
    return charts(data)[0][1]


def dump(path: str) -> str:
    with open(path, "rb") as fh:
        data = fh.read()
    streams = read_streams(data)
    all_charts = charts(data)
    declared = declared_chart_count(streams)
    lines = [f"{os.path.basename(path)}  ({len(data)} bytes, {len(streams)} streams)",
             f"  charts: {len(all_charts)} "
             f"({'declared ' + str(declared) if declared is not None else 'tab count not declared'}; "
             f"showing {', '.join(repr(n) for n, _ in all_charts)})"]
    for name, rows in all_charts:
        if len(all_charts) > 1:
            lines.append(f"  -- chart {name!r}")
        for depth, off, kind, desc in rows:
            where = f"@{off:<6}" if off is not None else " " * 7
            lines.append(f"  {where} {'  ' * depth}{kind}{': ' + desc if desc else ''}")
    lines.append(f"  nodes: {sum(1 for _, rows in all_charts for r in rows if r[1] is not None)}")
    return "\n".join(lines)


if __name__ == "__main__":
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    try:
        print("\n\n".join(dump(p) for p in sys.argv[1:]))
    except ValueError as e:
        sys.exit(f"rap_dump: {e}")
