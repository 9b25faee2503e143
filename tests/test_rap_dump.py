"""tools/rap_dump.py against a small MS-NRBF document assembled in this file.

The document is synthetic (a 4-node chart: Start, input, assign, End) so that the test
needs no reference data and runs on any machine.
"""
import struct

import pytest

from tools import rap_dump


def lps(s: str) -> bytes:
    b = s.encode("utf-8")
    assert len(b) < 128
    return bytes([len(b)]) + b


def i32(v: int) -> bytes:
    return struct.pack("<i", v)


HEADER = lambda root: b"\x00" + i32(root) + i32(-1) + i32(1) + i32(0)  # noqa: E731
END = b"\x0b"


def obj_string(oid, s):
    return b"\x06" + i32(oid) + lps(s)


def ref(oid):
    return b"\x09" + i32(oid)


def class_with_types(oid, name, members, lib):
    """members: [(name, binary type kind, primitive type or None)]"""
    out = b"\x05" + i32(oid) + lps(name) + i32(len(members))
    out += b"".join(lps(m[0]) for m in members)
    out += bytes(m[1] for m in members)
    out += bytes(m[2] for m in members if m[1] == 0)
    return out + i32(lib)


def int32_stream(value):
    """Like streams 0 and 2 of a .rap file: one boxed System.Int32."""
    body = b"\x04" + i32(1) + lps("System.Int32") + i32(1) + lps("m_value") + b"\x00" + b"\x08"
    return HEADER(1) + body + i32(value) + END


def chart_stream():
    shape = [("_text_str", 1, None), ("_Successor", 2, None)]
    io = shape + [("_is_input", 0, 1), ("_prompt", 1, None)]
    s = HEADER(1)
    s += b"\x0c" + i32(2) + lps("raptor, Version=4.1.0.1")
    s += class_with_types(1, "raptor.Oval", shape, 2) + obj_string(3, "Start") + ref(4)
    s += (class_with_types(4, "raptor.Parallelogram", io, 2) + obj_string(5, "h") + ref(6)
          + b"\x01" + obj_string(7, '"Enter h"'))
    s += class_with_types(6, "raptor.Rectangle", shape, 2) + obj_string(8, "x:=h*2") + ref(9)
    s += b"\x01" + i32(9) + i32(1) + obj_string(10, "End") + b"\x0a"   # ClassWithId -> Oval
    return s + END


DOC = int32_stream(17) + chart_stream() + int32_stream(1)


def test_streams_are_split_at_message_end():
    streams = rap_dump.read_streams(DOC)
    assert len(streams) == 3
    assert streams[0][0] == 0 and streams[-1][1] == len(DOC)
    assert [a[1] for a in streams[:-1]] == [b[0] for b in streams[1:]]
    assert streams[0][2].top[0].fields["m_value"] == 17
    assert streams[1][2].libraries == {2: "raptor, Version=4.1.0.1"}


def test_chart_is_walked_in_control_flow_order():
    rows = rap_dump.chart_rows(DOC)
    assert [(r[0], r[2], r[3]) for r in rows] == [
        (0, "Oval", "Start"),
        (0, "Parallelogram", "input  prompt='\"Enter h\"' -> h"),
        (0, "Rectangle", "assign x:=h*2"),
        (0, "Oval", "End"),
    ]
    # offsets point at the record byte of each node inside the whole document
    for _, off, _, _ in rows:
        assert DOC[off] in (1, 5)


def test_document_without_a_chart_is_rejected():
    with pytest.raises(ValueError, match="not a RAPTOR chart"):
        rap_dump.chart_rows(int32_stream(17))


def test_unknown_record_type_fails_loudly():
    with pytest.raises(ValueError, match="record type 7"):
        rap_dump.read_streams(HEADER(1) + b"\x07")


def test_dump_text(tmp_path):
    p = tmp_path / "toy.rap"
    p.write_bytes(DOC)
    text = rap_dump.dump(str(p))
    assert text.splitlines()[0] == f"toy.rap  ({len(DOC)} bytes, 3 streams)"
    assert text.splitlines()[-1] == "  nodes: 4"
    assert "Rectangle: assign x:=h*2" in text
