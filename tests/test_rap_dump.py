"""tools/rap_dump.py against small MS-NRBF documents assembled in this file.

Everything here is synthetic, so the tests need no reference data and run on any
machine (the GPU box has no /root/reference). The main document mirrors the layout of
a real ``.rap`` file (version, flag, tab count, tab name, chart)
and its chart has the two shapes the reference charts use: nested ``IF_Control`` nodes
as in BMI.rap@4246 and a ``Loop`` with both child lists as in
``multipication in reverse order.rap``@2790.
"""
import struct

import pytest

from tools import rap_dump


# -- NRBF byte builders ---------------------------------------------------------------
def lps(s: str) -> bytes:
    b = s.encode("utf-8")
    assert len(b) < 128
    return bytes([len(b)]) + b


def i32(v: int) -> bytes:
    return struct.pack("<i", v)


HEADER = lambda root: b"\x00" + i32(root) + i32(-1) + i32(1) + i32(0)  # noqa: E731
END = b"\x0b"
NULL = b"\x0a"
LIB = b"\x0c" + i32(2) + lps("raptor, Version=4.1.0.1")


def obj_string(oid, s):
    return b"\x06" + i32(oid) + lps(s)


def ref(oid):
    return b"\x09" + i32(oid)


def class_with_types(oid, name, members, lib):
    """members: [(name, binary type kind, extra)]; extra is the primitive type for kind 0,
    (class name, library id) for kind 4 (Class), None otherwise."""
    out = b"\x05" + i32(oid) + lps(name) + i32(len(members))
    out += b"".join(lps(m[0]) for m in members)
    out += bytes(m[1] for m in members)
    for _, kind, extra in members:
        if kind == 0:
            out += bytes([extra])
        elif kind == 4:
            out += lps(extra[0]) + i32(extra[1])
    return out + i32(lib)


def class_with_id(oid, metadata_oid):
    return b"\x01" + i32(oid) + i32(metadata_oid)


def int32_stream(value):
    """Like streams 0 and 2 of a .rap file: one boxed System.Int32."""
    body = b"\x04" + i32(1) + lps("System.Int32") + i32(1) + lps("m_value") + b"\x00" + b"\x08"
    return HEADER(1) + body + i32(value) + END


def bool_stream(value):
    """Like stream 1 of a .rap file: one boxed System.Boolean."""
    body = b"\x04" + i32(1) + lps("System.Boolean") + i32(1) + lps("m_value") + b"\x00" + b"\x01"
    return HEADER(1) + body + bytes([value]) + END


def string_stream(s):
    """Like stream 3 of a .rap file: the tab name."""
    return HEADER(1) + obj_string(1, s) + END


SHAPE = [("_text_str", 1, None), ("_Successor", 2, None)]
IO = SHAPE + [("_is_input", 0, 1), ("_prompt", 1, None)]
IF = SHAPE + [("_left_Child", 2, None), ("_right_Child", 2, None)]
LOOP = SHAPE + [("_before_Child", 2, None), ("_after_Child", 2, None)]


def linear_chart_stream():
    """Start -> input h -> x:=h*2 -> End."""
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(3, "Start") + ref(4)
    s += (class_with_types(4, "raptor.Parallelogram", IO, 2) + obj_string(5, "h") + ref(6)
          + b"\x01" + obj_string(7, '"Enter h"'))
    s += class_with_types(6, "raptor.Rectangle", SHAPE, 2) + obj_string(8, "x:=h*2") + ref(9)
    s += class_with_id(9, 1) + obj_string(10, "End") + NULL
    return s + END


def branching_chart_stream():
    """Start -> if a<1 {yes: out "A"; no: if a<2 {yes: out "B"; no: (empty)}}
    -> loop(exit i<0) {before: i:=i-1; after: m:=n*i, out m} -> End.

    Child objects are written after their parents and linked by MemberReference, the
    way BinaryFormatter writes the reference charts.
    """
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(20, "Start") + ref(2)
    # outer IF (id 2): successor = loop (8), yes = output A (3), no = inner IF (4)
    s += class_with_types(2, "raptor.IF_Control", IF, 2) + obj_string(21, "a<1") + ref(8) + ref(3) + ref(4)
    s += (class_with_types(3, "raptor.Parallelogram", IO, 2) + obj_string(22, '"A"') + NULL
          + b"\x00" + NULL)
    # inner IF (id 4): no successor, yes = output B (5), no branch empty
    s += class_with_id(4, 2) + obj_string(23, "a<2") + NULL + ref(5) + NULL
    s += class_with_id(5, 3) + obj_string(24, '"B"') + NULL + b"\x00" + NULL
    # loop (id 8): successor = End (12), before = i:=i-1 (9), after = m:=n*i (10) -> out m (11)
    s += class_with_types(8, "raptor.Loop", LOOP, 2) + obj_string(25, "i<0") + ref(12) + ref(9) + ref(10)
    s += class_with_types(9, "raptor.Rectangle", SHAPE, 2) + obj_string(26, "i:=i-1") + NULL
    s += class_with_id(10, 9) + obj_string(27, "m:=n*i") + ref(11)
    s += class_with_id(11, 3) + obj_string(28, "m") + NULL + b"\x00" + NULL
    s += class_with_id(12, 1) + obj_string(29, "End") + NULL
    return s + END


def rap_document(*chart_streams, tabs=None, names=None):
    """version, flag, tab count, then (tab name, chart) per chart, then a trailing flag."""
    tabs = len(chart_streams) if tabs is None else tabs
    names = names or ["main"] + [f"sub{i}" for i in range(1, len(chart_streams))]
    doc = int32_stream(17) + bool_stream(0) + int32_stream(tabs)
    for name, chart in zip(names, chart_streams):
        doc += string_stream(name) + chart
    return doc + bool_stream(0)


DOC = rap_document(linear_chart_stream())
BRANCHY = rap_document(branching_chart_stream())


# -- container and linear chart -------------------------------------------------------
def test_streams_are_split_at_message_end():
    streams = rap_dump.read_streams(DOC)
    assert len(streams) == 6
    assert streams[0][0] == 0 and streams[-1][1] == len(DOC)
    assert [a[1] for a in streams[:-1]] == [b[0] for b in streams[1:]]
    assert streams[0][2].top[0].fields["m_value"] == 17
    assert streams[1][2].top[0].fields["m_value"] is False
    assert streams[4][2].libraries == {2: "raptor, Version=4.1.0.1"}
    assert rap_dump.declared_chart_count(streams) == 1


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
    text = rap_dump.dump(str(p)).splitlines()
    assert text[0] == f"toy.rap  ({len(DOC)} bytes, 6 streams)"
    assert text[1] == "  charts: 1 (declared 1; showing 'main')"
    assert text[-1] == "  nodes: 4"
    assert any("Rectangle: assign x:=h*2" in ln for ln in text)


# -- (e) branch and loop walking, without the reference --------------------------------
def test_if_and_loop_children_are_walked_with_depth_and_labels():
    rows = rap_dump.chart_rows(BRANCHY)
    assert [(r[0], r[2], r[3]) for r in rows] == [
        (0, "Oval", "Start"),
        (0, "IF_Control", "if a<1"),
        (1, "yes", ""),
        (2, "Parallelogram", 'output "A"'),
        (1, "no", ""),
        (2, "IF_Control", "if a<2"),
        (3, "yes", ""),
        (4, "Parallelogram", 'output "B"'),      # the empty "no" branch prints no label
        (0, "Loop", "loop, exit when i<0"),
        (1, "before test", ""),
        (2, "Rectangle", "assign i:=i-1"),
        (1, "after test", ""),
        (2, "Rectangle", "assign m:=n*i"),
        (2, "Parallelogram", "output m"),
        (0, "Oval", "End"),
    ]
    # label rows carry no offset; node rows point at a class record
    assert all((off is None) == (kind in ("yes", "no", "before test", "after test"))
               for _, off, kind, _ in rows)
    assert sum(1 for r in rows if r[1] is not None) == 10


# -- (a) truncated input ---------------------------------------------------------------
def test_every_truncation_is_a_value_error_with_an_offset():
    """Cut the document at every byte: a clean ValueError each time, never
    struct.error / IndexError / KeyError."""
    complete = 0
    for cut in range(len(BRANCHY)):
        try:
            rap_dump.chart_rows(BRANCHY[:cut])
            complete += 1          # cut fell on a stream boundary after the chart
        except ValueError as e:
            assert "byte" in str(e) or "not a RAPTOR chart" in str(e), (cut, e)
    assert complete == 1           # only the cut just before the trailing flag stream


def test_truncation_message_names_the_byte():
    with pytest.raises(ValueError, match=r"truncated NRBF stream at byte 5"):
        rap_dump.read_streams(DOC[:5])
    with pytest.raises(ValueError, match=r"truncated NRBF stream at byte 400"):
        rap_dump.read_streams(DOC[:400])


def test_class_with_id_naming_an_unknown_class_is_a_value_error():
    with pytest.raises(ValueError, match="unknown class id 99"):
        rap_dump.read_streams(HEADER(1) + class_with_id(1, 99) + END)


# -- (b) cycles ------------------------------------------------------------------------
def test_successor_cycle_raises_instead_of_looping():
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(3, "Start") + ref(4)
    s += class_with_types(4, "raptor.Rectangle", SHAPE, 2) + obj_string(5, "x:=1") + ref(1)
    with pytest.raises(ValueError, match="revisits the node"):
        rap_dump.chart_rows(rap_document(s + END))


def test_child_cycle_raises_instead_of_recursing():
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(3, "Start") + ref(2)
    s += class_with_types(2, "raptor.IF_Control", IF, 2) + obj_string(4, "a<1") + NULL + ref(2) + NULL
    with pytest.raises(ValueError, match="revisits the node"):
        rap_dump.chart_rows(rap_document(s + END))


# -- (c) null runs and array records (types 13-17) --------------------------------------
def array_stream(body_records, n):
    """One ArraySingleObject of declared length ``n`` as the stream root."""
    return HEADER(1) + b"\x10" + i32(1) + i32(n) + body_records + END


def test_null_runs_expand_inside_arrays_only():
    run256 = b"\x0d" + bytes([3])                 # ObjectNullMultiple256, 3 nulls
    run = b"\x0e" + i32(2)                        # ObjectNullMultiple, 2 nulls
    empty = b"\x0d" + bytes([0])                  # zero-length run: contributes nothing
    body = obj_string(2, "a") + run256 + empty + obj_string(3, "b") + run
    (_, _, st), = rap_dump.read_streams(array_stream(body, 7))
    assert st.top[0] == ["a", None, None, None, "b", None, None]


def test_nested_array_starting_with_null_is_not_flattened():
    inner = b"\x10" + i32(2) + i32(2) + NULL + obj_string(3, "x")
    (_, _, st), = rap_dump.read_streams(array_stream(inner + obj_string(4, "y"), 2))
    assert st.top[0] == [[None, "x"], "y"]


def test_null_run_longer_than_the_array_is_rejected():
    with pytest.raises(ValueError, match="overflows"):
        rap_dump.read_streams(array_stream(b"\x0d" + bytes([5]), 2))


def test_primitive_and_string_arrays():
    prim = HEADER(1) + b"\x0f" + i32(1) + i32(3) + b"\x08" + i32(7) + i32(-1) + i32(0) + END
    (_, _, st), = rap_dump.read_streams(prim)
    assert st.top[0] == [7, -1, 0]
    strs = HEADER(1) + b"\x11" + i32(1) + i32(2) + obj_string(2, "u") + NULL + END
    (_, _, st), = rap_dump.read_streams(strs)
    assert st.top[0] == ["u", None]


def test_null_run_covers_consecutive_object_members():
    """MS-NRBF lets one ObjectNullMultiple record stand for several null members."""
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(3, "Start") + ref(2)
    s += (class_with_types(2, "raptor.IF_Control", IF, 2) + obj_string(4, "a<1")
          + b"\x0d" + bytes([3]))                 # _Successor, _left_Child, _right_Child
    rows = rap_dump.chart_rows(rap_document(s + END))
    assert [(r[0], r[2]) for r in rows] == [(0, "Oval"), (0, "IF_Control")]


# -- (d) tab count ----------------------------------------------------------------------
def test_multi_tab_document_lists_every_chart(tmp_path):
    doc = rap_document(linear_chart_stream(), branching_chart_stream())
    found = rap_dump.charts(doc)
    assert [name for name, _ in found] == ["main", "sub1"]
    assert [sum(1 for r in rows if r[1] is not None) for _, rows in found] == [4, 10]
    p = tmp_path / "two.rap"
    p.write_bytes(doc)
    text = rap_dump.dump(str(p)).splitlines()
    assert text[1] == "  charts: 2 (declared 2; showing 'main', 'sub1')"
    assert "  -- chart 'sub1'" in text and text[-1] == "  nodes: 14"


def test_declared_tab_count_must_match_charts_found():
    """A document that declares 2 tabs but holds 1 chart is refused, not listed as if
    'main' were the whole program."""
    with pytest.raises(ValueError, match="declares 2 chart tab"):
        rap_dump.chart_rows(rap_document(linear_chart_stream(), tabs=2))
    with pytest.raises(ValueError, match="declares 1 chart tab"):
        rap_dump.charts(rap_document(linear_chart_stream(), linear_chart_stream(), tabs=1))


def test_document_without_a_tab_count_is_still_listed():
    doc = string_stream("main") + linear_chart_stream()
    assert rap_dump.declared_chart_count(rap_dump.read_streams(doc)) is None
    assert len(rap_dump.chart_rows(doc)) == 4


# -- Rectangle kinds and MemberPrimitiveTyped (record type 8) ----------------------------
def test_rectangle_kind_enum_distinguishes_assignment_from_call():
    """In the reference charts every Rectangle carries a ``_kind`` enum object with
    ``value__`` 0 (assignment). A non-zero kind (a procedure call) must be shown as such,
    because "no call nodes" is part of the scalar-toy finding."""
    kind_cls = "raptor.Rectangle+Kind_Of"
    rect = SHAPE + [("_kind", 4, (kind_cls, 2))]
    s = HEADER(1) + LIB
    s += class_with_types(1, "raptor.Oval", SHAPE, 2) + obj_string(20, "Start") + ref(2)
    s += (class_with_types(2, "raptor.Rectangle", rect, 2) + obj_string(21, "x:=1") + ref(4)
          + class_with_types(3, kind_cls, [("value__", 0, 8)], 2) + i32(0))
    s += (class_with_id(4, 2) + obj_string(22, "draw_it(x)") + NULL
          + class_with_id(5, 3) + i32(1))
    rows = rap_dump.chart_rows(rap_document(s + END))
    assert [r[3] for r in rows] == ["Start", "assign x:=1", "call(kind=1) draw_it(x)"]


def test_member_primitive_typed_records_inside_an_object_array():
    body = b"\x08\x08" + i32(42) + b"\x08\x01\x01" + b"\x08\x06" + struct.pack("<d", 18.5)
    (_, _, st), = rap_dump.read_streams(array_stream(body, 3))
    assert st.top[0] == [42, True, 18.5]


# -- crafted inputs that once escaped as RecursionError or a huge allocation -----------
def array_of_objects(oid, n):
    return b"\x10" + i32(oid) + i32(n)


def test_a_long_run_of_library_records_is_read_in_a_loop():
    """5 000 BinaryLibrary records in a row used to recurse once each."""
    libs = b"".join(b"\x0c" + i32(2 + k) + lps("lib") for k in range(5000))
    s = rap_dump._Stream(HEADER(1) + libs + obj_string(1, "x") + END, 0).parse()
    assert s.top == ["x"] and len(s.libraries) == 5000


def test_deep_array_nesting_is_a_value_error_naming_the_byte():
    """5 000 arrays each holding the next. Depth is capped, so this is never RecursionError."""
    depth = 5000
    body = b"".join(array_of_objects(10 + k, 1) for k in range(depth)) + NULL
    data = HEADER(10) + body + b"\x00" * depth + END     # padding keeps each length plausible
    with pytest.raises(ValueError, match=rf"nested deeper than {rap_dump.MAX_DEPTH} at byte \d+"):
        rap_dump._Stream(data, 0).parse()


def test_nesting_up_to_the_cap_is_accepted():
    depth = rap_dump.MAX_DEPTH - 1
    body = b"".join(array_of_objects(10 + k, 1) for k in range(depth)) + obj_string(9, "leaf")
    s = rap_dump._Stream(HEADER(10) + body + b"\x00" * depth + END, 0)
    s.record()                                           # the header
    v = s.record()
    for _ in range(depth):
        assert isinstance(v, list) and len(v) == 1
        v = v[0]
    assert v == "leaf"


@pytest.mark.parametrize("n", [2 ** 28, 2 ** 31 - 1])
def test_a_giant_declared_array_is_refused_before_anything_is_allocated(n):
    """The 31-byte stream from VERDICT.md round 2: an array of 2^28 whose only element is
    a null run of 2^28. It used to allocate gigabytes before any check ran."""
    import tracemalloc
    data = HEADER(1) + array_of_objects(1, n) + b"\x0e" + i32(n) + END
    assert len(data) == 17 + 9 + 5 + 1
    tracemalloc.start()
    try:
        with pytest.raises(ValueError, match=rf"array at byte 17 declares {n} elements"):
            rap_dump._Stream(data, 0).parse()
        peak = tracemalloc.get_traced_memory()[1]
    finally:
        tracemalloc.stop()
    assert peak < 1 << 20


def test_a_null_run_longer_than_the_array_has_room_for_is_refused_before_extending():
    n = 8
    data = HEADER(1) + array_of_objects(1, n) + NULL + b"\x0e" + i32(n) + b"\x00" * n + END
    with pytest.raises(ValueError, match=r"null run of 8 overflows the 8-element array at byte 17"):
        rap_dump._Stream(data, 0).parse()


def test_a_giant_primitive_array_is_refused():
    data = HEADER(1) + b"\x0f" + i32(1) + i32(2 ** 30) + b"\x08" + END
    with pytest.raises(ValueError, match="declares 1073741824 elements"):
        rap_dump._Stream(data, 0).parse()


def test_message_end_as_a_member_is_refused():
    s = HEADER(1) + LIB + class_with_types(1, "raptor.Oval", SHAPE, 2) + END + NULL + END
    with pytest.raises(ValueError, match="MessageEnd as member '_text_str'"):
        rap_dump._Stream(s, 0).parse()


def test_a_null_run_past_the_last_member_is_refused():
    """A run of 3 nulls over an object with 2 reference members leaves 1 unaccounted for."""
    s = HEADER(1) + LIB + class_with_types(1, "raptor.Oval", SHAPE, 2) + b"\x0d\x03" + END
    with pytest.raises(ValueError, match="null run overflows the members of the object"):
        rap_dump._Stream(s, 0).parse()
    ok = HEADER(1) + LIB + class_with_types(1, "raptor.Oval", SHAPE, 2) + b"\x0d\x02" + END
    assert rap_dump._Stream(ok, 0).parse().top[0].fields == {"_text_str": None, "_Successor": None}
