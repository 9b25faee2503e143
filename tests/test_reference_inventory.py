"""Re-check the non-graftable finding whenever the reference is mounted.

If any of these fail, the reference has changed (for example it was re-pointed to the
upstream AMG library) and SURVEY.md / DESIGN.md must be redone before anything else.

File names, sizes and hashes come from ``__graft_entry__.REFERENCE_EXPECTED``: the same
table ``build()`` and ``STATUS.json`` use, so the three cannot disagree.
"""
import os

import __graft_entry__ as graft
from tools import rap_dump

# flowchart nodes per document.
NODES = {
    "RAPTOR/BMI.rap": 13,
    "RAPTOR/multipication in reverse order.rap": 10,
    "RAPTOR/summation of digits in 4 digited number.rap": 9,
}


def test_node_table_covers_the_same_files_as_the_hash_table():
    assert set(NODES) == set(graft.REFERENCE_EXPECTED)


def test_reference_is_exactly_the_three_surveyed_files(reference_dir):
    state, problems = graft.reference_state()
    assert (state, problems) == ("unchanged", [])


def test_check_reference_passes_on_the_mounted_reference(reference_dir, capsys):
    assert graft.check_reference() == "unchanged"
    assert "reference unchanged (3 files, hashes ok)" in capsys.readouterr().out


def test_reference_charts_are_scalar_toys(reference_dir):
    """Node counts match SURVEY.md section 2, each document declares and holds exactly
    one chart tab, and the only loop is the 11-step one."""
    loops = 0
    root = os.path.dirname(reference_dir)
    for rel, nodes in NODES.items():
        with open(os.path.join(root, rel), "rb") as fh:
            data = fh.read()
        streams = rap_dump.read_streams(data)
        assert len(streams) == 10, rel
        assert rap_dump.declared_chart_count(streams) == 1, rel
        found = rap_dump.charts(data)
        assert [name for name, _ in found] == ["main"], rel
        rows = [r for r in found[0][1] if r[1] is not None]
        assert len(rows) == nodes, rel
        assert {r[2] for r in rows} <= {"Oval", "Parallelogram", "Rectangle", "IF_Control", "Loop"}
        loops += sum(1 for r in rows if r[2] == "Loop")
        # no subchart / procedure call nodes: every Rectangle is a plain assignment
        assert all(r[3].startswith("assign ") for r in rows if r[2] == "Rectangle"), rel
    assert loops == 1
