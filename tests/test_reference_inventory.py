"""Re-check the non-graftable finding whenever the reference is mounted.

If any of these fail, the reference has changed (for example it was re-pointed to the
upstream AMG library) and SURVEY.md / DESIGN.md must be redone before anything else.
"""
import hashlib
import os

from tools import rap_dump

# file name -> (bytes, sha256 prefix, flowchart nodes)
EXPECTED = {
    "BMI.rap": (11378, "658d4628b43a", 13),
    "multipication in reverse order.rap": (8789, "bcd3016a53cb", 10),
    "summation of digits in 4 digited number.rap": (6711, "aa32ddcc02ca", 9),
}

SOURCE_SUFFIXES = (".c", ".cc", ".cpp", ".h", ".hpp", ".cu", ".cuh", ".rs", ".go", ".java",
                   ".py", ".js", ".ts", ".cs", ".f90", ".toml", ".txt", ".cmake", ".md")


def _all_files(root):
    return sorted(os.path.join(d, f) for d, _, fs in os.walk(root) for f in fs)


def test_reference_is_exactly_three_rap_files(reference_dir):
    files = _all_files(os.path.dirname(reference_dir))
    assert [os.path.basename(f) for f in files] == sorted(EXPECTED)
    assert not [f for f in files if f.lower().endswith(SOURCE_SUFFIXES)]


def test_reference_hashes_match_survey(reference_dir):
    for name, (size, sha, _) in EXPECTED.items():
        with open(os.path.join(reference_dir, name), "rb") as fh:
            data = fh.read()
        assert len(data) == size, name
        assert hashlib.sha256(data).hexdigest().startswith(sha), name


def test_reference_charts_are_scalar_toys(reference_dir):
    """Node counts match SURVEY.md section 2 and the only loop is the 11-step one."""
    loops = 0
    for name, (_, _, nodes) in EXPECTED.items():
        with open(os.path.join(reference_dir, name), "rb") as fh:
            data = fh.read()
        assert len(rap_dump.read_streams(data)) == 10, name
        rows = [r for r in rap_dump.chart_rows(data) if r[1] is not None]
        assert len(rows) == nodes, name
        assert {r[2] for r in rows} <= {"Oval", "Parallelogram", "Rectangle", "IF_Control", "Loop"}
        loops += sum(1 for r in rows if r[2] == "Loop")
        # no subchart / procedure call nodes: every Rectangle is a plain assignment
        assert all(r[3].startswith("assign ") for r in rows if r[2] == "Rectangle"), name
    assert loops == 1
