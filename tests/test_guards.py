"""The guards that must do the right thing on the day the entry is re-pointed.

1. ``__graft_entry__.build()`` raises when the mounted reference is no longer the three
   surveyed documents, and when a native source appears without a build recipe.
2. ``tests/conftest.py`` never rewrites pytest's "no tests collected" exit code: an
   empty ``-m gpu`` tier exits 5 and prints which of the two reasons applies.
3. ``STATUS.json`` equals what ``__graft_entry__.status()`` generates.

None of these needs /root/reference or a GPU: the reference is a temp directory and
the pytest runs are subprocesses inside a scratch copy of the repo skeleton.
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys

import pytest

import __graft_entry__ as graft

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# -- 1. reference tripwire -------------------------------------------------------------
@pytest.fixture
def fake_reference(tmp_path, monkeypatch):
    """A stand-in reference: three files, with REFERENCE_EXPECTED pointed at them."""
    root = tmp_path / "reference"
    expected = {}
    for i, rel in enumerate(graft.REFERENCE_EXPECTED):
        data = f"stand-in document {i}".encode() * (i + 1)
        path = root / rel
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_bytes(data)
        expected[rel] = (len(data), hashlib.sha256(data).hexdigest())
    monkeypatch.setattr(graft, "REFERENCE_EXPECTED", expected)
    monkeypatch.setattr(graft, "REFERENCE_ROOT", str(root))
    return root


def test_unchanged_reference_passes(fake_reference, capsys):
    assert graft.reference_state(str(fake_reference)) == ("unchanged", [])
    assert graft.check_reference(str(fake_reference)) == "unchanged"
    assert "reference unchanged (3 files, hashes ok)" in capsys.readouterr().out


def test_absent_reference_is_reported_not_failed(tmp_path, capsys):
    missing = str(tmp_path / "nowhere")
    assert graft.reference_state(missing) == ("absent", [])
    assert graft.check_reference(missing) == "absent"
    assert "inventory check skipped" in capsys.readouterr().out


def test_a_fourth_file_trips_the_wire(fake_reference):
    (fake_reference / "CMakeLists.txt").write_text("project(raptor)\n")
    state, problems = graft.reference_state(str(fake_reference))
    assert state == "changed" and problems == ["unexpected file: CMakeLists.txt"]
    with pytest.raises(RuntimeError, match="reference changed — re-survey before building"):
        graft.check_reference(str(fake_reference))


def test_an_edited_file_trips_the_wire(fake_reference):
    rel = next(iter(graft.REFERENCE_EXPECTED))
    with open(fake_reference / rel, "ab") as fh:
        fh.write(b"!")
    state, problems = graft.reference_state(str(fake_reference))
    assert state == "changed" and problems == [f"content differs: {rel}"]


def test_a_missing_file_trips_the_wire(fake_reference):
    rel = next(iter(graft.REFERENCE_EXPECTED))
    os.remove(fake_reference / rel)
    assert graft.reference_state(str(fake_reference)) == ("changed", [f"missing file: {rel}"])


def test_a_symlinked_source_tree_trips_the_wire(fake_reference, tmp_path):
    """A re-point done by linking a source tree beside RAPTOR/. os.walk would not follow
    the link and would see no new file."""
    src = tmp_path / "elsewhere" / "core"
    src.mkdir(parents=True)
    (src / "par_matrix.cpp").write_text("// upstream RAPtor\n")
    os.symlink(src.parent, fake_reference / "raptor")
    state, problems = graft.reference_state(str(fake_reference))
    assert state == "changed" and problems == ["unexpected entry (symlink): raptor"]
    with pytest.raises(RuntimeError, match="reference changed"):
        graft.check_reference(str(fake_reference))


def test_an_empty_directory_tree_trips_the_wire(fake_reference):
    (fake_reference / "raptor" / "core").mkdir(parents=True)
    state, problems = graft.reference_state(str(fake_reference))
    assert state == "changed"
    assert problems == ["unexpected entry (directory): raptor",
                        "unexpected entry (directory): raptor/core"]


def test_a_symlink_in_place_of_a_document_trips_the_wire(fake_reference, tmp_path):
    """Same bytes behind a link is still not the surveyed tree."""
    rel = next(iter(graft.REFERENCE_EXPECTED))
    moved = tmp_path / "moved.rap"
    shutil.move(fake_reference / rel, moved)
    os.symlink(moved, fake_reference / rel)
    state, problems = graft.reference_state(str(fake_reference))
    assert state == "changed"
    assert problems == [f"unexpected entry (symlink): {rel}", f"missing file: {rel}"]


def test_a_symlinked_root_trips_the_wire(fake_reference, tmp_path):
    link = tmp_path / "reference_link"
    os.symlink(fake_reference, link)
    assert graft.reference_state(str(link))[0] == "changed"


def test_the_root_is_read_at_call_time(fake_reference):
    """``fake_reference`` patches ``graft.REFERENCE_ROOT``; the no-argument calls must
    see that, not a default bound when the module was imported."""
    assert graft.reference_state() == ("unchanged", [])
    (fake_reference / "CMakeLists.txt").write_text("project(raptor)\n")
    assert graft.reference_state()[0] == "changed"
    with pytest.raises(RuntimeError, match="reference changed"):
        graft.check_reference()


def _build_with_root(monkeypatch, root):
    """Run build() against ``root``. ``fake_reference`` has already pointed
    ``graft.REFERENCE_ROOT`` there, and ``check_reference()`` reads it at call time."""
    assert graft.REFERENCE_ROOT == str(root)
    graft.build()


def test_build_runs_the_tripwire(fake_reference, monkeypatch, capsys):
    _build_with_root(monkeypatch, fake_reference)
    out = capsys.readouterr().out
    assert "reference unchanged" in out and "no CUDA extensions" in out
    (fake_reference / "raptor" / "core").mkdir(parents=True)
    (fake_reference / "raptor" / "core" / "par_matrix.cpp").write_text("// re-pointed\n")
    with pytest.raises(RuntimeError, match="reference changed"):
        graft.build()


def test_build_refuses_native_sources_without_a_recipe(fake_reference, monkeypatch):
    _build_with_root(monkeypatch, fake_reference)          # passes while the tree is clean
    for fake in ("raptor_b200/csrc/spmv.cu", "raptor_b200/csrc/abi.cpp", "setup.py"):
        monkeypatch.setattr(graft, "native_sources", lambda f=fake: [os.path.join(ROOT, f)])
        with pytest.raises(RuntimeError, match="no recipe"):
            graft.build()


def test_native_sources_sees_more_than_cuda(tmp_path, monkeypatch):
    """The widened guard. C++, CMake, setuptools, PTX, prebuilt binaries, header-only code
    and Python files that import a GPU package all count. Only scratch trees
    (``gpurun_out/``, ``baseline/_ref/``) are skipped; ``build/`` and the rest of
    ``baseline/`` are not."""
    monkeypatch.setattr(graft, "_HERE", str(tmp_path))
    files = {
        "pkg/csrc/k.cu": "", "pkg/csrc/abi.cpp": "", "include/abi.h": "", "oracle/ref.c": "",
        "CMakeLists.txt": "", "setup.py": "", "pkg/k.ptx": "", "pkg/k.cubin": "",
        "pkg/k.fatbin": "", "pkg/k.inl": "", "pkg/k.hh": "", "pkg/libk.so": "",
        "build/k.cu": "", "baseline/k.cu": "",
        "pkg/spmv_triton.py": "import triton\nimport triton.language as tl\n",
        "pkg/spmv_torch.py": "import os\n\n    from torch.utils import cpp_extension\n",
        "pkg/spmv_cupy.py": "import cupy as cp\n",
        # not counted:
        "gpurun_out/x.cu": "", "baseline/_ref/y.cpp": "", "baseline/_ref/z.py": "import torch\n",
        "pkg/host.py": "import os\n# import torch is only mentioned in a comment\n",
        "pkg/torchless.py": "import torchlike\n",
        "tests/test_x.py": "import torch\n", "tools/helper.py": "import torch\n",
    }
    for rel, text in files.items():
        path = tmp_path / rel
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_text(text)
    rel = lambda ps: sorted(os.path.relpath(p, tmp_path) for p in ps)  # noqa: E731
    assert rel(graft.cuda_sources()) == ["baseline/k.cu", "build/k.cu", "pkg/csrc/k.cu"]
    assert rel(graft.gpu_python_sources()) == ["pkg/spmv_cupy.py", "pkg/spmv_torch.py",
                                               "pkg/spmv_triton.py"]
    assert rel(graft.native_sources()) == sorted([
        "CMakeLists.txt", "baseline/k.cu", "build/k.cu", "include/abi.h", "oracle/ref.c",
        "pkg/csrc/abi.cpp", "pkg/csrc/k.cu", "pkg/k.cubin", "pkg/k.fatbin", "pkg/k.hh",
        "pkg/k.inl", "pkg/k.ptx", "pkg/libk.so", "pkg/spmv_cupy.py", "pkg/spmv_torch.py",
        "pkg/spmv_triton.py", "setup.py"])


def test_a_tool_config_pyproject_is_not_a_native_source(tmp_path, monkeypatch):
    """pytest or ruff settings in pyproject.toml must not make build() raise; a
    [build-system] table or ext_modules must."""
    monkeypatch.setattr(graft, "_HERE", str(tmp_path))
    cfg = tmp_path / "pyproject.toml"
    cfg.write_text("[tool.pytest.ini_options]\nmarkers = ['gpu']\n")
    assert graft.native_sources() == []
    cfg.write_text("[build-system]\nrequires = ['setuptools']\n")
    assert graft.native_sources() == [str(cfg)]
    cfg.write_text("[tool.setuptools]\next_modules = [{name = 'k', sources = ['k.c']}]\n")
    assert graft.native_sources() == [str(cfg)]


def test_the_real_tree_has_nothing_to_build():
    assert graft.native_sources() == [] and graft.gpu_python_sources() == []


# -- 2. the empty-GPU-tier hook in conftest.py ------------------------------------------
def _scratch_repo(tmp_path, extra_files=()):
    """A copy of the repo skeleton: __graft_entry__.py, tests/conftest.py and one CPU-only
    test file. ``__graft_entry__`` globs relative to its own location, so sources added
    here are seen by the copy and never touch the real tree."""
    repo = tmp_path / "repo"
    (repo / "tests").mkdir(parents=True)
    shutil.copy(os.path.join(ROOT, "__graft_entry__.py"), repo / "__graft_entry__.py")
    shutil.copy(os.path.join(ROOT, "tests", "conftest.py"), repo / "tests" / "conftest.py")
    (repo / "tests" / "test_cpu_only.py").write_text("def test_cpu():\n    assert True\n")
    for rel, text in extra_files:
        path = repo / rel
        path.parent.mkdir(parents=True, exist_ok=True)
        path.write_text(text)
    return repo


def _pytest(repo, *args):
    env = {k: v for k, v in os.environ.items() if not k.startswith("PYTEST_")}
    return subprocess.run([sys.executable, "-m", "pytest", "tests", "-q", "-p", "no:cacheprovider",
                           *args], cwd=repo, capture_output=True, text=True, timeout=120, env=env)


def test_empty_gpu_tier_reports_none_not_ok(tmp_path):
    """Zero GPU tests is exit 5 ("none"), with the reason printed. It is never exit 0."""
    p = _pytest(_scratch_repo(tmp_path), "-m", "gpu")
    assert p.returncode == 5, p.stdout + p.stderr
    assert "no gpu tests: the reference defines no hot path" in p.stdout


@pytest.mark.parametrize("rel", ["raptor_b200/csrc/spmv.cu", "raptor_b200/csrc/abi.cpp",
                                 "include/raptor_b200.h", "setup.py", "raptor_b200/k.ptx",
                                 "build/k.cu"])
def test_empty_gpu_tier_fails_once_native_sources_exist(tmp_path, rel):
    """The re-point day: code has landed but no test carries the gpu marker. That must
    be exit 5, not a green GPU tier with zero tests."""
    p = _pytest(_scratch_repo(tmp_path, [(rel, "// placeholder\n")]), "-m", "gpu")
    assert p.returncode == 5, p.stdout + p.stderr
    assert "EMPTY GPU TIER WITH NATIVE SOURCES" in p.stdout


def test_compound_gpu_expression_is_recognised(tmp_path):
    repo = _scratch_repo(tmp_path)
    p = _pytest(repo, "-m", "gpu and not slow")
    assert p.returncode == 5 and "no gpu tests" in p.stdout
    (repo / "k.cu").write_text("")
    p = _pytest(repo, "-m", "gpu and not slow")
    assert p.returncode == 5 and "EMPTY GPU TIER WITH NATIVE SOURCES" in p.stdout


def test_cpu_selection_is_never_rewritten(tmp_path):
    """``-m "not gpu"`` with nothing collected is exit 5 too, and the hook prints nothing."""
    repo = _scratch_repo(tmp_path)
    assert _pytest(repo, "-m", "not gpu").returncode == 0          # one CPU test runs
    os.remove(repo / "tests" / "test_cpu_only.py")
    p = _pytest(repo, "-m", "not gpu")
    assert p.returncode == 5
    assert "no gpu tests:" not in p.stdout and "EMPTY GPU TIER" not in p.stdout
    assert _pytest(repo).returncode == 5


def test_a_real_gpu_test_is_collected_not_swallowed(tmp_path):
    """With a gpu-marked test present the hook stays out of the way (pass stays pass,
    fail stays fail)."""
    body = "import pytest\n\n@pytest.mark.gpu\ndef test_g():\n    assert {}\n"
    ok = _scratch_repo(tmp_path / "a", [("tests/test_g.py", body.format("True"))])
    bad = _scratch_repo(tmp_path / "b", [("tests/test_g.py", body.format("False"))])
    assert _pytest(ok, "-m", "gpu").returncode == 0
    assert _pytest(bad, "-m", "gpu").returncode == 1


def test_selects_gpu_expression_matcher():
    from conftest import selects_gpu
    assert selects_gpu("gpu") and selects_gpu("gpu and not slow") and selects_gpu("slow or gpu")
    assert not selects_gpu("not gpu") and not selects_gpu("") and not selects_gpu(None)
    assert not selects_gpu("gpus") and not selects_gpu("not  gpu and slow")


# -- 3. STATUS.json ---------------------------------------------------------------------
def test_status_file_is_what_the_generator_writes(tmp_path):
    with open(graft.STATUS_PATH) as fh:
        committed = json.load(fh)
    assert committed == graft.status()
    out = tmp_path / "STATUS.json"
    graft.write_status(str(out))
    with open(graft.STATUS_PATH) as a, open(out) as b:
        assert a.read() == b.read()


def test_status_is_not_written_over_a_changed_reference(fake_reference, tmp_path):
    out = tmp_path / "STATUS.json"
    graft.write_status(str(out))                       # unchanged stand-in: written
    assert json.loads(out.read_text())["graftable"] is False
    out.unlink()
    (fake_reference / "CMakeLists.txt").write_text("project(raptor)\n")
    with pytest.raises(RuntimeError, match="reference changed"):
        graft.write_status(str(out))
    assert not out.exists()


def test_status_says_what_the_tests_check():
    st = graft.status()
    assert st["graftable"] is False and st["gpu_workload_built"] is False
    assert st["decision_needed"] == "drop | re-point to raptor-library/raptor"
    assert st["since_round"] == 1
    ev = st["evidence"]
    assert "VERDICT.md" in ev["see"] and not any("round" in x for x in ev["see"])
    assert ev["files"] == 3 and ev["bytes"] == 26878
    assert ev["sha256"] == {rel: sha for rel, (_, sha) in graft.REFERENCE_EXPECTED.items()}
    with open(os.path.join(ROOT, "BASELINE.json")) as fh:
        assert json.load(fh)["metric"].startswith("N/A")
