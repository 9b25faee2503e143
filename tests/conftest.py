"""Test configuration.

The ``gpu`` marker is registered because the driver selects on it. This repo has no
GPU code (the reference has no hot path), so ``-m gpu`` selects nothing.
pytest reports that as exit code 5 ("no tests collected"). The hook below turns that
one case into exit 0 and prints the reason, so an empty GPU tier is not misread as a
failing one.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIR = "/root/reference/RAPTOR"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (none exist in this repo)")


def pytest_sessionfinish(session, exitstatus):
    if exitstatus == pytest.ExitCode.NO_TESTS_COLLECTED and session.config.getoption("-m") == "gpu":
        print("\nno gpu tests: the reference defines no hot path, so no CUDA code exists "
              "(DESIGN.md, SURVEY.md section 8)")
        session.exitstatus = pytest.ExitCode.OK


@pytest.fixture
def reference_dir():
    if not os.path.isdir(REFERENCE_DIR):
        pytest.skip("/root/reference is not mounted on this machine")
    return REFERENCE_DIR
