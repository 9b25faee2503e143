#!/usr/bin/env python3
"""Driver bench contract for an entry that has NO hot path to measure.

The mounted reference (three RAPTOR flowchart documents) defines no data-parallel
path, no metric and no CPU implementation (SURVEY.md sections 0 and 8).
BASELINE.json's ``metric`` is "N/A". This script therefore accepts the driver's
command line, does no GPU work, and prints ONE JSON line that says so: ``value`` is
null, ``higher_is_better`` is null (a metric that does not exist has no direction),
``gpu_launches`` is 0 and ``unavailable`` gives the reason. It never prints a number
that could be mistaken for a measurement.

Under torchrun (N>1) only rank 0 prints. All ranks exit 0. No process group is
created because there is nothing to synchronise.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))

WHY = ("reference holds only three RAPTOR .rap flowchart documents (scalar teaching "
       "exercises): no hot path, API, CPU implementation, tests or metric exist; "
       "entry is non-graftable as mounted (BASELINE.json north_star, SURVEY.md s0/s8)")


def _metric() -> str:
    try:
        with open(os.path.join(_HERE, "BASELINE.json")) as fh:
            return json.load(fh)["metric"]
    except (OSError, KeyError, ValueError):
        return "N/A: no graftable hot path"


def line(args, ignored=()) -> dict:
    if args.impl == "reference":
        out = {"impl": "reference",
               "unavailable": "reference is not software: no setup.py/pyproject, no source "
                              "code, and the RAPTOR interpreter (.NET) is not in it"}
        if ignored:
            out["ignored_args"] = list(ignored)
        return out
    out = {
        "metric": _metric(), "value": None, "unit": None,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": None, "higher_is_better": None, "scaling": None,
        "vs_baseline": None, "dtype": None, "data": "none",
        "config": {"workload": "none: BASELINE.json configs 1-3 are single interactive "
                               "runs of a scalar flowchart; configs 4-5 apply only if "
                               "the entry is re-pointed"},
        "e2e": None, "roofline": None, "cpu_baseline": None, "clocks": None,
        "gpu_launches": 0,
        "unavailable": WHY,
    }
    if ignored:
        out["ignored_args"] = list(ignored)
    return out


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("b200", "reference"), default="b200")
    # Unknown driver flags are recorded, not fatal: an argparse exit 2 would read as a
    # bench crash instead of the ``unavailable`` finding.
    args, ignored = ap.parse_known_args(argv)
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(line(args, ignored)), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
