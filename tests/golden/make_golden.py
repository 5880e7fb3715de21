#!/usr/bin/env python3
"""Freeze the UNMODIFIED reference's output on the shared test cases.

Run in the build container (needs ``/root/reference/gpu.py``):

    python tests/golden/make_golden.py

Writes ``tests/golden/polls.json.gz``: for every case of ``tests/cases.all_cases()`` the
eight poll inputs plus what ``gpu.main()`` POSTed (``lines``) and the class of the
exception it died with (``err``), obtained through ``oracle/ref_harness.py``.  The GPU box
has no reference tree; tests there read this file.
"""
import gzip
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "k8s-gpu-monitor_b200"), os.path.join(ROOT, "tests")]

import cases  # noqa: E402
from oracle import ref_harness  # noqa: E402


def main():
    if not ref_harness.available():
        sys.exit("reference tree not found at %s" % ref_harness.REFERENCE_DIR)
    out = []
    for c in cases.all_cases():
        lines, err = ref_harness.run_reference(*cases.inputs_of(c))
        rec = {k: c[k] for k in ("name",) + cases.INPUT_KEYS}
        rec["lines"] = lines
        rec["err"] = err
        out.append(rec)
    path = os.path.join(HERE, "polls.json.gz")
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        f.write(json.dumps(out, sort_keys=True).encode())
    nerr = sum(1 for r in out if r["err"])
    print("wrote %s: %d cases (%d raising), %d lines" %
          (path, len(out), nerr, sum(len(r["lines"]) for r in out)))


if __name__ == "__main__":
    main()
