#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- makes the reference script available where /root/reference is not.

    python oracle/make_ref.py

Copies the UNMODIFIED `/root/reference/gpu.py` to `oracle/_ref/gpu.py` and records its SHA-256
in `oracle/_ref/SOURCE`.  `oracle/_ref/` is git-ignored (no reference source enters the
history) but not gpurun-ignored, so the copy travels to the GPU box like the built
`libkgm.so` does.  There `oracle/ref_harness.py` drives it, so that `bench.py --impl reference`
and the `cpu_baseline` leg time the reference's own `main()` (gpu.py:90-256) instead of the
port in `oracle/poll_oracle.py`.  The reference is one Python script: there is nothing to
compile.
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(os.environ.get("KGM_REFERENCE_DIR", "/root/reference"), "gpu.py")
DST_DIR = os.path.join(HERE, "_ref")


def main():
    if not os.path.isfile(SRC):
        print("make_ref: %s not found; keeping whatever oracle/_ref/ holds" % SRC)
        return 0
    os.makedirs(DST_DIR, exist_ok=True)
    dst = os.path.join(DST_DIR, "gpu.py")
    shutil.copyfile(SRC, dst)
    digest = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    with open(os.path.join(DST_DIR, "SOURCE"), "w") as f:
        f.write("%s sha256=%s\n" % (SRC, digest))
    print("make_ref: %s -> %s (sha256 %s)" % (SRC, dst, digest[:16]))
    return 0


if __name__ == "__main__":
    sys.exit(main())
