"""In-tree build of the CUDA library (and of the CPU test helper).

    libkgm.so          nvcc, sm_100a only, -lineinfo        -> k8s-gpu-monitor_b200/kgm/
    _hostcheck.so      g++ build of the shared leaf headers  -> tests/native/   (tests only)

nvcc cross-compiles without a GPU; the built .so files travel to the GPU box with the
snapshot (they are git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(os.path.dirname(PKG), "csrc")
ROOT = os.path.dirname(os.path.dirname(PKG))
LIB = os.path.join(PKG, "libkgm.so")
HOSTCHECK_SRC = os.path.join(ROOT, "tests", "native", "hostcheck.cpp")
HOSTCHECK = os.path.join(ROOT, "tests", "native", "_hostcheck.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]
SOURCES = ["kgm_kernels.cu", "kgm_api.cu", "kgm_window.cu", "kgm_replay.cu", "kgm_docker.cu"]
HEADERS = ["kgm_common.cuh", "kgm_d2s.cuh", "kgm_d2s_tables.h", "kgm_device.cuh", "kgm_skeleton.h",
           "kgm_poll2_leaf.cuh", "kgm_poll2.cuh",
           "kgm_fast.cuh", "kgm_fused.cuh", "kgm_window.cuh", "kgm_docker.cuh"]


def _stale(target, deps):
    if not os.path.isfile(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.isfile(d))


def _run(cmd):
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        sys.stderr.write("\n".join(proc.stdout.splitlines()[:40]) + "\n")
        raise RuntimeError("build failed: %s" % " ".join(cmd))
    return proc.stdout


def build_library(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS] + [os.path.join(ROOT, "include", "kgm.h")]
    if force or _stale(LIB, deps):
        nvcc = os.environ.get("NVCC", "nvcc")
        extra = os.environ.get("KGM_NVCC_DEFINES", "").split()      # e.g. -DKGM_PROFILE_STAGES
        out = _run([nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + srcs + ["-o", LIB])
        if verbose:
            print(out)
    return LIB


def build_hostcheck(force=False):
    deps = [HOSTCHECK_SRC] + [os.path.join(CSRC, h) for h in HEADERS]
    if force or _stale(HOSTCHECK, deps):
        _run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-x", "c++", HOSTCHECK_SRC, "-o", HOSTCHECK])
    return HOSTCHECK


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_hostcheck(force="--force" in sys.argv))
