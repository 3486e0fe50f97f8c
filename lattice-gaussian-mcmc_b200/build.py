"""Builds liblgs_b200.so (the C-ABI library, include/lgs_b200.h) in-tree with nvcc for sm_100a.

    python lattice-gaussian-mcmc_b200/build.py [--force]

nvcc cross-compiles without a GPU, so this also runs in the CPU-only build container.  The ordered path is
compiled with -fmad=false (reference rounding order); the blocked DMMA path keeps FMA contraction.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "liblgs_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]

# (source, extra flags)
UNITS = [
    ("lgs_ordered.cu", ["-fmad=false"]),
    ("lgs_blocked.cu", []),
    ("lgs_chain.cu", []),
    ("lgs_diag.cu", []),
    ("lgs_capi.cu", []),
    ("lgs_reduce.cpp", ["-Xcompiler", "-mfma"]),
    ("lgs_hostpack.cpp", ["-Xcompiler", "-mavx2"]),   # host-only: widening of the packed D2H transport   # host-only C++ (LLL / BKZ); nvcc hands it to g++
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "lgs_b200.h"))
    return hdrs


def build_native(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_time = max(os.path.getmtime(h) for h in _deps())
    objs = []
    rebuilt = False
    for src, extra in UNITS:
        s = os.path.join(CSRC, src)
        if not os.path.exists(s):
            continue
        o = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(s), hdr_time):
            cmd = [nvcc, *ARCH, *COMMON, *extra, "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd))
            subprocess.check_call(cmd)
            rebuilt = True
    if rebuilt or not os.path.exists(LIB):
        cmd = [nvcc, *ARCH, "-shared", "-o", LIB, *objs, "-Xcompiler", "-fPIC"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build_native(force="--force" in sys.argv, verbose=True))
