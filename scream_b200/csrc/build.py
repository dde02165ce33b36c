"""Builds libscream_render.so in-tree with nvcc for sm_100a (B200).  No torch, no pybind: a plain C-ABI
shared object (include/scream_render.h) that scream_b200/_lib.py loads through ctypes.

    python -m scream_b200.csrc.build [--force] [--verbose]

-fmad=false: no implicit FMA contraction anywhere (the reference rounds every ATen op separately);
the FMAs that the reference does perform (MKL sgemm) are written as explicit __fmaf_rn.
Denormals are kept (-ftz=false), division and sqrt are IEEE (-prec-div/-prec-sqrt=true).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
LIB = os.path.join(HERE, "libscream_render.so")
SOURCES = [os.path.join(HERE, "scream_render.cu")]
HEADERS = [os.path.join(HERE, f) for f in sorted(os.listdir(HERE)) if f.endswith(".cuh")] + \
          [os.path.join(ROOT, "include", "scream_render.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-ftz=false", "-prec-div=true", "-prec-sqrt=true",
    "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-cudart", "static",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libscream_render.so cannot be built (there is no CPU fallback)")


def up_to_date() -> bool:
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    return all(os.path.getmtime(p) <= t for p in SOURCES + HEADERS + [os.path.abspath(__file__)])


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and up_to_date():
        return LIB
    # SCREAM_DEFS: extra -D switches of an A/B build (tools/ab_build.sh); the shipped library is built without
    cmd = [find_nvcc()] + NVCC_FLAGS + os.environ.get("SCREAM_DEFS", "").split() + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
    res = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building libscream_render.so")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
