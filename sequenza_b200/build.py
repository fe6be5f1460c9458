"""Builds libsequenza_b200.so in-tree with nvcc for sm_100a (no other target).

    python -m sequenza_b200.build [--force]

The shared object is git-ignored but travels with the tree to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libsequenza_b200.so")
SOURCES = ["sqz_api.cu", "sqz_baf_grid.cu", "sqz_baf_split.cu", "sqz_baf_colwarp.cu",
           "sqz_mufreq.cu", "sqz_finish.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",  # faithful nmath code must not contract a*b+c; fast path calls fma()
    "-Xcompiler", "-fPIC", "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libsequenza_b200.so cannot be built")


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "sequenza_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return SO
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
    cmd += ["-o", SO] + [os.path.join(CSRC, s) for s in SOURCES]
    env = dict(os.environ)
    # nvcc's host compiler: the image's CC points at a gcc wrapper without libgomp specs
    cmd += ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    subprocess.check_call(cmd, env=env)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
