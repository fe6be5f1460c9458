"""Builds libsequenza_b200.so in-tree with nvcc for sm_100a (no other target).

    python -m sequenza_b200.build [--force] [-v]

Every translation unit is compiled to an object under sequenza_b200/build/ (in parallel,
only when it or a header changed) and the objects are linked into the shared object.
Both are git-ignored; the shared object travels with the tree to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
SO = os.path.join(HERE, "libsequenza_b200.so")
SOURCES = ["sqz_api.cu", "sqz_baf_grid.cu", "sqz_baf_split.cu", "sqz_baf_colwarp.cu",
           "sqz_mufreq.cu", "sqz_finish.cu", "sqz_segstats.cu", "sqz_peer.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",  # faithful nmath code must not contract a*b+c; fast path calls fma()
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libsequenza_b200.so cannot be built")


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "sequenza_b200.h"))
    return hs


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + _headers()
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return SO
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    # nvcc's host compiler: the image's CC points at a gcc wrapper without libgomp specs
    ccbin = ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    hdr_t = max(os.path.getmtime(h) for h in _headers())
    self_t = os.path.getmtime(os.path.abspath(__file__))

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src[:-3] + ".o")
        if (not force and os.path.exists(o)
                and os.path.getmtime(o) > max(os.path.getmtime(s), hdr_t, self_t)):
            return o
        cmd = [nvcc] + NVCC_FLAGS + ccbin + (["-Xptxas", "-v"] if verbose else [])
        subprocess.check_call(cmd + ["-c", "-o", o, s])
        return o

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                           "-Xcompiler", "-fPIC"] + ccbin + ["-o", SO] + objs)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
