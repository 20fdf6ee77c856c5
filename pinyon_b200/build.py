"""In-tree build of libpinyon_cuda.so for sm_100a (and only sm_100a).

`python -m pinyon_b200.build` or `pinyon_b200.build.build_library()`.  The .so lands
next to this file so it travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "libpinyon_cuda.so"
SOURCES = [CSRC / "pinyon_cuda.cu", CSRC / "mt_jump.cu", CSRC / "mt_poly.cpp"]
HEADERS = [CSRC / "kernels.cuh", CSRC / "stream.cuh", CSRC / "lane.cuh", CSRC / "seed.cuh", CSRC / "leaf.cuh", CSRC / "thread.cuh", CSRC / "warp2.cuh", CSRC / "mt19937.cuh", CSRC / "mt_jump.h", ROOT / "include" / "pinyon_cuda.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    # exact IEEE double arithmetic for the chance-sampling path: no fast math, no contraction
    "-fmad=false",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libpinyon_cuda cannot be built")


def is_stale() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    return any(p.stat().st_mtime > t for p in SOURCES + HEADERS)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    if not force and not is_stale():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, *map(str, SOURCES), "-o", str(LIB)]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        print(res.stdout)
        print(res.stderr, file=sys.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
