#!/usr/bin/env python
"""Build libia_b200.so (the C-ABI shared library) in-tree with nvcc for sm_100a.

    python inference-arena_b200/build.py [--force] [--verbose]

The library is written to ``inference-arena_b200/lib/libia_b200.so`` (git-ignored; it
travels to the GPU box with the working tree).  nvcc cross-compiles without a GPU.
"""

from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "lib" / "libia_b200.so"
SOURCES = ["ia_preprocess.cu", "ia_decode_nms.cu", "ia_crop.cu", "ia_float_crops.cu", "ia_classify.cu", "ia_jpeg.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> list[Path]:
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def needs_build() -> bool:
    if not LIB.exists():
        return True
    newest = max(p.stat().st_mtime for p in list(CSRC.glob("*")) + [HERE.parent / "include" / "ia_b200.h"])
    return newest > LIB.stat().st_mtime


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    LIB.parent.mkdir(parents=True, exist_ok=True)
    cmd = [nvcc_path(), "-shared", "-Xcompiler", "-fPIC", "-O3", "-std=c++17", "-lineinfo",
           "-fmad=false", *ARCH, "-I", str(HERE.parent / "include"), "-o", str(LIB),
           *[str(s) for s in sources()], "-lcudart"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed")
    if verbose:
        sys.stderr.write(res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
