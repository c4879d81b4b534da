#!/usr/bin/env python3
"""Build libbt_b200.so (C ABI + CUDA kernels) in-tree for sm_100a.

    python pytorch-blender_b200/build_native.py [--force] [--verbose]

nvcc cross-compiles without a GPU.  The library lands next to the Python host
layer (blendtorch/btt/_native/libbt_b200.so) so that it travels with the tree.
"""
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
CSRC = HERE / "csrc"
OUT_DIR = HERE / "blendtorch" / "btt" / "_native"
LIB = OUT_DIR / "libbt_b200.so"
SOURCES = ["scan.cpp", "rx.cpp", "zmtp.cpp", "push.cpp", "ring.cu", "shm.cu", "preprocess.cu", "project.cu", "raster.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; libbt_b200.so cannot be built")
    return cand


def _fingerprint():
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*")) + [ROOT / "include" / "bt_b200.h", Path(__file__)]):
        if f.is_file():
            h.update(f.name.encode())
            h.update(f.read_bytes())
    return h.hexdigest()


def build(force=False, verbose=False):
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    stamp = OUT_DIR / "libbt_b200.stamp"
    fp = _fingerprint()
    if not force and LIB.exists() and stamp.exists() and stamp.read_text() == fp:
        return LIB
    srcs = [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
    cmd = [_nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", "--threads", "4",
           "-Xcompiler", "-fPIC,-O3,-Wall", "-shared",
           "-I", str(ROOT / "include"), "-I", str(CSRC),
           "-o", str(LIB), *srcs, "-lpthread", "-lrt"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libbt_b200.so")
    if verbose:
        print(res.stdout + res.stderr)
    stamp.write_text(fp)
    return LIB


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(p)
