#!/usr/bin/env python3
"""Recipe for oracle/_ref (TEST / BASELINE INFRASTRUCTURE, never imported by the product).

The reference is pure Python: there is nothing to compile.  What `bench.py --impl reference` times is
the reference's own consumer code -- `blendtorch.btt.RemoteIterableDataset` and what it imports -- so
this recipe copies the unmodified package `pkg_pytorch/blendtorch` from the read-only reference tree
into `oracle/_ref/blendtorch` (git-ignored build output, not source of this repository; it travels to
the GPU box with the snapshot, where /root/reference does not exist).  Run by `__graft_entry__.build()`
whenever the reference tree is present; a no-op otherwise.

    python oracle/make_ref.py [/root/reference]
"""
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"


def make(reference_root="/root/reference"):
    src = Path(reference_root) / "pkg_pytorch" / "blendtorch"
    if not src.is_dir():
        return None
    dst = OUT / "blendtorch"
    if dst.exists():
        shutil.rmtree(dst)
    OUT.mkdir(parents=True, exist_ok=True)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    (OUT / "README").write_text("Copied unmodified from %s by oracle/make_ref.py; not part of this repository's sources.\n" % src)
    return dst


if __name__ == "__main__":
    print(make(*sys.argv[1:2]))
