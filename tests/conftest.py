"""pytest configuration: markers and import paths.

The product package lives in `pytorch-blender_b200/` (not an importable name), which
holds a drop-in `blendtorch.btt`; tests put that directory on sys.path exactly as
a user of the reference puts `pkg_pytorch/` there.
"""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "pytorch-blender_b200"
for p in (str(ROOT), str(PKG)):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = Path(__file__).resolve().parent / "golden"


def _ensure_native_library():
    """The C-ABI library is built in-tree and is not in git history: build it (nvcc, a few seconds; a
    no-op when it is up to date) so that a fresh checkout can run the suite."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bt_build_native", PKG / "build_native.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build()


_ensure_native_library()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
