"""The C-ABI library loads and exports every symbol include/bt_b200.h declares; the product never
imports the oracle; there is no CPU route around the kernels."""
import ctypes
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "pytorch-blender_b200"


def _declared():
    text = (ROOT / "include" / "bt_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from blendtorch.btt import native as N
    lib = ctypes.CDLL(str(N.LIB_PATH))
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), "libbt_b200.so lacks %s" % n
    assert sorted(N.EXPORTS) == sorted(x for x in names if x in N.EXPORTS)
    assert set(names) <= set(N.EXPORTS)
    assert N.lib.bt_abi_version() == 3
    out = subprocess.run(["nm", "-D", "--defined-only", str(N.LIB_PATH)], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (bt_[a-z0-9_]+)", out))
    assert set(names) <= exported


def test_struct_sizes_match_header(tmp_path):
    """ctypes mirrors vs. the C compiler's view of the header."""
    from blendtorch.btt import native as N
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "bt_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(bt_field),sizeof(bt_column),sizeof(bt_preprocess_params),sizeof(bt_project_params),'
                   'sizeof(bt_raster_params),sizeof(bt_zmq_api));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    from blendtorch.btt.raster import RasterParams
    from blendtorch.btt.rx import ZmqApi
    assert sizes == [ctypes.sizeof(N.Field), ctypes.sizeof(N.Column),
                     ctypes.sizeof(N.PreprocessParams), ctypes.sizeof(N.ProjectParams),
                     ctypes.sizeof(RasterParams), ctypes.sizeof(ZmqApi)]


def test_error_reporting():
    from blendtorch.btt import native as N
    assert N.lib.bt_preprocess_launch(None, None) == N.BT_E_INVALID
    assert "null" in N.last_error()
    p = N.PreprocessParams()
    p.src, p.dst = 16, 16
    p.batch = p.height = p.width = 1
    p.channels = 7
    assert N.lib.bt_preprocess_launch(ctypes.byref(p), None) == N.BT_E_INVALID
    assert "channels" in N.last_error()
    with pytest.raises(N.NativeError):
        N.gamma_lut(0.0)


def test_product_never_imports_oracle_and_has_no_cpu_kernels():
    for f in list(PKG.rglob("*.py")) + list(PKG.rglob("*.cu")) + list(PKG.rglob("*.cpp")) + list(PKG.rglob("*.h")):
        text = f.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
        assert "/root/reference" not in text, f
    # importing the product in a clean interpreter must not pull the oracle in
    code = ("import sys; sys.path.insert(0, %r); from blendtorch import btt; "
            "assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules)") % str(PKG)
    subprocess.run([sys.executable, "-c", code], check=True)


def test_device_paths_fail_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from blendtorch import btt
    with pytest.raises(RuntimeError):
        btt.preprocess_images(torch.zeros(1, 4, 4, 4, dtype=torch.uint8), btt.Preprocess())
    with pytest.raises(RuntimeError):
        btt.Camera(np.eye(4), np.eye(4), (4, 4), device="cpu")


def test_only_tests_smoke_and_bench_touch_the_oracle():
    """Diagnostic tools and examples use the product; the oracle is for tests, smoke() and bench.py's CPU legs."""
    for folder in ("tools", "examples"):
        for f in sorted((ROOT / folder).glob("*.py")):
            assert not re.search(r"^\s*(from|import)\s+oracle\b", f.read_text(), flags=re.M), f


def test_synthetic_generators_agree_with_oracle_copy():
    from blendtorch.btt import synthetic as S
    from oracle import synth as O
    for cfg in (1, 2, 3):
        a = S.message(S.rng_for(cfg, 4), cfg, 4, 9, 12, 16)
        b = O.message(O.rng_for(cfg, 4), cfg, 4, 9, 12, 16)
        assert a.keys() == b.keys()
        for k in a:
            np.testing.assert_array_equal(np.asarray(a[k]), np.asarray(b[k]))
    np.testing.assert_array_equal(S.gradient_frame(5, 7), O.gradient_frame(5, 7))


def _build_c_demo(out):
    from blendtorch.btt import native as N
    cuda = Path("/usr/local/cuda")
    cmd = ["gcc", "-O2", "-Wall", "-Werror", "-std=c99", "-I", str(ROOT / "include"), "-I", str(cuda / "include"),
           str(ROOT / "examples" / "c_abi_demo.c"), "-L", str(N._LIB_PATH.parent), "-lbt_b200",
           "-L", str(cuda / "lib64"), "-lcudart", "-lm", "-Wl,-rpath," + str(N._LIB_PATH.parent), "-o", str(out)]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return out


def test_header_is_plain_c_and_a_c_program_links(tmp_path):
    """include/bt_b200.h must be usable without C++, torch or Python: examples/c_abi_demo.c is compiled as
    C99 with -Werror and linked against the library (it is run on the GPU box by the gpu suite)."""
    exe = _build_c_demo(tmp_path / "c_abi_demo")
    assert exe.exists()
