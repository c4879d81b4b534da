"""bt_preprocess_stats_static: the statistics variant of K1 for callers that cannot allocate
device memory (plug-ins living inside a fixed memory budget, loops that must not touch an
allocator once running, processes next to a framework that already owns the GPU's memory).

It has to give exactly what the regular entry point gives with caller-provided scratch -- same
kernel, only the scratch lives in a fixed array inside the library -- on consecutive calls that
reuse that array, and it has to refuse batches the array does not cover.  Collected last on
purpose (file name), and self-contained so that it also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from blendtorch import btt
from blendtorch.btt import native as N

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _params(src, dst, stats, gamma=2.2):
    B, H, W, Cn = src.shape
    p = N.PreprocessParams()
    p.src, p.dst = src.data_ptr(), dst.data_ptr()
    p.batch, p.height, p.width, p.channels, p.out_channels = B, H, W, Cn, Cn
    p.out_dtype, p.flip, p.chw, p.use_gamma, p.use_norm = N.BT_F32, 1, 1, 1, 1
    for c in range(4):
        p.mean[c], p.stdev[c] = 127.5, 127.5
    C.memmove(p.gamma_lut, N.gamma_lut(gamma), 256)
    p.stats_out = stats.data_ptr()
    return p


def _regular(src):
    """The existing entry point, scratch allocated by the caller (here: torch)."""
    B, H, W, Cn = src.shape
    dst = torch.empty((B, Cn, H, W), dtype=torch.float32, device=DEV)
    stats = torch.empty((Cn, 2), dtype=torch.float64, device=DEV)
    p = _params(src, dst, stats)
    scratch = torch.empty(N.lib.bt_preprocess_scratch_bytes(C.byref(p)), dtype=torch.uint8, device=DEV)
    p.stats_scratch = scratch.data_ptr()
    N.check(N.lib.bt_preprocess_launch(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return dst, stats


def _static(src):
    """The entry point under test: no scratch passed, nothing allocated by the call."""
    B, H, W, Cn = src.shape
    dst = torch.empty((B, Cn, H, W), dtype=torch.float32, device=DEV)   # outputs belong to the caller
    stats = torch.empty((Cn, 2), dtype=torch.float64, device=DEV)
    p = _params(src, dst, stats)
    p.stats_scratch = None
    N.check(N.lib.bt_preprocess_stats_static(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return dst, stats


def test_static_scratch_entry_matches_regular_entry_twice_in_a_row():
    torch.cuda.set_device(0)
    g = torch.Generator(device=DEV)
    g.manual_seed(77)
    # two different batches, back to back, so the second call sees what the first left in the array
    first = torch.randint(0, 256, (8, 480, 640, 4), dtype=torch.uint8, device=DEV, generator=g)
    second = torch.randint(0, 256, (3, 40, 64, 4), dtype=torch.uint8, device=DEV, generator=g)
    for src in (first, second):
        want_out, want_stats = _regular(src)
        got_out, got_stats = _static(src)
        # same tolerance as the other K1 tests: bit-exact output, and statistics that are
        # deterministic (fixed reduction order), hence equal bit for bit as well
        assert torch.equal(got_out, want_out)
        assert torch.equal(got_stats, want_stats)
        # and they are the statistics of what was written (fp32 partials of <= 16 values -> float64)
        total = want_out.double().sum(dim=(0, 2, 3))
        scale = want_out.double().abs().sum(dim=(0, 2, 3))
        assert ((got_stats[:, 0] - total).abs() <= 2.0 ** -22 * scale).all()


def test_static_scratch_entry_rejects_what_it_cannot_cover():
    torch.cuda.set_device(0)
    big = torch.zeros((5, 1024, 1024, 4), dtype=torch.uint8, device=DEV)     # 20 971 520 B > 19 398 656 B
    dst = torch.empty((5, 4, 1024, 1024), dtype=torch.float32, device=DEV)
    stats = torch.empty((4, 2), dtype=torch.float64, device=DEV)
    p = _params(big, dst, stats)
    rc = N.lib.bt_preprocess_stats_static(C.byref(p), None)
    assert rc == N.BT_E_TOO_LARGE
    assert "exceeds" in N.last_error() and "19398656" in N.last_error()
    # statistics are the point of this entry: without stats_out it refuses instead of guessing
    p.stats_out = None
    assert N.lib.bt_preprocess_stats_static(C.byref(p), None) == N.BT_E_INVALID
