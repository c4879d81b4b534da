#!/usr/bin/env python3
"""Host-side cost of a DeviceLoader batch (diagnostic): cProfile of the producer's `finish` and of the
consumer's per-batch bookkeeping on a replayed recording small enough that copies and kernels are free."""
import cProfile
import pstats
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402
from blendtorch.btt import loader as L  # noqa: E402
from blendtorch.btt import synthetic as S  # noqa: E402

d = tempfile.mkdtemp(dir="/dev/shm")
S.write_recordings(d + "/r", 2, 1, 512, 48, 64, pool=16)
ds = btt.FileDataset(d + "/r")
dl = btt.DeviceLoader(ds, batch_size=8, device="cuda:0", num_receivers=1,
                      preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
for b in dl:
    pass
prof = cProfile.Profile()
orig = L._Stager.finish


def timed(self, *a, **k):
    prof.enable()
    try:
        return orig(self, *a, **k)
    finally:
        prof.disable()


L._Stager.finish = timed
t0 = time.perf_counter()
n = 0
for b in dl:
    n += 1
dt = time.perf_counter() - t0
print("batches", n, "us per batch (whole loop)", round(dt / n * 1e6, 1))
pstats.Stats(prof).sort_stats("cumulative").print_stats(14)
