#!/usr/bin/env python3
"""Replay of a recording that stays resident in HBM (DeviceLoader(cache='device')): epoch 1 goes
through the pinned pipeline (PCIe-bound), later epochs are gathered from the device by K1.
Prints one JSON object per case; diagnostic / report tool."""
import json
import os
import shutil
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402
from blendtorch.btt import synthetic as S  # noqa: E402

DEV = torch.device("cuda", 0)
HBM_PEAK = 6548.5
try:
    HBM_PEAK = float(json.load(open(ROOT / "MEASURED_PEAKS.json"))["hbm_gbs"])
except Exception:
    pass


def epoch(dl):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    n = 0
    for b in dl:
        n += b["image"].shape[0]
    e1.record()
    torch.cuda.synchronize()
    return n, time.perf_counter() - t0, e0.elapsed_time(e1) / 1e3


def case(name, config, h, w, batch, files, frames, spec, epochs=4):
    d = tempfile.mkdtemp(prefix="btrec_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        prefix = d + "/rec"
        S.write_recordings(prefix, config, files, frames, h, w, pool=min(frames, 32))
        ds = btt.FileDataset(prefix)
        out = {"case": name, "frames": [h, w], "batch": batch, "items": len(ds)}
        for mode in (None, "host", "device"):
            dl = btt.DeviceLoader(ds, batch_size=batch, shuffle=True, device=DEV, preprocess=spec, ingest_threads=4,
                                  ring_depth=3, prefetch=3, drop_last=False, cache=mode)
            rates = []
            for e in range(epochs):
                n, wall, dev = epoch(dl)
                rates.append({"images_per_s": round(n / wall, 1), "device_s": round(dev, 5), "wall_s": round(wall, 5)})
            key = {None: "uncached", "host": "pinned_host", "device": "resident"}[mode]
            out[key] = rates
            if mode == "host":
                out["host_cache_fill_s"] = round(dl.stats.get("cache_fill_s", 0), 3)
                best = max(r["images_per_s"] for r in rates[1:])
                out["pinned_host_GBps_h2d"] = round(best * os.path.getsize(ds.filenames[0]) / frames / 1e9, 1)
            if mode == "device":
                out["cache_bytes"] = dl.stats["cache_bytes"]
                out["resident_batches"] = dl.stats["resident_batches"]
                px_bytes = 4 + 4 * (4 if spec.dtype == torch.float32 else 2)
                best = max(r["images_per_s"] for r in rates[1:])
                out["resident_k1_GBps"] = round(best * h * w * px_bytes / 1e9, 1)
                out["resident_frac_of_hbm_peak"] = round(out["resident_k1_GBps"] / HBM_PEAK, 3)
            dl.close()
            del dl
            torch.cuda.empty_cache()
        return out
    finally:
        shutil.rmtree(d, ignore_errors=True)


def main():
    torch.cuda.set_device(0)
    spec = btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5)
    for args in (("C1 replay 640x480 b8, 1024 frames", 1, 480, 640, 8, 4, 256),
                 ("C1 replay 640x480 b64, 1024 frames", 1, 480, 640, 64, 4, 256),
                 ("C4 replay 1024x1024 b64, 512 frames", 4, 1024, 1024, 64, 2, 256),
                 ("C3 replay 1920x1080 + depth b32, 128 frames", 3, 1080, 1920, 32, 2, 64)):
        print(json.dumps(case(*args, spec)), flush=True)


if __name__ == "__main__":
    main()
