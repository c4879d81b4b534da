#!/usr/bin/env python3
"""End-to-end numbers for the BASELINE.json configurations that are not the headline bench line
(diagnostic / report tool): C1 replay, C3 1080p + depth + projection, C4 1024^2 batch 64, C5 small
frames.  Prints one JSON object per case."""
import json
import os
import shutil
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from blendtorch import btt  # noqa: E402
from blendtorch.btt import synthetic as S  # noqa: E402

DEV = torch.device("cuda", 0)


def timed(dl, warm, steps, after=None):
    it = iter(dl)
    for _ in range(warm):
        b = next(it)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 0
    for _ in range(steps):
        try:
            b = next(it)
        except StopIteration:
            break
        if after:
            after(b)
        n += b["image"].shape[0]
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    del it
    return n, dt


def replay_case(name, config, h, w, batch, files, frames, threads, steps, spec, extra=None, shuffle=True, producers=1):
    d = tempfile.mkdtemp(prefix="btrec_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        prefix = d + "/rec"
        S.write_recordings(prefix, config, files, frames, h, w, pool=min(frames, 32))
        ds = btt.FileDataset(prefix)
        per_item = os.path.getsize(ds.filenames[0]) / frames
        dl = btt.DeviceLoader(ds, batch_size=batch, shuffle=shuffle, device=DEV, preprocess=spec, ingest_threads=threads,
                              ring_depth=3, prefetch=3, drop_last=True, num_receivers=producers)
        epochs = max(1, -(-(steps + 3) * batch // len(ds)))

        class Multi:
            def __iter__(self_):
                for _ in range(epochs):
                    yield from dl
        n, dt = timed(Multi(), 3, steps, extra)
        return {"case": name, "images_per_s": n / dt, "GBps_h2d": n / dt * per_item / 1e9, "batch": batch,
                "ingest_threads": threads, "producers": producers, "frames": [h, w]}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def stream_case(name, config, h, w, batch, publishers, receivers, steps, spec, shm_slots=0):
    bench.H, bench.W, bench.CONFIG, bench.PUBLISHERS = h, w, config, publishers
    addrs, procs = bench.spawn_publishers("blendtorch.btt.synthetic", publishers, "tcp", tag=name,
                                          extra=("--shm-slots", str(shm_slots)) if shm_slots else ())
    try:
        per = -(-(steps + 3) // receivers) * batch
        ds = btt.RemoteIterableDataset(addrs, max_items=per * receivers, timeoutms=60000)
        dl = btt.DeviceLoader(ds, batch_size=batch, device=DEV, preprocess=spec, num_receivers=receivers,
                              ring_depth=3, prefetch=3)
        n, dt = timed(dl, 3, steps)
        return {"case": name, "images_per_s": n / dt, "GBps_h2d": n / dt * h * w * 4 / 1e9, "batch": batch,
                "publishers": publishers, "receivers": receivers, "shm_slots": shm_slots, "frames": [h, w]}
    finally:
        bench.stop_publishers(procs)


def main():
    torch.cuda.set_device(0)
    spec = btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5)
    out = []
    for th, pr in ((1, 1), (4, 1), (4, 2), (4, 3), (2, 4)):
        out.append(replay_case("C1 replay 640x480 b8 (.btr in page cache)", 1, 480, 640, 8, 4, 256, th, 400, spec, producers=pr))
    shape = (1080, 1920)
    cam = btt.Camera.from_lens(shape, lens=35.0, distance=0.0, device=DEV)

    def project(b):
        b["xy"], b["boxes"] = cam.bbox_world_to_pixel_boxes(b["corners"], 8)
    out.append(replay_case("C3 replay 1920x1080 + depth f32 + 64 corners, b32, + K2 boxes", 3, 1080, 1920, 32, 2, 64, 8, 40,
                           spec, extra=project))
    out.append(replay_case("C4 replay 1024x1024 b64", 4, 1024, 1024, 64, 2, 128, 8, 40, spec))
    out.append(stream_case("C4 stream 1024x1024 b64, 5 publishers", 4, 1024, 1024, 64, 5, 4, 40, spec))
    out.append(stream_case("C5 stream 128x128 b1024, 16 publishers", 5, 128, 128, 1024, 16, 4, 60, spec))
    out.append(stream_case("C5 stream 128x128 b1024, 16 publishers, shm wire", 5, 128, 128, 1024, 16, 2, 60, spec, shm_slots=256))
    for r in out:
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
