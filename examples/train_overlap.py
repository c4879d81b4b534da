#!/usr/bin/env python3
"""Does ingest stall training?  A small convolutional discriminator (the shape of
examples/densityopt/densityopt.py:139-190 in the reference) is trained on 640x480 RGBA frames from
five synthetic publishers, fed three ways:

  resident         the same device batch every step: what the model can do when data is free
  device_loader    btt.DeviceLoader: pinned ring -> H2D -> K1 (flip, gamma, normalise, NCHW) on the GPU
  reference_style  the reference's script structure (densityopt.py:210-268, generate.py:10-18):
                   RemoteIterableDataset(item_transform = numpy flip + gamma + normalise + CHW)
                   + torch DataLoader(num_workers) + batch.to(device)

Prints one JSON line per arm with steps/s and images/s.  No Blender needed.

    PYTHONPATH=pytorch-blender_b200 python examples/train_overlap.py --steps 300 --workers 8
"""
import argparse
import json
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import torch.utils.data as tud  # noqa: E402

from blendtorch import btt  # noqa: E402

BATCH, H, W = 8, 480, 640
GAMMA = 2.2


def numpy_transform(item):
    """What a reference script does per item on the CPU (offscreen.py:95-112, densityopt.py:117-119)."""
    img = np.flipud(item["image"])
    rgb = np.uint8(255.0 * (img[..., :3].astype(np.float32) / 255) ** (1 / GAMMA))
    img = np.concatenate((rgb, img[..., 3:4]), axis=-1)
    item["image"] = ((img.astype(np.float32) - 127.5) / 127.5).transpose(2, 0, 1)
    return item


def discriminator():
    def block(i, o):
        return [nn.Conv2d(i, o, 4, 2, 1, bias=False), nn.BatchNorm2d(o), nn.LeakyReLU(0.2, inplace=True)]
    return nn.Sequential(nn.Conv2d(4, 32, 4, 2, 1, bias=False), nn.LeakyReLU(0.2, inplace=True),
                         *block(32, 64), *block(64, 128), *block(128, 256), *block(256, 512),
                         nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(512, 1))


def spawn(n, tag):
    import socket
    addrs, procs = [], []
    for i in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
        s.close()
        addr = "tcp://127.0.0.1:%d" % port
        procs.append(subprocess.Popen([sys.executable, "-m", "blendtorch.btt.synthetic", "--bind", addr, "--btid", str(i),
                                       "--config", "2", "--height", str(H), "--width", str(W)],
                                      env={**__import__("os").environ, "PYTHONPATH": str(ROOT / "pytorch-blender_b200")},
                                      stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        addrs.append(addr)
    return addrs, procs


def train(model, opt, batches, steps, warmup, dev):
    loss_fn = nn.BCEWithLogitsLoss()
    target = torch.ones(BATCH, 1, device=dev)
    it = iter(batches)
    n = 0
    for step in range(steps + warmup):
        if step == warmup:
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
        x = next(it)
        if not x.is_cuda:
            x = x.to(dev)                      # the reference's pageable, synchronous copy (densityopt.py:261)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = loss_fn(model(x.contiguous(memory_format=torch.channels_last)).float(), target[:x.shape[0]])
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        if step >= warmup:
            n += x.shape[0]
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    return {"steps_per_s": round(steps / dt, 1), "images_per_s": round(n / dt, 1), "ms_per_step": round(dt / steps * 1e3, 3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workers", type=int, default=8, help="DataLoader workers of the reference-style arm")
    ap.add_argument("--receivers", type=int, default=4)
    ap.add_argument("--publishers", type=int, default=5)
    ap.add_argument("--arms", default="resident,device_loader,reference_style")
    ap.add_argument("--switch-interval", type=float, default=0.0,
                    help="sys.setswitchinterval for the run (seconds; 0 = leave Python's 5 ms): the loader's receiver "
                         "threads need the interpreter lock for ~0.1 ms per batch, the training loop holds it")
    a = ap.parse_args()
    if a.switch_interval > 0:
        sys.setswitchinterval(a.switch_interval)
    dev = torch.device("cuda", 0)
    torch.backends.cudnn.benchmark = True
    model = discriminator().to(dev).to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    total = a.steps + a.warmup
    for arm in a.arms.split(","):
        if arm == "resident":
            x = torch.randn(BATCH, 4, H, W, device=dev)

            def same():
                while True:
                    yield x
            res = train(model, opt, same(), a.steps, a.warmup, dev)
        else:
            addrs, procs = spawn(a.publishers, arm)
            try:
                if arm == "device_loader":
                    per = -(-total // a.receivers) * BATCH
                    ds = btt.RemoteIterableDataset(addrs, max_items=per * a.receivers, timeoutms=60000)
                    dl = btt.DeviceLoader(ds, batch_size=BATCH, device=dev, num_receivers=a.receivers,
                                          preprocess=btt.Preprocess(flip=True, gamma=GAMMA, mean=127.5, std=127.5))
                    res = train(model, opt, (b["image"] for b in dl), a.steps, a.warmup, dev)
                    res["h2d_GB_per_s"] = round(res["images_per_s"] * H * W * 4 / 1e9, 2)
                    # who waited for whom: the training loop for data, or the receivers for room in the queue
                    res["consumer_wait_s"] = round(dl.stats["consumer_wait_s"], 3)
                    res["producer_wait_s"] = round(dl.stats["producer_wait_s"], 3)
                else:
                    w = max(a.workers, 1)
                    per = (-(-total // w) + 2) * BATCH
                    ds = btt.RemoteIterableDataset(addrs, max_items=per * w, timeoutms=60000, item_transform=numpy_transform)
                    dl = tud.DataLoader(ds, batch_size=BATCH, num_workers=a.workers, prefetch_factor=4 if a.workers else None)
                    res = train(model, opt, (b["image"] for b in dl), a.steps, a.warmup, dev)
                    res["workers"] = a.workers
            finally:
                for p in procs:
                    p.terminate()
                for p in procs:
                    p.wait(timeout=10)
        print(json.dumps({"arm": arm, **res}), flush=True)


if __name__ == "__main__":
    main()
