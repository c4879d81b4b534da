#!/usr/bin/env python3
"""C5 (128x128 RGBA, batch 1024, 16 publishers): receivers x transport x receiver kind (diagnostic)."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200"), str(ROOT / "tools")]
import torch  # noqa: E402

import bench  # noqa: E402
from blendtorch import btt  # noqa: E402
from config_bench import timed  # noqa: E402

DEV = torch.device("cuda", 0)


def case(transport, receivers, rx, publishers=16, batch=1024, steps=80, h=128, w=128):
    bench.H, bench.W, bench.CONFIG, bench.PUBLISHERS = h, w, 5, publishers
    addrs, procs = bench.spawn_publishers("blendtorch.btt.synthetic", publishers, transport, tag="c5")
    try:
        per = -(-(steps + 3) // receivers) * batch
        ds = btt.RemoteIterableDataset(addrs, max_items=per * receivers, timeoutms=60000)
        dl = btt.DeviceLoader(ds, batch_size=batch, device=DEV, num_receivers=receivers, ring_depth=3, prefetch=3,
                              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5), transport=rx)
        n, dt = timed(dl, 3, steps)
        return {"transport": transport, "rx": rx, "receivers": receivers, "publishers": publishers,
                "images_per_s": round(n / dt), "GBps_h2d": round(n / dt * h * w * 4 / 1e9, 2)}
    finally:
        bench.stop_publishers(procs)


def main():
    torch.cuda.set_device(0)
    combos = [(t, r, x) for t in ("tcp", "ipc") for x in ("direct", "zmq") for r in (2, 4, 8)]
    if len(sys.argv) > 1:
        combos = [tuple(int(v) if v.isdigit() else v for v in a.split(",")) for a in sys.argv[1:]]
    for c in combos:
        print(json.dumps(case(*c)), flush=True)


if __name__ == "__main__":
    main()
