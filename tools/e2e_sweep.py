#!/usr/bin/env python3
"""Sweep receiver threads / ZMQ I/O threads / transport for the end-to-end leg (diagnostic)."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

import bench  # noqa: E402
from blendtorch import btt  # noqa: E402


class _S:
    active = False


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    res = []
    combos = [("tcp", 4, 1), ("tcp", 8, 1), ("tcp", 2, 5), ("tcp", 3, 5), ("tcp", 4, 3), ("ipc", 4, 1), ("ipc", 8, 1),
              ("ipc", 2, 5), ("ipc", 3, 5), ("ipc", 6, 2)]
    if len(sys.argv) > 1:
        combos = [tuple(int(x) if x.isdigit() else x for x in a.split(",")) for a in sys.argv[1:]]
    for combo in combos:
        transport, recv, io = combo[:3]
        shm_slots = combo[3] if len(combo) > 3 else 0
        bench.PUBLISHERS = combo[4] if len(combo) > 4 else 5
        rxmode = combo[5] if len(combo) > 5 else "auto"
        raw = len(combo) > 6 and combo[6] == "raw"
        images, sec, h2d, d2h = bench.e2e_leg(torch, btt, dev, 600, 10, _S(), transport, recv, io, shm_slots, rxmode,
                                              raw_parts=raw)
        r = {"transport": transport, "rx": rxmode, "receivers": recv, "io_threads": io, "shm_slots": shm_slots,
             "publishers": bench.PUBLISHERS, "raw_parts": raw, "images_per_s": images / sec,
             "GBps": images / sec * 1228800 / 1e9}
        print(json.dumps(r), flush=True)
        res.append(r)


if __name__ == "__main__":
    main()
