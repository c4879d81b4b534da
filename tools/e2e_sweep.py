#!/usr/bin/env python3
"""End-to-end rate of the headline stream leg for chosen core sets and thread splits (diagnostic):
    python tools/e2e_sweep.py --cpus 0-7 --split 2x4 3x4 --transport tcp ipc [--python-publishers]
Prints one JSON line per case."""
import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import bench  # noqa: E402


def parse_cpus(text):
    out = []
    for part in text.split(","):
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpus", default="")
    ap.add_argument("--split", nargs="+", default=["5x8"], help="SENDERSxRECEIVERS")
    ap.add_argument("--transport", nargs="+", default=["tcp"])
    ap.add_argument("--seconds", type=float, default=1.5)
    ap.add_argument("--shm-slots", type=int, default=0)
    ap.add_argument("--python-publishers", action="store_true")
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--in-place", action="store_true")
    ap.add_argument("--replay", nargs="*", default=None, metavar="THREADSxPRODUCERS",
                    help="replay legs (C1 recording) instead of stream legs, e.g. 4x2 8x2; add --cache-host")
    ap.add_argument("--cache-host", action="store_true")
    ap.add_argument("--group", type=int, nargs="+", default=[0], help="DeviceLoader(group=...), 0 = automatic")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    cpus = parse_cpus(a.cpus) if a.cpus else sorted(os.sched_getaffinity(0))
    os.sched_setaffinity(0, cpus)
    import numpy as np
    import torch
    from blendtorch import btt
    from blendtorch.btt import native as N, synthetic as S
    torch.set_num_threads(1)
    dev = torch.device("cuda", a.gpu)
    torch.cuda.set_device(a.gpu)
    ctx = {"torch": torch, "btt": btt, "np": np, "S": S, "N": N, "dev": dev, "rank": a.gpu, "local": a.gpu, "world": 1,
           "cpus": cpus, "steps": 20, "warmup": 5, "sampler": None, "barrier": lambda: torch.cuda.synchronize(dev)}
    c2 = {"name": "C2", "config": 2, "h": bench.H, "w": bench.W, "batch": bench.BATCH, "publishers": bench.PUBLISHERS}
    if a.replay is not None:
        import shutil
        import tempfile
        tmp = tempfile.mkdtemp(prefix="btrec_sw_", dir="/dev/shm")
        try:
            S.write_recordings(tmp + "/c1", 1, 2, 512, 480, 640, pool=16)
            c1 = {"name": "C1", "config": 1, "h": 480, "w": 640, "batch": 8, "pool": 16}
            for sp in a.replay or ["4x2"]:
                th, pr = (int(x) for x in sp.split("x"))
                for grp in a.group:
                    bench.GROUP = grp or None
                    leg = bench.replay_leg(ctx, c1, tmp + "/c1", cache="host" if a.cache_host else None, min_seconds=a.seconds,
                                           threads=th, producers=pr)
                    print(json.dumps({"replay": True, "cache": "host" if a.cache_host else None, "ingest_threads": th,
                                      "producers": pr, "group": grp, "images_per_s": leg["images"] / leg["sec"],
                                      "GBps": leg["images"] / leg["sec"] * 1228800 / 1e9, "check": leg["check"],
                                      "consumer_wait_s": leg["consumer_wait_s"]}), flush=True)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)
        return
    for tr in a.transport:
        for sp, grp in ((x, y) for x in a.split for y in a.group):
            s, r = (int(x) for x in sp.split("x"))
            bench.GROUP = grp or None
            if a.python_publishers:
                orig = bench.spawn_native_publisher

                def py(n, transport="tcp", threads=1, tag="n", extra=(), cpus=None, config=None, h=None, w=None):
                    addrs, procs = bench.spawn_publishers("blendtorch.btt.synthetic", n, transport, tag=tag, extra=extra)
                    for p in procs:
                        os.sched_setaffinity(p.pid, cpus)

                    class Group:
                        pid = procs[0].pid

                        def send_signal(self, sig):
                            [p.send_signal(sig) for p in procs]

                        def wait(self, timeout=None):
                            [p.wait(timeout) for p in procs]

                        def kill(self):
                            [p.kill() for p in procs]
                    return addrs, Group()
                bench.spawn_native_publisher = py
            t0 = time.time()
            leg = bench.stream_leg(ctx, c2, tr, r, s, shm_slots=a.shm_slots, min_seconds=a.seconds, in_place=a.in_place)
            if a.python_publishers:
                bench.spawn_native_publisher = orig
            print(json.dumps({"tag": a.tag, "gpu": a.gpu, "cpus": len(cpus), "transport": tr, "senders": s if not a.python_publishers else "5 python procs",
                              "receivers": r, "shm_slots": a.shm_slots, "in_place": a.in_place, "group": grp,
                              "images_per_s": leg["images"] / leg["sec"], "GBps": leg["images"] / leg["sec"] * 1228800 / 1e9,
                              "consumer_wait_s": leg["consumer_wait_s"], "check": leg["check"], "leg_s": time.time() - t0}), flush=True)


if __name__ == "__main__":
    main()
