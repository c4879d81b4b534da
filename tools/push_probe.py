#!/usr/bin/env python3
"""Host-side rate of the stream path without a GPU in the loop (diagnostic): publishers ->
direct receivers (bt_pull_fill into an unpinned ring), for Python/pyzmq publisher processes against the
native publisher (bt_push_*), over tcp and ipc.  Prints one JSON line per case."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import bench  # noqa: E402
from blendtorch.btt import native as N  # noqa: E402
from blendtorch.btt.rx import DirectReceiver  # noqa: E402
from blendtorch.btt.staging import Ring, Schema  # noqa: E402


def drain(addrs, seconds, receivers, batch=8):
    counts = [0] * receivers
    t_end = [0.0]

    def run(r):
        rx = DirectReceiver(addrs)
        while not rx.recv(1000):
            pass
        sch = Schema.from_message(rx.pending(), batch)
        ring = Ring(sch.slot_bytes, 2, None)
        stop = time.perf_counter() + seconds
        while time.perf_counter() < stop:
            slot = ring.acquire()
            rc, n = rx.fill(ring, slot, sch, 0, batch, 100)
            counts[r] += n
        rx.close()

    ts = [threading.Thread(target=run, args=(r,)) for r in range(receivers)]
    t0 = time.perf_counter()
    [t.start() for t in ts]
    [t.join() for t in ts]
    return sum(counts) / (time.perf_counter() - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=2.0)
    ap.add_argument("--receivers", type=int, nargs="+", default=[2, 4, 8])
    ap.add_argument("--threads", type=int, nargs="+", default=[1, 2, 3, 5])
    ap.add_argument("--transports", nargs="+", default=["tcp", "ipc"])
    ap.add_argument("--python-publishers", action="store_true")
    a = ap.parse_args()
    nbytes = bench.H * bench.W * bench.C
    for transport in a.transports:
        if a.python_publishers:
            addrs, procs = bench.spawn_publishers("blendtorch.btt.synthetic", bench.PUBLISHERS, transport, tag="pp")
            try:
                time.sleep(1.5)
                for r in a.receivers:
                    rate = drain(addrs, a.seconds, r)
                    print(json.dumps({"publishers": "5 python processes", "transport": transport, "receivers": r,
                                      "images_per_s": rate, "GBps": rate * nbytes / 1e9, "cpus": os.cpu_count()}), flush=True)
            finally:
                bench.stop_publishers(procs)
        for th in a.threads:
            addrs, proc = bench.spawn_native_publisher(bench.PUBLISHERS, transport, threads=th, tag="np%d" % th)
            try:
                time.sleep(1.5)
                for r in a.receivers:
                    rate = drain(addrs, a.seconds, r)
                    print(json.dumps({"publishers": "native, %d threads" % th, "transport": transport, "receivers": r,
                                      "images_per_s": rate, "GBps": rate * nbytes / 1e9, "cpus": os.cpu_count()}), flush=True)
            finally:
                bench.stop_publishers([proc])


if __name__ == "__main__":
    main()
