#!/usr/bin/env python3
"""Component ceilings of the host side of the ingest path on this box (diagnostic, not a bench):
raw ZMQ drain rate, scan+memcpy into pinned staging, pinned H2D.  Prints one JSON object."""
import argparse
import ctypes
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import numpy as np  # noqa: E402
import zmq  # noqa: E402

import bench  # noqa: E402


def drain(addrs, seconds, receivers, copy):
    counts = [0] * receivers
    stop = time.perf_counter() + seconds

    def run(r):
        ctx = zmq.Context()
        s = ctx.socket(zmq.PULL)
        s.setsockopt(zmq.RCVHWM, 10)
        for a in addrs:
            s.connect(a)
        while time.perf_counter() < stop:
            if s.poll(100):
                s.recv(copy=copy)
                counts[r] += 1
        s.close(0)
        ctx.term()

    ts = [threading.Thread(target=run, args=(r,)) for r in range(receivers)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    return sum(counts) / seconds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=3.0)
    a = ap.parse_args()
    out = {"cpus": os.cpu_count()}
    for transport in ("tcp", "ipc"):
        addrs, procs = bench.spawn_publishers("blendtorch.btt.synthetic", bench.PUBLISHERS, transport, tag="probe")
        try:
            time.sleep(1.0)
            for r in (1, 2, 4, 8):
                rate = drain(addrs, a.seconds, r, False)
                out["drain_%s_r%d_msgs_per_s" % (transport, r)] = rate
                out["drain_%s_r%d_GBps" % (transport, r)] = rate * 1228800 / 1e9
        finally:
            bench.stop_publishers(procs)
    # scan + memcpy into a pinned slot, per thread count
    import torch
    from blendtorch.btt import native as N, synthetic as S
    from blendtorch.btt.staging import Ring, Schema
    msgs = S.pickled_pool(2, 0, bench.H, bench.W, pool=16)
    sch = Schema.from_message(msgs[0], 8)
    for nthreads in (1, 2, 4, 8):
        rings = [Ring(sch.slot_bytes, 2, 0 if torch.cuda.is_available() else None) for _ in range(nthreads)]
        n_iter = 200

        def work(t):
            ring = rings[t]
            for it in range(n_iter):
                s = it & 1
                for i in range(8):
                    m = msgs[(it + i) % 16]
                    N.lib.bt_schema_ingest(sch.handle, ring.ptr(s), i, *N.buffer_address(m))

        ts = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
        t0 = time.perf_counter()
        [t.start() for t in ts]
        [t.join() for t in ts]
        dt = time.perf_counter() - t0
        out["ingest_t%d_imgs_per_s" % nthreads] = nthreads * n_iter * 8 / dt
        out["ingest_t%d_GBps" % nthreads] = nthreads * n_iter * 8 * 1228800 / dt / 1e9
        for r in rings:
            r.close()
    if torch.cuda.is_available():
        g = ctypes.c_double()
        for mb in (1, 10, 64, 256):
            N.check(N.lib.bt_measure_h2d(0, mb << 20, 20, ctypes.byref(g)))
            out["h2d_%dMB_GBps" % mb] = g.value
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
