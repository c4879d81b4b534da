#!/usr/bin/env python3
"""Benchmark of the blendtorch ingest hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1], the configuration the metric is quoted on):
RemoteIterableDataset fed by 5 synthetic ZMQ publishers, 640x480 RGBA + bbox annotations,
batch 8, fused flip + gamma 2.2 + normalise + NHWC->NCHW to fp32 on one B200.
A step is one pass of the hot path over one batch of 8 frames.

Printed JSON line (rank 0):
  value      images/s of the device leg: the fused kernel K1 over batches already resident in HBM
             (rotating over distinct batches so the working set exceeds L2), replayed from CUDA graphs
             only, timed with CUDA events over a window of >= 20 ms, max over ranks.
  e2e        images/s through the public API (btt.DeviceLoader over btt.RemoteIterableDataset) with HOST
             buffers: socket receive, framing, pinned staging, H2D copy, K1 and a D2H read of the batch
             statistics are all inside the timed region; after warm-up the batches that were already
             produced are discarded and at least `steps` batches AND 2 s are timed.  A sampled batch is
             checked against the publishers' frames outside the timed region (e2e_check).
  roofline   K1 algorithmic bytes per launch (H*W*C*(1 + 4) per image = 20 B/pixel)
             / average launch duration, against the measured HBM copy peak; `single_launch` is the same
             kernel on one stream, `e2e_kernel` the statistics variant + fold the e2e path runs.
  configs    the other BASELINE.json shapes end to end (C1 replay, C3 1080p + depth + projection,
             C4 1024^2 batch 64, C5 128^2 batch 1024), with their kernel times.
  cpu_baseline  the reference's CPU path timed on this box's host cores (see --impl reference).
With --impl reference only the reference's CPU path is run: its own btt.RemoteIterableDataset (the
unmodified files, copied to oracle/_ref by build(); the oracle port if they are absent) + the example
item_transforms + default_collate under torch DataLoader workers on every host core.
"""
import argparse
import json
import os
import signal
import socket
import statistics
import shutil
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "pytorch-blender_b200"
REF = ROOT / "oracle" / "_ref"

H, W, C, BATCH, PUBLISHERS, CONFIG = 480, 640, 4, 8, 5, 2
GAMMA, MEAN, STD = 2.2, 127.5, 127.5
WORKLOAD = "C2: 5 synthetic ZMQ publishers, 640x480 RGBA uint8 + bbox, batch 8, flip+gamma2.2+normalise+NCHW fp32"
METRIC = "images/sec into device tensors (640x480 RGBA, batch 8)"
CONFIG_KEYS = {"workload": WORKLOAD, "batch": BATCH, "publishers": PUBLISHERS}   # identical in both arms
DEVICE_STREAMS = 4  # independent batches in flight in the device leg (one stream per receiver in the pipeline)
ROTATE = 8  # distinct resident batches: 8 x (9.8 MB in + 39.3 MB out) = 393 MB > 126 MB L2


def _use_product():
    for p in (str(ROOT), str(PKG)):
        if p not in sys.path:
            sys.path.insert(0, p)


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def spawn_publishers(module, n, transport="tcp", tag="p", extra=()):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(ROOT), str(PKG), env.get("PYTHONPATH", "")])
    addrs, procs = [], []
    for i in range(n):
        if transport == "ipc":
            addr = "ipc:///tmp/bt_bench_%d_%s_%d" % (os.getpid(), tag, i)
        else:
            addr = "tcp://127.0.0.1:%d" % free_port()
        cmd = [sys.executable, "-m", module, "--bind", addr, "--btid", str(i), "--config", str(CONFIG),
               "--height", str(H), "--width", str(W), *extra]
        procs.append(subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        addrs.append(addr)
    return addrs, procs


def spawn_native_publisher(n, transport="tcp", threads=1, tag="n", extra=(), cpus=None, config=None, h=None, w=None):
    """One process hosting `n` publishers on the library's own PUSH endpoints (bt_push_*), `threads`
    sender threads.  Returns (addresses, process)."""
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([str(ROOT), str(PKG), env.get("PYTHONPATH", "")])
    addrs = []
    for i in range(n):
        if transport == "ipc":
            addrs.append("ipc:///tmp/bt_bench_%d_%s_%d" % (os.getpid(), tag, i))
        else:
            addrs.append("tcp://127.0.0.1:%d" % free_port())
    cmd = [sys.executable, "-m", "blendtorch.btt.synthetic", "--native", "--threads", str(threads), "--btid", "0",
           "--config", str(CONFIG if config is None else config), "--height", str(h or H), "--width", str(w or W), *extra]
    for a in addrs:
        cmd += ["--bind", a]
    if cpus:
        cmd += ["--cpus", ",".join(str(c) for c in cpus)]
    proc = subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return addrs, proc


def stop_publishers(procs):
    for p in procs:
        try:
            p.send_signal(signal.SIGTERM)
        except Exception:
            pass
    for p in procs:
        try:
            p.wait(timeout=5)
        except Exception:
            p.kill()
    import glob
    for p in procs:  # shared-memory segments of publishers that did not get to unlink them
        for f in glob.glob("/dev/shm/btshm_%d_*" % p.pid):
            try:
                os.unlink(f)
            except OSError:
                pass


class ClockSampler:
    """Polls NVML for SM clocks and throttle reasons while a region is marked active."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        self.samples, self.active, self._stop = [], False, False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)
        if self.nv:
            self.t.start()

    def _run(self):
        nv = self.nv
        while not self._stop:
            if self.active:
                try:
                    mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                    try:
                        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                    except Exception:
                        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    self.samples.append((mhz, r))
                except Exception:
                    pass
            time.sleep(0.002)

    def report(self):
        self._stop = True
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        reasons = set()
        for _, r in self.samples:
            for bit, name in self.REASONS.items():
                if r & bit:
                    reasons.add(name)
        return {"sm_mhz": statistics.median(m for m, _ in self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def measured_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the headline kernel, from the committed
    ncu --set full capture (profiles/); None if no capture is there."""
    for name in ("r2_k1_traffic.json", "r1_k1_traffic.json"):
        try:
            return float(json.loads((ROOT / "profiles" / name).read_text())["dram_bytes_read_plus_write_per_launch"])
        except Exception:
            continue
    return None


def measured_peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def rank_cpus(local_rank, world):
    """The host cores of this rank: the process's allowed cpus split into `world` equal, disjoint runs
    (one process per GPU, each with its own receivers, publishers and consumer thread on its own cores)."""
    cpus = sorted(os.sched_getaffinity(0))
    per = max(1, len(cpus) // max(world, 1))
    mine = cpus[local_rank * per:(local_rank + 1) * per] if world > 1 else cpus
    return mine or cpus


# --------------------------------------------------------------------------------------------
# CPU path of the reference: the cpu_baseline leg and the --impl reference arm.
# This is the only place bench.py touches oracle/ (and it never imports the product).
# --------------------------------------------------------------------------------------------
def _cpu_transform(item):
    from oracle import preprocess as opre
    img = opre.flip_vertical(item["image"])          # btb/offscreen.py:95-96
    img = opre.gamma_correct(img, GAMMA)             # examples/datagen/generate.py:10-14
    item["image"] = opre.normalise_chw(img)          # examples/densityopt/densityopt.py:117-119
    return item


def _worker_init(_):
    import torch
    torch.set_num_threads(1)


def _reference_dataset(addrs, max_items):
    """(dataset, kind): the reference's own RemoteIterableDataset when build() has copied the unmodified
    package to oracle/_ref ("reference"), else the oracle's restatement of it ("port")."""
    if (REF / "blendtorch" / "btt" / "dataset.py").exists():
        if str(REF) not in sys.path:
            sys.path.insert(0, str(REF))
        from blendtorch import btt as ref_btt
        assert str(REF) in ref_btt.__file__, "the product package shadows oracle/_ref"
        return ref_btt.RemoteIterableDataset(addrs, max_items=max_items, item_transform=_cpu_transform,
                                             timeoutms=60000), "reference"
    from oracle import stream
    return stream.torch_stream_dataset(addrs, max_items=max_items, item_transform=_cpu_transform, timeoutms=60000), "port"


def run_cpu_reference(steps, warmup, workers, min_seconds=5.0, max_seconds=30.0):
    """Reference CPU path on this box: synthetic publishers -> RemoteIterableDataset -> item_transform ->
    default_collate in `workers` DataLoader processes (prefetch_factor 2, torch's default and what
    benchmarks/benchmark.py:26-28 runs with).  After warm-up, workers x prefetch_factor batches are
    discarded (they were made before the clock starts); then at least `steps` batches and `min_seconds`
    are timed.  Returns (images/s, batches timed, seconds, kind)."""
    if str(ROOT) not in sys.path:
        sys.path.insert(0, str(ROOT))
    from torch.utils.data import DataLoader
    addrs, procs = spawn_publishers("oracle.synth", PUBLISHERS, tag="ref")
    try:
        ds, kind = _reference_dataset(addrs, max_items=10 ** 8)
        prefetch = 2
        dl = DataLoader(ds, batch_size=BATCH, num_workers=workers, worker_init_fn=_worker_init,
                        persistent_workers=False, prefetch_factor=prefetch if workers else None)
        it = iter(dl)
        for _ in range(warmup + (workers * prefetch if workers else 0)):
            next(it)
        t0 = time.perf_counter()
        done = 0
        while True:
            b = next(it)
            assert b["image"].shape == (BATCH, C, H, W)
            done += 1
            dt = time.perf_counter() - t0
            if (done >= steps and dt >= min_seconds) or dt >= max_seconds:
                break
        dt = time.perf_counter() - t0
        del it
        return done * BATCH / dt, done, dt, kind
    finally:
        stop_publishers(procs)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = len(os.sched_getaffinity(0))
    workers = max(1, min(cores, 64))  # all host threads: measured fastest (profiles/r1_cpu_baseline_sweep.jsonl)
    ips, done, dt, kind = run_cpu_reference(args.steps, args.warmup, workers, min_seconds=args.cpu_seconds)
    what = ("the reference's own btt.RemoteIterableDataset (unmodified files in oracle/_ref)" if kind == "reference"
            else "oracle port of btt.RemoteIterableDataset")
    line = {
        "impl": "reference", "metric": METRIC, "value": ips, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(done, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(CONFIG_KEYS),
        "config_detail": {"transform": "flipud + gamma_correct(2.2) + (x-127.5)/127.5 + CHW, default_collate",
                          "workers": workers, "prefetch_factor": 2, "timed_steps": done, "window_s": dt},
        "timed_steps": done,
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": workers, "kind": kind,
                         "sample": "%d batches of %d in %.1f s through %d DataLoader workers (%d host cpus); %s + "
                                   "flipud/gamma/normalise item_transform + default_collate" % (done, BATCH, dt, workers, cores, what)},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------
# Our arm
# --------------------------------------------------------------------------------------------
def _graph_of(torch, stream, enqueue):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        enqueue()
    return g


def _time_graphs(torch, stream, graphs, min_ms=20.0, sampler=None):
    """Replay the graph sequence R times so that the CUDA-event window is >= min_ms; (seconds, R)."""
    for g in graphs:
        g.replay()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for g in graphs:
        g.replay()
    e1.record(stream)
    e1.synchronize()
    once = max(e0.elapsed_time(e1), 1e-3)
    reps = max(1, int(min_ms / once) + 1)
    if sampler:
        sampler.active = True
    e0.record(stream)
    for _ in range(reps):
        for g in graphs:
            g.replay()
    e1.record(stream)
    e1.synchronize()
    if sampler:
        sampler.active = False
    return e0.elapsed_time(e1) * 1e-3, reps


def device_leg(torch, btt, dev, steps, warmup, sampler):
    """K1 over HBM-resident batches, no eager launch inside any timed region.  Returns a dict:
    overlapped (DEVICE_STREAMS streams, the pipeline's shape: `value`), single (one stream, back to back),
    e2e_kernel (statistics variant + fold, one stream: what the e2e path launches per batch)."""
    g = torch.Generator(device=dev)
    g.manual_seed(2000)
    srcs = [torch.randint(0, 256, (BATCH, H, W, C), dtype=torch.uint8, device=dev, generator=g) for _ in range(ROTATE)]
    outs = [torch.empty((BATCH, C, H, W), dtype=torch.float32, device=dev) for _ in range(ROTATE)]
    spec = btt.Preprocess(key="image", flip=True, gamma=GAMMA, mean=MEAN, std=STD)
    spec_stats = btt.Preprocess(key="image", flip=True, gamma=GAMMA, mean=MEAN, std=STD, stats=True)
    streams = [torch.cuda.Stream(dev) for _ in range(DEVICE_STREAMS)]
    stream = streams[0]
    stats_out = [torch.empty((C, 2), dtype=torch.float64, device=dev) for _ in range(ROTATE)]
    scratch = torch.empty(4 << 20, dtype=torch.uint8, device=dev)

    def fan(n):
        def enqueue():
            fork = torch.cuda.Event()
            fork.record(stream)
            for s in streams[1:]:
                s.wait_event(fork)
            for i in range(n):
                btt.preprocess_images(srcs[i % ROTATE], spec, out=outs[i % ROTATE], stream=streams[i % len(streams)])
            for s in streams[1:]:
                join = torch.cuda.Event()
                join.record(s)
                stream.wait_event(join)
        return enqueue

    def chain(n, sp, with_stats):
        def enqueue():
            for i in range(n):
                if with_stats:
                    btt.preprocess_images(srcs[i % ROTATE], sp, out=outs[i % ROTATE], stream=stream,
                                          stats_out=stats_out[i % ROTATE], _scratch=scratch)
                else:
                    btt.preprocess_images(srcs[i % ROTATE], sp, out=outs[i % ROTATE], stream=stream)
        return enqueue

    out = {}
    with torch.cuda.stream(stream):
        for i in range(max(warmup, 3)):
            btt.preprocess_images(srcs[i % ROTATE], spec, out=outs[i % ROTATE], stream=stream)
            btt.preprocess_images(srcs[i % ROTATE], spec_stats, out=outs[i % ROTATE], stream=stream,
                                  stats_out=stats_out[i % ROTATE], _scratch=scratch)
        stream.synchronize()
        # `steps` launches = whole rounds of ROTATE plus a tail, both as graphs
        n_round, tail = divmod(steps, ROTATE)
        graphs = [_graph_of(torch, stream, fan(ROTATE))] * n_round
        if tail:
            graphs.append(_graph_of(torch, stream, fan(tail)))
        sec, reps = _time_graphs(torch, stream, graphs, sampler=sampler)
        out["overlapped"] = {"sec_per_launch": sec / (reps * steps), "launches_timed": reps * steps, "replays": reps,
                             "window_ms": sec * 1e3}
        sec, reps = _time_graphs(torch, stream, [_graph_of(torch, stream, chain(ROTATE, spec, False))], sampler=sampler)
        out["single"] = {"sec_per_launch": sec / (reps * ROTATE), "launches_timed": reps * ROTATE, "window_ms": sec * 1e3}
        sec, reps = _time_graphs(torch, stream, [_graph_of(torch, stream, chain(ROTATE, spec_stats, True))], sampler=sampler)
        out["e2e_kernel"] = {"sec_per_launch": sec / (reps * ROTATE), "launches_timed": reps * ROTATE, "window_ms": sec * 1e3}
    out["algo_bytes"] = BATCH * H * W * C * (1 + 4)
    return out


def kernel_times(torch, btt, dev, h, w, batch, depth=False, corners=0):
    """CUDA-event time per launch of the kernels a config's batch runs (K1 with statistics + fold, the
    depth flip, K2), each over inputs larger than L2 or rotated so, from graphs."""
    stream = torch.cuda.Stream(dev)
    res = {}
    frame_bytes = batch * h * w * 4
    rot = max(2, min(8, -(-(200 << 20) // (frame_bytes * 5))))
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    with torch.cuda.stream(stream):
        srcs = [torch.randint(0, 256, (batch, h, w, 4), dtype=torch.uint8, device=dev, generator=g) for _ in range(rot)]
        outs = [torch.empty((batch, 4, h, w), dtype=torch.float32, device=dev) for _ in range(rot)]
        st = [torch.empty((4, 2), dtype=torch.float64, device=dev) for _ in range(rot)]
        spec = btt.Preprocess(flip=True, gamma=GAMMA, mean=MEAN, std=STD, stats=True)
        scratch = torch.empty(8 << 20, dtype=torch.uint8, device=dev)

        def k1():
            for i in range(rot):
                btt.preprocess_images(srcs[i], spec, out=outs[i], stream=stream, stats_out=st[i], _scratch=scratch)
        k1()
        stream.synchronize()
        sec, reps = _time_graphs(torch, stream, [_graph_of(torch, stream, k1)], min_ms=10.0)
        t = sec / (reps * rot)
        res["k1_stats_us"] = t * 1e6
        res["k1_gbs"] = frame_bytes * 5 / t / 1e9
        del srcs, outs
        if depth:
            d_in = [torch.rand((batch, h, w), dtype=torch.float32, device=dev) for _ in range(2)]
            d_out = [torch.empty_like(d_in[0]) for _ in range(2)]

            def fl():
                for i in range(2):
                    btt.flip_rows(d_in[i], out=d_out[i], stream=stream)
            fl()
            stream.synchronize()
            sec, reps = _time_graphs(torch, stream, [_graph_of(torch, stream, fl)], min_ms=10.0)
            res["flip_us"] = sec / (reps * 2) * 1e6
            res["flip_gbs"] = batch * h * w * 8 / (sec / (reps * 2)) / 1e9
            del d_in, d_out
        if corners:
            cam = btt.Camera.from_lens((h, w), lens=35.0, distance=0.0, device=dev)
            pts = torch.rand((batch, corners, 3), dtype=torch.float64, device=dev) * 4 + torch.tensor([0, 0, 6.0], device=dev)
            cam.bbox_world_to_pixel_boxes(pts, 8)
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(50):
                cam.bbox_world_to_pixel_boxes(pts, 8)
            e1.record(stream)
            e1.synchronize()
            res["k2_us_eager"] = e0.elapsed_time(e1) * 1e3 / 50   # launched from Python: host-paced
    return res
