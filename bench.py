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
GROUP = None  # DeviceLoader(group=...) of the e2e legs (None = automatic); tools override it
ROTATE = 8  # distinct resident batches: 8 x (9.8 MB in + 39.3 MB out) = 393 MB > 126 MB L2


def _use_product():
    for p in (str(ROOT), str(PKG)):
        if p not in sys.path:
            sys.path.insert(0, p)


def free_port():
    """A port nobody listens on, taken from outside the kernel's ephemeral range
    (/proc/sys/net/ipv4/ip_local_port_range): the address is handed to a process that binds it a moment
    later, and meanwhile no outgoing connection of ours can be given the same number as its source port
    (least of all a consumer's retry that connects to itself)."""
    import random
    import socket
    lo, hi = 32768, 60999
    try:
        with open("/proc/sys/net/ipv4/ip_local_port_range") as f:
            lo, hi = (int(x) for x in f.read().split()[:2])
    except (OSError, ValueError):
        pass
    spans = [(a, b) for a, b in ((10000, min(lo, 65536) - 1), (hi + 1, 65535)) if b - a >= 500] or [(20000, 32000)]
    for _ in range(400):
        a, b = random.choice(spans)
        port = random.randint(a, b)
        s = socket.socket()
        try:
            s.bind(("127.0.0.1", port))
        except OSError:
            continue
        finally:
            s.close()
        return port
    raise RuntimeError("no free port found")


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


_POOLS = {}


def _pool_messages(S, config, btid, h, w, pool=16):
    key = (config, btid, h, w)
    if key not in _POOLS:
        rng = S.rng_for(config, btid)
        _POOLS[key] = [S.message(rng, config, btid, i, h, w) for i in range(pool)]
    return _POOLS[key]


def check_batch(np, S, b, config, h, w, flipped_depth=False, pool=16):
    """Outside every timed region: a delivered batch against the frames its publishers generated --
    annotations exactly, and the fused statistics (per-channel sum and sum of squares of the flipped,
    gamma-corrected, normalised frames) against the same expression in numpy on those frames."""
    btid = b["btid"].cpu().numpy()
    fid = b["frameid"].cpu().numpy()
    lut = np.uint8(255.0 * (np.arange(256, dtype=np.uint8).astype(np.float32) / 255) ** (1 / GAMMA))
    s1, s2, mag = np.zeros(4), np.zeros(4), np.zeros(4)
    n_check = min(len(btid), 64)      # large batches: the statistics cover all items, annotations the first 64
    ann = {k: b[k][:n_check].cpu().numpy() for k in ("bbox", "xy", "corners") if k in b}
    for i in range(len(btid)):
        m = _pool_messages(S, config, int(btid[i]), h, w, pool)[int(fid[i]) % pool]
        img = m["image"]
        g = np.concatenate((lut[img[..., :3]], img[..., 3:4]), axis=-1)
        x = ((g.astype(np.float32) - np.float32(MEAN)) / np.float32(STD)).astype(np.float64)
        s1 += x.sum(axis=(0, 1))
        s2 += (x * x).sum(axis=(0, 1))
        mag += np.abs(x).sum(axis=(0, 1))
        if i >= n_check:
            continue
        for k, got in ann.items():
            if not np.array_equal(got[i], m[k]):
                return "field %s of item %d differs from the publisher's" % (k, i)
        if "depth" in m and i < 2:
            want = m["depth"][::-1] if flipped_depth else m["depth"]
            if not np.array_equal(b["depth"][i].cpu().numpy(), want):
                return "depth of item %d differs from the publisher's" % i
    st = b["image_stats"].cpu().numpy()
    if not (np.abs(st[:, 0] - s1) <= 1e-6 * mag + 1e-3).all() or not (np.abs(st[:, 1] - s2) <= 1e-6 * s2 + 1e-3).all():
        return "image_stats %s differ from the frames' %s" % (st.tolist(), [s1.tolist(), s2.tolist()])
    return "ok"


def _timed_batches(torch, it, steps, min_seconds, sampler, barrier=None):
    """Time batches from iterator `it` until >= steps batches and >= min_seconds; D2H read of the step's
    result (the statistics) every step.  (images, seconds, batches, last batch, stamps)"""
    if barrier:
        barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler:
        sampler.active = True
    t0 = time.perf_counter()
    e0.record()
    images = n = 0
    stamps = [t0]
    b = None
    while True:
        try:
            b = next(it)
        except StopIteration:
            break
        b["image_stats"].cpu()                     # D2H read of the step's result
        images += b["image"].shape[0]
        n += 1
        now = time.perf_counter()
        stamps.append(now)
        if n >= steps and now - t0 >= min_seconds:
            break
    e1.record()
    e1.synchronize()
    wall = time.perf_counter() - t0
    if sampler:
        sampler.active = False
    return images, max(e0.elapsed_time(e1) * 1e-3, wall), n, b, stamps


def _annotation_bytes(torch, b, batch):
    return sum(v[0].numel() * v.element_size() * batch for k, v in b.items()
               if isinstance(v, torch.Tensor) and k in ("btid", "frameid", "bbox", "xy", "corners"))


def stream_leg(ctx, cfg, transport, receivers, senders, shm_slots=0, min_seconds=2.0, project=None, depth_flip=False,
               in_place=False):
    """Public API with host buffers: native publishers -> DeviceLoader -> device batch (+ stats read back)."""
    torch, btt, np, S = ctx["torch"], ctx["btt"], ctx["np"], ctx["S"]
    h, w, batch, npub, config = cfg["h"], cfg["w"], cfg["batch"], cfg["publishers"], cfg["config"]
    extra = ("--shm-slots", str(shm_slots)) if shm_slots else ()
    if in_place:
        extra += ("--in-place",)
    senders = max(1, min(senders, npub))
    addrs, proc = spawn_native_publisher(npub, transport, threads=senders, tag="r%d%s" % (ctx["rank"], cfg["name"]),
                                         extra=extra, cpus=ctx["cpus"], config=config, h=h, w=w)
    dl = it = None
    try:
        ds = btt.RemoteIterableDataset(addrs, max_items=10 ** 9, timeoutms=60000)
        pre = [btt.Preprocess(key="image", flip=True, gamma=GAMMA, mean=MEAN, std=STD, stats=True)]
        if depth_flip:
            pre.append(btt.FlipRows(key="depth"))
        prefetch = 4 if batch <= 64 else 2
        dl = btt.DeviceLoader(ds, batch_size=batch, device=ctx["dev"], preprocess=pre, num_receivers=receivers,
                              ring_depth=4 if batch <= 64 else 3, prefetch=prefetch, project=project, group=GROUP)
        it = iter(dl)
        for _ in range(ctx["warmup"]):
            next(it)["image_stats"].cpu()
        for _ in range(prefetch + receivers):        # produced before the clock starts: not ours to count
            next(it)
        torch.cuda.synchronize(ctx["dev"])
        before = dict(dl.stats)
        images, sec, n, b, stamps = _timed_batches(torch, it, ctx["steps"], min_seconds, ctx["sampler"], ctx["barrier"])
        stats = {k: dl.stats[k] - before[k] for k in ("consumer_wait_s", "producer_wait_s")}   # of the timed window
        check = check_batch(np, S, next(it), config, h, w, flipped_depth=depth_flip)
        per_item = h * w * 4 * (2 if depth_flip else 1)
        h2d = _annotation_bytes(torch, b, batch) + batch * per_item
        win = max(1, min(20, len(stamps) // 5))
        rates = sorted(win * batch / (stamps[i + win] - stamps[i]) for i in range(0, len(stamps) - win, win)
                       if stamps[i + win] > stamps[i])
        spread = ({"window_batches": win, "p10": rates[len(rates) // 10], "p50": rates[len(rates) // 2],
                   "p90": rates[(len(rates) * 9) // 10]} if rates else None)
        return {"images": images, "sec": sec, "batches": n, "h2d": h2d, "d2h": 64, "check": check, "spread": spread,
                "bytes_per_image": per_item, "consumer_wait_s": stats["consumer_wait_s"],
                "producer_wait_s": stats["producer_wait_s"], "receivers": receivers, "senders": senders}
    finally:
        if it is not None:
            it.close()
        if dl is not None:
            dl.close()
        stop_publishers([proc])


def replay_leg(ctx, cfg, prefix, cache=None, min_seconds=2.0, project=None, depth_flip=False, threads=4, producers=2):
    """FileDataset -> DeviceLoader (shuffled, epochs back to back): mmap -> pinned -> DMA, or with
    cache='host' DMA straight from a page-locked copy of the recording."""
    torch, btt, np, S = ctx["torch"], ctx["btt"], ctx["np"], ctx["S"]
    h, w, batch, config = cfg["h"], cfg["w"], cfg["batch"], cfg["config"]
    ds = btt.FileDataset(prefix)
    pre = [btt.Preprocess(key="image", flip=True, gamma=GAMMA, mean=MEAN, std=STD, stats=True)]
    if depth_flip:
        pre.append(btt.FlipRows(key="depth"))
    dl = btt.DeviceLoader(ds, batch_size=batch, shuffle=True, drop_last=True, device=ctx["dev"], preprocess=pre,
                          ingest_threads=threads, ring_depth=3, prefetch=3, num_receivers=producers, cache=cache,
                          project=project, group=GROUP)
    it = None
    try:
        def epochs():
            while True:
                yield from dl
        it = epochs()
        for _ in range(ctx["warmup"] + 3 + producers):
            next(it)["image_stats"].cpu()
        torch.cuda.synchronize(ctx["dev"])
        waited = dl.stats["consumer_wait_s"]
        images, sec, n, b, _ = _timed_batches(torch, it, ctx["steps"], min_seconds, ctx["sampler"], ctx["barrier"])
        waited = dl.stats["consumer_wait_s"] - waited
        check = check_batch(np, S, next(it), config, h, w, flipped_depth=depth_flip, pool=cfg.get("pool", 16))
        per_item = h * w * 4 * (2 if depth_flip else 1)
        return {"images": images, "sec": sec, "batches": n, "check": check, "bytes_per_image": per_item,
                "cache_fill_s": dl.stats.get("cache_fill_s"), "consumer_wait_s": waited}
    finally:
        if it is not None:
            it.close()
        dl.close()


def host_probes(ctx):
    """What the host can do at most, measured with every rank busy at once: pinned H2D per GPU and
    plain memcpy on this rank's cores."""
    import ctypes
    N, np = ctx["N"], ctx["np"]
    out = {}
    g = ctypes.c_double(0.0)
    ctx["barrier"]()
    N.check(N.lib.bt_measure_h2d(ctx["local"], 64 << 20, 40, ctypes.byref(g)))
    out["pcie_h2d_gbs"] = g.value
    ctx["barrier"]()
    # memcpy: one thread per core of this rank but one, 64 MiB blocks (numpy releases the GIL)
    nthreads = max(1, len(ctx["cpus"]) - 1)
    bufs = [(np.ones(64 << 20, np.uint8), np.empty(64 << 20, np.uint8)) for _ in range(nthreads)]
    done = [0] * nthreads
    stop_at = time.perf_counter() + 0.6

    def work(t):
        a, b = bufs[t]
        while time.perf_counter() < stop_at:
            np.copyto(b, a)
            done[t] += 1
    ts = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    t0 = time.perf_counter()
    [t.start() for t in ts]
    [t.join() for t in ts]
    out["host_memcpy_gbs"] = sum(done) * (64 << 20) / (time.perf_counter() - t0) / 1e9
    out["host_memcpy_threads"] = nthreads
    ctx["barrier"]()
    return out


def socket_probe(ctx, transport, receivers, senders, seconds=1.0, cfg=None):
    """Images/s of the socket path alone (C2 messages unless `cfg` says otherwise): native publishers ->
    `receivers` direct receivers filling an unpinned ring (scan + the one copy), no GPU, no Python per
    message."""
    from blendtorch.btt.rx import DirectReceiver
    from blendtorch.btt.staging import Ring, Schema
    cfg = cfg or {"config": CONFIG, "h": H, "w": W, "publishers": PUBLISHERS}
    addrs, proc = spawn_native_publisher(cfg["publishers"], transport, threads=min(senders, cfg["publishers"]),
                                         tag="sp%d" % ctx["rank"], cpus=ctx["cpus"], config=cfg["config"], h=cfg["h"], w=cfg["w"])
    counts = [0] * receivers
    try:
        def run(r):
            rx = DirectReceiver(addrs)
            give_up = time.perf_counter() + 20
            while not rx.recv(500):
                if time.perf_counter() > give_up:
                    return
            sch = Schema.from_message(rx.pending(), BATCH)
            ring = Ring(sch.slot_bytes, 2, None)
            for phase in (0, 1):               # phase 0: let every receiver get going
                stop = time.perf_counter() + (0.3 if phase == 0 else seconds)
                n = 0
                while time.perf_counter() < stop:
                    _, k = rx.fill(ring, ring.acquire(), sch, 0, BATCH, 100)
                    n += k
                counts[r] = n
            rx.close()
        ts = [threading.Thread(target=run, args=(r,)) for r in range(receivers)]
        [t.start() for t in ts]
        [t.join() for t in ts]
    finally:
        stop_publishers([proc])
    return sum(counts) / seconds


def stream_threads(cores, endpoints):
    """(sender threads, receiver threads) of one rank for the large-frame stream legs.

    One sender thread per publisher endpoint, always: a sender sleeps in poll() while its consumers'
    sockets are full, so it costs little beyond its copies, whereas a thread that serves several
    endpoints keeps retrying the full sockets of one while another still takes data -- and in the
    sweeps behind this rule (profiles/r2_e2e_sweep_native.jsonl, r2_sender_threads.jsonl) exactly the
    unbalanced split of 4 threads over 5 endpoints collapsed now and then (5 k instead of 24 k images/s),
    while 5 threads on 4 cores carried 43 % more than 1.  Receivers: two thirds of the rank's cores but
    one (16 cores: 10 -> 26.0 k images/s against 24.0 k with 8; 12: 7; 8: 4; 4: 2), at most 10."""
    return endpoints, max(1, min(10, (cores - 1) * 2 // 3))


def split_threads(cores, sender_rate, receiver_rate, max_senders, max_receivers=10):
    """Sender and receiver thread counts for one rank from the measured per-thread rates: the split of
    (cores - 1) threads (one core stays with the consumer / interpreter) that maximises
    min(senders * sender_rate, receivers * receiver_rate)."""
    budget = max(2, cores - 1)
    best = (1, 1, 0.0)
    for s in range(1, min(max_senders, budget - 1) + 1):
        r = max(1, min(max_receivers, budget - s))
        rate = min(s * sender_rate, r * receiver_rate)
        if rate > best[2] * 1.001:
            best = (s, r, rate)
    return best


def ours(args):
    _use_product()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpus = rank_cpus(local, world)
    os.sched_setaffinity(0, cpus)      # before CUDA, torch and the pinned rings exist: first touch stays local
    import numpy as np
    import torch
    torch.set_num_threads(1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL announces its version on stdout at the first collective; stdout is reserved for the one
        # JSON line, so send everything to stderr until the communicator is up.
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    import importlib.util
    spec_ = importlib.util.spec_from_file_location("bt_build_native", PKG / "build_native.py")
    mod = importlib.util.module_from_spec(spec_)
    spec_.loader.exec_module(mod)
    if not mod.LIB.exists():
        mod.build()
    from blendtorch import btt
    from blendtorch.btt import native as N, synthetic as S

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def gather(x):
        """The value of every rank, on every rank."""
        if dist is None:
            return [float(x)]
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    sampler = ClockSampler(local)
    ctx = {"torch": torch, "btt": btt, "np": np, "S": S, "N": N, "dev": dev, "rank": rank, "local": local, "world": world,
           "cpus": cpus, "steps": args.steps, "warmup": args.warmup, "sampler": sampler, "barrier": barrier}

    # ---- device leg -----------------------------------------------------------------------------
    barrier()
    dleg = device_leg(torch, btt, dev, args.steps, args.warmup, sampler)
    barrier()
    sec_per_launch = max(gather(dleg["overlapped"]["sec_per_launch"]))
    algo_bytes = dleg["algo_bytes"]
    value = world * BATCH / sec_per_launch
    achieved = algo_bytes / sec_per_launch / 1e9
    c2 = {"name": "C2", "config": 2, "h": H, "w": W, "batch": BATCH, "publishers": PUBLISHERS}

    # ---- what the host can do at most, every rank busy at once ------------------------------------
    probes = host_probes(ctx)
    s_rate = r_rate = None
    if args.senders > 0 and args.receivers > 0:
        senders, receivers = args.senders, args.receivers
    else:
        # per-thread rates of the socket path on this rank's cores (reported, and the basis of the C5 split):
        # 1 sender feeding 3 receivers is sender-bound, 3 senders feeding 1 receiver is receiver-bound
        spare = max(1, len(cpus) - 2)
        barrier()
        s_rate = socket_probe(ctx, args.transport, min(3, spare), 1, 0.8)
        barrier()
        r_rate = socket_probe(ctx, args.transport, 1, min(3, PUBLISHERS, spare), 0.8)
        barrier()
        s_rate, r_rate = max(1.0, min(gather(s_rate))), max(1.0, min(gather(r_rate)))
        senders, receivers = stream_threads(len(cpus), PUBLISHERS)
        if args.senders > 0:
            senders = args.senders
        if args.receivers > 0:
            receivers = args.receivers
    barrier()
    sock_rate = socket_probe(ctx, args.transport, receivers, senders, 1.0)
    barrier()
    probes.update({"socket_path_images_per_s": sum(gather(sock_rate)), "senders": senders, "receivers": receivers})

    # ---- end to end: the headline shape over tcp, ipc and the shared-memory wire ---------------------
    legs = {}
    legs["tcp"] = stream_leg(ctx, c2, args.transport, receivers, senders)
    barrier()
    legs["ipc"] = stream_leg(ctx, c2, "ipc", receivers, senders)
    barrier()
    shm_senders = max(1, min(PUBLISHERS, len(cpus) - 2))
    legs["shm"] = stream_leg(ctx, c2, args.transport, min(2, receivers), shm_senders, shm_slots=args.shm_slots)
    barrier()
    legs["shm_in_place"] = stream_leg(ctx, c2, args.transport, min(2, receivers), 1, shm_slots=args.shm_slots, in_place=True)
    barrier()

    # ---- the other BASELINE.json shapes ---------------------------------------------------------------
    tmp = tempfile.mkdtemp(prefix="btrec_r%d_" % rank, dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    raw = {}
    try:
        # C1: replay of a recorded Cube-scene stream, 640x480 RGBA + xy, batch 8 (2 files x 512 frames)
        c1 = {"name": "C1", "config": 1, "h": 480, "w": 640, "batch": 8, "pool": 16}
        S.write_recordings(tmp + "/c1", 1, 2, 512, 480, 640, pool=16)
        rep = replay_leg(ctx, c1, tmp + "/c1", threads=max(1, min(8, len(cpus) - 3)))
        barrier()
        rep_host = replay_leg(ctx, c1, tmp + "/c1", cache="host")
        barrier()
        for f in Path(tmp).glob("c1_*.btr"):
            f.unlink()
        raw["C1"] = (rep, kernel_times(torch, btt, dev, 480, 640, 8),
                     "replay, shuffled, 2 x 512 frames, .btr in page cache: mmap -> pinned -> DMA, K1+stats")
        raw["C1_cache_host"] = (rep_host, None, "same from a page-locked copy of the recording (DMA only, no CPU copy per epoch)")
        if not args.no_configs:
            # C3: 1920x1080 RGBA + float32 depth (sent bottom-up), batch 32, 8 cubes x 8 corners -> pixel boxes
            c3 = {"name": "C3", "config": 3, "h": 1080, "w": 1920, "batch": 32, "pool": 8}
            S.write_recordings(tmp + "/c3", 3, 2, 64, 1080, 1920, pool=8)
            cam = btt.Camera.from_lens((1080, 1920), lens=35.0, distance=0.0, device=dev)
            pj = btt.Projection("corners", camera=cam, xy_key="corners_xy", bbox_key="boxes", corners_per_object=8)
            r3 = replay_leg(ctx, c3, tmp + "/c3", project=pj, depth_flip=True, threads=max(1, min(12, len(cpus) - 3)),
                            min_seconds=1.5)
            barrier()
            r3h = replay_leg(ctx, c3, tmp + "/c3", cache="host", project=pj, depth_flip=True, min_seconds=1.5)
            barrier()
            for f in Path(tmp).glob("c3_*.btr"):
                f.unlink()
            raw["C3"] = (r3, kernel_times(torch, btt, dev, 1080, 1920, 32, depth=True, corners=64),
                         "replay, shuffled, .btr in page cache: mmap -> pinned -> DMA, K1+stats, depth flip, K2 boxes")
            raw["C3_cache_host"] = (r3h, None, "same from a page-locked copy of the recording (DMA only)")
            # C4: 1024x1024 RGBA, batch 64 per GPU, own publishers and ring per GPU
            c4 = {"name": "C4", "config": 4, "h": 1024, "w": 1024, "batch": 64, "publishers": PUBLISHERS}
            r4 = stream_leg(ctx, c4, args.transport, receivers, senders, min_seconds=1.5)
            barrier()
            raw["C4"] = (r4, kernel_times(torch, btt, dev, 1024, 1024, 64),
                         "stream, %d native publishers over %s, %d receivers" % (PUBLISHERS, args.transport, receivers))
            # C5: 128x128 RGBA, batch 1024, 16 publishers
            c5 = {"name": "C5", "config": 5, "h": 128, "w": 128, "batch": 1024, "publishers": 16}
            # 64 KiB messages cost per message, not per byte: the split that suits C2 need not suit them.
            # Try the neighbours of the C2 rule on the socket path alone and keep the best (every rank
            # measures at once; all ranks take the same choice).
            s5, rcv5, _ = split_threads(len(cpus), s_rate or 1.0, r_rate or 1.0, 16)
            budget = s5 + rcv5
            best = None
            # (sender threads sleep while their sockets are full, so more of them than the budget allows can
            # pay when cores are few: 5 sender threads on 4 cores carried 43 % more C2 frames than one,
            # profiles/r2_sender_threads.jsonl)
            for cand in sorted({(s5, rcv5), (max(1, s5 - 1), min(10, budget - max(1, s5 - 1))), (min(16, s5 + 1), max(1, budget - min(16, s5 + 1))),
                                (min(16, 2 * s5 + 2), rcv5), (min(16, 4 * s5), rcv5)}):
                barrier()
                rate = min(gather(socket_probe(ctx, args.transport, cand[1], cand[0], 0.8, cfg=c5)))
                if best is None or rate > best[0] * 1.03:
                    best = (rate, cand)
            barrier()
            s5, rcv5 = best[1]
            r5 = stream_leg(ctx, c5, args.transport, rcv5, s5, min_seconds=1.5)
            barrier()
            raw["C5"] = (r5, kernel_times(torch, btt, dev, 128, 128, 1024),
                         "stream, 16 native publishers over %s, %d receivers, %d sender threads" % (args.transport, rcv5, s5))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    clocks = sampler.report()
    return _report(args, ctx, gather, dist, dleg, value, achieved, sec_per_launch, algo_bytes, probes, s_rate, r_rate,
                   senders, receivers, legs, raw, clocks)


def _report(args, ctx, gather, dist, dleg, value, achieved, sec_per_launch, algo_bytes, probes, s_rate, r_rate,
            senders, receivers, legs, raw, clocks):
    world, rank, cpus = ctx["world"], ctx["rank"], ctx["cpus"]
    h2d_conc = gather(probes["pcie_h2d_gbs"])
    memcpy_conc = gather(probes["host_memcpy_gbs"])
    pcie = sum(h2d_conc) / world   # mean per GPU with every GPU copying at once: what fractions are quoted against
    nbytes = H * W * C
    memcpy_total = sum(memcpy_conc)

    def leg_json(leg, note=None):
        rate = sum(gather(leg["images"])) / max(gather(leg["sec"]))     # whole job: all images / the slowest rank's window
        gbs = rate * leg["bytes_per_image"] / 1e9 / world
        all_ok = min(gather(1.0 if leg["check"] == "ok" else 0.0)) == 1.0
        d = {"value": rate, "unit": "images/s", "gbs_h2d_per_gpu": gbs, "frac_of_pcie_h2d": gbs / pcie,
             "timed_batches": leg["batches"], "window_s": leg["sec"],
             "e2e_check": leg["check"] if (all_ok or leg["check"] != "ok") else "failed on another rank",
             "consumer_wait_s": leg.get("consumer_wait_s")}
        for k in ("producer_wait_s", "receivers", "senders", "cache_fill_s"):
            if leg.get(k) is not None:
                d[k] = leg[k]
        if note:
            d["note"] = note
        return d

    configs = {}
    for name in sorted(raw):
        leg, kt, note = raw[name]
        configs[name] = leg_json(leg, note)
        if kt:
            configs[name]["kernels"] = kt
    e2e, ipc, shm = leg_json(legs["tcp"]), leg_json(legs["ipc"]), leg_json(legs["shm"])
    shm_ip = leg_json(legs["shm_in_place"])
    tcp = legs["tcp"]

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the reference's CPU path, in its own process (it imports the reference's `blendtorch`, not ours)
        try:
            out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "20",
                                  "--warmup", "3", "--cpu-seconds", "12"], capture_output=True, text=True, timeout=240)
            cpu = json.loads(out.stdout.strip().splitlines()[-1])["cpu_baseline"]
        except Exception as e:  # reported, not fatal: the reference arm is also run by the driver
            cpu = {"error": "%s: %s" % (type(e).__name__, e)}
    if rank == 0:
        peak, peak_src = measured_peaks()
        bound = ("the consumer waits for batches (the host receive side binds)" if tcp["consumer_wait_s"] > 0.5 * tcp["sec"]
                 else "the producers wait for the consumer")
        single = algo_bytes / dleg["single"]["sec_per_launch"] / 1e9
        e2ek = algo_bytes / dleg["e2e_kernel"]["sec_per_launch"] / 1e9
        k4 = configs.get("C4", {}).get("kernels", {}).get("k1_gbs")
        checks = [e2e, ipc, shm, shm_ip] + list(configs.values())
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec_per_launch * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(CONFIG_KEYS),
            "config_detail": {
                "l2": "device leg rotates over %d distinct resident batches (%.0f MB working set > 126 MB L2)"
                      % (ROTATE, ROTATE * BATCH * H * W * C * 5 / 1e6),
                "device_leg": "CUDA graphs only (rounds of %d launches over %d streams + a tail graph), replayed %d times: "
                              "%d launches in a %.1f ms CUDA-event window; value = batch / average launch time"
                              % (ROTATE, DEVICE_STREAMS, dleg["overlapped"]["replays"], dleg["overlapped"]["launches_timed"],
                                 dleg["overlapped"]["window_ms"]),
                "e2e_transport": args.transport, "e2e_receivers": receivers, "e2e_sender_threads": senders,
                "e2e_receiver": "direct ZMTP endpoint (bt_pull_*)",
                "e2e_publishers": "%d PUSH endpoints of one native publisher process per GPU (bt_push_*), reference pickle wire" % PUBLISHERS,
                "e2e_window": "after warm-up, prefetch + receivers batches are discarded; then >= steps batches and >= 2 s",
                "cpu_affinity": "rank k of N runs on cores [k*n/N, (k+1)*n/N) with its publishers, receivers and consumer",
                "parallelism": "dp%d, publishers sharded per GPU, no collective" % world},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": measured_traffic(), "peak_source": peak_src,
                         "traffic_note": "ncu, cold L2: 9.85 MB read (= algorithmic 9.83 MB) + <0.2 MB written back inside the "
                                         "kernel; the 39.3 MB of fp32 output of a batch-8 launch are absorbed by the 126 MB L2 "
                                         "and written back after it (C4-size launches write through: see profiles/)",
                         "kernel": "k1_planar<float,4,4,1> (C,PPT,UNROLL: one 4-pixel unit per thread at this batch size)",
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "frac_of_nominal_8TBs": achieved / 8000.0,
                         "overlapped_streams": DEVICE_STREAMS,
                         "single_launch": {"achieved": single, "frac": single / peak,
                                           "us": dleg["single"]["sec_per_launch"] * 1e6,
                                           "note": "same kernel, one stream, back-to-back graph nodes"},
                         "e2e_kernel": {"achieved": e2ek, "frac": e2ek / peak, "us": dleg["e2e_kernel"]["sec_per_launch"] * 1e6,
                                        "note": "statistics variant + fold kernel, one stream: what every e2e batch launches"},
                         "large_batch": {"workload": "C4 per-GPU batch: 64 x 1024x1024 RGBA -> fp32 NCHW (K1 + stats + fold)",
                                         "achieved": k4, "frac": (k4 or 0) / peak}},
            "e2e": dict(e2e, h2d_bytes_per_step=tcp["h2d"], d2h_bytes_per_step=tcp["d2h"],
                        pcie_h2d_measured_gbs=pcie, rank0_images_per_s_spread=tcp["spread"], binds=bound),
            "e2e_check": "ok" if all(v.get("e2e_check") == "ok" for v in checks) else "FAILED",
            "e2e_ipc": dict(ipc, note="as e2e, publishers bound to ipc:// instead of tcp://"),
            "e2e_shm_wire": dict(shm, note="opt-in wire for co-located producers: payload goes producer -> shared "
                                 "memory -> DMA, no consumer-side CPU copy; not the reference's pickle wire, reported beside "
                                 "e2e, not instead of it (%d slots per publisher)" % args.shm_slots),
            "e2e_shm_wire_in_place": dict(shm_ip, note="same wire, frames published from where they stand in shared memory "
                                          "(a producer that renders into the segment: no producer-side copy either); "
                                          "every payload byte crosses host memory once, as a DMA read"),
            "e2e_replay": configs.get("C1"),
            "e2e_replay_cache_host": configs.get("C1_cache_host"),
            "configs": configs,
            "host": {"cpus": len(cpus) * world if world > 1 else len(cpus), "cpus_per_rank": len(cpus),
                     "pcie_h2d_gbs_per_gpu_concurrent": h2d_conc, "host_memcpy_gbs_per_rank_concurrent": memcpy_conc,
                     "host_memcpy_threads_per_rank": probes["host_memcpy_threads"],
                     "socket_path_images_per_s": probes["socket_path_images_per_s"],
                     "socket_path_note": "native publishers -> direct receivers into an unpinned ring, same thread split, "
                                         "every rank at once, no GPU: the ceiling of the pickle wire on these cores",
                     "memory_ceilings": {
                         "host_memcpy_gbs_total": memcpy_total,
                         "pickle_wire_images_per_s": memcpy_total / 2.5 * 1e9 / nbytes,
                         "one_copy_images_per_s": memcpy_total / 1.5 * 1e9 / nbytes,
                         "dma_only_images_per_s": sum(h2d_conc) * 1e9 / nbytes,
                         "note": "host memory bandwidth is shared by all ranks.  In units of the concurrent memcpy rate "
                                 "(1 copy = 1 read + 1 write), a payload byte costs 2 copies + 1 DMA read = 2.5 on the socket "
                                 "wire (sender -> kernel, kernel -> pinned, DMA), 1.5 with one CPU copy (shared-memory wire "
                                 "with a copying producer; replay from the page cache), and only the DMA read when frames "
                                 "already stand in page-locked memory (in-place shared memory, cache='host'): those scale "
                                 "with the GPUs' concurrent PCIe rate instead"},
                     "sender_images_per_s_per_thread": s_rate, "receiver_images_per_s_per_thread": r_rate,
                     "thread_split": "senders %d (one per endpoint), receivers %d per rank" % (senders, receivers)},
            "gpu_launches": dleg["overlapped"]["launches_timed"],
            # e2e batches carry statistics (the step's result that is read back): K1 + the fold kernel of its rows
            "gpu_launches_e2e": 2 * tcp["batches"],
            "clocks": clocks,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--transport", default="tcp", choices=["tcp", "ipc"])
    ap.add_argument("--receivers", type=int, default=0, help="receiver threads per GPU (0 = from measured per-thread rates)")
    ap.add_argument("--senders", type=int, default=0, help="publisher sender threads per GPU (0 = from measured rates)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C3/C4/C5 legs")
    ap.add_argument("--shm-slots", type=int, default=32)
    ap.add_argument("--cpu-seconds", type=float, default=5.0, help="minimum timed window of the reference arm")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return reference_arm(args)
    return ours(args)


if __name__ == "__main__":
    sys.exit(main())
