#!/usr/bin/env python3
"""Micro-benchmark of K1 variants (diagnostic): per-launch time from CUDA-graph replays over rotating
buffers larger than L2.  Usage: python tools/k1_bench.py [quick]"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402

DEV = torch.device("cuda", 0)
ESZ = {torch.float32: 4, torch.bfloat16: 2, torch.uint8: 1}


def run(shape, dtype, layout="NCHW", stats=False, variant="auto", gamma=2.2, norm=True, streams=1, reps=200,
        smooth=False, tune=0):
    B, H, W, C = shape
    in_b = B * H * W * C
    out_b = in_b * ESZ[dtype]
    rot = max(2, int(400e6 // (in_b + out_b)) + 1)
    g = torch.Generator(device=DEV)
    g.manual_seed(1)
    if smooth:
        base = torch.arange(W, device=DEV).view(1, 1, W, 1) // 3 + torch.arange(H, device=DEV).view(1, H, 1, 1) // 2
        srcs = [((base + i) % 256).to(torch.uint8).expand(B, H, W, C).contiguous() for i in range(rot)]
    else:
        srcs = [torch.randint(0, 256, shape, dtype=torch.uint8, device=DEV, generator=g) for _ in range(rot)]
    oshape = (B, C, H, W) if layout == "NCHW" else (B, H, W, C)
    outs = [torch.empty(oshape, dtype=dtype, device=DEV) for _ in range(rot)]
    kw = dict(flip=True, gamma=gamma, layout=layout, dtype=dtype, stats=stats, variant=variant, tune=tune)
    if norm and dtype != torch.uint8:
        kw.update(mean=127.5, std=127.5)
    spec = btt.Preprocess(**kw)
    scratch = torch.empty(1 << 17, dtype=torch.uint8, device=DEV)
    st_out = torch.empty((C, 2), dtype=torch.float64, device=DEV)
    ss = [torch.cuda.Stream(DEV) for _ in range(streams)]
    main = ss[0]

    def launch(i, s):
        btt.preprocess_images(srcs[i], spec, out=outs[i], stream=s, stats_out=st_out, _scratch=scratch)

    with torch.cuda.stream(main):
        for i in range(3):
            launch(i % rot, main)
        main.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=main):
            if streams > 1:
                ev0 = torch.cuda.Event()
                ev0.record(main)
                for s in ss[1:]:
                    s.wait_event(ev0)
            for i in range(rot):
                launch(i, ss[i % streams])
            if streams > 1:
                for s in ss[1:]:
                    e = torch.cuda.Event()
                    e.record(s)
                    main.wait_event(e)
        graph.replay()
        main.synchronize()
        n = max(1, reps // rot)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        for _ in range(n):
            graph.replay()
        e1.record(main)
        e1.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (n * rot)
    return {"shape": list(shape), "dtype": str(dtype).split(".")[-1], "layout": layout, "stats": stats,
            "variant": variant, "tune": hex(tune), "streams": streams, "smooth": smooth, "us": round(us, 2),
            "GBps": round((in_b + out_b) / us / 1e3, 1), "frac_meas": round((in_b + out_b) / us / 1e3 / 6548.5, 3)}


def main():
    C2, C4, C5, C3 = (8, 480, 640, 4), (64, 1024, 1024, 4), (1024, 128, 128, 4), (32, 1080, 1920, 4)
    cases = [
        dict(shape=C2, dtype=torch.float32), dict(shape=C2, dtype=torch.float32, streams=2),
        dict(shape=C2, dtype=torch.float32, streams=4), dict(shape=C2, dtype=torch.float32, stats=True),
        dict(shape=C2, dtype=torch.float32, smooth=True),
        dict(shape=C2, dtype=torch.bfloat16), dict(shape=C2, dtype=torch.uint8, layout="NHWC"),
        dict(shape=C4, dtype=torch.float32), dict(shape=C4, dtype=torch.float32, stats=True),
        dict(shape=C4, dtype=torch.float32, smooth=True),
        dict(shape=C4, dtype=torch.bfloat16), dict(shape=C4, dtype=torch.uint8, layout="NHWC"),
        dict(shape=C4, dtype=torch.float32, layout="NHWC"),
        dict(shape=C3, dtype=torch.float32), dict(shape=C5, dtype=torch.float32),
        dict(shape=(8, 480, 640, 3), dtype=torch.float32), dict(shape=(8, 481, 641, 4), dtype=torch.float32),
    ]
    if len(sys.argv) > 1 and sys.argv[1] == "lf":
        cases = [dict(shape=sh, dtype=dt, tune=t, streams=st) for sh in (C2, (16, 480, 640, 4), C4, C5)
                 for dt in (torch.float32, torch.bfloat16) for t in (0, 0x40000, 0x50000)
                 for st in ((1, 4) if sh[0] <= 16 else (1,))]
    elif len(sys.argv) > 1 and sys.argv[1] == "stats":
        cases = [dict(shape=sh, dtype=dt, stats=st, streams=n) for sh in (C2, C4, C3, C5)
                 for dt in (torch.float32, torch.bfloat16) for st in (False, True)
                 for n in ((1, 4) if sh == C2 else (1,))]
    elif len(sys.argv) > 1 and sys.argv[1] == "statscap":
        cases = [dict(shape=sh, dtype=dt, stats=True, streams=n, tune=cap << 8) for sh in (C2, C4, C5)
                 for dt in (torch.float32, torch.bfloat16) for cap in (0, 4, 8, 16, 32, 64)
                 for n in ((1, 4) if sh == C2 else (1,))]
    elif len(sys.argv) > 1 and sys.argv[1] == "wave":
        cases = [dict(shape=(b, 480, 640, 4), dtype=dt, streams=n, tune=cap << 8) for b in (8, 12, 16, 24)
                 for dt in (torch.float32, torch.bfloat16) for cap in (0, 3, 4, 8) for n in (1, 4)]
    elif len(sys.argv) > 1 and sys.argv[1] == "retune":
        cases = [dict(shape=C2, dtype=torch.float32, streams=n, tune=t) for n in (1, 2, 3, 4, 8)
                 for t in (0, 0x40000, 2, 8, 1)]
        cases += [dict(shape=sh, dtype=torch.float32, tune=t) for sh in (C4, C5) for t in (0, 0x40000, 2, 8)]
        cases += [dict(shape=sh, dtype=torch.bfloat16, tune=t, streams=n) for sh in (C2, C4) for t in (0, 0x40000, 4, 1)
                  for n in ((1, 2, 4) if sh == C2 else (1,))]
        cases += [dict(shape=C2, dtype=torch.float32, variant=v, streams=n) for v in ("bulk",) for n in (1, 2, 4)]
        cases += [dict(shape=C4, dtype=torch.float32, variant="bulk")]
    elif len(sys.argv) > 1 and sys.argv[1] == "table":
        # the table built by every CTA (default) against a host-built table in kernel parameters (bit 19)
        cases = [dict(shape=sh, dtype=dt, tune=t, streams=n, stats=st) for sh in (C2, (16, 480, 640, 4), C5, C4)
                 for dt in (torch.float32, torch.bfloat16) for st in (False, True) for t in (0, 0x80000)
                 for n in ((1, 4) if sh[0] <= 16 else (1,))] * 2
        cases += [dict(shape=C4, dtype=torch.uint8, layout="NHWC", tune=t) for t in (0, 0x80000)]
        cases += [dict(shape=sh, dtype=torch.float32, layout="NHWC", tune=t) for sh in (C2, C4) for t in (0, 0x80000)]
    elif len(sys.argv) > 1 and sys.argv[1] == "ncu":
        # a short run for `ncu --set full`: the headline launch, its statistics variant, C4, and an odd width
        cases = [dict(shape=C2, dtype=torch.float32, reps=16), dict(shape=C2, dtype=torch.float32, stats=True, reps=16),
                 dict(shape=C4, dtype=torch.float32, reps=4), dict(shape=(64, 481, 641, 4), dtype=torch.float32, reps=4)]
    elif len(sys.argv) > 1 and sys.argv[1] == "statsunroll":
        # statistics variant: units in flight per thread (registers vs occupancy) x grid cap
        cases = [dict(shape=sh, dtype=torch.float32, stats=True, streams=n, tune=u | (cap << 8))
                 for sh in (C2, (16, 480, 640, 4), C5, C4) for u in (0, 1, 2) for cap in (0, 4, 8, 16)
                 for n in ((1, 4) if sh[0] <= 16 else (1,))]
    elif len(sys.argv) > 1 and sys.argv[1] == "single":
        # one launch alone at the headline batch: units per thread x whole-wave grids, and the bulk variant
        cases = [dict(shape=C2, dtype=torch.float32, streams=n, tune=t) for t in (0, 0x2, 0x802, 0x402, 0x808, 0x408, 0x801, 0x1001)
                 for n in (1, 4)] * 2
        cases += [dict(shape=C2, dtype=torch.float32, streams=n, variant="bulk", tune=t) for t in (0, 0x800, 0x600) for n in (1, 4)]
    elif len(sys.argv) > 1 and sys.argv[1] == "bulksmall":
        # where the bulk variant (8 CTAs per SM) beats the register kernel: small batches, alone and overlapped
        cases = [dict(shape=sh, dtype=dt, streams=n, variant=v, tune=t)
                 for sh in ((4, 480, 640, 4), (8, 480, 640, 4), (12, 480, 640, 4), (16, 480, 640, 4), (24, 480, 640, 4),
                            (32, 480, 640, 4), (4, 1024, 1024, 4), (8, 1024, 1024, 4), (2, 1080, 1920, 4), (32, 256, 256, 4))
                 for dt in (torch.float32, torch.bfloat16) for v, t in (("auto", 0), ("bulk", 0x800)) for n in (1, 4)] * 2
    elif len(sys.argv) > 1 and sys.argv[1] == "waves":
        # grids trimmed to whole resident waves (cap = CTAs per SM) at the headline batch and around it
        cases = [dict(shape=(b, 480, 640, 4), dtype=dt, streams=n, tune=cap << 8, stats=st)
                 for b in (8, 16) for dt in (torch.float32, torch.bfloat16) for st in (False, True)
                 for cap in (0, 8, 16, 24, 32) for n in (1, 4)] * 2
    elif len(sys.argv) > 1 and sys.argv[1] == "odd":
        # odd widths: aligned plane groups (auto, k1_window) against one pixel per thread (k1_pixel)
        cases = [dict(shape=sh, dtype=dt, variant=v, streams=n, stats=st)
                 for sh in ((8, 481, 641, 4), (64, 481, 641, 4), (32, 1080, 1918, 4), (64, 1023, 1022, 4))
                 for dt in (torch.float32, torch.bfloat16) for st in (False, True) for v in ("auto", "pixel")
                 for n in ((1, 4) if sh[0] == 8 else (1,))]
        cases += [dict(shape=(64, 481, 641, 4), dtype=torch.uint8, variant=v) for v in ("auto", "pixel")]
    elif len(sys.argv) > 1 and sys.argv[1] == "nhwc_lf":
        cases = [dict(shape=sh, dtype=dt, layout="NHWC", tune=t, streams=n) for sh in (C2, (64, 480, 640, 4), C5, C4)
                 for dt in (torch.float32, torch.uint8, torch.bfloat16) for t in (0, 0x40000)
                 for n in ((1, 4) if sh == C2 else (1,))] * 2
    elif len(sys.argv) > 1 and sys.argv[1] == "unroll2":
        cases = [dict(shape=sh, dtype=dt, tune=t) for sh in (C4, C3, C5, (256, 480, 640, 4))
                 for dt in (torch.float32, torch.bfloat16) for t in ((0, 1, 2) if dt == torch.float32 else (0, 1, 4))] * 2
    elif len(sys.argv) > 1 and sys.argv[1] == "unroll":
        cases = [dict(shape=(b, 480, 640, 4), dtype=torch.float32, streams=n, tune=t) for b in (8, 16, 32, 64)
                 for n in (1, 4, 8) for t in (0, 1, 2)] * 2
    elif len(sys.argv) > 1 and sys.argv[1] == "bulk":
        cases = []
        for shape in (C2, C4, C3, C5):
            for variant, tune in (("vector", 0), ("bulk", 0), ("bulk", 0x200), ("bulk", 0x600), ("bulk", 0x800)):
                cases.append(dict(shape=shape, dtype=torch.float32, variant=variant, tune=tune))
        cases += [dict(shape=C2, dtype=torch.float32, variant=v, streams=4) for v in ("vector", "bulk")]
        cases += [dict(shape=C4, dtype=torch.bfloat16, variant=v) for v in ("vector", "bulk")]
    elif len(sys.argv) > 1 and sys.argv[1] == "bf16":
        cases = [dict(shape=sh, dtype=torch.bfloat16, tune=t, streams=st) for sh in (C4, C2) for t in (0, 4, 1, 44)
                 for st in ((1, 4) if sh == C2 else (1,))]
        cases += [dict(shape=(64, 1024, 1024, 1), dtype=torch.float32)]
    elif len(sys.argv) > 1 and sys.argv[1] == "extra":
        cases = [dict(shape=C4, dtype=torch.uint8, layout="NHWC"), dict(shape=C4, dtype=torch.uint8, layout="NHWC", gamma=None),
                 dict(shape=C4, dtype=torch.float32, layout="NHWC"), dict(shape=C4, dtype=torch.bfloat16, layout="NHWC"),
                 dict(shape=(64, 1024, 1024, 3), dtype=torch.float32), dict(shape=(64, 1024, 1024, 3), dtype=torch.uint8, layout="NHWC"),
                 dict(shape=(64, 1024, 1024, 1), dtype=torch.float32), dict(shape=C4, dtype=torch.uint8),
                 dict(shape=C2, dtype=torch.bfloat16, streams=4), dict(shape=C2, dtype=torch.uint8, layout="NHWC", streams=4)]
    elif len(sys.argv) > 1 and sys.argv[1] == "pdl":
        cases = []
        for shape in (C2, (16, 480, 640, 4), C4):
            for dt in (torch.float32, torch.bfloat16):
                for tune in (0, 0x10000):
                    for streams in (1, 4):
                        cases.append(dict(shape=shape, dtype=dt, tune=tune, streams=streams))
    elif len(sys.argv) > 1 and sys.argv[1] == "tune2":
        cases = []
        for shape in (C4, C3, C5, (16, 480, 640, 4), (32, 480, 640, 4)):
            for unroll in (0, 8):
                for ctas in (0, 32, 64, 128, 255):
                    cases.append(dict(shape=shape, dtype=torch.float32, tune=unroll | (ctas << 8)))
    elif len(sys.argv) > 1 and sys.argv[1] == "tune":
        cases = []
        for shape in (C2, C4):
            for unroll in (0, 1, 2, 8):
                for ctas in (0, 4, 6, 16, 32):
                    cases.append(dict(shape=shape, dtype=torch.float32, tune=unroll | (ctas << 8)))
        cases += [dict(shape=C2, dtype=torch.float32, tune=t, streams=4) for t in (0, 1, 2, 0x2001, 0x2002)]
    for c in cases:
        print(json.dumps(run(**c)), flush=True)


if __name__ == "__main__":
    main()
