#!/usr/bin/env python3
"""Closing the loop: a training loop that steers its producers and learns from what comes back.

The structure of examples/densityopt/densityopt.py in the reference (lines 95-119 and 257-316): sample
simulation parameters from a distribution, send every producer its share over `btt.DuplexChannel`
together with the indices of the samples (`update_simulations`), receive the frames those parameters
produced -- each carrying the `shape_id` that ties it to its sample --, score them, and move the
distribution with a score-function gradient (the renderer is a black box).  What differs is where
the frames go: `btt.DeviceLoader` delivers them as device batches (frames flipped and normalised by K1,
the drawn cube's corners projected to a pixel box by K2), so scoring runs on the GPU without a
host-side transform, collate or copy in the loop.

Task (small enough to verify): producers draw a cube at the sampled position; the distribution over
positions is moved until the cube's projected box centre sits at a target pixel.

    python examples/steer_producers.py [--producers 2] [--steps 30]
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))
from blendtorch import btt  # noqa: E402


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


def launch_producers(n, height, width, device):
    """`n` steered producers: (data addresses, control addresses, processes)."""
    data = ["tcp://127.0.0.1:%d" % free_port() for _ in range(n)]
    ctrl = ["tcp://127.0.0.1:%d" % free_port() for _ in range(n)]
    procs = [subprocess.Popen([sys.executable, str(ROOT / "examples" / "steered_producer.py"), "--data", d, "--ctrl", c,
                               "--btid", str(i), "--height", str(height), "--width", str(width), "--device", device])
             for i, (d, c) in enumerate(zip(data, ctrl))]
    return data, ctrl, procs


def update_simulations(remotes, params, repeat=False):
    """densityopt.py:95-107: split the N samples over the R producers, send each its chunk and the
    indices of its samples; the indices come back with the frames."""
    ids = torch.arange(params.shape[0]).long()
    R = len(remotes)
    for remote, subset, subset_ids in zip(remotes, torch.chunk(params, R), torch.chunk(ids, R)):
        remote.send(shape_params=subset.cpu().numpy(), shape_ids=subset_ids.numpy(), repeat=repeat)


def run(producers=2, steps=30, batch=8, height=240, width=320, device="cuda:0", target=(200.0, 90.0), lr=0.05,
        steer=True, log=None):
    """Returns (history of distribution means and errors, seconds).  With steer=False the producers get
    their parameters once and then run free (they keep drawing those samples), and the same loop consumes
    and scores what arrives: the rate steering is compared with."""
    dev = torch.device(device)
    data, ctrl, procs = launch_producers(producers, height, width, device)
    remotes = [btt.DuplexChannel(c) for c in ctrl]
    cam = btt.Camera.from_lens((height, width), device=dev)
    mean = torch.tensor([0.0, 0.0], requires_grad=True)        # distribution over the cube's (x, y)
    log_std = torch.tensor([-0.5, -0.5], requires_grad=True)
    opt = torch.optim.Adam([mean, log_std], lr=lr)
    tgt = torch.tensor(target, device=dev, dtype=torch.float64)
    history, baseline, t0, sec = [], None, None, 0.0
    try:
        ds = btt.RemoteIterableDataset(data, max_items=batch * steps, timeoutms=30000)
        dl = btt.DeviceLoader(ds, batch_size=batch, device=dev, group=1,
                              preprocess=btt.Preprocess(flip=True, mean=127.5, std=127.5),
                              project=btt.Projection("corners", camera=cam, xy_key="corners_xy", bbox_key="box"))
        dist = torch.distributions.Normal(mean, log_std.exp())
        samples = dist.sample((batch,))
        params = torch.cat([samples, torch.full((batch, 1), 0.5)], dim=1)     # (cx, cy, half extent)
        update_simulations(remotes, params, repeat=not steer)
        for step, b in enumerate(dl):
            if t0 is None:
                t0 = time.perf_counter()            # the clock starts with the first batch (producers are up)
            # score on the device: squared pixel distance of the projected box centre from the target
            box = b["box"][:, 0]                                              # (B, 4): xmin, ymin, xmax, ymax
            centre = torch.stack(((box[:, 0] + box[:, 2]) / 2, (box[:, 1] + box[:, 3]) / 2), dim=1)
            err = ((centre - tgt) ** 2).sum(dim=1).sqrt().float().cpu()       # the step's result comes to the host
            shape_id = b["shape_id"].cpu()
            if log is not None:
                log(step, samples, shape_id, centre.cpu(), b)
            # score-function gradient (densityopt.py:291-302): the frame with shape_id i scores sample i
            base = err.mean() if baseline is None else baseline
            opt.zero_grad()
            dist = torch.distributions.Normal(mean, log_std.exp())
            logp = dist.log_prob(samples).sum(dim=1)
            (logp[shape_id] * (err - base)).mean().backward()
            opt.step()
            baseline = err.mean() if baseline is None else 0.9 * baseline + 0.1 * err.mean()
            history.append({"step": step, "mean": mean.detach().tolist(), "err_px": float(err.mean())})
            if steer and step + 1 < steps:
                dist = torch.distributions.Normal(mean.detach(), log_std.detach().exp())
                samples = dist.sample((batch,))
                params = torch.cat([samples, torch.full((batch, 1), 0.5)], dim=1)
                update_simulations(remotes, params)
        sec = time.perf_counter() - t0
        dl.close()
    finally:
        for r in remotes:
            try:
                r.send(quit=True)
            except Exception:
                pass
            r.close()
        for p in procs:
            try:
                p.wait(2)
            except Exception:      # a free-running producer sits in a blocking send: nobody reads any more
                p.terminate()
                p.wait(10)
    return history, sec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--producers", type=int, default=2)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--batch", type=int, default=8)
    a = ap.parse_args()
    hist, sec = run(a.producers, a.steps, a.batch)
    _, sec_fixed = run(a.producers, a.steps, a.batch, steer=False)
    print(json.dumps({"steps": a.steps, "producers": a.producers, "batch": a.batch,
                      "first_err_px": hist[0]["err_px"], "last_err_px": float(np.mean([h["err_px"] for h in hist[-5:]])),
                      "final_mean": hist[-1]["mean"], "steps_per_s_steered": (a.steps - 1) / sec,
                      "steps_per_s_free_running_producers": (a.steps - 1) / sec_fixed}))


if __name__ == "__main__":
    main()
