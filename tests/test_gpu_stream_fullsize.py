"""The timed path is the tested path: BASELINE-size messages (C2 640x480x4 = 1.2 MB, C4 1024x1024x4 = 4 MB)
through `DeviceLoader` exactly as bench.py drives it -- several publishers, 1/4/8 receivers, tcp and ipc,
pickle protocols 3 and 4, the direct ZMTP endpoint and the pyzmq socket, Python/pyzmq publishers
(unmodified PUSH sockets, pkg_blender/blendtorch/btb/publisher.py:21-43) and the native publisher, the
raw-parts and shared-memory wires, and replay from a page-locked copy at C4 -- with every delivered
frame, annotation and K1 output compared (semantics: pkg_pytorch/blendtorch/btt/dataset.py:97-106;
protocol of tests/test_dataset.py:11-34).  At these sizes a payload spans many recv() calls into the
pinned slot (csrc/zmtp.cpp receive_streaming), which the small-frame tests never exercise.

Comparison: raw frames and annotations byte for byte against the generator's pool (on the GPU), K1's
output bit for bit against the oracle's 256-entry table gathered by torch for every batch, and against
the oracle itself (numpy, per frame) for the first batch of each case."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest
import torch

from blendtorch import btt
from blendtorch.btt import synthetic as S
from oracle import preprocess as opre

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
POOL = 8
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
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


def _addresses(n, transport, tag):
    if transport == "ipc":
        return ["ipc:///tmp/bt_fs_%d_%s_%d" % (os.getpid(), tag, i) for i in range(n)]
    return ["tcp://127.0.0.1:%d" % _free_port() for _ in range(n)]


def _spawn(addrs, config, h, w, protocol, count, native=False, extra=()):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([ROOT, os.path.join(ROOT, "pytorch-blender_b200"), env.get("PYTHONPATH", "")])
    base = [sys.executable, "-m", "blendtorch.btt.synthetic", "--config", str(config), "--height", str(h), "--width", str(w),
            "--pool", str(POOL), "--protocol", str(protocol), "--count", str(count), *extra]
    if native:
        cmd = base + ["--native", "--threads", "2", "--btid", "0"]
        for a in addrs:
            cmd += ["--bind", a]
        return [subprocess.Popen(cmd, env=env)]
    return [subprocess.Popen(base + ["--bind", a, "--btid", str(i)], env=env) for i, a in enumerate(addrs)]


def _stop(procs):
    for p in procs:
        if p.poll() not in (None, 0):      # a publisher that died on its own: say so next to whatever failed
            print("publisher %s exited with %s before the test ended" % (p.args[-6:], p.returncode))
        p.terminate()
    for p in procs:
        try:
            p.wait(10)
        except Exception:
            p.kill()


class _Expected:
    """The publishers' pools on the GPU, and the oracle's per-byte table for the fused transform."""

    def __init__(self, config, pubs, h, w):
        imgs, ann = [], []
        self.ann_key = {1: "xy", 2: "bbox"}.get(config)
        for btid in range(pubs):
            rng = S.rng_for(config, btid)
            msgs = [S.message(rng, config, btid, i, h, w) for i in range(POOL)]
            imgs.append(np.stack([m["image"] for m in msgs]))
            if self.ann_key:
                ann.append(np.stack([m[self.ann_key] for m in msgs]))
        self.images = torch.from_numpy(np.stack(imgs)).to(DEV)            # (pubs, POOL, H, W, 4)
        self.ann = torch.from_numpy(np.stack(ann)).to(DEV) if ann else None
        ramp = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(4, axis=2)
        lut = opre.preprocess_frame(ramp, gamma=2.2, mean=127.5, std=127.5)[:, 0, :]   # oracle on all 256 bytes: (4, 256)
        self.lut = torch.from_numpy(lut).to(DEV)

    def transformed(self, raw):
        idx = raw.long().flip(1)
        return torch.stack([self.lut[c][idx[..., c]] for c in range(4)], dim=1)


SPEC = dict(flip=True, gamma=2.2, mean=127.5, std=127.5)


def _consume(dl, exp, total, oracle_batches=1, exactly_once=None):
    seen = {}
    n = 0
    for k, b in enumerate(dl):
        btid, fid = b["btid"], b["frameid"]
        raw = b["image_raw"]
        want_raw = exp.images[btid, fid % POOL]
        assert torch.equal(raw, want_raw), "batch %d: a delivered frame differs from the one published" % k
        if exp.ann is not None:
            assert torch.equal(b[exp.ann_key], exp.ann[btid, fid % POOL])
        assert torch.equal(b["image"], exp.transformed(raw)), "batch %d: K1 output differs from the oracle's table" % k
        if k < oracle_batches:
            want = opre.preprocess_batch(list(raw.cpu().numpy()), **SPEC)
            np.testing.assert_array_equal(b["image"].cpu().numpy().view(np.uint32), want.view(np.uint32))
        for key in zip(btid.tolist(), fid.tolist()):
            seen[key] = seen.get(key, 0) + 1
        n += len(btid)
    assert n == total
    if exactly_once is not None:        # finite publishers, one receiver: every message exactly once
        pubs, per_pub = exactly_once
        assert seen == {(p, f): per_pub // POOL for p in range(pubs) for f in range(POOL)}
    return seen


CASES = [
    # name, (h, w), publishers, messages per publisher, batch, receivers, transport, rx, protocol, native publisher
    ("C2-tcp-direct-r1-p3", (480, 640), 5, 200, 8, 1, "tcp", "direct", 3, False),
    ("C2-tcp-direct-r4-p4", (480, 640), 5, 200, 8, 4, "tcp", "direct", 4, False),
    ("C2-tcp-direct-r8-p3", (480, 640), 5, 200, 8, 8, "tcp", "direct", 3, False),
    ("C2-ipc-direct-r4-p3", (480, 640), 5, 200, 8, 4, "ipc", "direct", 3, False),
    ("C2-ipc-direct-r8-p4", (480, 640), 5, 200, 8, 8, "ipc", "direct", 4, False),
    ("C2-tcp-zmq-r1-p4", (480, 640), 5, 200, 8, 1, "tcp", "zmq", 4, False),
    ("C2-tcp-zmq-r4-p3", (480, 640), 5, 200, 8, 4, "tcp", "zmq", 3, False),
    ("C2-ipc-zmq-r8-p3", (480, 640), 5, 200, 8, 8, "ipc", "zmq", 3, False),
    ("C2-tcp-direct-r4-p3-native", (480, 640), 5, 200, 8, 4, "tcp", "direct", 3, True),
    ("C2-ipc-direct-r8-p4-native", (480, 640), 5, 200, 8, 8, "ipc", "direct", 4, True),
    ("C2-tcp-zmq-r1-p3-native", (480, 640), 5, 200, 8, 1, "tcp", "zmq", 3, True),
    ("C4-tcp-direct-r1-p3", (1024, 1024), 3, 200, 64, 1, "tcp", "direct", 3, False),
    ("C4-tcp-direct-r4-p4", (1024, 1024), 3, 200, 64, 4, "tcp", "direct", 4, False),
    ("C4-ipc-direct-r8-p3-native", (1024, 1024), 3, 200, 64, 8, "ipc", "direct", 3, True),
    ("C4-tcp-zmq-r4-p3", (1024, 1024), 3, 200, 64, 4, "tcp", "zmq", 3, False),
]


@pytest.mark.parametrize("name,shape,pubs,per_pub,batch,receivers,transport,rx,protocol,native", CASES, ids=[c[0] for c in CASES])
def test_full_size_stream_parity(name, shape, pubs, per_pub, batch, receivers, transport, rx, protocol, native):
    h, w = shape
    config = 2
    exp = _Expected(config, pubs, h, w)
    addrs = _addresses(pubs, transport, name.replace("-", ""))
    one = receivers == 1
    # one receiver: finite publishers, every message must arrive exactly once.  Several receivers each
    # take max_items // receivers (dataset.py:97) from whichever publisher serves them, so publishers
    # run on (messages queued to a receiver that has its share are dropped with its socket, as in the
    # reference) and what is checked is every delivered item.
    procs = _spawn(addrs, config, h, w, protocol, per_pub if one else -1, native)
    try:
        total = pubs * per_pub
        total -= total % (receivers * batch) if not one else 0
        ds = btt.RemoteIterableDataset(addrs, max_items=total, timeoutms=30000)
        dl = btt.DeviceLoader(ds, batch_size=batch, device=DEV, num_receivers=receivers, keep_raw=True, transport=rx,
                              ring_depth=3, prefetch=2, preprocess=btt.Preprocess(**SPEC))
        _consume(dl, exp, total, exactly_once=(pubs, per_pub) if one else None)
        assert dl.stats["compat_items"] == 0          # everything went through the native scanner
        dl.close()
    finally:
        _stop(procs)


@pytest.mark.parametrize("wire", ["raw-parts", "shm", "shm-native"])
@pytest.mark.parametrize("shape,pubs,batch", [((480, 640), 5, 8), ((1024, 1024), 3, 64)], ids=["C2", "C4"])
def test_full_size_opt_in_wires(wire, shape, pubs, batch):
    h, w = shape
    exp = _Expected(2, pubs, h, w)
    addrs = _addresses(pubs, "tcp", wire.replace("-", ""))
    extra = ("--raw-parts",) if wire == "raw-parts" else ("--shm-slots", "24")
    procs = _spawn(addrs, 2, h, w, 3, -1, native=wire.endswith("native"), extra=extra)
    try:
        total = 200 * pubs
        total -= total % (2 * batch)
        ds = btt.RemoteIterableDataset(addrs, max_items=total, timeoutms=30000)
        dl = btt.DeviceLoader(ds, batch_size=batch, device=DEV, num_receivers=2, keep_raw=True, ring_depth=3, prefetch=2,
                              preprocess=btt.Preprocess(**SPEC))
        _consume(dl, exp, total)
        assert dl.stats["compat_items"] == 0
        dl.close()
    finally:
        _stop(procs)


def test_full_size_replay_from_page_locked_copy_c4(tmp_path):
    """cache='host' at C4: every epoch assembled by DMA gather from the pinned recording; same frames as
    the recording, same K1 bits, for a shuffled order."""
    h, w, batch, frames = 1024, 1024, 64, 96
    S.write_recordings(str(tmp_path / "c4"), 4, 2, frames, h, w, pool=POOL)
    exp = _Expected(4, 2, h, w)
    ds = btt.FileDataset(str(tmp_path / "c4"))
    for cache in (None, "host"):
        dl = btt.DeviceLoader(ds, batch_size=batch, shuffle=True, device=DEV, keep_raw=True, cache=cache, seed=5,
                              preprocess=btt.Preprocess(**SPEC))
        for epoch in range(2):
            seen = _consume(dl, exp, 2 * frames, oracle_batches=0)
            assert sum(seen.values()) == 2 * frames
        dl.close()
