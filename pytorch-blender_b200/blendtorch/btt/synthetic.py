"""Synthetic producers and recordings in the reference's wire/record format.

No Blender is needed (or available) to exercise the ingest path: frames are random uint8 RGBA
images with the annotation fields of the reference's example scenes
(examples/datagen/cube.blend.py:20-24: image, xy, frameid + btid;
examples/datagen/falling_cubes.blend.py:28-32), generated deterministically as SURVEY.md 8(d)
prescribes: ``rng = default_rng(1000*config + publisher_id)``.

    python -m blendtorch.btt.synthetic --bind tcp://127.0.0.1:12000 --btid 0 --config 2 \
        --height 480 --width 640 --count 512

runs one publisher process; `write_recordings` writes `.btr` files with `FileRecorder`.
"""
import argparse
import pickle
import sys
import time

import numpy as np

from .file import FileRecorder
from .publisher import DataPublisher

# BASELINE.json configs -> (height, width, channels)
CONFIG_SHAPES = {1: (480, 640, 4), 2: (480, 640, 4), 3: (1080, 1920, 4), 4: (1024, 1024, 4), 5: (128, 128, 4)}


def rng_for(config, publisher):
    return np.random.default_rng(1000 * config + publisher)


def frame(rng, h, w, c=4):
    return rng.integers(0, 256, (h, w, c), dtype=np.uint8)


def gradient_frame(h, w, c=4):
    y, x, k = np.meshgrid(np.arange(h), np.arange(w), np.arange(c), indexing="ij")
    return ((y * w + x + k) % 256).astype(np.uint8)


def xy_points(rng, h, w, n=8):
    return rng.random((n, 2)) * np.array([w, h], dtype=np.float64)


def bboxes(rng, n=8):
    a = np.sort(rng.random((n, 2, 2)), axis=1)
    return np.concatenate((a[:, 0, :], a[:, 1, :]), axis=-1).astype(np.float32)


def depth(rng, h, w):
    return rng.random((h, w), dtype=np.float32) * np.float32(99.9) + np.float32(0.1)


def cube_corners(rng, n_cubes=8):
    centres = rng.uniform((-3, -3, 6), (3, 3, 12), size=(n_cubes, 3))
    eul = rng.uniform(-np.pi, np.pi, size=(n_cubes, 3))
    base = np.array([[x, y, z] for x in (1.0, -1.0) for y in (1.0, -1.0) for z in (1.0, -1.0)])
    out = np.empty((n_cubes, 8, 3))
    for i in range(n_cubes):
        cx, cy, cz = np.cos(eul[i])
        sx, sy, sz = np.sin(eul[i])
        rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
        ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
        rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
        out[i] = base @ (rz @ ry @ rx).T + centres[i]
    return out.reshape(-1, 3)


def message(rng, config, btid, frameid, h, w, c=4):
    msg = {"btid": btid, "image": frame(rng, h, w, c), "frameid": frameid}
    if config == 1:
        msg["xy"] = xy_points(rng, h, w)
    elif config == 2:
        msg["bbox"] = bboxes(rng)
    elif config == 3:
        msg["depth"] = depth(rng, h, w)
        msg["corners"] = cube_corners(rng)
    return msg


def pickled_pool(config, btid, h, w, c=4, pool=16, protocol=3):
    """`pool` distinct pre-pickled messages (publishers cycle through them so that the producer
    side costs a send, not a render + pickle -- the producer is not what is measured)."""
    rng = rng_for(config, btid)
    return [pickle.dumps(message(rng, config, btid, i, h, w, c), protocol=protocol) for i in range(pool)]


def write_recordings(prefix, config, n_files, frames_per_file, h=None, w=None, c=4, protocol=3, pool=None):
    """Write `n_files` recordings ``{prefix}_{k:02d}.btr`` the way btt/dataset.py:100-103 does
    (raw pickled messages, ``is_pickled=True``).  With `pool`, only that many distinct frames are
    generated per file and cycled (frameid still counts up in the first `pool` messages)."""
    if h is None:
        h, w, c = CONFIG_SHAPES[config]
    names = []
    for k in range(n_files):
        name = FileRecorder.filename(prefix, k)
        rng = rng_for(config, k)
        cache = []
        with FileRecorder(name, max_messages=frames_per_file) as rec:
            for i in range(frames_per_file):
                if pool and i >= pool:
                    data = cache[i % pool]
                else:
                    data = pickle.dumps(message(rng, config, k, i, h, w, c), protocol=protocol)
                    if pool:
                        cache.append(data)
                rec.save(data, is_pickled=True)
        names.append(name)
    return names


def run_publisher(bind, btid, config, h, w, c=4, count=-1, pool=16, protocol=3, send_hwm=10,
                  lingerms=5000, rate=0.0, shm_slots=0, raw_parts=False):
    """Publish `count` messages (-1: until killed) from a pre-generated pool: pre-pickled for the
    reference wire, arrays copied into shared-memory slots for the shared-memory wire, arrays sent as
    their own frames for the raw-parts wire."""
    pub = DataPublisher(bind, btid=btid, send_hwm=send_hwm, lingerms=lingerms, protocol=protocol,
                        shm_slots=shm_slots, raw_parts=raw_parts)
    by_object = shm_slots > 0 or raw_parts
    if by_object:
        rng = rng_for(config, btid)
        objs = [message(rng, config, btid, i, h, w, c) for i in range(pool)]
        for o in objs:
            o.pop("btid")
    else:
        msgs = pickled_pool(config, btid, h, w, c, pool, protocol)
    i = 0
    try:
        while count < 0 or i < count:
            if by_object:
                pub.publish(**objs[i % pool])
            else:
                pub.publish_raw(msgs[i % len(msgs)], copy=False)
            i += 1
            if rate > 0:
                time.sleep(1.0 / rate)
    finally:
        pub.close()
    return i


_TICKET_SENTINEL = 0x5A5A5A5A


def shm_templates(objs, segment, protocol=3):
    """For the native publisher's shared-memory mode: per message the pickled dict in which the largest
    array is replaced by its shm.py marker, the array itself, and the byte offset of the marker's
    ticket (a 4-byte BININT the publisher patches per send)."""
    from . import shm
    tmpls, payloads, offs = [], [], []
    needle = b"J" + _TICKET_SENTINEL.to_bytes(4, "little")
    for o in objs:
        o = dict(o)
        key = max((k for k, v in o.items() if isinstance(v, np.ndarray)), key=lambda k: o[k].nbytes)
        a = np.ascontiguousarray(o[key])
        o[key] = (shm.MARKER, segment.name, _TICKET_SENTINEL, a.dtype.str, tuple(a.shape))
        blob = pickle.dumps(o, protocol=protocol)
        at = blob.find(needle)
        assert at >= 0 and blob.find(needle, at + 1) < 0
        tmpls.append(blob)
        payloads.append(a)
        offs.append(at + 1)
    return tmpls, payloads, offs


def run_native_publishers(binds, btids, config, h, w, c=4, count=-1, pool=16, protocol=3, threads=1,
                          shm_slots=0, ready=None, stop=None, in_place=False):
    """All publishers of one consumer in this process: one `NativePublisher` endpoint per bind address,
    cycling the same deterministic pools as `run_publisher`, on `threads` sender threads.  Blocks until
    every endpoint has sent `count` messages, or (count < 0) until `stop` (a threading.Event) is set.
    in_place (with shm_slots): the frames are written into the segment's slots once, before the first
    message, and published from where they stand -- a producer that renders straight into shared
    memory, so that no producer-side copy is part of what a consumer benchmark measures."""
    from .publisher import NativePublisher
    pub = NativePublisher(list(binds))
    segs = []
    try:
        if shm_slots > 0:
            from . import shm
            pools, payloads, offs = [], [], []
            for btid in btids:
                rng = rng_for(config, btid)
                objs = [message(rng, config, btid, i, h, w, c) for i in range(pool)]
                if in_place:   # slot k holds frame k % pool for good; one template per slot
                    objs = [objs[k % pool] for k in range(shm_slots)]
                big = max(v.nbytes for v in objs[0].values() if isinstance(v, np.ndarray))
                seg = shm.Segment(big, shm_slots)
                seg.set_lease(0)
                segs.append(seg)
                t, p, o = shm_templates(objs, seg, protocol=protocol)
                if in_place:
                    for k, a in enumerate(p):
                        seg.view(k)[:a.nbytes] = a.reshape(-1).view(np.uint8)
                pools.append(t)
                payloads.append(p)
                offs.append(o)
            pub.start(pools, messages=count, threads=threads, shm=(segs, None if in_place else payloads, offs))
        else:
            pools = [pickled_pool(config, btid, h, w, c, pool, protocol) for btid in btids]
            pub.start(pools, messages=count, threads=threads)
        if ready is not None:
            ready(pub.addresses)
        while not pub.join(0.2):
            if stop is not None and stop.is_set():
                break
        sent = pub.sent
        pub.stop()
        return sent
    finally:
        pub.close()
        for seg in segs:
            seg.drain(2000)
            seg.close()


def main(argv=None):
    ap = argparse.ArgumentParser(description="synthetic blendtorch publisher")
    ap.add_argument("--bind", required=True, action="append",
                    help="address to bind; several (with --native) = several publishers in this process")
    ap.add_argument("--native", action="store_true", help="the library's own PUSH endpoints instead of pyzmq")
    ap.add_argument("--threads", type=int, default=1, help="sender threads (--native)")
    ap.add_argument("--cpus", default="", help="comma-separated cpu ids to run on (sched_setaffinity)")
    ap.add_argument("--btid", type=int, default=0)
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--channels", type=int, default=4)
    ap.add_argument("--count", type=int, default=-1)
    ap.add_argument("--pool", type=int, default=16)
    ap.add_argument("--protocol", type=int, default=3)
    ap.add_argument("--send-hwm", type=int, default=10)
    ap.add_argument("--shm-slots", type=int, default=0)
    ap.add_argument("--raw-parts", action="store_true")
    ap.add_argument("--in-place", action="store_true", help="--native --shm-slots: publish frames from where they stand")
    a = ap.parse_args(argv)
    h, w, c = CONFIG_SHAPES.get(a.config, (480, 640, 4))
    h, w, c = a.height or h, a.width or w, a.channels or c
    import signal
    signal.signal(signal.SIGTERM, lambda *_: sys.exit(0))  # run `finally` blocks (unlink shared memory)
    if a.cpus:
        import os
        os.sched_setaffinity(0, {int(x) for x in a.cpus.split(",")})
    if a.native:
        btids = list(range(a.btid, a.btid + len(a.bind)))
        run_native_publishers(a.bind, btids, a.config, h, w, c, a.count, a.pool, a.protocol, a.threads, a.shm_slots,
                              in_place=a.in_place)
        return 0
    if len(a.bind) != 1:
        ap.error("several --bind addresses need --native")
    a.bind = a.bind[0]
    run_publisher(a.bind, a.btid, a.config, h, w, c, a.count, a.pool, a.protocol, a.send_hwm, shm_slots=a.shm_slots,
                  raw_parts=a.raw_parts)


if __name__ == "__main__":
    sys.exit(main())
