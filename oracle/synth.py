"""Deterministic synthetic inputs for the oracle side (TEST INFRASTRUCTURE).

Follows SURVEY.md 8(d): rng = default_rng(1000*config + publisher_id), uniform
uint8 RGBA frames, `xy` float64, bboxes float32 from sorted pairs, depth float32
in [0.1, 100), cube centres in the range of
examples/datagen/falling_cubes.blend.py:21.  The product package carries its own
generator (blendtorch/btt/synthetic.py) because it may not import the oracle;
tests/test_abi_and_rules.py::test_synthetic_generators_agree_with_oracle_copy checks that the two emit
identical values.
"""

import pickle

import numpy as np


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
    a = np.sort(rng.random((n, 2, 2)), axis=1)  # [lo, hi] per axis
    return np.concatenate((a[:, 0, :], a[:, 1, :]), axis=-1).astype(np.float32)


def depth(rng, h, w):
    return rng.random((h, w), dtype=np.float32) * np.float32(99.9) + np.float32(0.1)


def cube_corners(rng, n_cubes=8):
    """(n_cubes*8, 3) float64 world corners of randomly rotated unit-half-extent cubes."""
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
    """One wire message dict in the schema of examples/datagen/cube.blend.py:20-24."""
    msg = {"btid": btid, "image": frame(rng, h, w, c), "frameid": frameid}
    if config == 1:
        msg["xy"] = xy_points(rng, h, w)
    elif config == 2:
        msg["bbox"] = bboxes(rng)
    elif config == 3:
        msg["depth"] = depth(rng, h, w)
        msg["corners"] = cube_corners(rng)
    return msg


def pickled(msg, protocol=3):
    return pickle.dumps(msg, protocol=protocol)


def run_publisher(bind, btid, config, h, w, c=4, count=-1, pool=16, protocol=3):
    """Endless (count < 0) synthetic publisher for the CPU-baseline / reference arm of bench.py:
    the restated btb publisher (oracle/stream.py) fed from a pre-pickled pool so that the
    producer costs a send, not a render."""
    from .stream import Publisher

    rng = rng_for(config, btid)
    blobs = [pickled(message(rng, config, btid, i, h, w, c), protocol) for i in range(pool)]
    pub = Publisher(bind, btid=btid, send_hwm=10, lingerms=0)
    i = 0
    try:
        while count < 0 or i < count:
            pub.sock.send(blobs[i % pool], copy=False)
            i += 1
    finally:
        pub.close()


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser()
    ap.add_argument("--bind", required=True)
    ap.add_argument("--btid", type=int, default=0)
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--count", type=int, default=-1)
    ap.add_argument("--pool", type=int, default=16)
    a = ap.parse_args()
    run_publisher(a.bind, a.btid, a.config, a.height, a.width, 4, a.count, a.pool)
