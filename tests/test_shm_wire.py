"""Shared-memory wire (opt-in) on the host: segment protocol, scanner support for
the reference marker, in-line resolution for consumers without a GPU, and transparency of the
datasets' host iteration.  No counterpart in the reference; the oracle for values is simply the
array that was published."""
import multiprocessing as mp
import pickle
import socket

import numpy as np
import pytest

from blendtorch import btt
from blendtorch.btt import native as N, shm
from blendtorch.btt.staging import Ring, Schema


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_segment_put_fetch_and_backpressure():
    seg = shm.Segment(slot_bytes=1000, nslots=3)
    assert seg.slot_bytes == 4096       # rounded to pages
    arrs = [np.arange(i, i + 200, dtype=np.float32).reshape(10, 20) for i in range(3)]
    marks = [seg.put(a) for a in arrs]
    assert [m[2] & 0xFFFF for m in marks] == [0, 1, 2] and all(m[2] >> 16 for m in marks) and marks[0][3] == "<f4" and marks[0][4] == (10, 20)
    with pytest.raises(N.NativeError) as e:   # every slot is held by a consumer that has not released
        seg.put(arrs[0], timeout_ms=50)
    assert e.value.code == N.BT_E_TIMEOUT
    np.testing.assert_array_equal(shm.fetch(marks[0]), arrs[0])      # releases slot 0
    m = seg.put(arrs[2], timeout_ms=50)
    assert m[2] & 0xFFFF == 0 and m[2] != marks[0][2]          # same slot, next generation
    for mk, a in ((marks[1], arrs[1]), (marks[2], arrs[2]), (m, arrs[2])):
        np.testing.assert_array_equal(shm.fetch(mk), a)
    with pytest.raises(ValueError):
        seg.put(np.zeros(5000, np.uint8))
    seg.close()
    with pytest.raises(N.NativeError):
        shm._attached.clear()
        shm.fetch(("__btshm__", seg.name + "x", 0, "|u1", (4,)))


def test_scanner_and_schema_understand_references():
    seg = shm.Segment(slot_bytes=6 * 10 * 4, nslots=4)
    rng = np.random.default_rng(3)
    frames = [rng.integers(0, 256, (6, 10, 4), dtype=np.uint8) for _ in range(3)]
    blobs = [pickle.dumps({"btid": 1, "image": seg.put(f), "frameid": i, "xy": np.full((2, 2), float(i))}, protocol=3)
             for i, f in enumerate(frames)]
    f = {x.key.decode(): x for x in N.scan_message(blobs[1])}
    assert f["image"].kind == N.BT_KIND_SHMREF and f["image"].dtype == N.BT_U8 and f["image"].ival & 0xFFFF == 1
    assert [f["image"].shape[d] for d in range(3)] == [6, 10, 4]
    assert bytes(blobs[1][f["image"].payload_off:f["image"].payload_off + f["image"].payload_len]).decode() == seg.name
    sch = Schema.from_message(blobs[0], 3)
    assert sch.native and sch.column("image").shape == (6, 10, 4) and sch.column("image").dtype == np.uint8
    ring = Ring(sch.slot_bytes, depth=1, device=None)          # no GPU: references are resolved in-line
    s = ring.acquire()
    for i, b in enumerate(blobs):
        ring.ingest(s, sch, i, b)
    c = sch.column("image")
    got = ring.view(s)[c.slot_off:c.slot_off + 3 * c.item_bytes].reshape(3, 6, 10, 4)
    np.testing.assert_array_equal(got, np.stack(frames))
    # the slots were released by the in-line copy: the producer can reuse all of them at once
    for _ in range(4):
        seg.put(frames[0], timeout_ms=50)
    ring.close()
    seg.close()


def _publish(addr, n, slots):
    with btt.DataPublisher(addr, btid=2, lingerms=5000, shm_slots=slots) as pub:
        for i in range(n):
            pub.publish(frameid=i, image=np.full((128, 256, 4), i, np.uint8), small=np.arange(4))


def test_host_iteration_is_transparent(tmp_path):
    addr = "tcp://127.0.0.1:%d" % _free_port()
    p = mp.get_context("fork").Process(target=_publish, args=(addr, 6, 2), daemon=True)
    p.start()
    try:
        ds = btt.RemoteIterableDataset([addr], max_items=6, record_path_prefix=str(tmp_path / "r"))
        items = list(ds)           # only 2 slots: each item must be released before the 3rd is sent
    finally:
        p.join(timeout=20)
    assert [it["frameid"] for it in items] == list(range(6))
    for i, it in enumerate(items):
        assert isinstance(it["image"], np.ndarray) and it["image"].shape == (128, 256, 4)
        assert (it["image"] == i).all() and it["btid"] == 2
        np.testing.assert_array_equal(it["small"], np.arange(4))       # below the threshold: in-line
    # recordings made from this wire are self-contained and reference-readable
    from oracle import btr
    r = btr.Reader(tmp_path / "r_00.btr")
    assert len(r) == 6 and (r[4]["image"] == 4).all()
