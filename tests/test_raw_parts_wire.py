"""The opt-in raw-parts wire (wire.py, DataPublisher(raw_parts=True)): large arrays travel as their own
zero-copy frames of a multipart message.  Host-side behaviour: framing, scanner, the direct receiver
landing frames in their columns, mixed framings on one socket, the pyzmq paths, recordings."""
import ctypes as C
import pickle
import threading
import time

import numpy as np
import pytest
import zmq
from torch.utils.data import default_collate

from blendtorch import btt
from blendtorch.btt import native as N
from blendtorch.btt import wire
from blendtorch.btt.rx import DirectReceiver, Receiver
from blendtorch.btt.staging import Ring, Schema

from test_direct_receiver import _drain, _free_tcp, _msgs


def _raw_publisher(addr, msgs, raw=True, delay=0.0, protocol=None):
    def run():
        time.sleep(delay)
        with btt.DataPublisher(addr, btid=None, lingerms=5000, raw_parts=raw, protocol=protocol) as pub:
            for m in msgs:
                m = dict(m)
                m.pop("btid", None)
                pub.btid = msgs[0]["btid"]
                pub.publish(**m)
    t = threading.Thread(target=run)
    t.start()
    return t


def test_framing_and_scanner():
    m = _msgs(1, extra=True)[0]
    m["small"] = np.arange(6, dtype=np.int16)           # below the threshold: stays in the header
    frames = wire.pack(m)
    assert len(frames) == 3                              # header + image + depth (bbox is 128 bytes: in the header)
    frames = wire.pack(m, threshold=100)
    assert [wire.is_raw(v) for v in pickle.loads(frames[0]).values()].count(True) == 3
    back = wire.decode([frames[0]] + [np.ascontiguousarray(f).tobytes() for f in frames[1:]])
    assert set(back) == set(m)
    for k in m:
        np.testing.assert_array_equal(np.asarray(back[k]), np.asarray(m[k]))
    assert back["image"].flags.writeable                 # fresh arrays, as after pickle.loads
    # the scanner reports the references
    fields = (N.Field * 16)()
    hdr = frames[0]
    n = N.lib.bt_scan_message((C.c_uint8 * len(hdr)).from_buffer_copy(hdr), len(hdr), fields, 16)
    kinds = {fields[i].key.decode(): fields[i].kind for i in range(n)}
    assert kinds["image"] == N.BT_KIND_RAWREF and kinds["depth"] == N.BT_KIND_RAWREF and kinds["btid"] == N.BT_KIND_INT
    img = next(fields[i] for i in range(n) if fields[i].key == b"image")
    assert img.ival == 1 and img.ndim == 3 and list(img.shape[:3]) == [120, 160, 4] and img.payload_len == 0
    # a header alone cannot be ingested: the frame is missing
    sch = Schema.from_message(hdr, 2)
    slot = np.zeros(sch.slot_bytes, np.uint8)
    rc = N.lib.bt_schema_ingest(sch.handle, slot.ctypes.data_as(C.c_void_p), 0,
                                (C.c_uint8 * len(hdr)).from_buffer_copy(hdr), len(hdr))
    assert rc == N.BT_E_SCHEMA and "missing" in N.last_error()


@pytest.mark.parametrize("first_is_raw", [True, False])
def test_direct_receiver_lands_frames_and_mixes_framings(first_is_raw):
    a1, a2 = _free_tcp(), _free_tcp()
    m1, m2 = _msgs(9, seed=1, extra=True), _msgs(8, seed=2, extra=True)
    odd = dict(m1[4], image=np.zeros((60, 80, 4), np.uint8))                   # wrong shape in a raw-parts message
    t1 = _raw_publisher(a1, [odd if i == 4 else m for i, m in enumerate(m1)], raw=True,
                        delay=0.0 if first_is_raw else 0.6)
    t2 = _raw_publisher(a2, m2, raw=False, delay=0.6 if first_is_raw else 0.0)  # the reference framing
    rx = DirectReceiver([a1, a2])
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    sizes = rx.pending_parts()
    assert (len(sizes) == 3) == first_is_raw             # header + image + depth (bbox is small)
    sch = Schema.from_message(rx.pending(), 4, sizes)
    assert sch.native and sch.column("image").shape == (120, 160, 4)
    ring = Ring(sch.slot_bytes, depth=2, device=None)
    rows, odd_seen, placed = {}, [], 0
    while placed < 17:
        slot, filled = ring.acquire(), 0
        want = min(4, 17 - placed)
        waited = 0
        while filled < want:
            rc, n = rx.fill(ring, slot, sch, filled, want - filled, 200)
            filled += n
            if rc == N.BT_OK:
                break
            if rc == N.BT_E_TIMEOUT:
                waited += 200
                assert waited < 20000
                continue
            obj = wire.loads(rx.pending(), rx.pending_parts())
            odd_seen.append(obj["image"].shape)
            obj["image"] = m1[4]["image"]
            sch.ingest_object(ring.view(slot), filled, obj)
            rx.drop()
            filled += 1
        v = ring.view(slot)
        b = {c.key: v[c.slot_off:c.slot_off + filled * c.item_bytes].view(c.dtype).reshape((filled,) + c.shape).copy()
             for c in sch.columns}
        for i in range(filled):
            rows[(int(b["btid"][i]), int(b["frameid"][i]))] = {k: b[k][i] for k in b}
        placed += filled
    t1.join()
    t2.join()
    rx.close()
    ring.close()
    assert odd_seen == [(60, 80, 4)]
    assert len(rows) == 17
    for seed, ms in ((1, m1), (2, m2)):
        for m in ms:
            got = rows[(seed, m["frameid"])]
            for k in ("image", "depth", "bbox"):
                np.testing.assert_array_equal(got[k], m[k])
            assert got["score"] == m["score"]


def test_host_iteration_and_recording_of_a_raw_parts_stream(tmp_path):
    """`RemoteIterableDataset` (the reference's per-item path) reads the framing too, and the recording
    tee stores self-contained messages that the unmodified FileReader format replays."""
    addr = _free_tcp()
    msgs = _msgs(6, seed=4)
    t = _raw_publisher(addr, msgs, raw=True)
    ds = btt.RemoteIterableDataset([addr], max_items=6, record_path_prefix=str(tmp_path / "rec"))
    got = list(ds)
    t.join()
    assert [g["frameid"] for g in got] == list(range(6))
    for g, m in zip(got, msgs):
        np.testing.assert_array_equal(g["image"], m["image"])
        np.testing.assert_array_equal(g["bbox"], m["bbox"])
    replay = btt.FileDataset(str(tmp_path / "rec"))
    assert len(replay) == 6
    np.testing.assert_array_equal(replay[3]["image"], msgs[3]["image"])
    exp = default_collate(msgs)
    rep = default_collate([replay[i] for i in range(6)])
    assert set(rep) == set(exp)


def test_libzmq_loop_hands_raw_parts_messages_to_the_python_decode():
    addr = _free_tcp()
    msgs = _msgs(5, seed=6)
    t = _raw_publisher(addr, msgs, raw=True)
    ctx = zmq.Context()
    sock = ctx.socket(zmq.PULL)
    sock.setsockopt(zmq.RCVHWM, 10)
    sock.connect(addr)
    rx = Receiver(sock)
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    assert rx.more()
    sch = Schema.from_message(rx.pending(), 5, rx.pending_parts())
    ring = Ring(sch.slot_bytes, depth=1, device=None)
    slot, filled = ring.acquire(), 0
    while filled < 5:
        rc, n = rx.fill(ring, slot, sch, filled, 5 - filled, 200)
        filled += n
        if rc in (N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
            obj = wire.decode([bytes(rx.pending())] + rx.recv_rest())
            sch.ingest_object(ring.view(slot), filled, obj)
            rx.drop()
            filled += 1
        else:
            assert rc in (N.BT_OK, N.BT_E_TIMEOUT) and time.time() < deadline
    c = sch.column("image")
    img = ring.view(slot)[c.slot_off:c.slot_off + 5 * c.item_bytes].reshape((5,) + c.shape)
    for i, m in enumerate(msgs):
        np.testing.assert_array_equal(img[i], m["image"])
    t.join()
    rx.close()
    ring.close()
    sock.close()
    ctx.term()
