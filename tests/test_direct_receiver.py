"""The direct ZMTP receiver (bt_pull_*) against unmodified pyzmq PUSH sockets: handshake, framing,
streaming placement of payloads (large ones never enter the message buffer), fallback to the caller's
pickle.loads for messages outside the subset, several publishers, late publishers, ipc://."""
import os
import pickle
import threading
import time

import numpy as np
import pytest
import zmq
from torch.utils.data import default_collate

from blendtorch import btt
from blendtorch.btt import native as N
from blendtorch.btt.rx import DirectReceiver
from blendtorch.btt.staging import Ring, Schema


def _msgs(n, shape=(120, 160, 4), seed=0, extra=False):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        m = {"btid": seed, "image": rng.integers(0, 256, shape, dtype=np.uint8), "frameid": i,
             "bbox": rng.random((8, 4)).astype(np.float32)}
        if extra:
            m["depth"] = rng.random(shape[:2], dtype=np.float32)      # a second large payload after the first
            m["score"] = float(i) / 3
        out.append(m)
    return out


def _publisher(addr, blobs, delay=0.0):
    def run():
        time.sleep(delay)
        ctx = zmq.Context()
        s = ctx.socket(zmq.PUSH)
        s.setsockopt(zmq.SNDHWM, 10)
        s.setsockopt(zmq.LINGER, 5000)
        s.setsockopt(zmq.IMMEDIATE, 1)
        s.bind(addr)
        for b in blobs:
            s.send(b)
        s.close()
        ctx.term()
    t = threading.Thread(target=run)
    t.start()
    return t


def _free_tcp():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return "tcp://127.0.0.1:%d" % p


def _drain(rx, sch, total, B, on_odd=None):
    ring = Ring(sch.slot_bytes, depth=2, device=None)
    batches, placed, waited = [], 0, 0
    while placed < total:
        slot, filled = ring.acquire(), 0
        want = min(B, total - placed)
        while filled < want:
            rc, n = rx.fill(ring, slot, sch, filled, want - filled, 200)
            filled += n
            if rc == N.BT_OK:
                break
            if rc == N.BT_E_TIMEOUT:
                waited += 200
                assert waited < 20000, "publishers went silent"
                continue
            obj = pickle.loads(rx.pending())       # outside the subset: the reference's own decode
            if on_odd:
                obj = on_odd(obj)
            sch.ingest_object(ring.view(slot), filled, obj)
            rx.drop()
            filled += 1
        v = ring.view(slot)
        batches.append({c.key: v[c.slot_off:c.slot_off + filled * c.item_bytes].view(c.dtype).reshape((filled,) + c.shape).copy()
                        for c in sch.columns})
        placed += filled
    ring.close()
    return batches


@pytest.mark.parametrize("protocol", [3, 4, 5])
@pytest.mark.parametrize("extra", [False, True])
def test_streams_from_an_unmodified_push_socket(protocol, extra):
    msgs = _msgs(7, extra=extra)
    addr = _free_tcp()
    t = _publisher(addr, [pickle.dumps(m, protocol=protocol) for m in msgs])
    rx = DirectReceiver([addr])
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    sch = Schema.from_message(rx.pending(), 3)

    def no_fallback(obj):
        pytest.fail("protocol %d message left the streaming path: %s" % (protocol, N.last_error()))

    got = _drain(rx, sch, 7, 3, on_odd=no_fallback)
    t.join()
    rx.close()
    exp = default_collate(msgs)
    assert [len(b["frameid"]) for b in got] == [3, 3, 1]
    for key in exp:
        cat = np.concatenate([b[key] for b in got])
        assert cat.dtype == exp[key].numpy().dtype, key
        np.testing.assert_array_equal(cat, exp[key].numpy())


def test_odd_messages_small_messages_and_two_publishers():
    a1, a2 = _free_tcp(), "ipc:///tmp/bt_direct_test_%d" % os.getpid()
    m1, m2 = _msgs(6, seed=1), _msgs(5, seed=2)
    odd = dict(m1[3], image=np.zeros((60, 80, 4), np.uint8))          # wrong shape: rejected while streaming
    blobs1 = [pickle.dumps(odd if i == 3 else m, protocol=3) for i, m in enumerate(m1)]
    t1 = _publisher(a1, blobs1)
    t2 = _publisher(a2, [pickle.dumps(m, protocol=4) for m in m2], delay=0.5)     # comes up late: retried
    rx = DirectReceiver([a1, a2])
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    sch = Schema.from_message(rx.pending(), 4)
    seen_odd = []

    def fix(obj):
        seen_odd.append(obj["image"].shape)
        obj["image"] = m1[3]["image"]
        return obj

    got = _drain(rx, sch, 11, 4, on_odd=fix)
    t1.join()
    t2.join()
    rx.close()
    assert seen_odd == [(60, 80, 4)]
    rows = {}
    for b in got:
        for i in range(len(b["frameid"])):
            rows[(int(b["btid"][i]), int(b["frameid"][i]))] = (b["image"][i], b["bbox"][i])
    assert len(rows) == 11                                   # every message exactly once
    for seed, ms in ((1, m1), (2, m2)):
        for m in ms:
            img, bb = rows[(seed, m["frameid"])]
            np.testing.assert_array_equal(img, m["image"])
            np.testing.assert_array_equal(bb, m["bbox"])
    # tiny messages (no payload worth streaming) and the reference's own test message
    addr = _free_tcp()
    small = [{"btid": 0, "frameid": i, "img": np.zeros((64, 64), np.uint8)} for i in range(5)]
    t = _publisher(addr, [pickle.dumps(m, protocol=3) for m in small])
    rx = DirectReceiver([addr])
    while not rx.recv(200):
        pass
    sch = Schema.from_message(rx.pending(), 5)
    got = _drain(rx, sch, 5, 5)
    t.join()
    rx.close()
    np.testing.assert_array_equal(got[0]["frameid"], np.arange(5))
    assert got[0]["img"].shape == (5, 64, 64)


def test_rejects_what_it_does_not_speak():
    with pytest.raises(N.NativeError):
        DirectReceiver(["inproc://x"])
    rx = DirectReceiver([_free_tcp()])        # nobody there: time-out, not an error
    assert rx.recv(150) is False
    rx.close()


def test_publisher_restart_is_picked_up_again():
    """A publisher that goes away and comes back on the same address (a restarted Blender instance) is
    reconnected without the consumer doing anything, as with a ZMQ socket."""
    addr = _free_tcp()
    first, second = _msgs(3, shape=(40, 64, 4), seed=4), _msgs(4, shape=(40, 64, 4), seed=5)
    t = _publisher(addr, [pickle.dumps(m, protocol=3) for m in first])
    rx = DirectReceiver([addr])
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    sch = Schema.from_message(rx.pending(), 7)
    got = _drain(rx, sch, 3, 3)
    t.join()                                   # the first publisher has closed its socket
    t = _publisher(addr, [pickle.dumps(m, protocol=3) for m in second], delay=0.3)
    got += _drain(rx, sch, 4, 4)
    t.join()
    rx.close()
    ids = [(int(b["btid"][i]), int(b["frameid"][i])) for b in got for i in range(len(b["frameid"]))]
    assert ids == [(4, 0), (4, 1), (4, 2), (5, 0), (5, 1), (5, 2), (5, 3)]
    np.testing.assert_array_equal(got[1]["image"][3], second[3]["image"])


@pytest.mark.parametrize("shape", [(8, 8, 4), (45, 47, 4), (90, 91, 4), (128, 128, 4)])
def test_read_ahead_keeps_every_byte_in_place(shape):
    """Frames from far below to far above the endpoint's read-ahead buffer (8 KiB): several messages
    per socket read, messages that straddle a read, payloads that start in the buffer and finish in the
    slot, an odd message in the middle of a burst, three publishers served in turn.  Every message
    arrives once, bit for bit."""
    per_pub, pubs, B = 40, 3, 16
    addrs = [_free_tcp() for _ in range(pubs)]
    all_msgs, threads = {}, []
    for k, addr in enumerate(addrs):
        ms = _msgs(per_pub, shape=shape, seed=10 + k)
        blobs = [pickle.dumps(m, protocol=3 + (k % 3)) for m in ms]
        if k == 1:   # a message outside the schema in the middle of the stream
            blobs[17] = pickle.dumps(dict(ms[17], image=ms[17]["image"][:-1]), protocol=3)
        for m in ms:
            all_msgs[(10 + k, m["frameid"])] = m
        threads.append(_publisher(addr, blobs, delay=0.2 * k))
    rx = DirectReceiver(addrs)
    deadline = time.time() + 20
    while not rx.recv(200):
        assert time.time() < deadline
    sch = Schema.from_message(rx.pending(), B)
    odd = []

    def fix(obj):
        odd.append(obj["image"].shape)
        obj["image"] = all_msgs[(obj["btid"], obj["frameid"])]["image"]
        return obj

    got = _drain(rx, sch, per_pub * pubs, B, on_odd=fix)
    for t in threads:
        t.join()
    rx.close()
    assert odd == [(shape[0] - 1,) + shape[1:]]
    seen = set()
    for b in got:
        for i in range(len(b["frameid"])):
            key = (int(b["btid"][i]), int(b["frameid"][i]))
            assert key not in seen
            seen.add(key)
            np.testing.assert_array_equal(b["image"][i], all_msgs[key]["image"])
            np.testing.assert_array_equal(b["bbox"][i], all_msgs[key]["bbox"])
    assert len(seen) == per_pub * pubs
