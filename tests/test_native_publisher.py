"""The native publisher (bt_push_*, csrc/push.cpp): a ZMTP PUSH endpoint without libzmq, checked against
unmodified pyzmq PULL sockets (the reference's consumer side, btt/dataset.py:73-78), against the
library's own direct receiver, and for the socket semantics it keeps from btb/publisher.py:21-27
(messages only to completed connections, load balancing over consumers, blocking without one)."""
import os
import pickle
import subprocess
import sys
import threading
import time

import numpy as np
import pytest
import zmq

from blendtorch import btt
from blendtorch.btt import native as N, shm, synthetic as S, wire
from blendtorch.btt.rx import DirectReceiver
from blendtorch.btt.staging import Ring, Schema


def _pull(ctx, addrs, hwm=10):
    s = ctx.socket(zmq.PULL)
    s.setsockopt(zmq.RCVHWM, hwm)
    s.setsockopt(zmq.LINGER, 0)
    for a in addrs:
        s.connect(a)
    return s


def _wait_peers(pub, endpoint, n, seconds=5.0):
    t0 = time.time()
    while pub.peers(endpoint) < n:
        assert time.time() - t0 < seconds, "consumer did not connect"
        time.sleep(0.01)


@pytest.mark.parametrize("bind", ["tcp://127.0.0.1:*", "ipc:///tmp/bt_np_test_%d" % os.getpid()])
def test_pyzmq_pull_receives_identical_messages_in_order(bind):
    rng = np.random.default_rng(0)
    blobs = [pickle.dumps({"btid": 0, "frameid": i, "image": rng.integers(0, 256, (60 + i, 80, 4), dtype=np.uint8)},
                          protocol=3 + i % 2) for i in range(12)]
    blobs += [b"", b"x", bytes(255), bytes(256)]          # short and long frame headers, empty message
    with btt.NativePublisher(bind) as pub:
        assert pub.addresses[0].startswith(bind[:6]) and not pub.addresses[0].endswith("*")
        ctx = zmq.Context()
        sock = _pull(ctx, pub.addresses)
        _wait_peers(pub, 0, 1)
        for b in blobs:
            pub.publish_raw(b, timeoutms=5000)
        got = [sock.recv() for _ in blobs]                # one consumer: in order (examples/datagen/Readme.md:175)
        assert got == blobs
        sock.close()
        ctx.term()


def test_blocks_without_a_consumer_and_only_completed_connections_get_messages():
    with btt.NativePublisher("tcp://127.0.0.1:*") as pub:
        assert pub.peers() == 0
        t0 = time.time()
        with pytest.raises(TimeoutError):                  # IMMEDIATE=1 + blocking send of publisher.py:24-26
            pub.publish_raw(b"nobody listens", timeoutms=100)
        assert 0.09 < time.time() - t0 < 2.0
        ctx = zmq.Context()
        sock = _pull(ctx, pub.addresses)
        pub.publish_raw(b"now somebody does", timeoutms=5000)   # waits for the handshake, then sends
        assert sock.recv() == b"now somebody does"
        sock.close()
        ctx.term()


def test_raw_parts_messages_are_multipart_for_libzmq():
    frame = np.arange(48 * 64 * 4, dtype=np.uint8).reshape(48, 64, 4)
    parts = wire.pack({"btid": 3, "image": frame, "frameid": 7})
    assert len(parts) == 2
    with btt.NativePublisher("tcp://127.0.0.1:*") as pub:
        ctx = zmq.Context()
        sock = _pull(ctx, pub.addresses)
        _wait_peers(pub, 0, 1)
        pub.publish_parts(parts, timeoutms=5000)
        got = sock.recv_multipart()
        assert len(got) == 2
        obj = wire.decode(got)
        assert obj["btid"] == 3 and obj["frameid"] == 7
        np.testing.assert_array_equal(obj["image"], frame)
        sock.close()
        ctx.term()


def test_pools_are_balanced_over_consumers_exactly_once_and_survive_a_consumer_leaving():
    pools = [[pickle.dumps({"btid": e, "frameid": i, "image": np.full((64, 64, 4), 16 * e + i, np.uint8)}, protocol=3)
              for i in range(8)] for e in range(3)]
    per_ep = 40
    with btt.NativePublisher(["tcp://127.0.0.1:*"] * 3) as pub:
        ctx = zmq.Context()
        a, b = _pull(ctx, pub.addresses, hwm=2), _pull(ctx, pub.addresses, hwm=2)
        for e in range(3):
            _wait_peers(pub, e, 2)
        pub.start(pools, messages=per_ep, threads=2)
        seen = {e: [] for e in range(3)}
        counts = [0, 0]
        poller = zmq.Poller()
        poller.register(a, zmq.POLLIN)
        poller.register(b, zmq.POLLIN)
        total = 0
        while total < 3 * per_ep:
            ready = dict(poller.poll(5000))
            assert ready, "publisher stalled after %d messages" % total
            for k, s in enumerate((a, b)):
                if s in ready:
                    m = pickle.loads(s.recv())
                    assert (m["image"] == 16 * m["btid"] + m["frameid"]).all()
                    seen[m["btid"]].append(m["frameid"])
                    counts[k] += 1
                    total += 1
        assert pub.join(5) and pub.sent == 3 * per_ep
        for e in range(3):                                  # every message of every publisher exactly once
            assert sorted(seen[e]) == sorted(i % 8 for i in range(per_ep))
        assert min(counts) > 0                              # both consumers were served
        pub.stop()
        # one consumer leaves: an endless stream keeps flowing to the other
        b.close()
        pub.start(pools, messages=-1, threads=1)
        for _ in range(30):
            pickle.loads(a.recv())
        pub.stop()
        a.close()
        ctx.term()


def test_direct_receiver_takes_native_publishers_messages():
    """Library to library: bt_push_* -> bt_pull_fill, payloads streamed into the staging slot."""
    H, W, n_ep, per_ep, B = 120, 160, 2, 12, 4
    pools = [S.pickled_pool(2, e, H, W, pool=6, protocol=3 + e) for e in range(n_ep)]
    with btt.NativePublisher(["ipc:///tmp/bt_np_dr_%d_%d" % (os.getpid(), e) for e in range(n_ep)]) as pub:
        pub.start(pools, messages=per_ep, threads=2)
        rx = DirectReceiver(pub.addresses)
        assert rx.recv(5000)
        sch = Schema.from_message(rx.pending(), B)
        ring = Ring(sch.slot_bytes, depth=1, device=None)
        got, placed = {}, 0
        while placed < n_ep * per_ep:
            slot, filled = ring.acquire(), 0
            while filled < B:
                rc, n = rx.fill(ring, slot, sch, filled, B - filled, 2000)
                assert rc in (N.BT_OK,), N.last_error()
                filled += n
            view = ring.view(slot)
            cols = {c.key: view[c.slot_off:c.slot_off + B * c.item_bytes].view(c.dtype).reshape((B,) + c.shape) for c in sch.columns}
            for i in range(B):
                got.setdefault(int(cols["btid"][i]), []).append((int(cols["frameid"][i]), cols["image"][i].copy(), cols["bbox"][i].copy()))
            placed += B
        rx.close()
        for e in range(n_ep):
            rng = S.rng_for(2, e)
            want = [S.message(rng, 2, e, i, H, W) for i in range(6)]
            assert [f for f, _, _ in got[e]] == [i % 6 for i in range(per_ep)]      # in order per publisher
            for f, img, bbox in got[e]:
                np.testing.assert_array_equal(img, want[f]["image"])
                np.testing.assert_array_equal(bbox, want[f]["bbox"])


@pytest.mark.parametrize("in_place", [False, True])
def test_shared_memory_mode_and_the_dataset_host_path(in_place):
    """run_native_publishers(shm_slots=N): payloads through shared memory, tickets patched into the pickled
    template; the datasets' host iteration resolves them like any shared-memory message.  in_place: the
    frames stand in their slots (a producer that renders into shared memory), no copy per message."""
    H, W = 64, 96
    binds = ["tcp://127.0.0.1:*"] * 2
    addrs, stop = [], threading.Event()
    th = threading.Thread(target=S.run_native_publishers, args=(binds, [0, 1], 2, H, W),
                          kwargs=dict(count=-1, pool=4, threads=2, shm_slots=6, ready=addrs.extend, stop=stop, in_place=in_place))
    th.start()
    try:
        t0 = time.time()
        while len(addrs) < 2:
            assert time.time() - t0 < 10
            time.sleep(0.01)
        ds = btt.RemoteIterableDataset(addrs, max_items=24, timeoutms=10000)
        n = 0
        want = {e: [S.message(r, 2, e, i, H, W) for r in [S.rng_for(2, e)] for i in range(4)] for e in (0, 1)}
        for item in ds:
            w = want[item["btid"]][item["frameid"]]
            np.testing.assert_array_equal(item["image"], w["image"])
            np.testing.assert_array_equal(item["bbox"], w["bbox"])
            n += 1
        assert n == 24
    finally:
        stop.set()
        th.join(10)


def test_slot_generations_turn_a_reclaimed_slot_into_an_error_not_into_wrong_data():
    """a slot taken back by the producer (held past the lease) must not be read, or freed,
    through the stale reference."""
    seg = shm.Segment(slot_bytes=4096, nslots=1)
    seg.set_lease(30)
    a, b = np.full(1000, 1, np.uint8), np.full(1000, 2, np.uint8)
    old = seg.put(a)
    assert old[2] & 0xFFFF == 0 and old[2] >> 16 >= 1
    with pytest.raises(N.NativeError):
        seg.put(b, timeout_ms=5)                            # within the lease: the slot is the consumer's
    time.sleep(0.06)
    new = seg.put(b, timeout_ms=100)                        # past it: taken back, new generation
    assert new[2] & 0xFFFF == 0 and new[2] != old[2]
    with pytest.raises(N.NativeError) as e:
        shm.fetch(old)                                      # the slow consumer finds out
    assert e.value.code == N.BT_E_IO and "taken back" in str(e.value)
    assert N.lib.bt_shm_release(seg.handle, old[2]) == N.BT_E_IO     # and cannot free it under the new owner
    with pytest.raises(N.NativeError):
        seg.put(a, timeout_ms=5)
    np.testing.assert_array_equal(shm.fetch(new), b)        # the rightful holder is unaffected
    seg.set_lease(0)                                        # 0: never take slots back
    held = seg.put(a)
    time.sleep(0.05)
    with pytest.raises(N.NativeError):
        seg.put(b, timeout_ms=5)
    np.testing.assert_array_equal(shm.fetch(held), a)
    seg.close()


def test_command_line_native_publisher_feeds_the_reference_style_dataset(tmp_path):
    addrs = ["ipc://%s/p%d" % (tmp_path, i) for i in range(2)]
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join(sys.path)
    cmd = [sys.executable, "-m", "blendtorch.btt.synthetic", "--native", "--threads", "2", "--config", "1",
           "--height", "32", "--width", "48", "--count", "10", "--pool", "5"]
    for a in addrs:
        cmd += ["--bind", a]
    proc = subprocess.Popen(cmd, env=env)
    try:
        ds = btt.RemoteIterableDataset(addrs, max_items=20, timeoutms=20000)
        items = list(ds)
        assert sorted((it["btid"], it["frameid"]) for it in items) == sorted((e, i % 5) for e in (0, 1) for i in range(10))
        assert items[0]["image"].shape == (32, 48, 4) and items[0]["xy"].shape == (8, 2)
        assert proc.wait(10) == 0
    finally:
        proc.kill()


def test_raw_parts_from_the_native_publisher_land_in_the_direct_receivers_columns():
    """Library to library over the raw-parts wire: `publish_parts` frames (pickled header + array frames)
    are received straight into the schema's columns by bt_pull_fill; values equal what was published."""
    rng = np.random.default_rng(5)
    B = 4
    items = [{"btid": 7, "image": rng.integers(0, 256, (90, 120, 4), dtype=np.uint8), "frameid": i,
              "depth": rng.random((90, 120), dtype=np.float32)} for i in range(2 * B)]
    with btt.NativePublisher("ipc:///tmp/bt_np_raw_%d" % os.getpid()) as pub:
        rx = DirectReceiver(pub.addresses)
        sender = threading.Thread(target=lambda: [pub.publish_parts(wire.pack(it), timeoutms=10000) for it in items])
        sender.start()
        assert rx.recv(5000)
        sizes = rx.pending_parts()
        assert len(sizes) == 3                                    # header + image + depth
        sch = Schema.from_message(rx.pending(), B, sizes)
        assert sch.native and sch.column("image").shape == (90, 120, 4) and sch.column("depth").dtype == np.float32
        ring = Ring(sch.slot_bytes, depth=1, device=None)
        for first in (0, B):
            slot, filled = ring.acquire(), 0
            while filled < B:
                rc, n = rx.fill(ring, slot, sch, filled, B - filled, 2000)
                assert rc == N.BT_OK, N.last_error()
                filled += n
            view = ring.view(slot)
            cols = {c.key: view[c.slot_off:c.slot_off + B * c.item_bytes].view(c.dtype).reshape((B,) + c.shape) for c in sch.columns}
            for i in range(B):
                want = items[first + i]
                assert int(cols["frameid"][i]) == first + i and int(cols["btid"][i]) == 7
                np.testing.assert_array_equal(cols["image"][i], want["image"])
                np.testing.assert_array_equal(cols["depth"][i], want["depth"])
        sender.join(10)
        rx.close()


def test_a_message_whose_consumer_vanished_before_it_started_goes_to_the_next_consumer():
    """Finite pools: every message must reach somebody even when a consumer disconnects while idle."""
    pools = [[pickle.dumps({"btid": 0, "frameid": i, "image": np.full((32, 32, 4), i, np.uint8)}, protocol=3) for i in range(6)]]
    with btt.NativePublisher("tcp://127.0.0.1:*") as pub:
        ctx = zmq.Context()
        a, b = _pull(ctx, pub.addresses), _pull(ctx, pub.addresses)
        _wait_peers(pub, 0, 2)
        b.close()                                                  # leaves before anything is sent
        time.sleep(0.1)
        pub.start(pools, messages=30, threads=1)
        got = [pickle.loads(a.recv())["frameid"] for _ in range(30)]
        assert got == [i % 6 for i in range(30)]                   # all of them, in order, at the remaining consumer
        assert pub.join(5)
        a.close()
        ctx.term()
