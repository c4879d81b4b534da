"""The C receive loop (bt_rx_*) on the host: a pyzmq PULL socket drained straight into an (unpinned)
staging slot, a batch per call, compared with what pickle.loads + default_collate give for the
same messages (reference: btt/dataset.py:97-106 + torch collate)."""
import pickle
import threading

import numpy as np
import zmq
from torch.utils.data import default_collate

from blendtorch.btt import native as N
from blendtorch.btt.rx import Receiver
from blendtorch.btt.staging import Ring, Schema


def test_fill_batches_from_socket():
    ctx = zmq.Context()
    push, pull = ctx.socket(zmq.PUSH), ctx.socket(zmq.PULL)
    port = push.bind_to_random_port("tcp://127.0.0.1")
    pull.connect("tcp://127.0.0.1:%d" % port)
    rng = np.random.default_rng(5)
    B, total = 4, 10
    msgs = [{"btid": 0, "image": rng.integers(0, 256, (8, 12, 4), dtype=np.uint8), "frameid": i,
             "xy": rng.random((3, 2))} for i in range(total)]
    odd = dict(msgs[6], note="hello")          # a message outside the schema -> held for the caller

    def send():
        for i, m in enumerate(msgs):
            push.send(pickle.dumps(odd if i == 6 else m, protocol=3 + (i % 2)))

    t = threading.Thread(target=send)
    t.start()
    rx = Receiver(pull)
    assert rx.recv(5000)
    sch = Schema.from_message(rx.pending(), B)
    ring = Ring(sch.slot_bytes, depth=2, device=None)
    got, placed = [], 0
    while placed < total:
        slot, filled = ring.acquire(), 0
        want = min(B, total - placed)
        while filled < want:
            rc, n = rx.fill(ring, slot, sch, filled, want - filled, 2000)
            filled += n
            if rc == N.BT_OK:
                break
            assert rc == N.BT_E_SCHEMA, N.last_error()        # the odd message: caller's decode
            obj = pickle.loads(rx.pending())
            assert obj["note"] == "hello"
            obj.pop("note")
            sch.ingest_object(ring.view(slot), filled, obj)
            rx.drop()
            filled += 1
        view = ring.view(slot)
        batch = {c.key: view[c.slot_off:c.slot_off + filled * c.item_bytes].view(c.dtype).reshape((filled,) + c.shape).copy()
                 for c in sch.columns}
        got.append(batch)
        placed += filled
    t.join()
    assert [len(b["frameid"]) for b in got] == [4, 4, 2]
    exp = default_collate(msgs)
    for key in ("btid", "image", "frameid", "xy"):
        cat = np.concatenate([b[key] for b in got])
        assert cat.dtype == exp[key].numpy().dtype
        np.testing.assert_array_equal(cat, exp[key].numpy())
    # nothing left: a further fill reports a timeout, not an error
    rc, n = rx.fill(ring, ring.acquire(), sch, 0, 1, 50)
    assert (rc, n) == (N.BT_E_TIMEOUT, 0)
    assert rx.release_queued() == 0
    rx.close()
    ring.close()
    push.close(0)
    pull.close(0)
    ctx.term()
