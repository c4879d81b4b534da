"""Staging copies use non-temporal stores (bt::stream_copy, csrc/scan.cpp): same bytes as memcpy for
every size, source offset and destination alignment -- through the entry points that use it
(bt_schema_ingest for a message in memory, bt_schema_ingest_many on the worker pool)."""
import pickle

import numpy as np
import pytest

from blendtorch.btt.staging import Ring, Schema


@pytest.mark.parametrize("protocol", [3, 4])
def test_payloads_of_every_size_and_alignment_are_copied_exactly(protocol):
    rng = np.random.default_rng(7)
    for n in (4095, 4096, 4097, 8191, 65535, 65536, 65537, 100003, 262144, 262145, 300001, 1228800):
        for key in ("a", "abcdefg", "k" * 13):            # shifts the payload's offset inside the pickle
            frames = [rng.integers(0, 256, n, dtype=np.uint8) for _ in range(3)]
            msgs = [pickle.dumps({"btid": 1, key: f, "frameid": i}, protocol=protocol) for i, f in enumerate(frames)]
            sch = Schema.from_message(msgs[0], 3)
            assert sch.native
            ring = Ring(sch.slot_bytes, depth=1, device=None)
            slot = ring.acquire()
            ring.ingest(slot, sch, 0, msgs[0])                                   # one message: copied by the caller
            addrs = np.array([np.frombuffer(m, np.uint8).ctypes.data for m in msgs[1:]], dtype=np.uint64)
            lens = np.array([len(m) for m in msgs[1:]], dtype=np.uint64)
            assert sch.ingest_many(ring.ptr(slot), 1, addrs, lens, nthreads=3) == -1   # two more: worker pool, 256 KiB tasks
            c = sch.column(key)
            got = ring.view(slot)[c.slot_off:c.slot_off + 3 * n].reshape(3, n)
            for i in range(3):
                np.testing.assert_array_equal(got[i], frames[i])
            ring.close()
            sch.close()
