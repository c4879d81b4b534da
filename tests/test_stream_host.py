"""Live-stream host iteration with synthetic publishers (no Blender): the protocol of the
reference's tests/test_dataset.py:11-34 and the record tee of btt/dataset.py:87-103."""
import multiprocessing as mp
import pickle
import socket

import numpy as np
import pytest
from torch.utils.data import DataLoader

from blendtorch import btt


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _publish_frames(addr, btid, n):
    # same message as tests/blender/dataset.blend.py:8-9
    # (the Blender script plays endlessly; n < 0 does the same and the test kills the process)
    with btt.DataPublisher(addr, btid, lingerms=5000 if n >= 0 else 0) as pub:
        i = 0
        while n < 0 or i < n:
            pub.publish(frameid=i, img=np.zeros((64, 64), dtype=np.uint8))
            i += 1


def test_dataset():
    addr = "tcp://127.0.0.1:%d" % _free_port()
    ctx = mp.get_context("fork")
    p = ctx.Process(target=_publish_frames, args=(addr, 0, -1), daemon=True)
    p.start()
    try:
        ds = btt.RemoteIterableDataset([addr], max_items=16)
        dl = DataLoader(ds, batch_size=4, num_workers=4, drop_last=False, shuffle=False)
        count = 0
        for item in dl:
            assert item["img"].shape == (4, 64, 64)
            assert item["frameid"].shape == (4,)
            count += 1
        assert count == 4
    finally:
        p.terminate()
        p.join(timeout=10)


def test_timeout_message():
    ds = btt.RemoteIterableDataset(["tcp://127.0.0.1:%d" % _free_port()], timeoutms=200, max_items=2)
    with pytest.raises(AssertionError, match="No response within timeout interval."):
        next(iter(ds))


def test_record_tee_then_replay(tmp_path):
    addr = "tcp://127.0.0.1:%d" % _free_port()
    ctx = mp.get_context("fork")
    p = ctx.Process(target=_publish_frames, args=(addr, 5, 6), daemon=True)
    p.start()
    try:
        ds = btt.RemoteIterableDataset([addr], max_items=6, item_transform=lambda it: it["frameid"])
        ds.enable_recording(str(tmp_path / "rec"))
        ds.stream_length(6)
        assert list(ds) == list(range(6))
    finally:
        p.join(timeout=10)
    rep = btt.FileDataset(str(tmp_path / "rec"))
    assert len(rep) == 6 and rep[4]["frameid"] == 4 and rep[4]["btid"] == 5
    # the oracle (reference algorithm) reads our recording as well
    from oracle import btr
    assert btr.Reader(tmp_path / "rec_00.btr")[2]["frameid"] == 2
