"""End-to-end ingest through the drop-in surface on the GPU: DeviceLoader over FileDataset and
RemoteIterableDataset, compared with what the unmodified reference produced for the same
recordings (tests/golden/rec_small_expected.npz) and with the oracle for live streams."""
import multiprocessing as mp
import pickle
import socket

import numpy as np
import pytest
import torch

from blendtorch import btt
from blendtorch.btt import synthetic as S
from oracle import preprocess as opre

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


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


def test_replay_equals_reference_collate(golden_dir):
    exp = np.load(golden_dir / "rec_small_expected.npz")
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    dl = btt.DeviceLoader(ds, batch_size=4, device=DEV)
    assert len(dl) == 3
    batches = list(dl)
    assert len(batches) == 3
    for k, b in enumerate(batches):
        assert set(b) == {"btid", "image", "frameid", "xy"}
        for key in ("image", "xy", "frameid", "btid"):
            t = b[key]
            assert t.is_cuda
            e = exp["raw_b%d_%s" % (k, key)]
            got = t.cpu().numpy()
            assert got.dtype == e.dtype and got.shape == e.shape, key     # collated dtypes/shapes are the contract
            np.testing.assert_array_equal(got, e)
    assert dl.stats["compat_items"] == 0 and dl.stats["items"] == 11


def test_replay_with_fused_preprocess_equals_reference(golden_dir):
    exp = np.load(golden_dir / "rec_small_expected.npz")
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, keep_raw=True,
                          preprocess=btt.Preprocess(gamma=2.2, mean=127.5, std=127.5, stats=True))
    for k, b in enumerate(dl):
        want = exp["gamma_norm_b%d_image" % k]
        got = b["image"].cpu().numpy()
        np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))
        np.testing.assert_array_equal(b["image_raw"].cpu().numpy(), exp["raw_b%d_image" % k])
        s, q = opre.batch_stats(want)
        np.testing.assert_allclose(b["image_stats"].cpu().numpy()[:, 0], s, rtol=1e-6, atol=1e-4)
    assert dl.stats["kernels"] == 6          # 3 batches x (K1 + the fold of its statistics rows)


def test_shuffle_drop_last_and_item_transform(golden_dir):
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    dl = btt.DeviceLoader(ds, batch_size=4, shuffle=True, drop_last=True, device=DEV, seed=3)
    seen = []
    for b in dl:
        assert b["image"].shape == (4, 12, 16, 4)
        seen += list(zip(b["btid"].tolist(), b["frameid"].tolist()))
    assert len(seen) == 8 and len(set(seen)) == 8
    first_epoch = list(seen)
    seen2 = [x for b in dl for x in zip(b["btid"].tolist(), b["frameid"].tolist())]
    assert seen2 != first_epoch                                   # a new permutation per epoch
    # a user item_transform still applies (compatibility decode), then the same device path
    def t(item):
        item["image"] = item["image"][::2, ::2].copy()
        item["tag"] = "f%d" % item["frameid"]
        return item
    ds2 = btt.FileDataset(str(golden_dir / "rec_small"), item_transform=t)
    dl2 = btt.DeviceLoader(ds2, batch_size=4, device=DEV, preprocess=btt.Preprocess(mean=127.5, std=127.5))
    b = next(iter(dl2))
    assert b["image"].shape == (4, 4, 6, 8) and b["tag"] == ["f0", "f1", "f2", "f3"]
    raw = ds[1]["image"][::2, ::2]
    np.testing.assert_array_equal(b["image"][1].cpu().numpy(), opre.preprocess_frame(raw, mean=127.5, std=127.5))


def _publisher(addr, btid, count, pool, h, w, cfg, proto):
    S.run_publisher(addr, btid, cfg, h, w, 4, count=count, pool=pool, protocol=proto, lingerms=10000)


@pytest.mark.parametrize("receivers", [1, 2])
def test_stream_exactly_once_and_values(receivers, tmp_path):
    H, W, per_pub, pubs = 24, 32, 12, 2
    addrs = ["tcp://127.0.0.1:%d" % _free_port() for _ in range(pubs)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_publisher, args=(a, i, -1 if receivers > 1 else per_pub, per_pub, H, W, 2, 3 + i), daemon=True)
             for i, a in enumerate(addrs)]
    for p in procs:
        p.start()
    try:
        total = per_pub * pubs
        ds = btt.RemoteIterableDataset(addrs, max_items=total, timeoutms=20000)
        ds.enable_recording(str(tmp_path / "tee"))
        dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, num_receivers=receivers, keep_raw=True,
                              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
        got = {}
        n = 0
        for b in dl:
            assert b["image"].dtype == torch.float32 and b["bbox"].dtype == torch.float32
            assert b["btid"].dtype == torch.int64
            raw = b["image_raw"].cpu().numpy()
            img = b["image"].cpu().numpy()
            np.testing.assert_array_equal(img, opre.preprocess_batch(list(raw), flip=True, gamma=2.2, mean=127.5, std=127.5))
            for i in range(len(raw)):
                key = (int(b["btid"][i]), int(b["frameid"][i]))
                if receivers == 1:
                    assert key not in got
                got[key] = (raw[i], b["bbox"][i].cpu().numpy())
                n += 1
        assert n == total
        # every delivered frame is byte-identical to what the publisher generated
        for btid in range(pubs):
            rng = S.rng_for(2, btid)
            msgs = [S.message(rng, 2, btid, i, H, W) for i in range(per_pub)]
            keys = [k for k in got if k[0] == btid]   # may be empty: endless publishers race
            for (_, i) in keys:
                raw, bbox = got[(btid, i)]
                np.testing.assert_array_equal(raw, msgs[i]["image"])
                np.testing.assert_array_equal(bbox, msgs[i]["bbox"])
        if receivers == 1:
            assert len(got) == total                      # exactly once
        # the tee recorded every raw message; replaying it gives the same frames
        rep = btt.FileDataset(str(tmp_path / "tee"))
        assert len(rep) == total
        it = rep[0]
        np.testing.assert_array_equal(it["image"], got[(it["btid"], it["frameid"])][0])
    finally:
        for p in procs:
            p.terminate()
            p.join(timeout=10)


def test_stream_timeout_and_string_fields():
    ds = btt.RemoteIterableDataset(["tcp://127.0.0.1:%d" % _free_port()], timeoutms=300, max_items=4)
    with pytest.raises(AssertionError, match="No response within timeout interval."):
        next(iter(btt.DeviceLoader(ds, batch_size=2, device=DEV)))
    addr = "tcp://127.0.0.1:%d" % _free_port()
    def pub():
        with btt.DataPublisher(addr, 7, lingerms=10000) as p:
            for i in range(4):
                p.publish(image=np.full((4, 8, 4), i, np.uint8), name="n%d" % i, frameid=i)
    pr = mp.get_context("fork").Process(target=pub, daemon=True)
    pr.start()
    try:
        ds = btt.RemoteIterableDataset([addr], max_items=4)
        bs = list(btt.DeviceLoader(ds, batch_size=2, device=DEV))
        assert [b["name"] for b in bs] == [["n0", "n1"], ["n2", "n3"]]
        assert bs[1]["image"].shape == (2, 4, 8, 4) and int(bs[1]["image"][1, 0, 0, 0]) == 3
    finally:
        pr.join(timeout=10)


def _shm_publisher(addr, btid, count, slots, h, w):
    S.run_publisher(addr, btid, 2, h, w, 4, count=count, pool=8, protocol=3, lingerms=10000, shm_slots=slots)


@pytest.mark.parametrize("slots,receivers", [(16, 2), (2, 1)])
def test_stream_shared_memory_wire(slots, receivers, tmp_path):
    """Opt-in shared-memory wire: payloads reach the GPU by DMA from the producer's segment.  With
    only 2 slots and batches of 4 the consumer must spill held slots when the stream stalls, or
    producer and consumer would wait for each other forever."""
    H, W, pubs = 64, 128, 2
    per_pub = 16
    addrs = ["tcp://127.0.0.1:%d" % _free_port() for _ in range(pubs)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_shm_publisher, args=(a, i, -1, slots, H, W), daemon=True) for i, a in enumerate(addrs)]
    for p in procs:
        p.start()
    try:
        total = per_pub * pubs
        ds = btt.RemoteIterableDataset(addrs, max_items=total, timeoutms=30000)
        dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, num_receivers=receivers, keep_raw=True,
                              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
        expected = {}
        for btid in range(pubs):
            rng = S.rng_for(2, btid)
            for i in range(8):
                expected[(btid, i)] = S.message(rng, 2, btid, i, H, W)
        n = 0
        for b in dl:
            raw = b["image_raw"].cpu().numpy()
            np.testing.assert_array_equal(b["image"].cpu().numpy(),
                                          opre.preprocess_batch(list(raw), flip=True, gamma=2.2, mean=127.5, std=127.5))
            for i in range(len(raw)):
                m = expected[(int(b["btid"][i]), int(b["frameid"][i]))]
                np.testing.assert_array_equal(raw[i], m["image"])
                np.testing.assert_array_equal(b["bbox"][i].cpu().numpy(), m["bbox"])
                n += 1
        assert n == total and dl.stats["compat_items"] == 0
    finally:
        for p in procs:
            p.terminate()
            p.join(timeout=10)
        import glob, os
        for p in procs:
            for f in glob.glob("/dev/shm/btshm_%d_*" % p.pid):
                os.unlink(f)


def test_config3_depth_and_projection_in_loader(tmp_path):
    """BASELINE config 3 in small: RGBA + float32 depth per frame, 8 cubes x 8 corners projected to
    pixel boxes per frame inside the loader (K1 + K2 per batch), replayed from a recording."""
    from oracle import btr, camera as ocam
    H, W = 54, 96
    prefix = str(tmp_path / "c3")
    S.write_recordings(prefix, config=3, n_files=2, frames_per_file=5, h=H, w=W)
    V, P = ocam.view_looking_down_z(0.0), ocam.blender_perspective((H, W), lens=35.0)
    cam = btt.Camera(V, P, (H, W), device=DEV)
    dl = btt.DeviceLoader(btt.FileDataset(prefix), batch_size=4, device=DEV, group=1,
                          preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5),
                          project=btt.Projection("corners", cam, xy_key="xy", depth_key="z", bbox_key="bbox"))
    ref = btr.ReplayDataset(prefix)
    n = 0
    for b in dl:
        k = len(b["image"])
        assert b["depth"].dtype == torch.float32 and b["depth"].shape == (k, H, W)
        for i in range(k):
            it = ref[n + i]
            np.testing.assert_array_equal(b["depth"][i].cpu().numpy(), it["depth"])          # pass-through, exact
            np.testing.assert_array_equal(b["image"][i].cpu().numpy(),
                                          opre.preprocess_frame(it["image"], flip=True, gamma=2.2, mean=127.5, std=127.5))
            exy, ez = ocam.world_to_pixel(it["corners"], V, P, (H, W), return_depth=True)
            np.testing.assert_allclose(b["xy"][i].cpu().numpy(), exy, rtol=1e-11, atol=1e-8)
            np.testing.assert_allclose(b["z"][i].cpu().numpy(), ez, rtol=1e-12, atol=1e-12)
            np.testing.assert_array_equal(b["bbox"][i].cpu().numpy(), ocam.pixel_bboxes(b["xy"][i].cpu().numpy()))
        n += k
    assert n == 10 and dl.stats["kernels"] == 6          # 3 batches x (K1 + K2)
    # a second epoch reuses the cached pinned ring and gives the same data
    again = [b["frameid"].tolist() for b in dl]
    assert sum(len(x) for x in again) == 10
    one_by_one = [{k: v.cpu() for k, v in b.items()} for b in dl]
    dl.close()
    # the default groups small batches (one copy per column, one K1 and one K2 launch for all three
    # batches here): same batches, fewer launches
    dl = btt.DeviceLoader(btt.FileDataset(prefix), batch_size=4, device=DEV,
                          preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5),
                          project=btt.Projection("corners", cam, xy_key="xy", depth_key="z", bbox_key="bbox"))
    grouped = [{k: v.cpu() for k, v in b.items()} for b in dl]
    assert dl.stats["kernels"] == 2 and len(grouped) == len(one_by_one) == 3
    for a, b in zip(one_by_one, grouped):
        assert a.keys() == b.keys()
        for k in a:
            assert torch.equal(a[k], b[k]), k
    dl.close()


@pytest.mark.parametrize("producers", [2, 3, 5])
def test_replay_with_several_producers_keeps_order(golden_dir, producers):
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    one = [(b["btid"].tolist(), b["frameid"].tolist(), b["image"].cpu()) for b in
           btt.DeviceLoader(ds, batch_size=2, device=DEV, shuffle=True, seed=5)]
    dl = btt.DeviceLoader(ds, batch_size=2, device=DEV, shuffle=True, seed=5, num_receivers=producers)
    many = [(b["btid"].tolist(), b["frameid"].tolist(), b["image"].cpu()) for b in dl]
    assert len(one) == len(many) == 6
    for a, b in zip(one, many):
        assert a[0] == b[0] and a[1] == b[1] and torch.equal(a[2], b[2])
    # drop_last and a second epoch
    dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, drop_last=True, num_receivers=producers)
    assert [len(b["frameid"]) for b in dl] == [4, 4]
    assert [b["frameid"].tolist() for b in dl] == [[0, 1, 2, 3], [4, 5, 0, 1]]
    dl.close()


def test_stream_through_the_direct_receiver():
    """transport='direct': the library's own ZMTP endpoint instead of pyzmq's socket; same batches."""
    H, W, pubs, per_pub = 48, 64, 2, 8
    addrs = ["tcp://127.0.0.1:%d" % _free_port() for _ in range(pubs)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_publisher, args=(a, i, per_pub, per_pub, H, W, 2, 3 + i), daemon=True)
             for i, a in enumerate(addrs)]
    for p in procs:
        p.start()
    try:
        ds = btt.RemoteIterableDataset(addrs, max_items=per_pub * pubs, timeoutms=20000)
        dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, keep_raw=True, transport="direct",
                              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
        got = {}
        for b in dl:
            raw = b["image_raw"].cpu().numpy()
            np.testing.assert_array_equal(b["image"].cpu().numpy(),
                                          opre.preprocess_batch(list(raw), flip=True, gamma=2.2, mean=127.5, std=127.5))
            for i in range(len(raw)):
                key = (int(b["btid"][i]), int(b["frameid"][i]))
                assert key not in got
                got[key] = (raw[i], b["bbox"][i].cpu().numpy())
        assert len(got) == per_pub * pubs and dl.stats["compat_items"] == 0
        for btid in range(pubs):
            rng = S.rng_for(2, btid)
            for i in range(per_pub):
                m = S.message(rng, 2, btid, i, H, W)
                np.testing.assert_array_equal(got[(btid, i)][0], m["image"])
                np.testing.assert_array_equal(got[(btid, i)][1], m["bbox"])
    finally:
        for p in procs:
            p.join(timeout=10)


def _epochs(dl, n):
    out = []
    for _ in range(n):
        out.append([{k: (v.cpu() if torch.is_tensor(v) else v) for k, v in b.items()} for b in dl])
    return out


def _same_batches(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert set(x) == set(y)
        for k in x:
            if torch.is_tensor(x[k]):
                assert x[k].dtype == y[k].dtype and x[k].shape == y[k].shape, k
                assert torch.equal(x[k].view(torch.uint8), y[k].view(torch.uint8)), k
            else:
                assert x[k] == y[k], k


@pytest.mark.parametrize("shuffle,drop_last", [(False, False), (True, False), (True, True)])
def test_resident_recording_serves_later_epochs_identically(golden_dir, shuffle, drop_last):
    """cache='device': epoch 1 goes through the pinned pipeline and leaves the raw columns in HBM; epochs
    2.. are gathered from there by K1 -- same batches, bit for bit, as the uncached loader with the same
    seed (which re-reads the files every epoch like the reference, file.py:100-108)."""
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    kw = dict(batch_size=4, device=DEV, shuffle=shuffle, drop_last=drop_last, seed=11, keep_raw=True,
              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True))
    plain = btt.DeviceLoader(ds, **kw)
    cached = btt.DeviceLoader(ds, cache="device", **kw)
    want, got = _epochs(plain, 4), _epochs(cached, 4)
    for e in range(4):
        _same_batches(want[e], got[e])
    st = cached.stats
    assert st["cache_bytes"] == 11 * (12 * 16 * 4 + 8 * 2 * 8 + 8 + 8)
    if drop_last and shuffle:
        # 3 of 11 frames are dropped each epoch: epochs stay on the pipeline until every frame an
        # epoch needs has been seen
        assert 0 <= st["resident_batches"] <= 6
    else:
        assert st["resident_batches"] == 3 * len(want[0])
        assert st["bytes_h2d"] == plain.stats["bytes_h2d"] // 4     # only the first epoch crossed PCIe
    plain.close()
    cached.close()


def test_resident_recording_limits_and_projection(tmp_path):
    names = S.write_recordings(str(tmp_path / "c3"), 3, n_files=2, frames_per_file=5, h=24, w=32)
    assert len(names) == 2
    ds = btt.FileDataset(str(tmp_path / "c3"))
    from oracle import camera as ocam
    cam = btt.Camera(ocam.view_looking_down_z(0.0), ocam.blender_perspective((24, 32), lens=35.0), (24, 32), device=DEV)
    kw = dict(batch_size=3, device=DEV, shuffle=True, seed=2,
              preprocess=[btt.Preprocess(mean=127.5, std=127.5, dtype=torch.bfloat16)],
              project=btt.Projection("corners", cam, xy_key="xy", bbox_key="bbox", depth_key="z"))
    plain, cached = btt.DeviceLoader(ds, **kw), btt.DeviceLoader(ds, cache="device", **kw)
    want, got = _epochs(plain, 3), _epochs(cached, 3)
    for e in range(3):
        _same_batches(want[e], got[e])
    assert cached.stats["resident_batches"] == 2 * len(want[0])
    # a budget the recording does not fit: replay stays uncached, and says so
    small = btt.DeviceLoader(ds, cache="device", cache_bytes=1000, **kw)
    with pytest.warns(UserWarning, match="replaying uncached"):
        first = _epochs(small, 1)
    _same_batches(_epochs(btt.DeviceLoader(ds, **kw), 1)[0], first[0])
    assert small.stats["resident_batches"] == 0 and small.stats["cache_bytes"] == 0
    # not for streams, not with an item_transform
    with pytest.raises(ValueError):
        btt.DeviceLoader(btt.FileDataset(str(tmp_path / "c3"), item_transform=lambda x: x), device=DEV, cache="device")
    with pytest.raises(ValueError):
        btt.DeviceLoader(btt.RemoteIterableDataset(["tcp://127.0.0.1:1"]), device=DEV, cache="device")


def _raw_publisher(addr, btid, count, h, w):
    S.run_publisher(addr, btid, 2, h, w, 4, count=count, pool=8, protocol=3, lingerms=10000, raw_parts=True)


@pytest.mark.parametrize("transport,receivers", [("direct", 2), ("zmq", 1)])
def test_stream_raw_parts_wire(transport, receivers):
    """Opt-in raw-parts wire (wire.py): frames go from the socket straight into the staging slot with the
    direct receiver (no compatibility decodes), and through the Python decode with pyzmq's socket."""
    H, W, pubs, per_pub = 64, 128, 2, 16
    addrs = ["tcp://127.0.0.1:%d" % _free_port() for _ in range(pubs)]
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_raw_publisher, args=(a, i, -1, H, W), daemon=True) for i, a in enumerate(addrs)]
    for p in procs:
        p.start()
    try:
        total = per_pub * pubs
        ds = btt.RemoteIterableDataset(addrs, max_items=total, timeoutms=30000)
        dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, num_receivers=receivers, keep_raw=True, transport=transport,
                              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
        expected = {}
        for btid in range(pubs):
            rng = S.rng_for(2, btid)
            for i in range(8):
                expected[(btid, i)] = S.message(rng, 2, btid, i, H, W)
        n = 0
        for b in dl:
            raw = b["image_raw"].cpu().numpy()
            np.testing.assert_array_equal(b["image"].cpu().numpy(),
                                          opre.preprocess_batch(list(raw), flip=True, gamma=2.2, mean=127.5, std=127.5))
            for i in range(len(raw)):
                m = expected[(int(b["btid"][i]), int(b["frameid"][i]))]
                np.testing.assert_array_equal(raw[i], m["image"])
                np.testing.assert_array_equal(b["bbox"][i].cpu().numpy(), m["bbox"])
                n += 1
        assert n == total
        assert (dl.stats["compat_items"] == 0) == (transport == "direct")
    finally:
        for p in procs:
            p.terminate()
            p.join(timeout=10)


@pytest.mark.parametrize("shuffle,drop_last", [(False, False), (True, True)])
def test_pinned_host_recording_serves_every_epoch_identically(golden_dir, shuffle, drop_last):
    """cache='host': the recording is read into page-locked memory once and every batch is gathered to
    the device by DMA (bt_gather_h2d) -- same batches, bit for bit, as the uncached loader."""
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    kw = dict(batch_size=4, device=DEV, shuffle=shuffle, drop_last=drop_last, seed=7, keep_raw=True,
              preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True))
    plain, cached = btt.DeviceLoader(ds, **kw), btt.DeviceLoader(ds, cache="host", **kw)
    want, got = _epochs(plain, 3), _epochs(cached, 3)
    for e in range(3):
        _same_batches(want[e], got[e])
    assert cached.stats["cache_bytes"] > 0 and cached.stats["cache_fill_s"] > 0
    assert cached.stats["bytes_h2d"] == plain.stats["bytes_h2d"]      # every epoch crosses PCIe, none touches the file
    plain.close()
    cached.close()
    small = btt.DeviceLoader(ds, cache="host", cache_bytes=100, **kw)
    with pytest.warns(UserWarning, match="replaying uncached"):
        first = _epochs(small, 1)
    _same_batches(want[0], first[0])
    small.close()


def test_gather_h2d_merges_runs_and_rejects_bad_rows():
    import ctypes as C
    from blendtorch.btt import native as N
    src = torch.arange(64 * 37, dtype=torch.int32).reshape(64, 37).pin_memory()
    idx = np.array([5, 6, 7, 0, 63, 62, 10, 11], dtype=np.int64)
    dst = torch.empty((len(idx), 37), dtype=torch.int32, device=DEV)
    st = torch.cuda.current_stream()
    N.check(N.lib.bt_gather_h2d(dst.data_ptr(), src.data_ptr(), 37 * 4, idx.ctypes.data, len(idx), C.c_void_p(st.cuda_stream)))
    st.synchronize()
    assert torch.equal(dst.cpu(), src[torch.from_numpy(idx)])
    bad = np.array([1, -2], dtype=np.int64)
    assert N.lib.bt_gather_h2d(dst.data_ptr(), src.data_ptr(), 37 * 4, bad.ctypes.data, 2, C.c_void_p(st.cuda_stream)) == N.BT_E_INVALID


def test_allow_pickle_false_never_unpickles():
    """The reference unpickles whatever arrives (dataset.py:103,105).  With allow_pickle=False ordinary
    streams run entirely on the scanner; a message that would need pickle.loads raises instead."""
    def pub(addr, with_string):
        with btt.DataPublisher(addr, 3, lingerms=10000) as p:
            for i in range(4):
                extra = {"name": "n%d" % i} if with_string else {}
                p.publish(image=np.full((8, 16, 4), i, np.uint8), frameid=i, **extra)
    for with_string in (False, True):
        addr = "tcp://127.0.0.1:%d" % _free_port()
        pr = mp.get_context("fork").Process(target=pub, args=(addr, with_string), daemon=True)
        pr.start()
        try:
            ds = btt.RemoteIterableDataset([addr], max_items=4)
            dl = btt.DeviceLoader(ds, batch_size=2, device=DEV, allow_pickle=False)
            if with_string:
                with pytest.raises(RuntimeError, match="allow_pickle=False"):
                    list(dl)
            else:
                bs = list(dl)
                assert [b["frameid"].tolist() for b in bs] == [[0, 1], [2, 3]] and dl.stats["compat_items"] == 0
        finally:
            pr.join(timeout=10)
    with pytest.raises(ValueError, match="allow_pickle=False"):
        btt.DeviceLoader(btt.RemoteIterableDataset(["tcp://127.0.0.1:1"], item_transform=lambda x: x), device=DEV,
                         allow_pickle=False)


def test_depth_field_is_flipped_with_fliprows(tmp_path):
    """BASELINE config 3 carries a float32 depth map next to the frame: a FlipRows spec flips it on the GPU
    like the frame (both are bottom-up after a GL read-back), in the pipeline and from resident caches."""
    from oracle import btr
    prefix = str(tmp_path / "c3")
    S.write_recordings(prefix, config=3, n_files=1, frames_per_file=6, h=24, w=32)
    ref = btr.ReplayDataset(prefix)
    ds = btt.FileDataset(prefix)
    specs = [btt.Preprocess(flip=True, mean=127.5, std=127.5), btt.FlipRows("depth")]
    for cache in (None, "device", "host"):
        dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, preprocess=specs, cache=cache, keep_raw=True)
        for epoch in range(2):
            n = 0
            for b in dl:
                for i in range(len(b["depth"])):
                    np.testing.assert_array_equal(b["depth"][i].cpu().numpy(), ref[n + i]["depth"][::-1])
                    np.testing.assert_array_equal(b["depth_raw"][i].cpu().numpy(), ref[n + i]["depth"])
                    np.testing.assert_array_equal(b["image"][i].cpu().numpy(),
                                                  opre.preprocess_frame(ref[n + i]["image"], flip=True, mean=127.5, std=127.5))
                n += len(b["depth"])
            assert n == 6
        dl.close()
    with pytest.raises(TypeError, match="uint8"):
        list(btt.DeviceLoader(ds, batch_size=4, device=DEV, preprocess=btt.Preprocess(key="depth")))


def _slow_publisher(addr, count, rate, h, w):
    S.run_publisher(addr, 0, 2, h, w, 4, count=count, pool=count, protocol=3, lingerms=10000, rate=rate)


@pytest.mark.parametrize("drop_last", [False, True])
def test_groups_of_batches_hand_out_what_is_complete_when_the_stream_pauses(drop_last):
    """group=4 stages 4 batches per pinned slot.  A slow publisher must not make the consumer wait for a
    whole group: when the stream pauses on a batch boundary the complete batches are handed out; the
    batches are the same as with group=1, in order, and the last, short one follows drop_last."""
    H, W, count = 24, 32, 10
    out = {}
    for group in (1, 4):
        addr = "tcp://127.0.0.1:%d" % _free_port()
        ctx = mp.get_context("spawn")
        p = ctx.Process(target=_slow_publisher, args=(addr, count, 6.0, H, W), daemon=True)   # 167 ms apart: every wait times out once
        p.start()
        try:
            ds = btt.RemoteIterableDataset([addr], max_items=count, timeoutms=20000)
            dl = btt.DeviceLoader(ds, batch_size=4, device=DEV, group=group, drop_last=drop_last,
                                  preprocess=btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True))
            got = [(b["frameid"].tolist(), b["image"].cpu(), b["image_stats"].cpu()) for b in dl]
            dl.close()
            out[group] = got
        finally:
            p.join(timeout=15)
    sizes = [len(f) for f, _, _ in out[4]]
    assert sizes == ([4, 4] if drop_last else [4, 4, 2])
    assert [f for f, _, _ in out[4]] == [f for f, _, _ in out[1]] == ([[0, 1, 2, 3], [4, 5, 6, 7]] + ([] if drop_last else [[8, 9]]))
    for (_, img1, st1), (_, img4, st4) in zip(out[1], out[4]):
        assert torch.equal(img1, img4) and torch.equal(st1, st4)       # per-batch statistics stay per batch
