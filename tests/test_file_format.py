"""`.btr` writer/reader: byte compatibility with the reference in both directions
(btt/file.py), the reference's own test protocol (tests/test_file.py:5-13), and the
zero-unpickle raw/span access."""
import pickle

import numpy as np
import pytest

from blendtorch import btt
from blendtorch.btt.file import FileReader, FileRecorder
from oracle import btr as obtr


def test_file_recorder_reader(tmp_path):
    # same statements as the reference's tests/test_file.py:5-13
    with FileRecorder(outpath=tmp_path / "record.mpkl", max_messages=10) as rec:
        for i in range(7):
            rec.save({"value": i}, is_pickled=False)
    r = FileReader(tmp_path / "record.mpkl")
    assert len(r) == 7
    for i in range(7):
        assert r[i]["value"] == i
    # random access works too (the reference's shared-memo files only replay sequentially)
    assert [r[i]["value"] for i in (5, 0, 3)] == [5, 0, 3]
    # and the oracle (= reference algorithm) reads the file we wrote
    o = obtr.Reader(tmp_path / "record.mpkl")
    assert [o[i]["value"] for i in range(7)] == list(range(7))


def test_reads_reference_written_files(golden_dir):
    r = FileReader(golden_dir / "file_roundtrip.btr")   # written by the reference with its shared memo
    assert len(r) == 7
    assert [r[i]["value"] for i in range(7)] == list(range(7))
    for pub, n in enumerate((6, 5)):
        r = FileReader(golden_dir / ("rec_small_%02d.btr" % pub))
        o = obtr.Reader(golden_dir / ("rec_small_%02d.btr" % pub))
        assert len(r) == n
        for i in (n - 1, 0, 2):
            a, b = r[i], o[i]
            assert a.keys() == b.keys()
            np.testing.assert_array_equal(a["image"], b["image"])
            assert pickle.loads(r.raw(i)).keys() == a.keys()
            np.testing.assert_array_equal(pickle.loads(r.raw(i))["xy"], b["xy"])
        r.close()


def test_writes_reference_identical_bytes(golden_dir, tmp_path):
    for pub in range(2):
        src = FileReader(golden_dir / ("rec_small_%02d.btr" % pub))
        out = tmp_path / ("copy_%02d.btr" % pub)
        with FileRecorder(out, max_messages=8) as rec:
            for i in range(len(src)):
                rec.save(src.raw(i), is_pickled=True)
        assert out.read_bytes() == (golden_dir / ("rec_small_%02d.btr" % pub)).read_bytes()


def test_capacity_and_errors(tmp_path):
    with FileRecorder(tmp_path / "cap.btr", max_messages=3) as rec:
        for i in range(5):
            rec.save({"v": i})              # beyond capacity: silently dropped (file.py:46)
    r = FileReader(tmp_path / "cap.btr")
    assert len(r) == 3 and r[2]["v"] == 2
    with pytest.raises(AssertionError, match="Cannot open .* for reading"):
        FileReader(tmp_path / "missing.btr")
    with pytest.raises(AssertionError, match="Found no recording files with prefix"):
        btt.FileDataset(str(tmp_path / "nothing"))
    assert FileRecorder.filename("a/b", 3) == "a/b_03.btr"
    # empty recording
    with FileRecorder(tmp_path / "empty.btr", max_messages=4):
        pass
    assert len(FileReader(tmp_path / "empty.btr")) == 0


def test_filedataset_matches_reference_replay(golden_dir):
    """FileDataset + torch DataLoader (host iteration) reproduces the reference's collated
    batches recorded in rec_small_expected.npz."""
    from torch.utils.data import DataLoader
    exp = np.load(golden_dir / "rec_small_expected.npz")
    ds = btt.FileDataset(str(golden_dir / "rec_small"))
    assert len(ds) == 11
    for k, b in enumerate(DataLoader(ds, batch_size=4, shuffle=False, num_workers=0)):
        for key in ("image", "xy", "frameid", "btid"):
            got = b[key].numpy()
            assert got.dtype == exp["raw_b%d_%s" % (k, key)].dtype
            np.testing.assert_array_equal(got, exp["raw_b%d_%s" % (k, key)])
    # with workers: a fresh dataset per loader (SURVEY defect 4)
    ds = btt.FileDataset(str(golden_dir / "rec_small"), item_transform=lambda it: {"f": it["frameid"]})
    got = sorted(int(x) for b in DataLoader(ds, batch_size=4, num_workers=2) for x in b["f"])
    assert got == sorted(list(range(6)) + list(range(5)))


def test_spans_point_at_messages(golden_dir):
    import ctypes
    r = FileReader(golden_dir / "rec_small_00.btr")
    addrs, lens = r.spans([0, 3, 5])
    for a, l, i in zip(addrs, lens, (0, 3, 5)):
        blob = ctypes.string_at(int(a), int(l))
        assert blob == bytes(r.raw(i))
        assert pickle.loads(blob)["frameid"] == i


def test_async_recorder_writes_identical_files(golden_dir, tmp_path):
    src = FileReader(golden_dir / "rec_small_00.btr")
    for mode in (False, True):
        with FileRecorder(tmp_path / ("m%d.btr" % mode), max_messages=8, async_write=mode) as rec:
            for i in range(len(src)):
                rec.save(src.raw(i), is_pickled=True)           # memoryviews, as the receive loop passes them
            rec.save({"v": 1})
    assert (tmp_path / "m0.btr").read_bytes() == (tmp_path / "m1.btr").read_bytes()
    assert len(FileReader(tmp_path / "m1.btr")) == len(src) + 1


def test_concurrent_first_use_of_the_mapping(golden_dir):
    """Several replay producers may touch a reader at once; its lazily created mapping must be set up once."""
    import ctypes
    import threading
    for _ in range(20):
        r = FileReader(golden_dir / "rec_small_00.btr")
        out, errs = [], []

        def work():
            try:
                a, l = r.spans([0, 5])
                out.append((int(a[1]), ctypes.string_at(int(a[1]), int(l[1]))))
            except Exception as e:
                errs.append(e)

        ts = [threading.Thread(target=work) for _ in range(8)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        assert not errs and len({x[0] for x in out}) == 1 and len({x[1] for x in out}) == 1
        r.close()


def test_index_sidecar_skips_the_capacity_sized_header(tmp_path):
    """N1's optional `.btr.idx`: the offsets without unpickling the 800 KB header of a default-capacity
    recording; ignored as soon as it no longer matches the recording; the recording itself unchanged."""
    import os
    import pickle
    import time
    from blendtorch import btt
    name = str(tmp_path / "rec_00.btr")
    msgs = [pickle.dumps({"btid": 0, "frameid": i, "image": np.full((4, 6, 4), i, np.uint8)}, protocol=3) for i in range(9)]
    with btt.FileRecorder(name) as rec:                         # default capacity: 100000 offsets in the header
        for m in msgs:
            rec.save(m, is_pickled=True)
    before = open(name, "rb").read()
    plain = btt.FileReader.read_offsets(name)
    idx = btt.FileReader.write_index(name)
    assert os.path.getsize(idx) == 8 + 24 + 8 * 9 and open(name, "rb").read() == before
    np.testing.assert_array_equal(btt.FileReader._read_index(name), plain)
    r = btt.FileReader(name)
    assert len(r) == 9 and r[4]["frameid"] == 4 and bytes(r.raw(8)) == msgs[8]
    r.close()
    t0 = time.perf_counter()
    for _ in range(20):
        btt.FileReader.read_offsets(name)
    with_idx = time.perf_counter() - t0
    os.rename(idx, idx + ".off")
    t0 = time.perf_counter()
    for _ in range(20):
        btt.FileReader.read_offsets(name)
    without = time.perf_counter() - t0
    assert with_idx < without                                   # 104 bytes against an 800 KB pickle
    os.rename(idx + ".off", idx)
    # a recording that changed after the sidecar was written: the sidecar is ignored, not trusted
    with btt.FileRecorder(name, max_messages=16) as rec:
        for m in msgs[:3]:
            rec.save(m, is_pickled=True)
    assert btt.FileReader._read_index(name) is None and len(btt.FileReader(name)) == 3
    # garbage sidecars are ignored too
    open(idx, "wb").write(b"BTRIDX1\n" + b"\xff" * 40)
    assert btt.FileReader._read_index(name) is None and len(btt.FileReader.read_offsets(name)) == 3
