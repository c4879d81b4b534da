"""Oracle: the `.btr` recording format (TEST INFRASTRUCTURE, see oracle/__init__.py).

Restates pkg_pytorch/blendtorch/btt/file.py.

File layout (file.py:56-74):
    [pickle(protocol 3) of np.int64[capacity] absolute offsets, -1 = unused]
    [message 0 bytes][message 1 bytes]...
Every message is one self-contained pickle (the `is_pickled=True` path of
file.py:35-53, which is the one btt/dataset.py:100-103 uses).  The
`is_pickled=False` path shares one Pickler memo across messages (reference
defect 3 in SURVEY.md) and is only safe for memo-free payloads read
sequentially; `Recorder.save` reproduces the reference behaviour for it by
keeping a single Pickler as the reference does.
"""

import io
import pickle
from pathlib import Path

import numpy as np

HEADER_PROTOCOL = 3  # file.py:63,72


def filename(prefix, worker_idx):
    """file.py:76-79."""
    return "%s_%02d.btr" % (prefix, worker_idx)


class Recorder:
    """file.py:10-79 (FileRecorder)."""

    def __init__(self, outpath, max_messages=100000):
        self.outpath = Path(outpath)
        self.outpath.parent.mkdir(parents=True, exist_ok=True)  # file.py:30
        self.capacity = max_messages

    def __enter__(self):  # file.py:56-67
        self.fh = io.open(self.outpath, "wb", buffering=0)
        self.offsets = np.full(self.capacity, -1, dtype=np.int64)
        self.count = 0
        self._pickler = pickle.Pickler(self.fh, protocol=HEADER_PROTOCOL)
        self._pickler.dump(self.offsets)  # placeholder, rewritten on exit
        return self

    def save(self, data, is_pickled=False):  # file.py:35-53
        if self.count >= self.capacity:
            return  # silently dropped, file.py:46
        self.offsets[self.count] = self.fh.tell()
        self.count += 1
        if is_pickled:
            self.fh.write(data)
        else:
            self._pickler.dump(data)

    def __exit__(self, *exc):  # file.py:69-74
        self.fh.seek(0)
        pickle.Pickler(self.fh, protocol=HEADER_PROTOCOL).dump(self.offsets)
        self.fh.close()
        self.fh = None


def read_offsets(fname):
    """file.py:120-132: header -> offsets trimmed at the first -1."""
    assert Path(fname).exists(), "Cannot open %s for reading." % fname
    with io.open(fname, "rb") as fh:
        offs = pickle.Unpickler(fh).load()
    unused = np.flatnonzero(offs == -1)
    n = int(unused[0]) if unused.size else len(offs)
    return offs[:n]


class Reader:
    """file.py:81-118 (FileReader): lazy fd, seek to offset, unpickle one message."""

    def __init__(self, path):
        self.path = path
        self.offsets = read_offsets(path)
        self._fh = None

    def __len__(self):
        return len(self.offsets)

    def __getitem__(self, idx):  # file.py:100-108
        if self._fh is None:
            self._fh = io.open(self.path, "rb", buffering=0)
            self._unpickler = pickle.Unpickler(self._fh)
        self._fh.seek(self.offsets[idx])
        return self._unpickler.load()

    def close(self):
        if self._fh is not None:
            self._fh.close()
            self._fh = None


class ReplayDataset:
    """btt/dataset.py:119-153 (SingleFileDataset + FileDataset over ConcatDataset):
    sorted glob of `{prefix}_*.btr`, indices concatenated in file order, then
    `item_transform`.  Subclasses torch Dataset lazily so that importing the
    oracle does not need torch for the pure-numpy parts."""

    def __init__(self, record_path_prefix, item_transform=None):
        from glob import glob

        fnames = sorted(glob("%s_*.btr" % record_path_prefix))  # dataset.py:142
        assert len(fnames) > 0, "Found no recording files with prefix %s" % record_path_prefix
        self.readers = [Reader(f) for f in fnames]
        self.cum = np.cumsum([len(r) for r in self.readers])
        self.item_transform = item_transform or (lambda x: x)

    def __len__(self):
        return int(self.cum[-1]) if len(self.cum) else 0

    def __getitem__(self, idx):
        if idx < 0:
            idx += len(self)
        k = int(np.searchsorted(self.cum, idx, side="right"))  # ConcatDataset bisect
        local = idx - (int(self.cum[k - 1]) if k else 0)
        return self.item_transform(self.readers[k][local])
