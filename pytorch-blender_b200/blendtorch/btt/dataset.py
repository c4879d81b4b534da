"""Stream and replay datasets with the reference's surface.

Drop-in for pkg_pytorch/blendtorch/btt/dataset.py: `RemoteIterableDataset` and `FileDataset`
keep their constructor signatures and defaults, their error texts and their per-item contract
(one dict per message, `item_transform` / overridable `_item`), so existing scripts that wrap
them in ``torch.utils.data.DataLoader`` keep working unchanged.

The accelerated path does not go through ``DataLoader`` workers at all: hand either dataset to
``btt.DeviceLoader`` (loader.py), which lands payloads in pinned staging, copies batches to the
GPU asynchronously and runs the fused preprocessing kernel there.
"""
import pickle
from contextlib import ExitStack
from glob import glob

import torch.utils.data as tud
import zmq

from .constants import DEFAULT_TIMEOUTMS
from . import shm
from . import wire
from .file import FileReader, FileRecorder


def _identity_item_transform(x):
    return x


def open_pull_socket(ctx, addresses, queue_size):
    """PULL socket connected to *all* publishers with a bounded receive queue: ZMQ fair-queues
    between them and stalls senders once `queue_size` messages wait (dataset.py:73-78)."""
    sock = ctx.socket(zmq.PULL)
    sock.setsockopt(zmq.RCVHWM, queue_size)
    for addr in addresses:
        sock.connect(addr)
    poller = zmq.Poller()
    poller.register(sock, zmq.POLLIN)
    return sock, poller


class RemoteIterableDataset(tud.IterableDataset):
    """Receives items published by remote Blender instances (``btb.DataPublisher``).

    Params (identical to the reference, dataset.py:45-51)
    ------
    addresses: list-like of ZMQ addresses to connect to.
    queue_size: receive queue size before publishers get stalled.
    timeoutms: max wait per item before ``AssertionError('No response within timeout interval.')``.
    max_items: artificial length of the dataset; also the capacity of any recording.
    item_transform: callable applied to each received item dict.
    record_path_prefix: when given, every consumer records raw messages to
        ``{prefix}_{worker_idx:02d}.btr``.
    """

    def __init__(self, addresses, queue_size=10, timeoutms=DEFAULT_TIMEOUTMS, max_items=100000,
                 item_transform=None, record_path_prefix=None):
        self.addresses = addresses
        self.queue_size = queue_size
        self.timeoutms = timeoutms
        self.max_items = max_items
        self.record_path_prefix = record_path_prefix
        self.item_transform = item_transform or _identity_item_transform

    def enable_recording(self, fname):
        """Record to prefix `fname`; must be set before iteration starts."""
        self.record_path_prefix = fname

    def stream_length(self, max_items):
        """Set the artificial dataset length."""
        self.max_items = max_items

    def __iter__(self):
        return self._stream()

    def has_custom_item(self):
        """True when items must pass through Python (a transform or an overridden `_item`)."""
        return (self.item_transform is not _identity_item_transform
                or type(self)._item is not RemoteIterableDataset._item)

    def _stream(self):
        # Per-item host iteration, as the reference does it; the socket is created here, not in
        # __init__, so that every DataLoader worker process owns its own (dataset.py:17-18).
        wi = tud.get_worker_info()
        worker_id, num_workers = (wi.id, wi.num_workers) if wi is not None else (0, 1)
        ctx = zmq.Context()
        sock = None
        try:
            sock, poller = open_pull_socket(ctx, self.addresses, self.queue_size)
            with ExitStack() as es:
                rec = None
                if self.record_path_prefix is not None:
                    rec = es.enter_context(FileRecorder(
                        FileRecorder.filename(self.record_path_prefix, worker_id), self.max_items))
                for _ in range(self.max_items // num_workers):
                    ready = dict(poller.poll(self.timeoutms))
                    assert sock in ready, "No response within timeout interval."
                    data = sock.recv()
                    if sock.getsockopt(zmq.RCVMORE):
                        # raw-parts wire (wire.py): the arrays follow as frames of their own
                        obj = shm.resolve(wire.decode([data] + sock.recv_multipart(copy=False)))
                        data = None
                    else:
                        obj = shm.resolve(pickle.loads(data))
                    if rec is not None:
                        if data is None or shm.MARKER_BYTES in data[:4096]:
                            # opt-in wires: record the message with its payload in-line
                            data = pickle.dumps(obj, protocol=3)
                        rec.save(data, is_pickled=True)
                    yield self._item(obj)
        finally:
            if sock is not None:
                sock.close()
            ctx.term()

    def _item(self, item):
        """Transform one item; override or pass `item_transform`."""
        return self.item_transform(item)


class SingleFileDataset(tud.Dataset):
    """Replays one recording file (dataset.py:119-132)."""

    def __init__(self, path, item_transform=None):
        self.reader = FileReader(path)
        self.item_transform = item_transform or _identity_item_transform

    def __len__(self):
        return len(self.reader)

    def __getitem__(self, idx):
        return self._item(self.reader[idx])

    def _item(self, item):
        return self.item_transform(item)


class FileDataset(tud.ConcatDataset):
    """Replays all recordings ``{record_path_prefix}_*.btr`` as one dataset (dataset.py:134-153)."""

    def __init__(self, record_path_prefix, item_transform=None, *, rank=None, world_size=None):
        fnames = sorted(glob(f"{record_path_prefix}_*.btr"))
        assert len(fnames) > 0, f"Found no recording files with prefix {record_path_prefix}"
        if world_size is not None:
            # one process per GPU: rank k replays files k::G of the sorted glob (no collective)
            fnames = fnames[(rank or 0)::world_size]
            assert len(fnames) > 0, f"No recording files left for rank {rank} of {world_size}"
        super().__init__([SingleFileDataset(f) for f in fnames])
        self.filenames = fnames
        self.item_transform = item_transform or _identity_item_transform

    def __getitem__(self, idx):
        return self._item(super().__getitem__(idx))

    def has_custom_item(self):
        return (self.item_transform is not _identity_item_transform
                or type(self)._item is not FileDataset._item)

    def _item(self, item):
        return self.item_transform(item)
