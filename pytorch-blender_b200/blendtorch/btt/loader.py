"""DeviceLoader: batches of training-ready *device* tensors from a btt dataset.

Takes the place of ``torch.utils.data.DataLoader(ds, batch_size, num_workers) + item_transform
+ default_collate + batch.to(device)`` in a reference script (benchmarks/benchmark.py:26-41,
examples/datagen/generate.py:79-85, examples/densityopt/densityopt.py:210-268):

    ds = btt.RemoteIterableDataset(addresses, max_items=N)          # or btt.FileDataset(prefix)
    dl = btt.DeviceLoader(ds, batch_size=8, device='cuda:0',
                          preprocess=btt.Preprocess(gamma=2.2, mean=127.5, std=127.5))
    for batch in dl:       # dict of CUDA tensors, same keys/dtypes/shapes default_collate gives
        ...

Pipeline per batch: receiver thread(s) -> `bt_schema_ingest` (one memcpy per payload byte, GIL
released) into a pinned slot -> one `cudaMemcpyAsync` on the loader's copy stream -> K1 on the
same stream -> CUDA event; the consumer's stream waits on the event, never the host.  Annotation
columns (xy, bbox, frameid, btid, depth ...) are views into the per-batch device buffer, so they
cost no kernel.  Ordering and delivery follow the reference: in order per publisher with one
receiver, unordered across receivers, every message delivered once
(examples/datagen/Readme.md:172-177); back-pressure is the bounded queue + ZMQ high-water marks.

Recordings that are replayed for many epochs can be kept in HBM (``cache='device'``): the first epoch
goes through the pipeline above and leaves the *raw* columns (uint8 frames, 1/4 of their fp32 size)
on the device; later epochs never touch the file, the host or PCIe -- the sampler's indices go to K1,
which gathers its frames straight from the resident set.
"""
import logging
import pickle
import queue
import threading
import time
import warnings
import ctypes as C

import numpy as np
import torch
import zmq

from . import native as N
from . import shm
from . import wire
from .dataset import FileDataset, RemoteIterableDataset, open_pull_socket
from .file import FileRecorder
from .preprocess import FlipRows, Preprocess, prepare
from .rx import DirectReceiver, Receiver, direct_capable
from .staging import Ring, Schema, refuse_pickle

_DONE = object()
_log = logging.getLogger("blendtorch")      # the reference's logger name (file.py:8)


def shard(seq, rank, world_size):
    """Round-robin shard of publishers / recording files for GPU `rank` of `world_size`: the
    reference's worker model (btt/dataset.py:77-85,97) with GPU = worker; no collective."""
    return list(seq)[rank::world_size]


def _torch_dtype(np_dtype):
    return torch.from_numpy(np.empty(0, dtype=np_dtype)).dtype


class _ResidentCache:
    """A recording kept in HBM: per column one dense device tensor whose rows are dataset indices.

    Filled as a side effect of the first pass over the data (a device-to-device copy per batch on the
    producer's stream); a row counts as present once the consumer's stream has waited for the batch
    that carried it, so later reads on that stream are ordered after the writes."""

    def __init__(self, schema, n, device, specs):
        self.n, self.device = n, device
        self.plan, self.cols = [], {}
        base = torch.cuda.default_stream(device)
        base.synchronize()                    # whatever used these blocks before is done
        with torch.cuda.stream(base):         # not tied to a producer's stream
            self._allocate(schema, n, device, specs)
        self.host = {k: [None] * n for k in schema.host_keys}
        self.present = np.zeros(n, dtype=bool)
        self.nbytes = sum(t.numel() * t.element_size() for t in self.cols.values())

    def _allocate(self, schema, n, device, specs):
        for c in schema.columns:
            dt = _torch_dtype(c.dtype)
            prep = prepare(specs[c.key], c.shape, c.dtype, dt, device) if c.key in specs else None
            self.plan.append((c.key, prep))
            self.cols[c.key] = torch.empty((n,) + tuple(c.shape), dtype=dt, device=device)

    @staticmethod
    def bytes_needed(schema, n):
        return n * sum(c.item_bytes for c in schema.columns)

    def store(self, idx, rows):
        """Enqueue (on the current stream) the copy of one batch's raw columns into rows `idx`."""
        n = len(idx)
        dense = n > 0 and int(idx[-1]) - int(idx[0]) == n - 1 and bool(np.all(np.diff(idx) == 1))
        where = None if dense else torch.from_numpy(np.ascontiguousarray(idx, dtype=np.int64)).to(self.device)
        for key, t in rows.items():
            if dense:
                self.cols[key][int(idx[0]):int(idx[0]) + n].copy_(t, non_blocking=True)
            else:
                self.cols[key].index_copy_(0, where, t)

    def commit(self, idx, batch):
        for k, rows in self.host.items():
            for i, v in zip(idx, batch.get(k, ())):
                rows[int(i)] = v
        self.present[idx] = True

    def covers(self, idx):
        return bool(self.present[idx].all())


class _Stager:
    """One pinned ring + one CUDA stream; turns filled slots into device batches.

    Everything that is the same for every batch (column views, kernel parameter blocks) is planned
    once; `finish` is then a handful of torch calls and two ctypes calls.  The owning thread makes
    `self.stream` its current torch stream, so allocations belong to that stream."""

    def __init__(self, schema, device, depth, specs, keep_raw, projections=()):
        self.schema = schema
        self.device = device
        self.projections = projections
        self.ring = Ring(schema.slot_bytes, depth, device.index) if depth > 0 else None   # 0: batches come by DMA gather
        self.stream = torch.cuda.Stream(device)
        torch.cuda.set_stream(self.stream)      # thread-local: this receiver thread only
        self._stream_handle = C.c_void_p(self.stream.cuda_stream)
        self.keep_raw = keep_raw
        self.kernels = 0
        self.plan = []
        for c in schema.columns:
            spec = specs.get(c.key)
            prep = prepare(spec, c.shape, c.dtype, _torch_dtype(c.dtype), device) if spec is not None else None
            self.plan.append((c.key, c.slot_off, c.item_bytes, _torch_dtype(c.dtype), c.shape, prep))
        self.item_bytes = sum(c.item_bytes for c in schema.columns)

    def begin(self):
        return self.ring.acquire()

    def ingest(self, slot, item, msg, host_rows):
        self.ring.ingest(slot, self.schema, item, msg, host_rows)

    def finish(self, slot, n, host_rows, cache=None, idx=None, fill=None, batch_size=None):
        """Enqueue, on this stager's stream, everything that turns the `n` staged items into device
        batches: the column copies, the kernels, an event.  The items are handed out as batches of
        `batch_size` (default: one batch) that share one set of allocations, one copy per column and
        -- unless per-batch statistics are wanted -- one launch per kernel: the per-batch cost of the
        host side is a handful of slices.  Returns (list of batch dicts, event, tensors to keep, idx)."""
        sch = self.schema
        B = n if not batch_size else batch_size
        used = 0
        for _, off, ib, _, _, _ in self.plan:
            used = max(used, off + n * ib)
        batch, per_batch = {}, {}
        raw = torch.empty(max(used, 1), dtype=torch.uint8, device=self.device)
        if used and fill is not None:
            fill(raw, self._stream_handle)      # e.g. rows gathered from a pinned recording
        elif used:
            self.ring.submit_batch(slot, sch, n, raw.data_ptr(), self._stream_handle)
        keep = [raw]                       # distinct storages the consumer stream must know about
        base = raw.data_ptr()
        if cache is not None and n:
            cache.store(idx, {key: raw[off:off + n * ib].view(dt).view((n,) + shape)
                              for key, off, ib, dt, shape, _ in self.plan})
        for key, off, ib, dt, shape, prep in self.plan:
            if prep is None or self.keep_raw:
                t = raw[off:off + n * ib].view(dt).view((n,) + shape)
            if prep is not None:
                out, stats = prep.run_group(base + off, n, B, self._stream_handle)
                self.kernels += prep.last_kernels
                keep.append(out)
                if stats is not None:
                    per_batch[key + "_stats"] = stats
                    keep.append(stats)
                if self.keep_raw:
                    batch[key + "_raw"] = t
                t = out
            batch[key] = t
        for pj in self.projections:
            keep.extend(pj.apply(batch, self.device))
            self.kernels += 1
        ev = torch.cuda.Event()
        ev.record(self.stream)
        for k in sch.host_keys:
            batch[k] = host_rows.get(k, [])
        # what the resident cache needs once the consumer has seen the group: rows and host-side fields
        tag = (idx, {k: batch[k] for k in sch.host_keys}) if cache is not None else None
        if n <= B:
            for k, v in per_batch.items():
                batch[k] = v[0]
            return [batch], ev, keep, tag
        parts = []
        for j, first in enumerate(range(0, n, B)):
            d = {k: v[first:first + B] for k, v in batch.items()}
            for k, v in per_batch.items():
                d[k] = v[j]
            parts.append(d)
        return parts, ev, keep, tag

    def close(self):
        self.stream.synchronize()
        if self.ring is not None:
            self.ring.close()
        self.schema.close()


class _HostCache:
    """A recording kept in page-locked host memory: the batch layout with batch = len(dataset), i.e. one
    column per field whose rows are dataset indices.  Filled once, straight from the mmap'ed files (one
    CPU copy per byte, on the worker pool); after that every batch of every epoch is assembled on the
    device by DMA (`bt_gather_h2d`): no file access, no scan, no CPU copy."""

    def __init__(self, first_msg, n, device):
        self.n, self.device, self.small = n, device, {}
        self.schema = Schema.from_message(first_msg, n)
        if not self.schema.native or self.schema.host_keys:
            self.schema.close()
            raise ValueError("messages need the Python decode")
        self.mem = torch.empty(self.schema.slot_bytes, dtype=torch.uint8, pin_memory=True)
        self.ptr = self.mem.data_ptr()
        self.nbytes = self.schema.slot_bytes

    @staticmethod
    def bytes_needed(first_msg, n):
        sch = Schema.from_message(first_msg, n)
        try:
            return sch.slot_bytes
        finally:
            sch.close()

    def fill(self, ds, threads):
        """Every message of every file to its row; False if one of them needs the Python decode."""
        base = 0
        for d in ds.datasets:
            reader, count = d.reader, len(d.reader)
            for c0 in range(0, count, 512):
                local = np.arange(c0, min(c0 + 512, count))
                addrs, lens = reader.spans(local)
                if self.schema.ingest_many(self.ptr, base + c0, addrs, lens, threads) >= 0:
                    return False
            base += count
        # annotation columns (a few bytes per item) are also kept on the device, whole
        view = self.mem.numpy()
        for c in self.schema.columns:
            if c.item_bytes < self.kSmall:
                rows = view[c.slot_off:c.slot_off + self.n * c.item_bytes].view(c.dtype).reshape((self.n,) + tuple(c.shape))
                self.small[c.key] = torch.from_numpy(rows).to(self.device)
        return True

    kSmall = 16 * 1024   # columns with smaller items also live on the device (see `fill`)

    def gather(self, plan, idx, raw, stream_handle, order_dev=None, start=0):
        """Rows `idx` of every column to the device batch `raw` (the batch layout of `plan`): frames by
        DMA from the page-locked recording, small annotation columns by an index_select from their
        device copy (a DMA per 8-byte item would cost more than it moves)."""
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        n = len(idx)
        where = None
        for col, (key, off, ib, dt, shape, _) in zip(self.schema.columns, plan):
            small = self.small.get(key)
            if small is not None:
                if where is None:
                    where = order_dev[start:start + n] if order_dev is not None else torch.from_numpy(idx).to(raw.device)
                torch.index_select(small, 0, where, out=raw[off:off + n * ib].view(dt).view((n,) + tuple(shape)))
            else:
                N.check(N.lib.bt_gather_h2d(raw.data_ptr() + off, self.ptr + col.slot_off, ib, idx.ctypes.data, n, stream_handle))

    def close(self):
        self.schema.close()
        self.mem = None


class DeviceLoader:
    """Iterate a `RemoteIterableDataset` or `FileDataset` as collated CUDA batches.

    batch_size, shuffle, drop_last: as in ``torch.utils.data.DataLoader`` (shuffle only for
        `FileDataset`; streams cannot be shuffled, generate.py:71-72).
    device: CUDA device of this process (one process per GPU; shard publishers/files with `shard`).
    preprocess: a `Preprocess` or list of them (one per image field) run on the GPU; None leaves
        the uint8 frames as they are -- exactly default_collate's output, on the device.  A `FlipRows`
        in the list flips a field of any dtype vertically (depth maps sent bottom-up).
    project: a `Projection` or list of them: world points carried by the messages are projected to
        pixels (and per-object boxes) on the GPU, one K2 launch per batch.
    num_receivers: receiver threads for streams (default 1); each owns a PULL socket connected to all
        addresses and takes ``max_items // num_receivers`` items, like the reference's DataLoader
        workers (dataset.py:80-85,97).  For replay: producer threads (default 2) that fill alternate
        batches; the sampler's batch order is preserved.
    ring_depth: pinned slots per receiver.  prefetch: finished batches queued ahead of the consumer.
    ingest_threads: host threads `bt_schema_ingest_many` uses per replay batch.
    transport: how streams are received.  'direct' (= 'auto' for tcp:// and ipc:// addresses): the
        library's own ZMTP endpoint, no libzmq: payloads go from the kernel socket straight into pinned
        staging -- one copy and one thread per message instead of two of each; give it about twice
        the receivers you would give pyzmq.  'zmq': pyzmq's socket drained by the C loop (its I/O thread
        overlaps the kernel read with our copy).  Streams with a recording tee or an item_transform
        always use pyzmq with the per-message Python loop.
    io_threads: ZMQ I/O threads per receiver; a receiver's connections (one per publisher) are
        spread over them, so socket reads of different publishers proceed in parallel.
    seed: shuffle seed (epoch number is added on every new iteration).
    keep_raw: also return the untouched uint8 frames as ``<key>_raw``.
    allow_pickle: False = never unpickle.  The reference unpickles every message it receives
        (dataset.py:103,105), and unpickling data from a socket can run code; the native scanner only
        locates payloads.  With False, a message outside its subset raises instead of being decoded by
        ``pickle.loads`` (so no item_transform, no recording tee, no string fields).
    group: how many batches are staged, copied and preprocessed together (one pinned slot, one
        allocation, one copy per column and one launch per kernel for `group` batches, which are then
        handed out one by one as slices): the host cost per batch of small batches all but disappears
        and the copies and launches get longer.  None = as many as make about 32 MiB, at most 8; 1 =
        every batch by itself.  Batches of a group share their storage; results are the same.
    cache: 'device' keeps a replayed recording resident in HBM (raw columns, i.e. uint8 frames): the
        first epoch fills it on the way through, every later epoch is served from it -- no file, host
        or PCIe traffic; K1 gathers the sampler's frames straight from the resident set.  Only for
        `FileDataset` without an item_transform (a transform may differ from epoch to epoch).
        'host' keeps it in page-locked host memory instead (for recordings larger than HBM): it is
        read from the files once, before the first batch, and every batch of every epoch is then
        assembled on the device by DMA from there -- no CPU copy per epoch.
    cache_bytes: memory budget of the cache (default: 80 % of the free device memory when the layout of
        the recording becomes known; half of the available host memory for 'host').  A recording that
        does not fit is replayed uncached, with a warning.
    """

    def __init__(self, dataset, batch_size=1, shuffle=False, drop_last=False, device="cuda",
                 preprocess=None, num_receivers=None, ring_depth=3, prefetch=2, ingest_threads=4,
                 seed=0, keep_raw=False, io_threads=1, project=None, transport="auto", cache=None,
                 cache_bytes=None, allow_pickle=True, group=None):
        self.dataset = dataset
        self.batch_size = int(batch_size)
        if self.batch_size < 1:
            raise ValueError("batch_size must be >= 1")
        self.shuffle, self.drop_last = shuffle, drop_last
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DeviceLoader produces CUDA batches; for host iteration wrap the dataset in "
                               "torch.utils.data.DataLoader exactly as with the reference")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if preprocess is None:
            preprocess = []
        elif isinstance(preprocess, (Preprocess, FlipRows)):
            preprocess = [preprocess]
        self.specs = {p.key: p for p in preprocess}
        if project is None:
            project = []
        elif not isinstance(project, (list, tuple)):
            project = [project]
        self.projections = tuple(project)
        self._num_receivers_arg = num_receivers
        self.num_receivers = max(1, int(num_receivers or 1))
        self.ring_depth, self.prefetch = int(ring_depth), int(prefetch)
        self.ingest_threads = int(ingest_threads)
        self.seed, self.keep_raw = seed, keep_raw
        if group is not None and int(group) < 1:
            raise ValueError("group must be >= 1")
        self.group = None if group is None else int(group)
        self.io_threads = max(1, int(io_threads))
        if transport not in ("auto", "direct", "zmq"):
            raise ValueError("transport must be 'auto', 'direct' or 'zmq'")
        self.transport = transport
        self._epoch = 0
        self._stagers = {}
        if cache not in (None, "device", "host"):
            raise ValueError("cache must be None, 'device' or 'host'")
        self.cache_mode, self.cache_bytes = cache, cache_bytes
        self._cache, self._cache_off, self._cache_lock = None, False, threading.Lock()
        self._host = None
        # counters: who waits for whom tells which side is the bottleneck (consumer_wait_s: the training loop
        # waited for a batch; producer_wait_s: receivers waited for room in the prefetch queue)
        self.stats = {"batches": 0, "items": 0, "bytes_h2d": 0, "kernels": 0, "compat_items": 0,
                      "resident_batches": 0, "cache_bytes": 0, "consumer_wait_s": 0.0, "producer_wait_s": 0.0}
        if isinstance(dataset, RemoteIterableDataset):
            if shuffle:
                raise ValueError("iterable datasets do not support shuffle")
            self._mode = "stream"
        elif isinstance(dataset, FileDataset):
            self._mode = "replay"
        else:
            raise TypeError("DeviceLoader takes a btt.RemoteIterableDataset or a btt.FileDataset")
        self.allow_pickle = bool(allow_pickle)
        if not self.allow_pickle and (dataset.has_custom_item() or getattr(dataset, "record_path_prefix", None)):
            raise ValueError("allow_pickle=False excludes item transforms and the recording tee (both need unpickled items)")
        if cache is not None and (self._mode != "replay" or dataset.has_custom_item()):
            raise ValueError("cache=%r is for FileDataset replay without an item_transform" % cache)

    def __len__(self):
        if self._mode == "replay":
            n = len(self.dataset)
        else:
            n = (self.dataset.max_items // self.num_receivers) * self.num_receivers
            per = self.dataset.max_items // self.num_receivers
            full, rem = divmod(per, self.batch_size)
            return (full + (1 if rem and not self.drop_last else 0)) * self.num_receivers
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    # ---- producers --------------------------------------------------------------------------
    _GROUP_BYTES = 32 << 20

    def _group_for(self, item_bytes):
        """Batches per group for items of `item_bytes`."""
        if self.group is not None:
            return self.group
        per_batch = max(1, item_bytes * self.batch_size)
        return int(max(1, min(8, -(-self._GROUP_BYTES // per_batch))))

    def _group_of_message(self, first_msg, sizes=None):
        if self.group is None and shm.MARKER_BYTES in bytes(memoryview(first_msg)[:4096]):
            # payloads by shared-memory reference are copied per item anyway, and a group would hold
            # its publishers' slots until it is full: with few slots that starves them
            return 1
        try:
            probe = Schema.from_message(first_msg, 1, sizes, self.allow_pickle)
        except Exception:
            return 1
        try:
            return self._group_for(sum(c.item_bytes for c in probe.columns))
        finally:
            probe.close()

    def _new_stager(self, first_msg, rid=0, sizes=None, depth=None, group=1):
        """Stager for the stream that starts with `first_msg`.  Pinned rings are expensive to create
        (page-locking hundreds of MB takes longer than an epoch of small recordings), so a stager
        is kept per producer and reused by later epochs while the layout stays the same."""
        schema = Schema.from_message(first_msg, self.batch_size * group, sizes, self.allow_pickle)
        for key, spec in self.specs.items():
            c = schema.column(key)  # KeyError: the named image field is not in the messages
            if c.dtype != np.uint8 and not isinstance(spec, FlipRows):
                raise TypeError("field %r is %s; Preprocess takes uint8 frames (FlipRows flips anything)" % (key, c.dtype))
        depth = self.ring_depth if depth is None else depth
        sig = (schema.batch, schema.native, tuple((c.key, c.kind, c.dtype.str, c.shape, c.slot_off) for c in schema.columns),
               schema.host_keys, depth)
        cached = self._stagers.get(rid)
        if cached is not None and cached[0] == sig:
            schema.close()
            stager = cached[1]
            torch.cuda.set_stream(stager.stream)   # a new thread runs this epoch
            return stager
        if cached is not None:
            cached[1].close()
        stager = _Stager(schema, self.device, depth, self.specs, self.keep_raw, self.projections)
        self._stagers[rid] = (sig, stager)
        return stager

    @staticmethod
    def _rid():
        name = threading.current_thread().name          # "bt-recv-<k>"
        return int(name.rsplit("-", 1)[-1]) if name.startswith("bt-recv-") else 0

    def _retire(self, stager):
        """End of an epoch: keep cached stagers (their ring), close the others."""
        if stager is None:
            return
        self.stats["kernels"] += stager.kernels
        stager.kernels = 0
        if not any(v[1] is stager for v in self._stagers.values()):
            stager.close()

    def close(self):
        """Release pinned rings and streams kept between epochs, and the resident recording."""
        for _, st in self._stagers.values():
            st.close()
        self._stagers = {}
        if self._cache is not None:
            torch.cuda.synchronize(self.device)    # readers of the resident rows may still be running
            self._cache = None
        if self._host is not None:
            torch.cuda.synchronize(self.device)    # copies out of the pinned recording may still be running
            self._host.close()
            self._host = None

    def _resident(self, schema):
        """The cache the producers fill, created when the first producer knows the layout."""
        if self.cache_mode != "device" or self._cache_off:
            return None
        with self._cache_lock:
            if self._cache is None and not self._cache_off:
                n = len(self.dataset)
                need = _ResidentCache.bytes_needed(schema, n)
                budget = self.cache_bytes
                if budget is None:
                    budget = int(torch.cuda.mem_get_info(self.device)[0] * 0.8)
                if need > budget:
                    self._cache_off = True
                    warnings.warn("recording needs %d bytes on the device, cache budget is %d: replaying uncached"
                                  % (need, budget))
                    return None
                self._cache = _ResidentCache(schema, n, self.device, self.specs)
                self.stats["cache_bytes"] = self._cache.nbytes
            return self._cache

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _fill_host_cache(self):
        """cache='host': read the whole recording into page-locked memory (once)."""
        ds = self.dataset
        n = len(ds)
        if n == 0:
            self._cache_off = True
            return
        first = ds.datasets[0].reader.raw(0)
        t0 = time.perf_counter()
        try:
            need = _HostCache.bytes_needed(first, n)
            budget = self.cache_bytes
            if budget is None:
                import psutil
                budget = psutil.virtual_memory().available // 2
            if need > budget:
                raise ValueError("recording needs %d bytes of page-locked memory, cache budget is %d" % (need, budget))
            host = _HostCache(first, n, self.device)
            if not host.fill(ds, max(self.ingest_threads, 1)):
                host.close()
                raise ValueError("a message needs the Python decode")
        except ValueError as e:
            self._cache_off = True
            warnings.warn("%s: replaying uncached" % e)
            return
        self._host = host
        self.stats["cache_bytes"] = host.nbytes
        self.stats["cache_fill_s"] = time.perf_counter() - t0

    def _emit(self, out_q, stop, stager, slot, n, host_rows, cache=None, idx=None, fill=None):
        item = stager.finish(slot, n, host_rows, cache, idx, fill, self.batch_size)
        self.stats["bytes_h2d"] += stager.item_bytes * n
        t0 = time.perf_counter()
        try:
            while not stop.is_set():
                try:
                    out_q.put(item, timeout=0.1)
                    return True
                except queue.Full:
                    continue
            return False
        finally:
            self.stats["producer_wait_s"] += time.perf_counter() - t0

    def _stream_producer(self, rid, out_q, stop):
        ds = self.dataset
        stager = None
        ctx = None
        sock = None
        rec = None
        try:
            torch.cuda.set_device(self.device)
            fast = ds.record_path_prefix is None and not ds.has_custom_item()
            # 'auto' prefers the direct endpoint: one thread and one copy per message instead of two of
            # each; with enough receivers it is the faster path on a busy host
            direct = fast and self.transport != "zmq" and direct_capable(ds.addresses)
            if self.transport == "direct" and not direct:
                raise ValueError("transport='direct' needs tcp:// or ipc:// addresses, no recording tee and no item_transform")
            if not direct:
                ctx = zmq.Context(io_threads=self.io_threads)
                sock, poller = open_pull_socket(ctx, ds.addresses, ds.queue_size)
            if ds.record_path_prefix is not None:
                rec = FileRecorder(FileRecorder.filename(ds.record_path_prefix, rid), ds.max_items, async_write=True)
                rec.__enter__()
            holder = []
            try:
                if direct:
                    self._stream_fast(ds, DirectReceiver(ds.addresses), out_q, stop, holder)
                elif fast:
                    self._stream_fast(ds, Receiver(sock), out_q, stop, holder)
                else:
                    self._stream_compat(ds, sock, poller, rec, out_q, stop, holder)
            finally:
                stager = holder[0] if holder else None
        except BaseException as e:  # delivered to the consumer thread
            out_q.put(e)
        finally:
            if rec is not None:
                rec.__exit__(None, None, None)
            if sock is not None:
                try:  # do not strand shared-memory slots announced by messages we will never read
                    tail = Receiver(sock)
                    tail.release_queued()
                    tail.close()
                except Exception:
                    pass
                sock.close()
            if ctx is not None:
                ctx.term()
            self._retire(stager)
            out_q.put(_DONE)

    def _decode_pending(self, rx):
        """The held message as the object the producer published (the reference's decode, for either
        framing): raw-parts frames are joined, shared-memory references fetched."""
        refuse_pickle(self.allow_pickle)
        msg, sizes = rx.pending(), rx.pending_parts()
        if len(sizes) > 1:
            obj = wire.loads(msg, sizes)
        elif rx.more():
            obj = wire.decode([bytes(msg)] + rx.recv_rest())
        else:
            obj = pickle.loads(msg)
        return shm.resolve(obj)

    def _stream_fast(self, ds, rx, out_q, stop, holder):
        """Batches filled by a C receive loop (`Receiver`: libzmq socket, `DirectReceiver`: own ZMTP
        endpoint): one call per batch, no GIL per message."""
        try:
            remaining = ds.max_items // self.num_receivers
            waited = 0
            # first message: defines the schema of the stream
            while remaining > 0 and not rx.recv(100):
                waited += 100
                if stop.is_set():
                    return
                assert waited < ds.timeoutms, "No response within timeout interval."
            if remaining <= 0:
                return
            group = self._group_of_message(rx.pending(), rx.pending_parts())
            stager = self._new_stager(rx.pending(), self._rid(), rx.pending_parts(), group=group)
            holder.append(stager)
            ring, schema = stager.ring, stager.schema
            while remaining > 0 and not stop.is_set():
                slot, filled, host_rows = stager.begin(), 0, {}
                want = min(self.batch_size * group, remaining)
                waited = 0
                while filled < want:
                    if schema.native:
                        rc, n = rx.fill(ring, slot, schema, filled, want - filled, 100)
                    else:  # schema needs the pickle.loads decode for every message
                        rc, n = (N.BT_E_UNSUPPORTED, 0) if rx.recv(100) else (N.BT_E_TIMEOUT, 0)
                    filled += n
                    if n:
                        waited = 0
                    if rc == N.BT_OK:
                        break
                    if rc == N.BT_E_TIMEOUT:
                        waited += 100
                        if stop.is_set():
                            return
                        if filled and filled % self.batch_size == 0:
                            break    # the stream paused on a batch boundary: hand out what is complete
                        # the stream stalled: do not sit on shared-memory slots of a partly filled
                        # batch (their publishers may be waiting for exactly those slots)
                        ring.spill(slot, schema)
                        ring.reap()
                        assert waited < ds.timeoutms, "No response within timeout interval."
                        continue
                    # this message is outside the native subset: the reference's own decode, same slot
                    stager.schema.ingest_object(ring.view(slot), filled, self._decode_pending(rx), host_rows)
                    rx.drop()
                    self.stats["compat_items"] += 1
                    filled += 1
                    waited = 0
                remaining -= filled
                if self.drop_last and filled % self.batch_size:      # only the very last batch can be short
                    filled -= filled % self.batch_size
                    remaining = 0
                    if not filled:
                        break
                if not self._emit(out_q, stop, stager, slot, filled, host_rows):
                    return
        finally:
            rx.close()

    def _stream_compat(self, ds, sock, poller, rec, out_q, stop, holder):
        """Per-message Python loop: needed when every message has to pass through Python anyway
        (a user item_transform / overridden _item, or the recording tee)."""
        stager = None
        custom = ds.has_custom_item()
        slot, filled, host_rows = None, 0, {}
        for _ in range(ds.max_items // self.num_receivers):
            if stop.is_set():
                return
            waited = 0
            while True:  # poll in slices so that an abandoned iterator stops promptly
                step = min(100, max(ds.timeoutms - waited, 0))
                if sock in dict(poller.poll(step)):
                    break
                waited += step
                if stop.is_set():
                    return
                if stager is not None:
                    # the stream stalled: do not sit on shared-memory slots of a partly filled
                    # batch (their publishers may be waiting for exactly those slots)
                    if slot is not None:
                        stager.ring.spill(slot, stager.schema)
                    stager.ring.reap()
                assert waited < ds.timeoutms, "No response within timeout interval."
            frame = sock.recv(copy=False)
            msg = frame.buffer
            parts = sock.recv_multipart(copy=False) if sock.getsockopt(zmq.RCVMORE) else None   # raw-parts wire
            if stager is not None:
                stager.ring.reap()
            by_ref = parts is not None or (shm.MARKER_BYTES in bytes(msg[:4096]) if (rec is not None or custom) else False)
            if rec is not None and not by_ref:
                rec.save(msg, is_pickled=True)
            if custom or parts is not None or (rec is not None and by_ref):
                obj = shm.resolve(wire.decode([msg] + parts) if parts is not None else pickle.loads(msg))
                if rec is not None and by_ref:
                    # recordings must be self-contained: store the message with its payload in-line
                    rec.save(pickle.dumps(obj, protocol=3), is_pickled=True)
                obj = ds._item(obj)
                if stager is None:
                    stager = _Stager(Schema.from_object(obj, self.batch_size), self.device,
                                     self.ring_depth, self.specs, self.keep_raw, self.projections)
                if slot is None:
                    slot, filled, host_rows = stager.begin(), 0, {}
                stager.schema.ingest_object(stager.ring.view(slot), filled, obj, host_rows)
                self.stats["compat_items"] += 1
            else:
                if stager is None:
                    stager = self._new_stager(msg, self._rid())
                    holder.append(stager)
                if slot is None:
                    slot, filled, host_rows = stager.begin(), 0, {}
                stager.ingest(slot, filled, msg, host_rows)
            filled += 1
            if filled == self.batch_size:
                if not self._emit(out_q, stop, stager, slot, filled, host_rows):
                    return
                slot = None
        if slot is not None and filled and not self.drop_last:
            self._emit(out_q, stop, stager, slot, filled, host_rows)

    def _replay_order(self):
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self._epoch)
            return torch.randperm(n, generator=g).numpy()
        return np.arange(n)

    def _replay_producer(self, rid, out_q, stop):
        ds = self.dataset
        stager = None
        try:
            torch.cuda.set_device(self.device)
            order = self._replay_order()
            cum = np.asarray(ds.cumulative_sizes)
            readers = [d.reader for d in ds.datasets]
            custom = ds.has_custom_item()
            B = self.batch_size
            n_prod = self._replay_producers()
            group = 1 if custom or not len(order) else self._group_of_message(readers[0].raw(0))
            G = B * group
            for start in range(rid * G, len(order), G * n_prod):   # producer `rid` takes groups rid, rid+n, ...
                if stop.is_set():
                    return
                idx = order[start:start + G]
                if self.drop_last and len(idx) % B:
                    idx = idx[:len(idx) // B * B]
                    if not len(idx):
                        break
                if self._host is not None:
                    # the recording is in page-locked memory: this batch is a handful of DMA copies
                    if stager is None:
                        stager = self._new_stager(readers[0].raw(0), rid, depth=0, group=group)
                        order_dev = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int64)).to(self.device)
                    host, plan = self._host, stager.plan
                    if not self._emit(out_q, stop, stager, None, len(idx), {}, None, None,
                                      lambda raw, st, idx=idx, start=start: host.gather(plan, idx, raw, st, order_dev, start)):
                        return
                    continue
                fidx = np.searchsorted(cum, idx, side="right")
                local = idx - np.where(fidx > 0, cum[np.maximum(fidx - 1, 0)], 0)
                if custom:
                    objs = [ds._item(readers[f][int(l)]) for f, l in zip(fidx, local)]
                    if stager is None:
                        stager = _Stager(Schema.from_object(objs[0], B), self.device, self.ring_depth,
                                         self.specs, self.keep_raw, self.projections)
                    slot, host_rows = stager.begin(), {}
                    for i, o in enumerate(objs):
                        stager.schema.ingest_object(stager.ring.view(slot), i, o, host_rows)
                    self.stats["compat_items"] += len(objs)
                else:
                    addrs = np.empty(len(idx), dtype=np.uint64)
                    lens = np.empty(len(idx), dtype=np.uint64)
                    for f in np.unique(fidx):
                        m = fidx == f
                        addrs[m], lens[m] = readers[f].spans(local[m])
                    if stager is None:
                        stager = self._new_stager(readers[fidx[0]].raw(int(local[0])), rid, group=group)
                    slot, host_rows = stager.begin(), {}
                    done = 0
                    if stager.schema.native:
                        bad = stager.schema.ingest_many(stager.ring.ptr(slot), 0, addrs, lens, self.ingest_threads)
                        done = len(idx) if bad < 0 else -1
                    if done != len(idx):  # at least one message needs the compatibility decode
                        for i, (f, l) in enumerate(zip(fidx, local)):
                            stager.ingest(slot, i, readers[f].raw(int(l)), host_rows)
                cache = None if custom else self._resident(stager.schema)
                if not self._emit(out_q, stop, stager, slot, len(idx), host_rows, cache, idx):
                    return
        except BaseException as e:
            out_q.put(e)
        finally:
            self._retire(stager)
            out_q.put(_DONE)

    def _resident_epoch(self, order):
        """An epoch served from HBM: per batch one K1 launch that gathers its frames from the resident
        set by index (other columns: index_select), all on the consumer's current stream."""
        cache, B, dev = self._cache, self.batch_size, self.device
        with torch.cuda.device(dev):
            order_dev = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int64)).to(dev)
            for _, st in self._stagers.values():   # the streams that wrote the resident rows
                torch.cuda.current_stream(dev).wait_stream(st.stream)
            # annotation columns are small: gather them for the whole epoch at once, batches are views
            whole = {}
            for key, prep in cache.plan:
                store = cache.cols[key]
                if (prep is None or self.keep_raw) and store[:1].numel() * store.element_size() * len(order) <= (64 << 20):
                    whole[key] = store.index_select(0, order_dev)
            seen_stream = torch.cuda.current_stream(dev)
        kernels = 0
        try:
            for start in range(0, len(order), B):
                idx = order_dev[start:start + B]
                n = idx.numel()
                if n < B and self.drop_last:
                    break
                with torch.cuda.device(dev):
                    cur = torch.cuda.current_stream(dev)
                    if cur != seen_stream:             # the consumer moved to another stream
                        seen_stream = cur
                        for t in (order_dev, *whole.values()):
                            t.record_stream(cur)
                    handle = C.c_void_p(cur.cuda_stream)
                    batch = {}
                    for key, prep in cache.plan:
                        store = cache.cols[key]
                        if prep is None or self.keep_raw:
                            t = whole[key][start:start + B] if key in whole else store.index_select(0, idx)
                        if prep is not None:
                            out, stats = prep.run(store.data_ptr(), n, handle, idx.data_ptr())
                            kernels += prep.last_kernels
                            if stats is not None:
                                batch[key + "_stats"] = stats
                            if self.keep_raw:
                                batch[key + "_raw"] = t
                            t = out
                        batch[key] = t
                    for pj in self.projections:
                        pj.apply(batch, dev)
                        kernels += 1
                host_idx = order[start:start + B]
                for k, rows in cache.host.items():
                    batch[k] = [rows[int(i)] for i in host_idx]
                self.stats["batches"] += 1
                self.stats["resident_batches"] += 1
                self.stats["items"] += n
                yield batch
        finally:
            self.stats["kernels"] += kernels

    # ---- consumer ---------------------------------------------------------------------------
    def _replay_producers(self):
        # two producers by default: one fills a pinned slot while the other submits its batch
        n = 2 if self._num_receivers_arg is None else self.num_receivers
        return max(1, min(n, 8))

    def __iter__(self):
        if self._cache is not None:
            self._epoch += 1                       # the order of the epoch that starts now
            order = self._replay_order()
            usable = len(order) - (len(order) % self.batch_size if self.drop_last else 0)
            if self._cache.covers(order[:usable]):
                yield from self._resident_epoch(order)
                return
            self._epoch -= 1
        if self.cache_mode == "host" and self._host is None and not self._cache_off:
            self._fill_host_cache()
        stop = threading.Event()
        if self._mode == "stream":
            n_prod, target = self.num_receivers, self._stream_producer
            queues = [queue.Queue(maxsize=max(1, self.prefetch))] * n_prod      # one shared queue: unordered
        else:
            # Replay keeps the sampler's order: producer r makes batches r, r+n, ... into its own
            # queue and the consumer reads the queues round-robin.
            n_prod, target = self._replay_producers(), self._replay_producer
            queues = [queue.Queue(maxsize=max(1, self.prefetch)) for _ in range(n_prod)]
        threads = [threading.Thread(target=target, args=(r, queues[r], stop), daemon=True, name="bt-recv-%d" % r)
                   for r in range(n_prod)]
        self._epoch += 1
        for t in threads:
            t.start()
        ordered = self._mode != "stream"
        alive = [True] * n_prod
        turn = 0
        try:
            while any(alive):
                t0 = time.perf_counter()
                if ordered:
                    if not alive[turn]:
                        break                      # the producer whose turn it is has no more batches
                    item = queues[turn].get()
                else:
                    item = queues[0].get()
                self.stats["consumer_wait_s"] += time.perf_counter() - t0
                if item is _DONE:
                    if ordered:
                        alive[turn] = False
                        # a finished producer ends the epoch only when its turn comes round again;
                        # later producers may still hold the batches in between
                        turn = (turn + 1) % n_prod
                        if not any(alive):
                            break
                        continue
                    alive[alive.index(True)] = False
                    continue
                if isinstance(item, BaseException):
                    raise item
                batches, ev, keep, tag = item
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                for t in keep:
                    t.record_stream(cur)
                if tag is not None and self._cache is not None:
                    self._cache.commit(tag[0], tag[1])
                for batch in batches:          # the batches of one group, in order
                    self.stats["batches"] += 1
                    self.stats["items"] += len(batch["btid"]) if "btid" in batch else next((len(v) for v in batch.values()), 0)
                    yield batch
                if ordered:
                    turn = (turn + 1) % n_prod
        finally:
            stop.set()
            _log.debug("DeviceLoader epoch %d: %s", self._epoch, self.stats)
            # unblock producers waiting on a full queue, then let them finish cleanly
            for t, q in zip(threads, queues):
                while t.is_alive():
                    try:
                        q.get_nowait()
                    except queue.Empty:
                        pass
                    t.join(timeout=0.05)
