"""Batch schema and pinned staging: messages -> batch-contiguous pinned slot -> device.

This is the host half of the new ingest path.  The reference moves every payload byte through
``recv -> bytes -> ndarray -> (transform temporaries) -> shared-memory batch -> pageable
staging`` (btt/dataset.py:100-106, torch collate.py:246-289, densityopt.py:261,268).  Here a
message's payloads are copied exactly once, by ``bt_schema_ingest`` in C with the GIL released,
straight into their final position inside a page-locked slot whose layout is the collated
batch; the slot then goes to the GPU with one ``cudaMemcpyAsync``.

A `Schema` says where every field of a message lives inside a slot.  It is normally derived in
C from the first message (`Schema.from_message`); messages the native scanner does not cover
(strings, nested containers, Fortran order, big-endian ...) use the same layout filled from
``pickle.loads`` -- the reference's own decode -- so they still end up on the device through the
same copy and the same kernels.
"""
import ctypes as C
import pickle

import numpy as np

from . import native as N
from . import shm
from . import wire

_ALIGN = 256


class ColumnSpec:
    __slots__ = ("key", "kind", "dtype", "shape", "item_bytes", "slot_off")

    def __init__(self, key, kind, dtype, shape, item_bytes, slot_off):
        self.key, self.kind, self.dtype = key, kind, np.dtype(dtype)
        self.shape, self.item_bytes, self.slot_off = tuple(shape), int(item_bytes), int(slot_off)

    def __repr__(self):
        return "ColumnSpec(%r, %s%s @%d)" % (self.key, self.dtype, list(self.shape), self.slot_off)


def refuse_pickle(allow_pickle):
    """The reference unpickles whatever arrives on the socket (dataset.py:103,105); pickle can run code.
    The native scanner never does; with allow_pickle=False nothing else is allowed to decode a message."""
    if not allow_pickle:
        raise RuntimeError("message outside the natively scanned subset (%s) and allow_pickle=False"
                           % (N.last_error() or "no reason recorded"))


def _align(v):
    return (v + _ALIGN - 1) // _ALIGN * _ALIGN


class Schema:
    """Column layout of one staging slot for batches of up to `batch` messages."""

    def __init__(self, batch, columns, host_keys=(), handle=None):
        self.batch = batch
        self.columns = columns
        self.host_keys = tuple(host_keys)  # fields kept on the host (str etc.), collated as lists
        self.handle = handle               # bt_schema* when the native fast path applies
        self.allow_pickle = True           # False: refuse messages that would need pickle.loads
        end = 0
        for c in columns:
            end = max(end, c.slot_off + c.item_bytes * batch)
        self.slot_bytes = max(_align(end), _ALIGN)

    @property
    def native(self):
        return self.handle is not None

    @classmethod
    def from_message(cls, msg, batch, sizes=None, allow_pickle=True):
        """Derive the schema from the first message of a stream (bytes-like pickle).  `sizes`: frame
        sizes when `msg` holds a raw-parts message (wire.py), header frame first.  `allow_pickle=False`:
        messages the native scanner does not cover are refused instead of unpickled."""
        addr, n = N.buffer_address(msg)
        if sizes is not None and len(sizes) > 1:
            n = int(sizes[0])
        h = C.c_void_p()
        rc = N.lib.bt_schema_create(addr, n, batch, C.byref(h))
        if rc == N.BT_OK:
            cols = []
            col = N.Column()
            for i in range(N.lib.bt_schema_num_columns(h)):
                N.check(N.lib.bt_schema_column(h, i, C.byref(col)))
                cols.append(ColumnSpec(col.key.decode(), col.kind, N.NUMPY_DTYPES[col.dtype],
                                       [col.shape[d] for d in range(col.ndim)], col.item_bytes, col.slot_off))
            out = cls(batch, cols, handle=h)
            out.allow_pickle = allow_pickle
            return out
        if rc not in (N.BT_E_UNSUPPORTED,):
            N.check(rc)
        refuse_pickle(allow_pickle)
        return cls.from_object(shm.resolve(wire.loads(msg, sizes)), batch)

    @classmethod
    def from_object(cls, obj, batch):
        """Same layout rules, from an unpickled message (the compatibility path)."""
        if not isinstance(obj, dict):
            raise TypeError("messages must be dicts (btb/publisher.py:42), got %s" % type(obj))
        cols, host, off = [], [], 0
        for key, v in obj.items():
            if isinstance(v, np.ndarray) or isinstance(v, np.generic):
                a = np.asarray(v)
                if a.dtype.hasobject or a.dtype.fields is not None:
                    host.append(key)
                    continue
                spec = (N.BT_KIND_NDARRAY, a.dtype.newbyteorder("="), a.shape, a.nbytes)
            elif isinstance(v, bool):
                spec = (N.BT_KIND_BOOL, np.bool_, (), 1)
            elif isinstance(v, int) and -2 ** 63 <= v < 2 ** 63:
                spec = (N.BT_KIND_INT, np.int64, (), 8)
            elif isinstance(v, float):
                spec = (N.BT_KIND_FLOAT, np.float64, (), 8)
            else:
                host.append(key)
                continue
            cols.append(ColumnSpec(key, spec[0], spec[1], spec[2], spec[3], off))
            off = _align(off + spec[3] * batch)
        return cls(batch, cols, host_keys=host)

    def column(self, key):
        for c in self.columns:
            if c.key == key:
                return c
        raise KeyError(key)

    def close(self):
        if self.handle is not None:
            N.lib.bt_schema_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- filling a slot ---------------------------------------------------------------------
    def ingest(self, slot_ptr, slot_view, item, msg, host_rows=None):
        """Place message `msg` (bytes-like) as item `item` of the slot.

        slot_ptr: integer address of the slot; slot_view: uint8 numpy view of the same memory.
        """
        if self.handle is not None:
            addr, n = N.buffer_address(msg)
            rc = N.lib.bt_schema_ingest(self.handle, slot_ptr, item, addr, n)
            if rc == N.BT_OK:
                return
            if rc not in (N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
                N.check(rc)
        refuse_pickle(self.allow_pickle)
        self.ingest_object(slot_view, item, shm.resolve(pickle.loads(msg)), host_rows)

    def ingest_object(self, slot_view, item, obj, host_rows=None):
        for c in self.columns:
            if c.key not in obj:
                raise KeyError("message lacks field %r present in the first message of the stream" % c.key)
            v = np.asarray(obj[c.key])
            if v.shape != c.shape:
                # same failure class as default_collate's torch.stack on ragged items
                raise RuntimeError("field %r: each element in a batch must have equal size, got %s and %s"
                                   % (c.key, list(c.shape), list(v.shape)))
            dst = slot_view[c.slot_off + item * c.item_bytes: c.slot_off + (item + 1) * c.item_bytes]
            dst.view(c.dtype).reshape(c.shape)[...] = v
        if host_rows is not None:
            for k in self.host_keys:
                host_rows.setdefault(k, []).append(obj.get(k))

    def ingest_many(self, slot_ptr, first_item, addrs, lens, nthreads=1):
        """Native bulk ingest of messages given as (address, length) arrays; returns the index of
        the first message that needs the compatibility path, or -1."""
        assert self.handle is not None
        n = len(addrs)
        bad = C.c_int(-1)
        addrs = np.ascontiguousarray(addrs, dtype=np.uint64)
        lens = np.ascontiguousarray(lens, dtype=np.uint64)
        rc = N.lib.bt_schema_ingest_many(self.handle, slot_ptr, first_item, n, addrs.ctypes.data,
                                         lens.ctypes.data, nthreads, C.byref(bad))
        if rc == N.BT_OK:
            return -1
        if rc in (N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
            return bad.value
        N.check(rc)


class Ring:
    """`depth` page-locked slots of `slot_bytes` (bt_ring_*); device=None gives an unpinned ring
    for host-only tools and tests."""

    def __init__(self, slot_bytes, depth=3, device=None):
        self.handle = C.c_void_p()
        dev = -1 if device is None else int(device)
        N.check(N.lib.bt_ring_create(dev, int(slot_bytes), int(depth), C.byref(self.handle)))
        self.depth = depth
        self.slot_bytes = N.lib.bt_ring_slot_bytes(self.handle)
        self._ptrs, self._views = [], []
        for s in range(depth):
            p = N.lib.bt_ring_slot_ptr(self.handle, s)
            self._ptrs.append(p)
            buf = (C.c_uint8 * self.slot_bytes).from_address(p)
            self._views.append(np.frombuffer(buf, dtype=np.uint8))

    def acquire(self):
        """Index of the next slot, free for writing (waits for its last H2D copy)."""
        s = C.c_int()
        N.check(N.lib.bt_ring_acquire(self.handle, C.byref(s)))
        return s.value

    def ptr(self, slot):
        return self._ptrs[slot]

    def view(self, slot):
        return self._views[slot]

    def submit(self, slot, dev_ptr, nbytes, stream_handle):
        N.check(N.lib.bt_ring_submit(self.handle, slot, dev_ptr, nbytes, stream_handle))

    def ingest(self, slot, schema, item, msg, host_rows=None):
        """Place message `msg` as item `item` of `slot`.  Payloads carried in-line are copied once
        into the pinned slot; payloads announced by shared-memory reference are only noted and go
        to the GPU by DMA in `submit_batch`."""
        if schema.handle is not None:
            addr, n = N.buffer_address(msg)
            rc = N.lib.bt_ring_ingest(self.handle, slot, schema.handle, item, addr, n)
            if rc == N.BT_OK:
                return
            if rc not in (N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
                N.check(rc)
        refuse_pickle(schema.allow_pickle)
        schema.ingest_object(self._views[slot], item, shm.resolve(pickle.loads(msg)), host_rows)

    def submit_batch(self, slot, schema, n_items, dev_ptr, stream_handle):
        """H2D of a filled slot: per column from pinned memory, per item from shared memory."""
        if schema.handle is not None:
            N.check(N.lib.bt_ring_submit_batch(self.handle, slot, schema.handle, n_items, dev_ptr, stream_handle))
        else:
            used = 0
            for c in schema.columns:
                used = max(used, c.slot_off + n_items * c.item_bytes)
            if used:
                self.submit(slot, dev_ptr, used, stream_handle)

    def spill(self, slot, schema):
        if schema.handle is not None:
            N.check(N.lib.bt_ring_spill(self.handle, slot, schema.handle))

    def reap(self):
        N.check(N.lib.bt_ring_reap(self.handle))

    def wait(self, slot):
        N.check(N.lib.bt_ring_wait(self.handle, slot))

    def close(self):
        if self.handle:
            self._views = []
            N.lib.bt_ring_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
