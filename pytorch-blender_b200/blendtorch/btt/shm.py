"""Shared-memory wire (opt-in): payloads travel through a POSIX shared-memory segment, the ZMQ
message only carries a reference.  No counterpart in the reference (its only transport is the
pickled ZMQ message, btb/publisher.py:43); see include/bt_b200.h `bt_shm_*`.

A message on this wire is the usual pickled dict in which a large array is replaced by

    ("__btshm__", segment_name, ticket, dtype_str, shape)

where ``ticket = slot | generation << 16``: every hand-out of a slot has its own generation, so a
consumer that holds a reference past the producer's lease (see `Segment.set_lease`) notices that the
slot has moved on -- before and after it copies -- and raises instead of delivering another frame.

Both ends must opt in: the producer with ``DataPublisher(..., shm_slots=N)``, the consumer by being
this package (``DeviceLoader`` DMAs the slot straight to the GPU; host iteration copies it out).
"""
import ctypes as C
import os

import numpy as np

from . import native as N

MARKER = "__btshm__"
MARKER_BYTES = b"__btshm__"


def is_marker(v):
    return isinstance(v, tuple) and len(v) == 5 and v[0] == MARKER


class Segment:
    """Producer side: a segment of `nslots` slots of `slot_bytes` each."""

    def __init__(self, slot_bytes, nslots, name=None):
        self.name = name or "/btshm_%d_%x" % (os.getpid(), int.from_bytes(os.urandom(4), "little"))
        self.handle = C.c_void_p()
        N.check(N.lib.bt_shm_create(self.name.encode(), int(slot_bytes), int(nslots), C.byref(self.handle)))
        self.nslots = nslots
        self.slot_bytes = N.lib.bt_shm_slot_bytes(self.handle)
        self._views = []
        for i in range(nslots):
            p = N.lib.bt_shm_slot_ptr(self.handle, i)
            self._views.append(np.frombuffer((C.c_uint8 * self.slot_bytes).from_address(p), dtype=np.uint8))

    def acquire(self, timeout_ms=10000):
        """Ticket (slot | generation << 16) of a free slot; waits up to `timeout_ms` for one."""
        s = C.c_int()
        N.check(N.lib.bt_shm_acquire(self.handle, int(timeout_ms), C.byref(s)))
        return s.value

    def view(self, ticket):
        return self._views[ticket & 0xFFFF]

    def set_lease(self, lease_ms):
        """How long a consumer may hold a slot before `acquire` takes it back (default 30 s, or
        ``BT_SHM_LEASE_MS``; 0 = never).  Taking slots back keeps consumers that died with messages queued
        from starving the producer, and is lossy: a live consumer slower than the lease loses those
        frames -- loudly (its copy fails the generation check), never silently."""
        N.check(N.lib.bt_shm_set_lease(self.handle, int(lease_ms)))

    def put(self, array, timeout_ms=10000):
        """Copy `array` into the next free slot and return the marker that announces it."""
        a = np.ascontiguousarray(array)
        if a.nbytes > self.slot_bytes:
            raise ValueError("array of %d bytes does not fit a %d-byte slot" % (a.nbytes, self.slot_bytes))
        ticket = self.acquire(timeout_ms)
        self.view(ticket)[:a.nbytes] = a.reshape(-1).view(np.uint8)
        return (MARKER, self.name, ticket, a.dtype.str, tuple(a.shape))

    def drain(self, timeout_ms):
        """Wait (up to `timeout_ms` per slot) until consumers have released every slot."""
        try:
            for _ in range(self.nslots):
                N.check(N.lib.bt_shm_release(self.handle, self.acquire(timeout_ms)))
        except N.NativeError:
            pass

    def close(self):
        if self.handle:
            self._views = []
            N.lib.bt_shm_close(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_attached = {}


def fetch(marker):
    """Consumer side without a GPU in the loop: copy the announced array out and free its slot."""
    _, name, slot, dtype_str, shape = marker
    h = _attached.get(name)
    if h is None:
        h = C.c_void_p()
        N.check(N.lib.bt_shm_attach(name.encode(), -1, C.byref(h)))
        _attached[name] = h
    dt = np.dtype(dtype_str)
    n = int(np.prod(shape, dtype=np.int64)) * dt.itemsize
    p = N.lib.bt_shm_slot_ptr(h, int(slot))
    if not p or n > N.lib.bt_shm_slot_bytes(h):
        raise N.NativeError(N.BT_E_INVALID, "shared-memory reference out of range")
    N.check(N.lib.bt_shm_check(h, int(slot)))      # still ours before the copy ...
    out = np.frombuffer((C.c_uint8 * n).from_address(p), dtype=np.uint8).copy().view(dt).reshape(shape)
    N.check(N.lib.bt_shm_release(h, int(slot)))    # ... and after it (the release compares generations)
    return out


def resolve(obj):
    """Replace every marker in a received message dict by the array it announces."""
    if isinstance(obj, dict):
        for k, v in obj.items():
            if is_marker(v):
                obj[k] = fetch(v)
    return obj
