"""bpy-free data publisher: the producer end of the ingest path.

Same class name, constructor and ``publish(**kwargs)`` contract as the reference's Blender-side
``btb.DataPublisher`` (pkg_blender/blendtorch/btb/publisher.py:21-43): a PUSH socket that binds,
a bounded send queue that blocks the caller when consumers are slow, and messages that are
``pickle({'btid': btid, **kwargs})``.  It exists on the btt side so that synthetic producers
(tests, benchmarks, simulators that are not Blender) speak the exact wire format.

Opt-in extensions (understood by this package, not by the reference): ``shm_slots=N``, for producers
on the same host as their consumers, sends every array of at least ``shm_threshold`` bytes through a
shared-memory segment and only a reference through ZMQ (see shm.py); ``raw_parts=True``, for any
producer, sends large arrays as separate zero-copy frames of a multipart message instead of pickling
them (see wire.py).
"""
import pickle

import numpy as np
import zmq


class DataPublisher:
    """
    bind_address: ZMQ address to bind to.
    btid: integer producer id added to every message (None is sent as-is, and like in the
        reference cannot be collated -- pass an int).
    send_hwm: send queue size before `publish` blocks.
    lingerms: how long unsent messages survive `close`.
    protocol: pickle protocol; None = pyzmq's default, 3 = what Blender's Python 3.7 emits.
    shm_slots: 0 = reference wire (everything pickled in-line); N > 0 = shared-memory wire with N
        slots (the segment is created on the first large array and sized for it).
    shm_threshold: smallest array, in bytes, that goes through shared memory.
    shm_timeoutms: how long `publish` waits for a free slot before raising (consumers gone).
    raw_parts: True = arrays of at least `raw_threshold` bytes travel as their own frames, sent
        straight from the array's memory (no pickling copy, no pyzmq copy); the caller must not
        modify a published array until it has been sent (pass a fresh array per frame, as renderers do).
    """

    def __init__(self, bind_address, btid=None, send_hwm=10, lingerms=0, protocol=None,
                 shm_slots=0, shm_threshold=65536, shm_timeoutms=10000, raw_parts=False, raw_threshold=4096):
        self.ctx = zmq.Context()
        self.sock = self.ctx.socket(zmq.PUSH)
        self.sock.setsockopt(zmq.SNDHWM, send_hwm)
        self.sock.setsockopt(zmq.LINGER, lingerms)
        self.sock.setsockopt(zmq.IMMEDIATE, 1)
        self.sock.bind(bind_address)
        self.btid = btid
        self.lingerms = lingerms
        self.protocol = pickle.DEFAULT_PROTOCOL if protocol is None else protocol  # == pyzmq default
        self.shm_slots, self.shm_threshold, self.shm_timeoutms = int(shm_slots), int(shm_threshold), int(shm_timeoutms)
        self.raw_parts, self.raw_threshold = bool(raw_parts), int(raw_threshold)
        self._segment = None

    def _by_reference(self, array):
        from . import shm
        if self._segment is None:
            self._segment = shm.Segment(array.nbytes, self.shm_slots)
        if array.nbytes > self._segment.slot_bytes:
            return array  # does not fit the segment sized by the first frame: send it in-line
        return self._segment.put(array, self.shm_timeoutms)

    def publish(self, **kwargs):
        """Send one message; `btid` is added implicitly."""
        data = {"btid": self.btid, **kwargs}
        if self.shm_slots > 0:
            for k, v in data.items():
                if isinstance(v, np.ndarray) and v.nbytes >= self.shm_threshold and not v.dtype.hasobject:
                    data[k] = self._by_reference(v)
        if self.raw_parts:
            from . import wire
            self.sock.send_multipart(wire.pack(data, self.raw_threshold, self.protocol), copy=False)
            return
        self.publish_raw(pickle.dumps(data, protocol=self.protocol))

    def publish_raw(self, data, copy=True):
        """Send an already pickled message (replaying recordings, pre-pickled synthetic frames)."""
        self.sock.send(data, copy=copy)

    def close(self):
        self.sock.close()
        self.ctx.term()
        if self._segment is not None:
            if self.lingerms > 0:
                self._segment.drain(self.lingerms)  # let consumers finish with the slots they hold
            self._segment.close()
            self._segment = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
