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


class NativePublisher:
    """Several publishers in one process without libzmq: the library's own ZMTP PUSH endpoints
    (``bt_push_*``, csrc/push.cpp), for producers that already hold pickled messages -- replaying
    recordings, synthetic load, simulators that are not Python.

    Each bound address is one publisher in the sense of btb/publisher.py:21-27 (PUSH, bind, messages only
    to completed connections, blocking when no consumer can take more); unmodified
    ``btt.RemoteIterableDataset`` consumers and the direct receiver connect to them alike.  A message
    goes from the caller's buffer to the kernel socket in one copy and the GIL is released while it
    does, so a few threads of one process do the work of a process (plus libzmq I/O thread) per
    publisher.

        pub = NativePublisher(["tcp://127.0.0.1:*"] * 5)
        pub.addresses                       # what consumers connect to
        pub.publish_raw(blob, endpoint=2)   # one pickled message
        pub.start(pools, threads=2)         # cycle pre-pickled pools on background threads
    """

    def __init__(self, bind_addresses):
        import ctypes as C
        from . import native as N
        self._C, self._N = C, N
        if isinstance(bind_addresses, str):
            bind_addresses = [bind_addresses]
        self.handle = C.c_void_p()
        N.check(N.lib.bt_push_create(C.byref(self.handle)))
        self.addresses = []
        try:
            for a in bind_addresses:
                ep = N.check(N.lib.bt_push_bind(self.handle, a.encode()))
                buf = C.create_string_buffer(256)
                N.check(N.lib.bt_push_endpoint(self.handle, ep, buf, 256))
                self.addresses.append(buf.value.decode())
        except Exception:
            self.close()
            raise
        self._threads, self._keep = [], []
        self._stop = C.c_int(0)
        self._sent = []

    def peers(self, endpoint=0):
        """Consumers connected to `endpoint` right now (not while `start`'s threads are running: an
        endpoint's connection list belongs to one thread at a time)."""
        return self._N.check(self._N.lib.bt_push_peers(self.handle, endpoint))

    def publish_raw(self, data, endpoint=0, timeoutms=-1):
        """Send one already pickled message (bytes-like); blocks while no consumer can take it."""
        self.publish_parts([data], endpoint, timeoutms)

    def publish_parts(self, frames, endpoint=0, timeoutms=-1):
        """Send one message of several frames (the raw-parts wire, wire.py: header, then arrays)."""
        C, N = self._C, self._N
        n = len(frames)
        ptrs, lens = (C.c_void_p * n)(), (C.c_uint64 * n)()
        for i, f in enumerate(frames):
            ptrs[i], lens[i] = N.buffer_address(f)
        rc = N.lib.bt_push_send(self.handle, endpoint, ptrs, lens, n, int(timeoutms))
        if rc == N.BT_E_TIMEOUT:
            raise TimeoutError("no consumer took the message within %d ms" % timeoutms)
        N.check(rc)

    def start(self, pools, messages=-1, threads=1, shm=None):
        """Cycle through pre-pickled `pools` (one list of bytes-like messages per endpoint, equal
        lengths) on `threads` background threads, `messages` per endpoint (-1: until `stop`).

        shm: None for the reference wire, or ``(segments, payloads, ticket_offsets)`` for the
        shared-memory wire (shm.py): per endpoint a `shm.Segment`, per message the array that travels
        through it and the byte offset of the 4-byte ticket inside the pickled template.  `payloads`
        None publishes in place: slot k of the segment already holds the array template k describes
        (one template per slot), as when a renderer draws straight into shared memory."""
        import threading
        C, N = self._C, self._N
        n_ep = len(self.addresses)
        if len(pools) != n_ep or len({len(p) for p in pools}) != 1:
            raise ValueError("one pool per endpoint, all of the same length")
        pool_n = len(pools[0])
        threads = max(1, min(int(threads), n_ep))
        bounds = [n_ep * t // threads for t in range(threads + 1)]
        self._stop.value = 0
        for t in range(threads):
            first, count = bounds[t], bounds[t + 1] - bounds[t]
            n = count * pool_n
            ptrs, lens = (C.c_void_p * n)(), (C.c_uint64 * n)()
            flat = [m for pool in pools[first:first + count] for m in pool]
            for i, m in enumerate(flat):
                ptrs[i], lens[i] = N.buffer_address(m)
            job = N.PushJob()
            job.first_endpoint, job.n_endpoints, job.pool_n, job.messages = first, count, pool_n, int(messages)
            job.pool, job.pool_len = C.cast(ptrs, C.c_void_p), C.cast(lens, C.c_void_p)
            keep = [ptrs, lens, flat]
            if shm is not None:
                segs, payloads, offs = shm
                sh = (C.c_void_p * count)(*[s.handle for s in segs[first:first + count]])
                po = (C.c_uint32 * n)(*[o for row in offs[first:first + count] for o in row])
                job.seg, job.tmpl_ticket_off = C.cast(sh, C.c_void_p), C.cast(po, C.c_void_p)
                keep += [sh, po, segs]
                if payloads is not None:     # None: frames already stand in their slots (published in place)
                    pp, pl = (C.c_void_p * n)(), (C.c_uint64 * n)()
                    fp = [a for row in payloads[first:first + count] for a in row]
                    for i, a in enumerate(fp):
                        pp[i], pl[i] = N.buffer_address(a)
                    job.payload, job.payload_len = C.cast(pp, C.c_void_p), C.cast(pl, C.c_void_p)
                    keep += [pp, pl, fp]
            sent = C.c_int64(0)
            job.stop = C.cast(C.pointer(self._stop), C.c_void_p)
            job.sent = C.cast(C.pointer(sent), C.c_void_p)
            self._sent.append(sent)
            self._keep.append((job, keep))
            th = threading.Thread(target=self._run, args=(job,), daemon=True, name="bt-push-%d" % t)
            th.start()
            self._threads.append(th)

    def _run(self, job):
        rc = self._N.lib.bt_push_run(self.handle, self._C.byref(job))
        if rc != self._N.BT_OK:
            self.error = self._N.last_error()

    error = None

    @property
    def sent(self):
        """Messages sent so far by `start`'s threads."""
        return sum(s.value for s in self._sent)

    def join(self, timeout=None):
        """Wait for `start`'s threads (they end by themselves when `messages` is finite)."""
        for th in self._threads:
            th.join(timeout)
        return not any(th.is_alive() for th in self._threads)

    def stop(self):
        self._stop.value = 1
        for th in self._threads:
            th.join()
        self._threads = []
        if self.error:
            raise RuntimeError("native publisher: " + self.error)

    def close(self):
        if getattr(self, "_threads", None):
            self._stop.value = 1
            for th in self._threads:
                th.join()
            self._threads = []
        if self.handle:
            self._N.lib.bt_push_destroy(self.handle)
            self.handle = None
        self._keep = []

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
