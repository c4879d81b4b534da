"""Python face of the C receive loop (``bt_rx_*``): drains a pyzmq PULL socket into a staging slot,
a whole batch per call, without the GIL.  Replaces the per-message Python of
pkg_pytorch/blendtorch/btt/dataset.py:97-106 on the accelerated path.

libzmq is not linked into libbt_b200.so; the entry points are taken from the libzmq that pyzmq
itself has loaded into this process, and the socket handle is pyzmq's (``socket.underlying``).
"""
import ctypes as C

import zmq

from . import native as N


class ZmqApi(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("msg_init", "msg_recv", "msg_data", "msg_size", "msg_close", "poll", "zmq_errno")]


_api = None


def _find_libzmq():
    with open("/proc/self/maps") as f:
        for line in f:
            path = line.rsplit(" ", 1)[-1].strip()
            if "libzmq" in path and path.startswith("/"):
                return path
    raise ImportError("the libzmq loaded by pyzmq was not found in this process")


def zmq_api():
    global _api
    if _api is None:
        lib = C.CDLL(_find_libzmq())
        api = ZmqApi()
        for field, sym in (("msg_init", "zmq_msg_init"), ("msg_recv", "zmq_msg_recv"), ("msg_data", "zmq_msg_data"),
                           ("msg_size", "zmq_msg_size"), ("msg_close", "zmq_msg_close"), ("poll", "zmq_poll"),
                           ("zmq_errno", "zmq_errno")):
            setattr(api, field, C.cast(getattr(lib, sym), C.c_void_p).value)
        _api = (api, lib)
    return _api[0]


class Receiver:
    """Owns nothing but a small C state; the socket stays pyzmq's and must be used from this thread only."""

    def __init__(self, sock):
        if sock.type != zmq.PULL:
            raise ValueError("the ingest receiver takes a PULL socket")
        self._sock = sock  # keep alive
        self.handle = C.c_void_p()
        N.check(N.lib.bt_rx_create(C.byref(zmq_api()), C.c_void_p(sock.underlying), C.byref(self.handle)))

    def recv(self, timeout_ms):
        """True when a message is held, False on timeout."""
        rc = N.lib.bt_rx_recv(self.handle, int(timeout_ms))
        if rc == N.BT_E_TIMEOUT:
            return False
        N.check(rc)
        return True

    def pending(self):
        """Zero-copy view of the held message."""
        p, n = C.c_void_p(), C.c_uint64()
        N.check(N.lib.bt_rx_pending(self.handle, C.byref(p), C.byref(n)))
        return memoryview((C.c_uint8 * n.value).from_address(p.value)).cast("B")

    def pending_parts(self):
        """Frame sizes of the held message as far as this receiver has them (always one frame: see `more`)."""
        return [len(self.pending())]

    def more(self):
        """True when the held frame is the first of a multipart message (raw-parts wire, wire.py) whose
        other frames are still queued on the socket."""
        return bool(self._sock.getsockopt(zmq.RCVMORE))

    def recv_rest(self):
        """The remaining frames of the message whose first frame is held."""
        return self._sock.recv_multipart(copy=False)

    def drop(self):
        N.check(N.lib.bt_rx_drop(self.handle))

    def fill(self, ring, slot, schema, first_item, max_items, slice_ms=100):
        """(status, n_placed): BT_OK batch part complete, BT_E_TIMEOUT nothing for slice_ms,
        BT_E_UNSUPPORTED / BT_E_SCHEMA the held message needs the pickle.loads path."""
        n = C.c_int(0)
        rc = N.lib.bt_rx_fill(self.handle, ring.handle, slot, schema.handle, first_item, max_items, slice_ms, C.byref(n))
        if rc not in (N.BT_OK, N.BT_E_TIMEOUT, N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
            N.check(rc)
        return rc, n.value

    def release_queued(self):
        """Give back the shared-memory slots referenced by messages still queued on the socket."""
        n = C.c_int(0)
        N.check(N.lib.bt_rx_release_queued(self.handle, C.byref(n)))
        return n.value

    def close(self):
        if self.handle:
            N.lib.bt_rx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def direct_capable(addresses):
    """True when every address is one the direct receiver can speak to (tcp:// or ipc://)."""
    return all(isinstance(a, str) and (a.startswith("tcp://") or a.startswith("ipc://")) for a in addresses)


class DirectReceiver:
    """Same interface as `Receiver`, without libzmq: a ZMTP 3 PULL endpoint in C (``bt_pull_*``) whose
    array payloads go from the kernel socket straight into the staging slot."""

    def __init__(self, addresses):
        self.handle = C.c_void_p()
        N.check(N.lib.bt_pull_create(C.byref(self.handle)))
        try:
            for a in addresses:
                N.check(N.lib.bt_pull_connect(self.handle, a.encode()))
        except Exception:
            self.close()
            raise

    def recv(self, timeout_ms):
        rc = N.lib.bt_pull_recv(self.handle, int(timeout_ms))
        if rc == N.BT_E_TIMEOUT:
            return False
        N.check(rc)
        return True

    def pending(self):
        p, n = C.c_void_p(), C.c_uint64()
        N.check(N.lib.bt_pull_pending(self.handle, C.byref(p), C.byref(n)))
        return memoryview((C.c_uint8 * n.value).from_address(p.value)).cast("B")

    def pending_parts(self):
        """Frame sizes of the held message: one for the reference wire, several for the raw-parts wire
        (`pending` then spans all frames back to back)."""
        sizes = (C.c_uint64 * 64)()
        n = N.lib.bt_pull_pending_parts(self.handle, sizes, 64)
        if n < 0:
            N.check(n)
        return [int(sizes[i]) for i in range(min(n, 64))]

    def more(self):
        return False     # this receiver always holds whole messages

    def recv_rest(self):
        return []

    def drop(self):
        N.check(N.lib.bt_pull_drop(self.handle))

    def fill(self, ring, slot, schema, first_item, max_items, slice_ms=100):
        n = C.c_int(0)
        rc = N.lib.bt_pull_fill(self.handle, ring.handle, slot, schema.handle, first_item, max_items, slice_ms, C.byref(n))
        if rc not in (N.BT_OK, N.BT_E_TIMEOUT, N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
            N.check(rc)
        return rc, n.value

    def release_queued(self):
        return 0  # nothing is queued in user space: unread messages stay with their publishers

    def close(self):
        if self.handle:
            N.lib.bt_pull_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
