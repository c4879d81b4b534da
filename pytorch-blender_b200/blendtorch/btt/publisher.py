"""bpy-free data publisher: the producer end of the ingest path.

Same class name, constructor and ``publish(**kwargs)`` contract as the reference's Blender-side
``btb.DataPublisher`` (pkg_blender/blendtorch/btb/publisher.py:21-43): a PUSH socket that binds,
a bounded send queue that blocks the caller when consumers are slow, and messages that are
``pickle({'btid': btid, **kwargs})``.  It exists on the btt side so that synthetic producers
(tests, benchmarks, simulators that are not Blender) speak the exact wire format.
"""
import pickle

import zmq


class DataPublisher:
    """
    bind_address: ZMQ address to bind to.
    btid: integer producer id added to every message (None is sent as-is, and like in the
        reference cannot be collated -- pass an int).
    send_hwm: send queue size before `publish` blocks.
    lingerms: how long unsent messages survive `close`.
    protocol: pickle protocol; None = pyzmq's default, 3 = what Blender's Python 3.7 emits.
    """

    def __init__(self, bind_address, btid=None, send_hwm=10, lingerms=0, protocol=None):
        self.ctx = zmq.Context()
        self.sock = self.ctx.socket(zmq.PUSH)
        self.sock.setsockopt(zmq.SNDHWM, send_hwm)
        self.sock.setsockopt(zmq.LINGER, lingerms)
        self.sock.setsockopt(zmq.IMMEDIATE, 1)
        self.sock.bind(bind_address)
        self.btid = btid
        self.protocol = pickle.DEFAULT_PROTOCOL if protocol is None else protocol  # == pyzmq default

    def publish(self, **kwargs):
        """Send one message; `btid` is added implicitly."""
        self.publish_raw(pickle.dumps({"btid": self.btid, **kwargs}, protocol=self.protocol))

    def publish_raw(self, data, copy=True):
        """Send an already pickled message (replaying recordings, pre-pickled synthetic frames)."""
        self.sock.send(data, copy=copy)

    def close(self):
        self.sock.close()
        self.ctx.term()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
