"""Control channel to one producer (next-row N4): the consumer-side ``DuplexChannel`` of the reference
(pkg_pytorch/blendtorch/btt/duplex.py:8-66) for the training loops that steer their producers
(examples/densityopt/densityopt.py:95-107).  Tiny dict messages over a ZMQ PAIR socket; no kernel
work.  ``bind=True`` gives the producer-side end (btb/duplex.py binds) for synthetic peers and tests.
"""
import os
import pickle

import zmq

from .constants import DEFAULT_TIMEOUTMS


class DuplexChannel:
    """Bidirectional messages with a single remote process.

    address: ZMQ address of the producer's PAIR socket.
    btid: identifier attached to every outgoing message (None on the consumer side, like the reference).
    lingerms / timeoutms: socket linger and send/receive time-outs.
    """

    def __init__(self, address, btid=None, lingerms=0, timeoutms=DEFAULT_TIMEOUTMS, bind=False):
        self.ctx = zmq.Context()
        self.sock = self.ctx.socket(zmq.PAIR)
        for opt, val in ((zmq.LINGER, lingerms), (zmq.RCVHWM, 10), (zmq.SNDHWM, 10),
                         (zmq.SNDTIMEO, timeoutms), (zmq.RCVTIMEO, timeoutms)):
            self.sock.setsockopt(opt, val)
        (self.sock.bind if bind else self.sock.connect)(address)
        self.btid = btid

    def recv(self, timeoutms=None):
        """Next message dict, or None when nothing arrives within `timeoutms` (None = wait)."""
        if self.sock.poll(timeoutms, zmq.POLLIN):
            return pickle.loads(self.sock.recv())
        return None

    def send(self, **kwargs):
        """Send `kwargs` with `btid` and a fresh random message id `btmid` attached; returns the id."""
        mid = int.from_bytes(os.urandom(4), "little")
        self.sock.send(pickle.dumps({"btid": self.btid, "btmid": mid, **kwargs}, protocol=pickle.DEFAULT_PROTOCOL))
        return mid

    def close(self):
        self.sock.close()
        self.ctx.term()
