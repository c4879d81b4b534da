"""Oracle: live-stream receive loop and the producer's publish call
(TEST INFRASTRUCTURE).

Restates
  pkg_pytorch/blendtorch/btt/dataset.py:68-111  (RemoteIterableDataset._stream)
  pkg_blender/blendtorch/btb/publisher.py:21-43 (DataPublisher, minus `import bpy`)
"""

import pickle

import zmq

from . import btr

DEFAULT_TIMEOUTMS = 10000  # btt/constants.py:4


class Publisher:
    """publisher.py:21-43: PUSH socket bound to `bind_address`, SNDHWM, LINGER,
    IMMEDIATE; `publish(**kw)` sends pickle({'btid': btid, **kw})."""

    def __init__(self, bind_address, btid=None, send_hwm=10, lingerms=0, protocol=None):
        self.ctx = zmq.Context()
        self.sock = self.ctx.socket(zmq.PUSH)
        self.sock.setsockopt(zmq.SNDHWM, send_hwm)
        self.sock.setsockopt(zmq.LINGER, lingerms)
        self.sock.setsockopt(zmq.IMMEDIATE, 1)
        self.sock.bind(bind_address)
        self.btid = btid
        # send_pyobj pickles at pyzmq's DEFAULT_PROTOCOL (4 here, 3 on Blender's
        # Python 3.7); `protocol` lets tests pick the Blender-faithful one.
        self.protocol = protocol

    def publish(self, **kwargs):
        msg = {"btid": self.btid, **kwargs}
        if self.protocol is None:
            self.sock.send_pyobj(msg)
        else:
            self.sock.send(pickle.dumps(msg, protocol=self.protocol))

    def close(self):
        self.sock.close()
        self.ctx.term()


def stream_items(addresses, queue_size=10, timeoutms=DEFAULT_TIMEOUTMS, max_items=100000,
                 item_transform=None, record_path_prefix=None, worker_id=0, num_workers=1):
    """dataset.py:68-111 as a generator for one consumer ("worker")."""
    transform = item_transform or (lambda x: x)
    ctx = zmq.Context()
    sock = ctx.socket(zmq.PULL)
    try:
        sock.setsockopt(zmq.RCVHWM, queue_size)  # dataset.py:74
        poller = zmq.Poller()
        poller.register(sock, zmq.POLLIN)
        for a in addresses:  # every consumer connects to ALL publishers, dataset.py:77-78
            sock.connect(a)
        rec = None
        if record_path_prefix is not None:  # dataset.py:87-95
            rec = btr.Recorder(btr.filename(record_path_prefix, worker_id), max_items)
            rec.__enter__()
        try:
            for _ in range(max_items // num_workers):  # dataset.py:97
                ready = dict(poller.poll(timeoutms))
                assert sock in ready, "No response within timeout interval."  # dataset.py:99
                raw = sock.recv()
                if rec is not None:
                    rec.save(raw, is_pickled=True)  # dataset.py:100-103
                yield transform(pickle.loads(raw))
        finally:
            if rec is not None:
                rec.__exit__(None, None, None)
    finally:
        sock.close()
        ctx.term()


def torch_stream_dataset(addresses, **kw):
    """The same loop wrapped as a torch IterableDataset so that DataLoader
    workers drive it the way they drive the reference class
    (dataset.py:80-85: worker id / count from get_worker_info)."""
    import torch.utils.data as tud

    class _DS(tud.IterableDataset):
        def __iter__(self):
            wi = tud.get_worker_info()
            wid, nw = (wi.id, wi.num_workers) if wi is not None else (0, 1)
            return stream_items(addresses, worker_id=wid, num_workers=nw, **kw)

    return _DS()
