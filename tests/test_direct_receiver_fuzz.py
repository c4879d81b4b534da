"""The direct ZMTP endpoint against a hostile peer: a raw-socket "publisher" that completes the
handshake and then sends frames whose pickles are mutated (flipped bytes, wrong lengths, truncated
frames, garbage after STOP, multipart flags, oversized size fields).  The endpoint may place a message,
hand it to the caller's decode, or drop the connection -- it must never crash, hang or leave the
receiver unusable for the well-formed messages that follow on a fresh connection."""
import pickle
import socket
import struct
import threading
import time

import numpy as np

from blendtorch.btt import native as N
from blendtorch.btt.rx import DirectReceiver
from blendtorch.btt.staging import Ring, Schema


def _greeting():
    g = bytearray(64)
    g[0], g[9], g[10], g[11] = 0xFF, 0x7F, 3, 0
    g[12:16] = b"NULL"
    return bytes(g)


def _ready():
    body = b"\x05READY" + b"\x0bSocket-Type" + struct.pack(">I", 4) + b"PUSH"
    return bytes([0x04, len(body)]) + body


def _frame(payload, flags=0):
    if len(payload) < 256:
        return bytes([flags, len(payload)]) + payload
    return bytes([flags | 0x02]) + struct.pack(">Q", len(payload)) + payload


class _Peer(threading.Thread):
    """Accepts connections one after another; each gets the handshake and the next script."""

    def __init__(self, scripts):
        super().__init__(daemon=True)
        self.srv = socket.socket()
        self.srv.bind(("127.0.0.1", 0))
        self.srv.listen(8)
        self.addr = "tcp://127.0.0.1:%d" % self.srv.getsockname()[1]
        self.scripts = scripts
        self.served = 0

    def run(self):
        for script in self.scripts:
            try:
                c, _ = self.srv.accept()
                c.settimeout(5)
                c.sendall(_greeting())
                got = b""
                while len(got) < 64:
                    got += c.recv(64 - len(got))
                c.sendall(_ready())
                c.recv(64)                     # the endpoint's READY
                for chunk in script:
                    c.sendall(chunk)
                self.served += 1
                time.sleep(0.05)
                c.close()
            except OSError:
                pass
        self.srv.close()


def _msg(i, shape=(40, 50, 4), proto=3):
    rng = np.random.default_rng(i)
    return {"btid": 0, "image": rng.integers(0, 256, shape, dtype=np.uint8), "frameid": i}


def _mutations(rng, good):
    b = bytearray(good)
    kind = int(rng.integers(0, 7))
    pos = int(rng.integers(0, len(b)))
    if kind == 0:
        b[pos] ^= 1 << int(rng.integers(0, 8))
        return [_frame(bytes(b))]
    if kind == 1:
        b[pos:pos + 4] = int(rng.integers(0, 2**32)).to_bytes(4, "little")
        return [_frame(bytes(b))]
    if kind == 2:   # frame announces more bytes than are sent, then the connection ends
        return [_frame(bytes(b))[:max(12, pos)]]
    if kind == 3:   # garbage after the pickle's STOP
        return [_frame(bytes(b) + bytes(rng.integers(0, 256, 40, dtype=np.uint8)))]
    if kind == 4:   # multipart flag: not part of the blendtorch wire
        return [_frame(bytes(b), flags=0x01), _frame(b"tail")]
    if kind == 5:   # absurd size field
        return [bytes([0x02]) + struct.pack(">Q", 1 << 50)]
    del b[pos:]     # pickle truncated inside a well-formed frame
    return [_frame(bytes(b))]


def _raw_parts_mutations(rng, k):
    """Raw-parts framing (wire.py) gone wrong."""
    from blendtorch.btt import wire
    frames = wire.pack(_msg(k))
    hdr, img = frames[0], frames[1].tobytes()
    kind = k % 5
    if kind == 0:   # the array frame is one byte short
        return [_frame(hdr, 1), _frame(img[:-1])]
    if kind == 1:   # the header promises a frame that never comes
        return [_frame(hdr, 0)]
    if kind == 2:   # more frames than the endpoint accepts
        return [_frame(hdr, 1)] + [_frame(b"x", 1)] * 80 + [_frame(b"x")]
    if kind == 3:   # header refers to frame 7 of 2
        bad = pickle.dumps({"btid": 0, "image": ("__btraw__", 7, "|u1", (40, 50, 4)), "frameid": k}, protocol=3)
        return [_frame(bad, 1), _frame(img)]
    # a well-formed one, followed by a frame header cut off in the middle
    return [_frame(hdr, 1), _frame(img), bytes([0x02, 0, 0])]


def test_hostile_peer_never_wedges_the_receiver():
    rng = np.random.default_rng(99)
    n_bad = 60
    scripts = []
    for k in range(n_bad):
        proto = 3 + k % 3
        good = pickle.dumps(_msg(k), protocol=proto)
        # a good message first, so that a schema exists; then the mutated one
        scripts.append([_frame(good)] + (_raw_parts_mutations(rng, k) if k % 4 == 3 else _mutations(rng, good)))
    final = [_frame(pickle.dumps(_msg(1000 + i), protocol=3)) for i in range(5)]
    scripts.append(final)
    peer = _Peer(scripts)
    peer.start()
    rx = DirectReceiver([peer.addr])
    deadline = time.time() + 60
    while not rx.recv(200):
        assert time.time() < deadline
    sch = Schema.from_message(rx.pending(), 4)
    ring = Ring(sch.slot_bytes, depth=2, device=None)
    seen, fallbacks = [], 0
    slot, filled = ring.acquire(), 0
    while time.time() < deadline and (peer.is_alive() or not seen or seen[-1] != 1004):
        rc, n = rx.fill(ring, slot, sch, filled, 4 - filled, 100)
        filled += n
        if rc in (N.BT_E_UNSUPPORTED, N.BT_E_SCHEMA):
            try:
                obj = pickle.loads(rx.pending())      # what the reference would do with it
                if isinstance(obj, dict) and isinstance(obj.get("frameid"), int):
                    seen.append(obj["frameid"])
            except Exception:
                pass
            fallbacks += 1
            rx.drop()
        if filled == 4 or rc == N.BT_E_TIMEOUT:
            col = sch.column("frameid")
            v = ring.view(slot)
            seen += v[col.slot_off:col.slot_off + filled * 8].view(np.int64).tolist()
            slot, filled = ring.acquire(), 0
    rx.close()
    ring.close()
    peer.join(5)
    assert peer.served == len(scripts)
    # the well-formed messages of the last connection all arrived, in order
    tail = [f for f in seen if f >= 1000]
    assert tail == [1000, 1001, 1002, 1003, 1004], (tail, fallbacks)
    # and the good first message of (nearly) every hostile connection did too
    assert len([f for f in seen if f < n_bad]) >= n_bad - 2


def test_peers_that_are_not_push_sockets_are_refused_and_a_socket_connected_to_itself_is_dropped():
    """A PULL endpoint takes messages from PUSH peers only.  This matters beyond protocol hygiene: a consumer
    that keeps retrying a local port nobody listens on yet can get connected to *itself* (the kernel picks
    the destination port as the source port); it then reads back its own greeting and its own READY -- a
    PULL socket -- and would sit on the port its publisher is about to bind.  The endpoint refuses a peer
    whose Socket-Type is not PUSH (and compares getsockname with getpeername before it even starts)."""
    good = pickle.dumps(_msg(1), protocol=3)

    class _PullPeer(_Peer):
        def run(self):
            c, _ = self.srv.accept()
            c.settimeout(5)
            try:
                c.sendall(_greeting())
                got = b""
                while len(got) < 64:
                    got += c.recv(64 - len(got))
                body = b"\x05READY" + b"\x0bSocket-Type" + struct.pack(">I", 4) + b"PULL"    # what a mirror would say
                c.sendall(bytes([0x04, len(body)]) + body)
                c.recv(64)
                c.sendall(_frame(good))
                self.closed_by_endpoint = c.recv(1) == b""        # the endpoint hangs up instead of reading on
            except OSError:
                self.closed_by_endpoint = True
            c.close()
            self.srv.close()

    peer = _PullPeer([])
    peer.start()
    rx = DirectReceiver([peer.addr])
    assert not rx.recv(1500)                                      # nothing is ever delivered from that peer
    peer.join(5)
    assert peer.closed_by_endpoint
    rx.close()
    # the mirror itself: a socket connected to its own port (bind, then connect to the bound address)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    s.connect(s.getsockname())
    assert s.getsockname() == s.getpeername()                     # what finish_connect tests for
    s.sendall(_greeting())
    assert s.recv(64) == _greeting()                              # it reads back what it wrote: a valid greeting
    s.close()
