"""DuplexChannel against a synthetic producer-side peer speaking the protocol of the reference's
tests/blender/duplex.blend.py (echo the first message, then say 'end'); assertions follow
tests/test_duplex.py:29-47."""
import socket
import threading

from blendtorch import btt


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _peer(addr, btid):
    ch = btt.DuplexChannel(addr, btid=btid, lingerms=5000, bind=True)     # producer side binds
    msg = ch.recv(timeoutms=10000)
    ch.send(echo=msg)
    ch.send(msg="end")
    ch.close()


def test_duplex():
    addrs = ["tcp://127.0.0.1:%d" % _free_port() for _ in range(2)]
    peers = [threading.Thread(target=_peer, args=(a, i)) for i, a in enumerate(addrs)]
    [t.start() for t in peers]
    duplex = [btt.DuplexChannel(a, lingerms=5000) for a in addrs]
    mids = [d.send(msg=f"hello {i}") for i, d in enumerate(duplex)]
    msgs = [[d.recv(timeoutms=5000), d.recv(timeoutms=5000)] for d in duplex]
    for i in range(2):
        assert msgs[i][0]["echo"]["msg"] == f"hello {i}"
        assert msgs[i][0]["echo"]["btid"] is None
        assert msgs[i][0]["echo"]["btmid"] == mids[i]
        assert msgs[i][0]["btid"] == i
        assert msgs[i][1]["msg"] == "end" and msgs[i][1]["btid"] == i
    assert duplex[0].recv(timeoutms=50) is None
    [t.join() for t in peers]
    [d.close() for d in duplex]
