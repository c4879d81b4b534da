"""Raw-parts wire: an opt-in framing for producers that are *not* on the consumer's host (for those,
see shm.py) and do not want to pay for pickling their frames.

The reference sends one frame per message, ``pickle({'btid': .., 'image': ndarray, ..})``
(pkg_blender/blendtorch/btb/publisher.py:43): the producer copies every image into the pickle and
pyzmq copies the pickle into a ZMQ message.  With ``DataPublisher(raw_parts=True)`` a message becomes a
ZMQ multipart message instead:

    frame 0   pickle of the same dict, with every large array replaced by the tuple
              ("__btraw__", k, dtype.str, shape)
    frame k   the C-contiguous bytes of that array, sent zero-copy from the array's own memory

The consumer side of this package understands both framings on the same socket: the direct receiver
(csrc/zmtp.cpp) receives frame k straight into its column of the pinned staging slot, the Python
paths rebuild the arrays with `decode`.  The unmodified reference does not (both ends opt in).
"""
import pickle

import numpy as np

RAW_MARKER = "__btraw__"


def is_raw(v):
    return isinstance(v, tuple) and len(v) == 4 and v[0] == RAW_MARKER


def pack(data, threshold=4096, protocol=pickle.DEFAULT_PROTOCOL):
    """dict -> list of frames (bytes header, then the arrays themselves)."""
    frames = [None]
    header = {}
    for k, v in data.items():
        if isinstance(v, np.ndarray) and v.nbytes >= threshold and not v.dtype.hasobject and v.dtype.fields is None:
            a = np.ascontiguousarray(v)          # a flipped view (offscreen.py:96) is materialised here
            header[k] = (RAW_MARKER, len(frames), a.dtype.str, tuple(a.shape))
            frames.append(a)
        else:
            header[k] = v
    frames[0] = pickle.dumps(header, protocol=protocol)
    return frames


def decode(frames):
    """list of frames (bytes-like) -> the dict the producer published; arrays are fresh copies, as after
    ``pickle.loads`` of the reference wire."""
    obj = pickle.loads(frames[0])
    if isinstance(obj, dict):
        for k, v in obj.items():
            if is_raw(v):
                _, part, dtype, shape = v
                if not 1 <= part < len(frames):
                    raise ValueError("message refers to frame %d but has %d frames" % (part, len(frames)))
                buf = frames[part]
                buf = buf.buffer if hasattr(buf, "buffer") else buf
                obj[k] = np.frombuffer(buf, dtype=np.dtype(dtype)).reshape(shape).copy()
    return obj


def split(buf, sizes):
    """One buffer holding all frames back to back (``DirectReceiver.pending``) -> list of memoryviews."""
    mv = memoryview(buf)
    out, at = [], 0
    for n in sizes:
        out.append(mv[at:at + n])
        at += n
    return out


def loads(buf, sizes=None):
    """Decode a held message of either framing."""
    if sizes is not None and len(sizes) > 1:
        return decode(split(buf, sizes))
    return pickle.loads(buf)
