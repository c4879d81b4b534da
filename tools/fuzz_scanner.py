#!/usr/bin/env python3
"""Mutation fuzzer for the message scanner (bt_scan_message / bt_schema_ingest): valid blendtorch
messages of every pickle protocol with random byte flips, truncations, splices and length-field edits.
The scanner must answer with a status code every time -- run it under AddressSanitizer to make memory
errors visible:

    LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 python tools/fuzz_scanner.py 200000
"""
import ctypes as C
import pickle
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import numpy as np  # noqa: E402

from blendtorch.btt import native as N  # noqa: E402
from blendtorch.btt.staging import Schema  # noqa: E402


def seeds(rng):
    out = []
    for proto in (2, 3, 4, 5):
        for shape in ((4, 6, 4), (70, 80, 4), (3, 5)):
            m = {"btid": 3, "image": rng.integers(0, 256, shape, dtype=np.uint8), "frameid": 17,
                 "xy": rng.random((8, 2)), "score": 0.5, "ok": True, "name": "cube", "none": None,
                 "depth": rng.random(shape[:2], dtype=np.float32), "big": 1 << 70,
                 "ref": ("__btshm__", "seg", 1, "|u1", (4, 6, 4))}
            out.append(pickle.dumps(m, protocol=proto))
    return out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    rng = np.random.default_rng(5)
    base = seeds(rng)
    fields = (N.Field * 64)()
    schema = Schema.from_message(pickle.dumps({"btid": 1, "image": np.zeros((4, 6, 4), np.uint8), "frameid": 2}, protocol=3), 4)
    slot = np.zeros(schema.slot_bytes, dtype=np.uint8)
    counts = {}
    for it in range(n):
        b = bytearray(base[it % len(base)])
        for _ in range(int(rng.integers(1, 6))):
            kind = int(rng.integers(0, 5))
            pos = int(rng.integers(0, len(b))) if len(b) else 0
            if kind == 0 and b:
                b[pos] ^= 1 << int(rng.integers(0, 8))
            elif kind == 1 and b:
                b[pos] = int(rng.integers(0, 256))
            elif kind == 2:
                del b[pos:]
            elif kind == 3 and b:
                other = base[int(rng.integers(0, len(base)))]
                cut = int(rng.integers(0, len(other)))
                b[pos:pos] = other[cut:cut + int(rng.integers(1, 64))]
            elif kind == 4 and len(b) > 8:
                b[pos:pos + 4] = int(rng.integers(0, 2**32)).to_bytes(4, "little")   # lengths, counts
        buf = (C.c_uint8 * max(len(b), 1)).from_buffer_copy(bytes(b) or b"\0")
        rc = N.lib.bt_scan_message(buf, len(b), fields, 64)
        counts[min(rc, 1)] = counts.get(min(rc, 1), 0) + 1
        rc = N.lib.bt_schema_ingest(schema.handle, slot.ctypes.data_as(C.c_void_p), it % 4, buf, len(b))
        counts[("ingest", rc)] = counts.get(("ingest", rc), 0) + 1
    print({str(k): v for k, v in sorted(counts.items(), key=str)})


if __name__ == "__main__":
    main()
