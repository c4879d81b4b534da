"""Probe: can a memory-mapped recording be page-locked (cudaHostRegister) and copied to the device
without a staging copy?  Prints one JSON line per (filesystem, mapping, register flags) case with
the registration time, the H2D rate from unaligned offsets inside the mapping and a content check.

    python tools/hostreg_probe.py [--mb 512]
"""
import argparse
import ctypes as C
import json
import mmap
import os
import sys
import time

import numpy as np
import torch


def cudart():
    with open("/proc/self/maps") as f:
        for line in f:
            path = line.rsplit(" ", 1)[-1].strip()
            if "libcudart" in path and path.startswith("/"):
                return C.CDLL(path)
    raise SystemExit("libcudart is not loaded")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--dirs", default="/tmp,/dev/shm," + os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gpurun_out"))
    a = ap.parse_args()
    torch.cuda.init()
    torch.zeros(1, device="cuda")
    rt = cudart()
    rt.cudaHostRegister.argtypes = [C.c_void_p, C.c_size_t, C.c_uint]
    rt.cudaHostUnregister.argtypes = [C.c_void_p]
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    rt.cudaGetLastError.restype = C.c_int
    nbytes = a.mb << 20
    rng = np.random.default_rng(1)
    payload = rng.integers(0, 256, nbytes, dtype=np.uint8)
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    chunk = 4 << 20
    off0 = 185  # where the frame of a protocol-3 message starts: not aligned to anything
    for d in a.dirs.split(","):
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "hostreg_probe.bin")
        try:
            with open(path, "wb") as f:
                f.write(payload.tobytes())
        except OSError as e:
            print(json.dumps({"dir": d, "error": str(e)}))
            continue
        fstype = "?"
        with open("/proc/mounts") as f:
            best = ""
            for line in f:
                parts = line.split()
                if os.path.realpath(d).startswith(parts[1]) and len(parts[1]) > len(best):
                    best, fstype = parts[1], parts[2]
        for mapname, mflags, prot in (("shared_ro", mmap.MAP_SHARED, mmap.PROT_READ),
                                      ("private_ro", mmap.MAP_PRIVATE, mmap.PROT_READ),
                                      ("private_rw", mmap.MAP_PRIVATE, mmap.PROT_READ | mmap.PROT_WRITE)):
            for regname, rflags in (("default", 0), ("readonly", 8)):
                fd = os.open(path, os.O_RDONLY)
                t0 = time.perf_counter()
                mm = mmap.mmap(fd, 0, flags=mflags | mmap.MAP_POPULATE, prot=prot)
                t_map = time.perf_counter() - t0
                os.close(fd)
                arr = np.frombuffer(mm, dtype=np.uint8)
                addr = arr.ctypes.data
                t0 = time.perf_counter()
                rc = rt.cudaHostRegister(addr, nbytes, rflags)
                t_reg = time.perf_counter() - t0
                rec = {"dir": d, "fs": fstype, "map": mapname, "register": regname, "mb": a.mb, "rc": rc,
                       "map_ms": round(t_map * 1e3, 2), "register_ms": round(t_reg * 1e3, 2)}
                if rc == 0:
                    st = torch.cuda.Stream()
                    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    for rep in range(3):
                        ev0.record(st)
                        pos = off0
                        while pos + chunk <= nbytes:
                            rt.cudaMemcpyAsync(dev.data_ptr() + pos, addr + pos, chunk, 1, st.cuda_stream)
                            pos += chunk
                        ev1.record(st)
                        st.synchronize()
                        ms = ev0.elapsed_time(ev1)
                    rec["h2d_GBps"] = round((pos - off0) / ms / 1e6, 2)
                    rec["content_ok"] = bool(torch.equal(dev[off0:pos].cpu(), torch.from_numpy(payload[off0:pos])))
                    t0 = time.perf_counter()
                    rec["unregister_rc"] = rt.cudaHostUnregister(addr)
                    rec["unregister_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
                else:
                    rt.cudaGetLastError()
                print(json.dumps(rec), flush=True)
                del arr
                mm.close()
        os.unlink(path)
    # reference point: the same copy from cudaHostAlloc'ed memory
    pin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    pin.copy_(torch.from_numpy(payload))
    st = torch.cuda.Stream()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(3):
        ev0.record(st)
        pos = off0
        while pos + chunk <= nbytes:
            rt.cudaMemcpyAsync(dev.data_ptr() + pos, pin.data_ptr() + pos, chunk, 1, st.cuda_stream)
            pos += chunk
        ev1.record(st)
        st.synchronize()
        ms = ev0.elapsed_time(ev1)
    print(json.dumps({"pinned_reference_GBps": round((pos - off0) / ms / 1e6, 2)}))


if __name__ == "__main__":
    sys.exit(main())
