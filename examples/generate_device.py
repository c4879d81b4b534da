#!/usr/bin/env python3
"""The reference's examples/datagen/generate.py on the accelerated path.

Reference (CPU):   ds = btt.RemoteIterableDataset(addresses, item_transform=gamma_correct_transform)
                   dl = DataLoader(ds, batch_size=4, num_workers=4)        # numpy gamma per item, collate, then .to(dev)
Here (GPU):        dl = btt.DeviceLoader(ds, batch_size=4, preprocess=btt.Preprocess(gamma=2.2, ...))

No Blender is needed to try it: `--synthetic N` starts N synthetic publishers that speak the same wire
format; `--replay PREFIX` replays `.btr` recordings; `--record PREFIX` tees the raw stream to disk.

    PYTHONPATH=pytorch-blender_b200 python examples/generate_device.py --synthetic 4 --record /tmp/cube
    PYTHONPATH=pytorch-blender_b200 python examples/generate_device.py --replay /tmp/cube
"""
import argparse
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))

import torch  # noqa: E402

from blendtorch import btt  # noqa: E402

BATCH = 4


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--addresses", nargs="*", default=[], help="tcp://... addresses of running btb.DataPublisher sockets")
    ap.add_argument("--synthetic", type=int, default=0, help="start this many synthetic publishers instead")
    ap.add_argument("--replay", default=None, help="replay recordings with this prefix instead of streaming")
    ap.add_argument("--record", default=None, help="record the raw stream to this prefix")
    ap.add_argument("--items", type=int, default=64)
    args = ap.parse_args()

    # what generate.py does per item on the CPU (gamma_correct) plus the usual normalise + CHW, on the GPU
    spec = btt.Preprocess(key="image", gamma=2.2, mean=127.5, std=127.5, layout="NCHW", dtype=torch.float32)
    procs = []
    try:
        if args.replay:
            ds = btt.FileDataset(args.replay)
            dl = btt.DeviceLoader(ds, batch_size=BATCH, shuffle=True, device="cuda", preprocess=spec)
        else:
            addresses = list(args.addresses)
            for i in range(args.synthetic):
                addr = "tcp://127.0.0.1:%d" % (11000 + i)       # the reference launcher's port range
                procs.append(subprocess.Popen([sys.executable, "-m", "blendtorch.btt.synthetic", "--bind", addr,
                                               "--btid", str(i), "--config", "1"],
                                              env={"PYTHONPATH": str(ROOT / "pytorch-blender_b200"), "PATH": "/usr/bin:/bin"}))
                addresses.append(addr)
            if not addresses:
                ap.error("give --addresses, --synthetic N or --replay PREFIX")
            ds = btt.RemoteIterableDataset(addresses)
            ds.stream_length(args.items)
            if args.record:
                ds.enable_recording(args.record)
            dl = btt.DeviceLoader(ds, batch_size=BATCH, device="cuda", preprocess=spec, num_receivers=2)
        t0 = time.time()
        n = 0
        for step, item in enumerate(dl):
            img, xy, btid, fid = item["image"], item["xy"], item["btid"], item["frameid"]
            n += len(img)
            print(f"batch {step}: image {tuple(img.shape)} {img.dtype} on {img.device}, xy {tuple(xy.shape)}, "
                  f"from Blender processes {btid.tolist()}, frames {fid.tolist()}")
        print(f"{n} images in {time.time() - t0:.2f} s")
    finally:
        for p in procs:
            p.terminate()


if __name__ == "__main__":
    main()
