#!/usr/bin/env python3
"""examples/datagen/generate.py of the reference, line for line, on the accelerated path.

What changes against the reference script (read the two side by side):

    reference (generate.py:10-18, 52-85)                     here
    ----------------------------------------------------    -----------------------------------------------------
    def gamma_correct(x): numpy, per item, on the CPU        GAMMA = btt.Preprocess(gamma=2.2, dtype=uint8, 'NHWC')
    def item_transform(item): item['image'] = gamma_...      (nothing: the kernel does it for the whole batch)
    btt.BlenderLauncher(scene=..., num_instances=4, ...)     launch_publishers(4): synthetic publishers, same wire
    ds = btt.RemoteIterableDataset(addresses,                ds = btt.RemoteIterableDataset(addresses)
             item_transform=item_transform)
    ds.stream_length(64); ds.enable_recording(prefix)        the same two calls
    ds = btt.FileDataset(prefix, item_transform=...)         ds = btt.FileDataset(prefix)
    dl = DataLoader(ds, batch_size=BATCH,                    dl = btt.DeviceLoader(ds, batch_size=BATCH, shuffle=shuffle,
             num_workers=WORKER_INSTANCES, shuffle=shuffle)          num_receivers=WORKER_INSTANCES, preprocess=GAMMA)
    iterate(dl): item['image'] uint8 (B,H,W,4) on the CPU    iterate(dl): the same tensor, same values, on the GPU

No Blender is needed (or available here): the publishers are `blendtorch.btt.synthetic` processes that
speak the wire format of `btb.DataPublisher`.  `scene` only names the recording, as in the reference.

    python examples/generate_device.py cube --record        # stream 64 items, record them to ./tmp/record_cube_*.btr
    python examples/generate_device.py cube --replay        # replay the recording, shuffled
"""
import argparse
import os
import socket
import subprocess
import sys
from contextlib import ExitStack
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))

import torch  # noqa: E402

import blendtorch.btt as btt  # noqa: E402

# generate.py:10-14 `gamma_correct` + :16-18 `item_transform`, as a description instead of numpy code:
# sRGB gamma on the colour channels, truncated to uint8, alpha untouched, image stays (H, W, 4) uint8.
GAMMA = btt.Preprocess(key="image", gamma=2.2, dtype=torch.uint8, layout="NHWC")


def iterate(dl):
    """generate.py:20-37 without the matplotlib figure: the same fields of the same batch."""
    for step, item in enumerate(dl):
        img, xy, btid, fid = item["image"], item["xy"], item["btid"], item["frameid"]
        print(f"Received batch from Blender processes {btid.cpu().numpy()}, frames {fid.cpu().numpy()}")
        H, W = img.shape[1], img.shape[2]
        assert img.dtype == torch.uint8 and img.shape[3] == 4 and img.is_cuda and xy.shape[1:] == (8, 2)
        print(f"  image {tuple(img.shape)} on {img.device}, {H}x{W}, mean {img.float().mean().item():.2f}")


BATCH = 4
BLENDER_INSTANCES = 4
WORKER_INSTANCES = 4


class launch_publishers:
    """Stands where `btt.BlenderLauncher(...)` stands in generate.py:56-63: a context manager whose
    `launch_info.addresses['DATA']` lists one address per producer."""

    def __init__(self, num_instances):
        self.num_instances = num_instances

    def __enter__(self):
        env = dict(os.environ, PYTHONPATH=str(ROOT / "pytorch-blender_b200"))
        self.procs, data = [], []
        for i in range(self.num_instances):
            s = socket.socket()
            s.bind(("127.0.0.1", 0))
            addr = "tcp://127.0.0.1:%d" % s.getsockname()[1]
            s.close()
            self.procs.append(subprocess.Popen([sys.executable, "-m", "blendtorch.btt.synthetic", "--bind", addr,
                                                "--btid", str(i), "--config", "1", "--height", "480", "--width", "640"], env=env))
            data.append(addr)
        self.launch_info = type("LaunchInfo", (), {"addresses": {"DATA": data}})()
        return self

    def __exit__(self, *exc):
        for p in self.procs:
            p.terminate()


def main():
    import logging
    logging.basicConfig(level=logging.INFO)

    parser = argparse.ArgumentParser()
    parser.add_argument("scene", help="name of the recording (the reference's Blender scene name)")
    parser.add_argument("--replay", action="store_true", help="Replay from disc instead of launching publishers")
    parser.add_argument("--record", action="store_true", help="Record raw data")
    args = parser.parse_args()

    with ExitStack() as es:
        if not args.replay:
            # Launch the producers.  Upon exit of this script all of them will be closed.
            bl = es.enter_context(launch_publishers(num_instances=BLENDER_INSTANCES))

            # Setup a streaming dataset
            ds = btt.RemoteIterableDataset(bl.launch_info.addresses["DATA"])
            # Iterable datasets do not support shuffle
            shuffle = False

            # Limit the total number of streamed elements
            ds.stream_length(64)

            # Setup raw recording if desired
            if args.record:
                os.makedirs("./tmp", exist_ok=True)
                ds.enable_recording(f"./tmp/record_{args.scene}")
        else:
            # Otherwise we replay from file.
            ds = btt.FileDataset(f"./tmp/record_{args.scene}")
            shuffle = True

        # Setup the loader and iterate
        dl = btt.DeviceLoader(ds, batch_size=BATCH, num_receivers=WORKER_INSTANCES, shuffle=shuffle, device="cuda",
                              preprocess=GAMMA)
        iterate(dl)


if __name__ == "__main__":
    main()
