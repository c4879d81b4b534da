#!/usr/bin/env python3
"""A producer that is steered over the control channel, without Blender.

Plays the part of examples/densityopt/supershape.blend.py in the reference: it listens on a
`DuplexChannel` for messages ``{'shape_params': (n, P), 'shape_ids': (n,)}``
(pkg_pytorch/../densityopt.py:95-107 sends them), produces one frame per parameter sample and
publishes ``image`` + ``shape_id`` (supershape.blend.py:24-41) -- here also the world corners of what
was drawn, so that the consumer can project them (K2).  The scene is one cube whose centre is
``(params[0], params[1], 0)`` and whose half-extent is ``params[2]``, drawn by the CUDA rasteriser K3
(`btt.render_cubes`); frames leave through `btt.DataPublisher` in the reference's wire format.

    python examples/steered_producer.py --data tcp://127.0.0.1:12000 --ctrl tcp://127.0.0.1:12001 --btid 0
"""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))

BASE = np.array([[x, y, z] for x in (1.0, -1.0) for y in (1.0, -1.0) for z in (1.0, -1.0)])   # default-cube corner order


def cube_corners(params):
    """World corners (8, 3) of the cube that parameter sample `params` = (cx, cy, half_extent) describes."""
    cx, cy, half = (float(v) for v in params[:3])
    return BASE * half + np.array([cx, cy, 0.0])


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--data", required=True)
    ap.add_argument("--ctrl", required=True)
    ap.add_argument("--btid", type=int, default=0)
    ap.add_argument("--height", type=int, default=240)
    ap.add_argument("--width", type=int, default=320)
    ap.add_argument("--device", default="cuda:0")
    a = ap.parse_args(argv)
    import torch
    from blendtorch import btt
    cam = btt.Camera.from_lens((a.height, a.width), device=a.device)
    pub = btt.DataPublisher(a.data, btid=a.btid, lingerms=2000, protocol=3)
    ctrl = btt.DuplexChannel(a.ctrl, btid=a.btid, bind=True)
    last = None
    try:
        while True:
            # supershape.blend.py:24-27 polls once per frame; a message sent with repeat=True is drawn over
            # and over until the next one arrives (a free-running producer)
            msg = ctrl.recv(timeoutms=0 if last is not None else 100)
            if msg is None:
                msg = last
            if msg is None:
                continue
            if msg.get("quit"):
                break
            last = msg if msg.get("repeat") else None
            for params, shape_id in zip(msg["shape_params"], msg["shape_ids"]):
                corners = cube_corners(params)
                img = btt.render_cubes(corners[None, None], cam.view_matrix, cam.proj_matrix, (a.height, a.width),
                                       bottom_up=True, with_depth=False, device=a.device)
                torch.cuda.synchronize()
                pub.publish(image=img[0].cpu().numpy(), shape_id=int(shape_id), corners=corners)
    finally:
        pub.close()
        ctrl.close()


if __name__ == "__main__":
    main()
