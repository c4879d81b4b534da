#!/usr/bin/env python3
"""Small, fast exercise of every kernel (all K1 variants incl. statistics, K2, K3) for
compute-sanitizer: `compute-sanitizer --tool memcheck python tools/sanitizer_case.py`."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402

dev = "cuda:0"
rng = np.random.default_rng(0)
n = 0
for shape in ((2, 12, 16, 4), (3, 7, 5, 4), (2, 9, 16, 3), (2, 8, 32, 3), (1, 5, 16, 1), (2, 6, 10, 2), (2, 33, 47, 4)):
    src = torch.from_numpy(rng.integers(0, 256, shape, dtype=np.uint8)).to(dev)
    for dt in (torch.float32, torch.bfloat16, torch.uint8):
        for layout in ("NCHW", "NHWC"):
            for stats in (False, True):
                for variant in ("auto", "generic"):
                    kw = dict(flip=True, gamma=2.2, layout=layout, dtype=dt, stats=stats, variant=variant)
                    if dt != torch.uint8:
                        kw.update(mean=127.5, std=127.5)
                    btt.preprocess_images(src, btt.Preprocess(**kw))
                    n += 1
cam = btt.Camera(np.eye(4), np.eye(4), (48, 64), device=dev)
pts = rng.random((3, 64, 3)) - 0.5
cam.world_to_pixel(pts, return_depth=True)
cam.bbox_world_to_pixel_boxes(pts, 8)
cam.ndc_to_pixel(pts)
V = np.eye(4)
V[2, 3] = -20
P = np.array([[2.0, 0, 0, 0], [0, 2.6, 0, 0], [0, 0, -1.002, -0.2], [0, 0, -1, 0]])
corners = rng.uniform(-3, 3, (2, 5, 8, 3))
btt.render_cubes(corners, V, P, (45, 70), device=dev)
torch.cuda.synchronize()
print("sanitizer case done:", n, "K1 launches + K2 + K3")
