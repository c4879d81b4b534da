#!/usr/bin/env python3
"""The headline K1 launch (batch 8 x 640x480 RGBA -> fp32 NCHW) twelve times on rotating inputs and
the C4-size launch three times, for one `ncu --set full -k regex:k1_planar` capture (diagnostic)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402

DEV = torch.device("cuda", 0)
spec = btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5)
g = torch.Generator(device=DEV)
g.manual_seed(1)
srcs = [torch.randint(0, 256, (8, 480, 640, 4), dtype=torch.uint8, device=DEV, generator=g) for _ in range(12)]
outs = [btt.preprocess_images(s, spec) for s in srcs]
torch.cuda.synchronize()
big = torch.randint(0, 256, (64, 1024, 1024, 4), dtype=torch.uint8, device=DEV, generator=g)
for _ in range(3):
    o = btt.preprocess_images(big, spec)
torch.cuda.synchronize()
print("ok")
