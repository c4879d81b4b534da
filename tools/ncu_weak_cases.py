#!/usr/bin/env python3
"""The K1 cases that sit furthest from the HBM roofline, three plain launches each, for one
`ncu --set full -k regex:k1_` capture (diagnostic)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402

DEV = torch.device("cuda", 0)
C2, C4 = (8, 480, 640, 4), (64, 1024, 1024, 4)
CASES = [
    ("u8 NHWC gamma C4", C4, dict(flip=True, gamma=2.2, layout="NHWC", dtype=torch.uint8)),
    ("fp32 NCHW stats C2", C2, dict(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True)),
    ("fp32 NCHW stats C4", C4, dict(flip=True, gamma=2.2, mean=127.5, std=127.5, stats=True)),
    ("fp32 NCHW odd width", (8, 481, 641, 4), dict(flip=True, gamma=2.2, mean=127.5, std=127.5)),
    ("fp32 NHWC C4", C4, dict(flip=True, gamma=2.2, mean=127.5, std=127.5, layout="NHWC")),
]
g = torch.Generator(device=DEV)
g.manual_seed(1)
for name, shape, kw in CASES:
    src = torch.randint(0, 256, shape, dtype=torch.uint8, device=DEV, generator=g)
    spec = btt.Preprocess(**kw)
    for _ in range(3):
        out = btt.preprocess_images(src, spec)
    torch.cuda.synchronize()
    print(name, "ok", flush=True)
    del src, out
