#!/usr/bin/env python3
"""GPU-resident producer: K3 (rasterise the synthetic Cube scene) -> K1 (fused preprocess), no PCIe
in the loop.  Diagnostic: prints frames/s of K3 alone and of the chain."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

from blendtorch import btt  # noqa: E402
from blendtorch.btt import synthetic as S  # noqa: E402


def main():
    dev = "cuda:0"
    spec = btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5)
    for (B, H, W) in ((8, 480, 640), (32, 1080, 1920), (64, 1024, 1024)):
        scenes = np.stack([S.cube_corners(S.rng_for(3, b)).reshape(8, 8, 3) for b in range(B)])
        cam = btt.Camera.from_lens((H, W), lens=35.0, distance=20.0, device=dev)
        pts = torch.as_tensor(scenes, device=dev)
        Vt, Pt = cam.view_matrix, cam.proj_matrix
        res = {}
        for name, chain in (("k3", False), ("k3+k1", True)):
            for _ in range(3):
                img = btt.render_cubes(pts, Vt, Pt, (H, W), with_depth=False, device=dev)
                if chain:
                    out = btt.preprocess_images(img, spec)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 30
            e0.record()
            for _ in range(reps):
                img = btt.render_cubes(pts, Vt, Pt, (H, W), with_depth=False, device=dev)
                if chain:
                    out = btt.preprocess_images(img, spec)
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1) / reps
            res[name] = {"ms_per_batch": round(ms, 3), "frames_per_s": round(B / ms * 1e3), "Gpixel_per_s": round(B * H * W / ms / 1e6, 2)}
        print(json.dumps({"batch": B, "frame": [H, W], "cubes": 8, **res}), flush=True)


if __name__ == "__main__":
    main()
