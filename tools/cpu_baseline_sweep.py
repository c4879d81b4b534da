#!/usr/bin/env python3
"""CPU baseline of the reference path over DataLoader worker counts (N in {0, 4,
cpu_count // n_gpus, cpu_count}), with and without the example item_transform.  Uses the oracle port
(test infrastructure), like bench.py's cpu_baseline leg.  Prints one JSON object per row."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "pytorch-blender_b200")]
import bench  # noqa: E402


def main():
    cores = os.cpu_count() or 1
    for workers in sorted({0, 4, max(1, cores - bench.PUBLISHERS - 1), cores}):
        for transform in (True, False):
            saved = bench._cpu_transform
            if not transform:
                bench._cpu_transform = _identity
            try:
                ips, done, dt = bench.run_cpu_reference(steps=100000, warmup=2, workers=workers, max_seconds=8.0,
                                                        check_shape=transform)
            finally:
                bench._cpu_transform = saved
            print(json.dumps({"workers": workers, "host_cpus": cores, "transform": "flipud+gamma+normalise+CHW" if transform
                              else "none (decode + collate only)", "images_per_s": round(ips, 1), "batches": done,
                              "seconds": round(dt, 1)}), flush=True)


def _identity(item):
    return item


if __name__ == "__main__":
    main()
