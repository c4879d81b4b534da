"""bench.py's host-side helpers (no GPU): core sets per rank, thread rules, ports, and the value check
that every end-to-end leg runs on a sampled batch."""
import importlib.util
import os
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", ROOT / "bench.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_ranks_get_disjoint_core_sets_and_thread_rules_hold():
    b = _bench()
    cpus = sorted(os.sched_getaffinity(0))
    for world in (1, 2, 4, 8):
        sets = [b.rank_cpus(r, world) for r in range(world)]
        assert all(sets)
        if world > 1 and len(cpus) >= world:
            flat = [c for s in sets for c in s]
            assert len(flat) == len(set(flat)) and set(flat) <= set(cpus)      # disjoint, inside the allowed set
    assert [b.stream_threads(c, 5) for c in (4, 8, 12, 16)] == [(5, 2), (5, 4), (5, 7), (5, 10)]
    s, r, _ = b.split_threads(16, 6500.0, 4800.0, 16)
    assert 1 <= s <= 16 and 1 <= r <= 10 and s + r <= 15
    lo, hi = (int(x) for x in open('/proc/sys/net/ipv4/ip_local_port_range').read().split())
    assert not lo <= b.free_port() <= hi                                        # outside the ephemeral range


def test_value_check_accepts_the_publishers_frames_and_rejects_anything_else():
    """check_batch recomputes a batch's statistics from the frames its (btid, frameid) pairs name, with the
    reference's expressions in numpy; a batch assembled from those frames passes, one wrong byte or a wrong
    annotation does not."""
    b = _bench()
    import sys
    sys.path.insert(0, str(ROOT / "pytorch-blender_b200"))
    from blendtorch.btt import synthetic as S
    from oracle import preprocess as opre
    h, w = 24, 32
    ids = [(0, 3), (1, 5), (0, 0), (1, 15)]
    msgs = [b._pool_messages(S, 2, bt, h, w)[f] for bt, f in ids]
    out = opre.preprocess_batch([m["image"] for m in msgs], flip=True, gamma=b.GAMMA, mean=b.MEAN, std=b.STD)
    s1, s2 = opre.batch_stats(out)
    batch = {"btid": torch.tensor([i[0] for i in ids]), "frameid": torch.tensor([i[1] for i in ids]),
             "bbox": torch.from_numpy(np.stack([m["bbox"] for m in msgs])),
             "image": torch.from_numpy(out), "image_stats": torch.from_numpy(np.stack([s1, s2], axis=1))}
    assert b.check_batch(np, S, batch, 2, h, w) == "ok"
    wrong = dict(batch, image_stats=batch["image_stats"] + torch.tensor([[0.0, 0.0], [0.0, 0.0], [0.5, 0.0], [0.0, 0.0]]))
    assert b.check_batch(np, S, wrong, 2, h, w) != "ok"
    wrong = dict(batch, bbox=batch["bbox"].roll(1, 0))
    assert "bbox" in b.check_batch(np, S, wrong, 2, h, w)
    wrong = dict(batch, frameid=torch.tensor([3, 5, 1, 15]))                       # another frame's statistics
    assert b.check_batch(np, S, wrong, 2, h, w) != "ok"
