"""Next-row N4: the loop of examples/densityopt/densityopt.py:95-119,257-316 closed over the device
path.  A training loop sends every producer its share of parameter samples over `btt.DuplexChannel`
(`update_simulations`), the producers draw what the parameters say (K3) and publish it with the
`shape_id` of the sample, `btt.DeviceLoader` delivers device batches (K1 + K2), and the loop scores
them and moves the distribution.  Checked: every frame comes back with the id of the sample that
produced it, its projected box is where that sample puts the cube (against the oracle's camera maths),
its pixels show the cube there, and the distribution moves towards the target."""
import importlib.util
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import camera as ocam

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _load(name):
    spec = importlib.util.spec_from_file_location(name, ROOT / "examples" / (name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_parameters_steer_producers_and_frames_come_back_with_their_sample_id():
    steer, producer = _load("steer_producers"), _load("steered_producer")
    H, W, B = 120, 160, 8
    from blendtorch import btt
    cam = btt.Camera.from_lens((H, W), device="cuda:0")
    V, P = cam.view_matrix.cpu().numpy(), cam.proj_matrix.cpu().numpy()
    seen = []

    def log(step, samples, shape_id, centre, batch):
        assert sorted(shape_id.tolist()) == list(range(B))          # every sample of this step, exactly once
        img = batch["image"].cpu().numpy()                           # (B, 4, H, W), flipped to top-down, normalised
        for i, sid in enumerate(shape_id.tolist()):
            corners = producer.cube_corners(list(samples[sid].tolist()) + [0.5])
            np.testing.assert_array_equal(batch["corners"][i].cpu().numpy(), corners)
            xy = ocam.world_to_pixel(corners, V, P, (H, W))
            want = np.array([(xy[:, 0].min() + xy[:, 0].max()) / 2, (xy[:, 1].min() + xy[:, 1].max()) / 2])
            np.testing.assert_allclose(centre[i].numpy(), want, rtol=1e-9, atol=1e-6)
            cx, cy = int(round(want[0])), int(round(want[1]))
            if 2 <= cx < W - 2 and 2 <= cy < H - 2:
                assert img[i, :3, cy, cx].max() > -1.0              # the cube is drawn where its sample puts it
        seen.append(step)

    torch.manual_seed(0)
    target = (110.0, 40.0)
    hist, _ = steer.run(producers=2, steps=40, batch=B, height=H, width=W, target=target, lr=0.08, log=log)
    assert seen == list(range(40))
    first = np.mean([h["err_px"] for h in hist[:5]])
    last = np.mean([h["err_px"] for h in hist[-5:]])
    assert last < 0.7 * first, (first, last)                          # the distribution moved towards the target
    assert hist[-1]["mean"][0] > 0.2 and hist[-1]["mean"][1] > 0.2    # right of and above the image centre


def test_free_running_producers_keep_publishing_the_last_parameters():
    steer = _load("steer_producers")
    torch.manual_seed(1)
    ids = []
    hist, _ = steer.run(producers=2, steps=6, batch=8, height=60, width=80, steer=False,
                        log=lambda step, samples, shape_id, centre, batch: ids.append(sorted(shape_id.tolist())))
    assert len(hist) == 6 and all(len(x) == 8 for x in ids)
