"""K3 (synthetic Cube-scene rasteriser) against its numpy oracle, bit for bit, and
against K2: annotation corners projected by K2 must land on the rasterised silhouettes.  There is
no reference output for this (the reference draws through OpenGL, unavailable on this box)."""
import numpy as np
import pytest
import torch

from blendtorch import btt
from oracle import camera as ocam, preprocess as opre, raster as oraster, synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _scene(seed, n_cubes):
    return synth.cube_corners(np.random.default_rng(seed), n_cubes).reshape(n_cubes, 8, 3)


@pytest.mark.parametrize("proj_kind", ["persp", "ortho"])
@pytest.mark.parametrize("bottom_up", [True, False])
def test_bit_exact_against_oracle(proj_kind, bottom_up):
    H, W = 90, 120
    V = ocam.view_looking_down_z(20.0)
    P = ocam.blender_perspective((H, W), lens=35.0) if proj_kind == "persp" else ocam.blender_orthographic((H, W), ortho_scale=16.0)
    scenes = np.stack([_scene(s, 5) for s in (1, 2, 3)])
    img, depth = btt.render_cubes(scenes, V, P, (H, W), bottom_up=bottom_up, background=(10, 20, 30, 255), device=DEV)
    for b in range(3):
        eimg, edepth = oraster.render(scenes[b], V, P, H, W, bottom_up=bottom_up, background=(10, 20, 30, 255))
        np.testing.assert_array_equal(img[b].cpu().numpy(), eimg)
        got = depth[b].cpu().numpy()
        np.testing.assert_array_equal(got.view(np.uint32), edepth.view(np.uint32))
        assert (eimg[..., :3] != (10, 20, 30)).any(axis=-1).mean() > 0.05      # something was drawn


def test_per_frame_cameras_colors_and_edge_cases():
    from blendtorch.btt import native as N
    H, W = 48, 64
    scenes = np.stack([_scene(s, 3) for s in (7, 8)])
    Vs = np.stack([ocam.view_looking_down_z(18.0), ocam.view_looking_down_z(25.0)])
    P = ocam.blender_perspective((H, W), lens=30.0)
    colors = np.array([[255, 0, 0, 255], [0, 255, 0, 255], [0, 0, 255, 255]], np.uint8)
    img, depth = btt.render_cubes(scenes, Vs, P, (H, W), bottom_up=False, colors=colors, light=(0, 0, 1), device=DEV)
    for b in range(2):
        eimg, edepth = oraster.render(scenes[b], Vs[b], P, H, W, bottom_up=False, colors=colors, light=(0, 0, 1))
        np.testing.assert_array_equal(img[b].cpu().numpy(), eimg)
        np.testing.assert_array_equal(depth[b].cpu().numpy(), edepth)
    # a cube behind the camera draws nothing; odd sizes are fine
    behind = _scene(1, 1) + np.array([0, 0, 40.0])
    out = btt.render_cubes(behind, ocam.view_looking_down_z(20.0), P, (33, 47), with_depth=False, device=DEV)
    assert out.shape == (1, 33, 47, 4) and (out[..., :3] == 0).all()
    with pytest.raises(N.NativeError):
        btt.render_cubes(np.zeros((40, 8, 3)), np.eye(4), P, (8, 8), device=DEV)


def test_projected_corners_lie_on_the_rasterised_silhouettes():
    """End-to-end consistency of frames and annotations: every corner K2 projects into the frame sits on
    (or within one pixel of) a pixel K3 covered, and nothing is drawn outside the K2 pixel boxes."""
    H, W = 270, 480
    V, P = ocam.view_looking_down_z(20.0), ocam.blender_perspective((H, W), lens=35.0)
    B = 4
    scenes = np.stack([_scene(100 + s, 8) for s in range(B)])
    img = btt.render_cubes(scenes, V, P, (H, W), bottom_up=False, with_depth=False, device=DEV).cpu().numpy()
    cam = btt.Camera(V, P, (H, W), device=DEV)
    xy, boxes = cam.bbox_world_to_pixel_boxes(scenes.reshape(B, 64, 3), corners_per_object=8)
    xy, boxes = xy.cpu().numpy(), boxes.cpu().numpy()
    drawn = (img[..., :3] != 0).any(axis=-1)
    for b in range(B):
        pad = np.pad(drawn[b], 1)
        near = np.zeros_like(drawn[b])
        for dy in range(3):
            for dx in range(3):
                near |= pad[dy:dy + H, dx:dx + W]
        for (x, y) in xy[b]:
            ix, iy = int(np.floor(x)), int(np.floor(y))
            if 1 <= ix < W - 1 and 1 <= iy < H - 1:
                assert near[iy, ix], "corner (%.2f, %.2f) is not on a drawn pixel" % (x, y)
        inside = np.zeros_like(drawn[b])
        for x0, y0, x1, y1 in boxes[b]:
            inside[max(int(np.floor(y0)), 0):int(np.ceil(y1)) + 1, max(int(np.floor(x0)), 0):int(np.ceil(x1)) + 1] = True
        assert not (drawn[b] & ~inside).any()


def test_gpu_resident_producer_feeds_k1():
    """raster (bottom-up, as a GL read-back) -> K1 with flip == preprocessing of the top-down frame."""
    H, W = 64, 96
    V, P = ocam.view_looking_down_z(20.0), ocam.blender_perspective((H, W), lens=35.0)
    scenes = np.stack([_scene(s, 6) for s in (11, 12)])
    up = btt.render_cubes(scenes, V, P, (H, W), bottom_up=True, with_depth=False, device=DEV)
    down = btt.render_cubes(scenes, V, P, (H, W), bottom_up=False, with_depth=False, device=DEV)
    out = btt.preprocess_images(up, btt.Preprocess(flip=True, gamma=2.2, mean=127.5, std=127.5))
    exp = opre.preprocess_batch(list(down.cpu().numpy()), gamma=2.2, mean=127.5, std=127.5)
    np.testing.assert_array_equal(out.cpu().numpy(), exp)
