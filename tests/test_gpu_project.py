"""K2 parity: bt_project_launch against the reference's own goldens (tests/test_camera.py:19-40),
against outputs of the reference's Camera code (tests/golden/camera_golden.npz) and against the
oracle on random batches.  Tolerance: 1e-6 relative is the north_star bar; float64 arithmetic
lands within 1e-12 (numpy's BLAS may fuse multiply-adds, so bit equality is not promised)."""
import numpy as np
import pytest
import torch
from numpy.testing import assert_allclose

from blendtorch import btt
from oracle import camera as ocam, synth

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-12, 1e-12


def test_reference_camera_goldens(golden_dir):
    g = np.load(golden_dir / "camera_golden.npz")
    shape = (480, 640)
    for tag, exp in (("ortho", g["expected_ortho_xy"]), ("persp", g["expected_proj_xy"])):
        cam = btt.Camera(g["V"], g["P_" + tag], shape)
        xy, z = cam.world_to_pixel(g["cube"], return_depth=True)
        assert_allclose(xy.cpu().numpy(), exp, atol=1e-2)                     # tests/test_camera.py:46-49
        assert_allclose(z.cpu().numpy(), g["expected_z"], atol=1e-2)
        ndc, z = cam.world_to_ndc(g["cube"], return_depth=True)
        assert_allclose(ndc.cpu().numpy(), g["ref_%s_ndc" % tag], rtol=RTOL, atol=ATOL)
        assert_allclose(z.cpu().numpy(), g["ref_%s_z" % tag], rtol=RTOL, atol=ATOL)
        assert_allclose(cam.world_to_ndc(g["cube"]).cpu().numpy(), g["ref_%s_ndc_nodepth" % tag], rtol=RTOL, atol=ATOL)
        for origin in ("upper-left", "lower-left"):
            px = cam.ndc_to_pixel(g["ref_%s_ndc" % tag], origin=origin)
            assert_allclose(px.cpu().numpy(), g["ref_%s_px_%s" % (tag, "ul" if origin[0] == "u" else "ll")],
                            rtol=RTOL, atol=1e-9)
    # the bpy-free constructor gives the matrices of the reference's test scene (cam.blend: 50 mm lens on a
    # 36 mm sensor at z = 7 looking down -Z; ortho_scale 4), hence the reference's own expected pixels
    for cam, exp in ((btt.Camera.from_lens(shape), g["expected_proj_xy"]),
                     (btt.Camera.from_lens(shape, ortho_scale=4.0), g["expected_ortho_xy"])):
        np.testing.assert_array_equal(cam.view_matrix.cpu().numpy(), g["V"])
        xy, z = cam.world_to_pixel(g["cube"], return_depth=True)
        assert_allclose(xy.cpu().numpy(), exp, atol=1e-2)
        assert_allclose(z.cpu().numpy(), g["expected_z"], atol=1e-2)
    cam = btt.Camera(g["rand_V"], g["rand_P"], (1080, 1920))
    xy, z = cam.world_to_pixel(g["rand_pts"], return_depth=True)
    assert_allclose(xy.cpu().numpy(), g["rand_px_ul"], rtol=1e-11, atol=1e-9)
    assert_allclose(z.cpu().numpy(), g["rand_z"], rtol=RTOL, atol=ATOL)
    assert_allclose(cam.world_to_pixel(g["rand_pts"], origin="lower-left").cpu().numpy(), g["rand_px_ll"],
                    rtol=1e-11, atol=1e-9)
    with pytest.raises(AssertionError):
        cam.ndc_to_pixel(g["rand_ndc"], origin="centre")


def test_batched_random_frames_and_boxes():
    """BASELINE config 3: batch 32, 8 cubes x 8 corners, per-frame view matrices."""
    B, shape = 32, (1080, 1920)
    pts = np.stack([synth.cube_corners(synth.rng_for(3, b)) for b in range(B)])
    rng = np.random.default_rng(33)
    Vs = np.tile(ocam.view_looking_down_z(0.0), (B, 1, 1))
    Vs[:, :3, 3] = rng.uniform(-0.5, 0.5, (B, 3)) + np.array([0, 0, -2.0])
    P = ocam.blender_perspective(shape, lens=35.0)
    cam = btt.Camera(Vs, P, shape)
    xy, z = cam.world_to_pixel(pts, return_depth=True)
    xy2, boxes = cam.bbox_world_to_pixel_boxes(pts, corners_per_object=8)
    xy, z, xy2, boxes = (t.cpu().numpy() for t in (xy, z, xy2, boxes))
    for b in range(B):
        exy, ez = ocam.world_to_pixel(pts[b], Vs[b], P, shape, return_depth=True)
        # 1e-6 relative is the bar; measured agreement is ~1e-13
        assert_allclose(xy[b], exy, rtol=1e-11, atol=1e-8)
        assert_allclose(z[b], ez, rtol=RTOL, atol=ATOL)
        exy_nd = ocam.world_to_pixel(pts[b], Vs[b], P, shape)
        assert_allclose(xy2[b], exy_nd, rtol=1e-10, atol=1e-8)
        np.testing.assert_array_equal(boxes[b], ocam.pixel_bboxes(xy2[b]))   # min/max of the same numbers: exact
    assert boxes.shape == (B, 8, 4)
    # float32 relative bar of the north_star on the same data
    rel = np.abs(xy.astype(np.float32) - np.stack([ocam.world_to_pixel(pts[b], Vs[b], P, shape) for b in range(B)]).astype(np.float32))
    assert (rel <= 1e-6 * np.abs(xy) + 1e-6).all()


def test_edge_cases():
    from blendtorch.btt import native as N
    cam = btt.Camera(np.eye(4), ocam.blender_orthographic((4, 4)), (4, 4))
    assert cam.world_to_pixel(np.zeros((0, 3))).shape == (0, 2)
    one = cam.world_to_pixel(np.array([[0.0, 0.0, -1.0]]))
    assert_allclose(one.cpu().numpy(), [[2.0, 2.0]])
    with pytest.raises(N.NativeError):
        cam.bbox_world_to_pixel_boxes(np.zeros((6, 3)), corners_per_object=3)
    with pytest.raises(ValueError):
        btt.Camera(np.eye(3), np.eye(4), (4, 4))
    # 10000 points, one launch, several blocks
    rng = np.random.default_rng(2)
    pts = rng.uniform(-1, 1, (10000, 3)) + np.array([0, 0, -5.0])
    P = ocam.blender_perspective((480, 640))
    cam = btt.Camera(np.eye(4), P, (480, 640))
    assert_allclose(cam.world_to_pixel(pts).cpu().numpy(), ocam.world_to_pixel(pts, np.eye(4), P, (480, 640)),
                    rtol=1e-11, atol=1e-8)
