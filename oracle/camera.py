"""Oracle: annotation projection world -> NDC -> pixel (TEST INFRASTRUCTURE).

Restates the bpy-free arithmetic of
  pkg_blender/blendtorch/btb/camera.py:84-112  (world_to_ndc)
  pkg_blender/blendtorch/btb/camera.py:115-136 (ndc_to_pixel)
  pkg_blender/blendtorch/btb/utils.py:112-121  (hom / dehom)
The bpy-dependent getters (camera.py:57-82, utils.py:30-109) only *fetch*
matrices and vertices; here those are plain arrays.  All float64, as numpy
computes them in the reference.

`blender_perspective` / `blender_orthographic` rebuild the matrices Blender's
`calc_matrix_camera` returns for the cameras of tests/blender/cam.blend (lens
50 mm, sensor 36 mm, clip 0.1..100, ortho_scale 4, horizontal sensor fit), which
is what makes tests/test_camera.py:19-40 reproducible without Blender.
"""

import numpy as np


def hom(x, v=1.0):
    """utils.py:112-117."""
    x = np.asarray(x)
    return np.concatenate((x, np.full((x.shape[0], 1), v, dtype=x.dtype)), axis=-1)


def dehom(x):
    """utils.py:119-121."""
    return x[..., :-1] / x[..., -1:]


def world_to_ndc(xyz_world, view_matrix, proj_matrix, return_depth=False):
    """camera.py:84-112.  With depth: two products (view, then projection) and
    d = -z_cam.  Without: one product with the pre-multiplied P @ V."""
    pts = hom(np.atleast_2d(xyz_world), 1.0)
    V = np.asarray(view_matrix)
    P = np.asarray(proj_matrix)
    if return_depth:
        cam = pts @ V.T
        depth = -cam[:, -2].copy()
        clip = cam @ P.T
        return dehom(clip), depth
    return dehom(pts @ (P @ V).T)


def ndc_to_pixel(ndc, shape, origin="upper-left"):
    """camera.py:115-136.  shape = (H, W)."""
    assert origin in ("upper-left", "lower-left")
    h, w = shape
    xy = np.atleast_2d(ndc)[:, :2]
    xy = (xy + 1) * 0.5
    if origin == "upper-left":
        xy[:, 1] = 1.0 - xy[:, 1]
    return xy * np.array([[w, h]])


def world_to_pixel(xyz_world, view_matrix, proj_matrix, shape, origin="upper-left",
                   return_depth=False):
    """camera.py:138-162 minus the bpy vertex fetch."""
    if return_depth:
        ndc, z = world_to_ndc(xyz_world, view_matrix, proj_matrix, True)
        return ndc_to_pixel(ndc, shape, origin), z
    return ndc_to_pixel(world_to_ndc(xyz_world, view_matrix, proj_matrix), shape, origin)


def pixel_bboxes(xy, corners_per_object=8):
    """(M*8, 2) projected box corners -> (M, 4) [xmin, ymin, xmax, ymax].

    NOT in the reference: camera.py:165-189 returns the 8 projected corners per
    object and stops.  BASELINE.json config 3 asks for pixel bboxes, so the
    oracle is the obvious min/max over each object's corners."""
    c = np.asarray(xy).reshape(-1, corners_per_object, 2)
    return np.concatenate((c.min(axis=1), c.max(axis=1)), axis=-1)


def blender_perspective(shape, lens=50.0, sensor=36.0, near=0.1, far=100.0):
    h, w = shape
    sx = 2.0 * lens / sensor
    sy = sx * w / h
    P = np.zeros((4, 4))
    P[0, 0], P[1, 1] = sx, sy
    P[2, 2] = -(far + near) / (far - near)
    P[2, 3] = -2.0 * far * near / (far - near)
    P[3, 2] = -1.0
    return P


def blender_orthographic(shape, ortho_scale=4.0, near=0.1, far=100.0):
    h, w = shape
    sx = 2.0 / ortho_scale
    sy = sx * w / h
    P = np.zeros((4, 4))
    P[0, 0], P[1, 1] = sx, sy
    P[2, 2] = -2.0 / (far - near)
    P[2, 3] = -(far + near) / (far - near)
    P[3, 3] = 1.0
    return P


def view_looking_down_z(distance=7.0):
    """Camera on +Z at `distance`, looking at the origin, Y up (cam.blend)."""
    V = np.eye(4)
    V[2, 3] = -distance
    return V


def default_cube_vertices():
    """Blender default cube vertex order: (+-1, +-1, +-1), z fastest, then y,
    x from +1 to -1 -- matches the row order of tests/test_camera.py:19-40."""
    return np.array([[x, y, z] for x in (1.0, -1.0) for y in (1.0, -1.0) for z in (1.0, -1.0)])
