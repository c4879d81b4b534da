"""Annotation projection on the device: the host face of kernel K2.

Mirrors the bpy-free arithmetic of the reference's ``btb.Camera``
(pkg_blender/blendtorch/btb/camera.py:84-189) with the same method names and argument
meaning, for batches of frames held in CUDA tensors.  The reference runs this per frame inside
Blender on the CPU; here one launch projects every point of every frame of a batch.  Matrices
are supplied by the caller (the bpy getters camera.py:57-82 have no place on the btt side).
"""
import ctypes as C

import numpy as np
import torch

from . import native as N


def _f64(t, device):
    t = torch.as_tensor(t, dtype=torch.float64, device=device)
    return t.contiguous()


class Camera:
    """Batched camera.

    view_matrix, proj_matrix: (4,4) shared by all frames, or (B,4,4) one per frame.
    shape: (H, W) of the image, as in the reference.
    """

    def __init__(self, view_matrix, proj_matrix, shape, device="cuda"):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("btt.Camera projects on the GPU; there is no CPU implementation here")
        self.view_matrix = _f64(view_matrix, self.device)
        self.proj_matrix = _f64(proj_matrix, self.device)
        for m in (self.view_matrix, self.proj_matrix):
            if m.shape[-2:] != (4, 4) or m.dim() not in (2, 3):
                raise ValueError("matrices must be (4,4) or (B,4,4)")
        self.shape = tuple(shape)

    @classmethod
    def from_lens(cls, shape, lens=50.0, sensor=36.0, near=0.1, far=100.0, distance=7.0, ortho_scale=None,
                  device="cuda"):
        """A camera without Blender: on the +Z axis at `distance`, looking at the origin, Y up, with the
        projection Blender computes for a perspective camera of focal length `lens` on a `sensor` mm wide
        sensor (horizontal fit) -- or an orthographic one of `ortho_scale`.  Stands in for the bpy-dependent
        getters of the reference (btb/camera.py:57-82) in synthetic scenes, benchmarks and examples."""
        h, w = shape
        view = np.eye(4)
        view[2, 3] = -float(distance)
        proj = np.zeros((4, 4))
        if ortho_scale is None:
            proj[0, 0] = 2.0 * lens / sensor
            proj[1, 1] = proj[0, 0] * w / h
            proj[2, 2] = -(far + near) / (far - near)
            proj[2, 3] = -2.0 * far * near / (far - near)
            proj[3, 2] = -1.0
        else:
            proj[0, 0] = 2.0 / ortho_scale
            proj[1, 1] = proj[0, 0] * w / h
            proj[2, 2] = -2.0 / (far - near)
            proj[2, 3] = -(far + near) / (far - near)
            proj[3, 3] = 1.0
        return cls(view, proj, shape, device=device)

    # -- internals ---------------------------------------------------------------------------
    def _points(self, xyz):
        pts = _f64(xyz, self.device)
        squeeze = pts.dim() == 2
        if pts.dim() == 1:
            pts = pts.view(1, 1, -1)
            squeeze = True
        elif squeeze:
            pts = pts.unsqueeze(0)
        if pts.dim() != 3 or pts.shape[-1] != 3:
            raise ValueError("points must be (N,3) or (B,N,3)")
        return pts, squeeze

    def _launch(self, pts, want_ndc, want_xy, want_depth, origin, bbox_corners=0, input_is_ndc=False):
        assert origin in ["upper-left", "lower-left"]
        B, Np, _ = pts.shape
        for m in (self.view_matrix, self.proj_matrix):
            if m.dim() == 3 and m.shape[0] != B:
                raise ValueError("per-frame matrices need one matrix per frame (%d != %d)" % (m.shape[0], B))
        p = N.ProjectParams()
        p.points = pts.data_ptr()
        p.view, p.proj = self.view_matrix.data_ptr(), self.proj_matrix.data_ptr()
        p.batch, p.npoints = B, Np
        p.view_per_frame = int(self.view_matrix.dim() == 3)
        p.proj_per_frame = int(self.proj_matrix.dim() == 3)
        p.height, p.width = self.shape
        p.origin_upper_left = int(origin == "upper-left")
        p.with_depth = int(want_depth)
        p.input_is_ndc = int(input_is_ndc)
        out = {}
        if want_ndc:
            out["ndc"] = torch.empty((B, Np, 3), dtype=torch.float64, device=self.device)
            p.ndc_out = out["ndc"].data_ptr()
        if want_xy:
            out["xy"] = torch.empty((B, Np, 2), dtype=torch.float64, device=self.device)
            p.xy_out = out["xy"].data_ptr()
        if want_depth:
            out["z"] = torch.empty((B, Np), dtype=torch.float64, device=self.device)
            p.depth_out = out["z"].data_ptr()
        if bbox_corners:
            out["bbox"] = torch.empty((B, Np // bbox_corners, 4), dtype=torch.float64, device=self.device)
            p.bbox_out = out["bbox"].data_ptr()
            p.corners_per_object = bbox_corners
        if B * Np == 0:  # nothing to project
            return out
        with torch.cuda.device(self.device):
            N.check(N.lib.bt_project_launch(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out

    # -- reference-named API -----------------------------------------------------------------
    def world_to_ndc(self, xyz_world, return_depth=False):
        """camera.py:84-112: NDC (and linear camera-space depth) of world points."""
        pts, squeeze = self._points(xyz_world)
        o = self._launch(pts, True, False, return_depth, "upper-left")
        ndc = o["ndc"][0] if squeeze else o["ndc"]
        if return_depth:
            return ndc, (o["z"][0] if squeeze else o["z"])
        return ndc

    def ndc_to_pixel(self, ndc, origin="upper-left"):
        """camera.py:115-136."""
        pts, squeeze = self._points(ndc)
        o = self._launch(pts, False, True, False, origin, input_is_ndc=True)
        return o["xy"][0] if squeeze else o["xy"]

    def world_to_pixel(self, xyz_world, return_depth=False, origin="upper-left"):
        """``ndc_to_pixel(world_to_ndc(xyz))`` in one launch -- what ``object_to_pixel``
        (camera.py:138-162) computes once the vertices are known."""
        pts, squeeze = self._points(xyz_world)
        o = self._launch(pts, False, True, return_depth, origin)
        xy = o["xy"][0] if squeeze else o["xy"]
        if return_depth:
            return xy, (o["z"][0] if squeeze else o["z"])
        return xy

    # the reference's convenience names, minus the bpy vertex fetch
    object_to_pixel = world_to_pixel
    bbox_object_to_pixel = world_to_pixel

    def bbox_world_to_pixel_boxes(self, corners_world, corners_per_object=8, origin="upper-left"):
        """Projected corners (B, M*k, 2) and their per-object pixel boxes (B, M, 4) =
        [xmin, ymin, xmax, ymax].  The box reduction is not in the reference (it stops at the
        corners, camera.py:165-189); BASELINE config 3 asks for it."""
        pts, squeeze = self._points(corners_world)
        o = self._launch(pts, False, True, False, origin, bbox_corners=corners_per_object)
        if squeeze:
            return o["xy"][0], o["bbox"][0]
        return o["xy"], o["bbox"]


class Projection:
    """Per-batch annotation projection inside `DeviceLoader`: world points carried by the messages
    (e.g. the 8 corners of every object's bounding box, as `btb.utils.bbox_world_coordinates`
    yields them, utils.py:84-109) become pixel coordinates, optional depths and optional per-object
    pixel boxes, by one K2 launch per batch on the loader's stream.

    key:        message field holding the points, shape (N, 3) per frame.
    camera:     a `Camera` with fixed matrices, or None to take them from the message fields
                `view_key` / `proj_key` ((4, 4) per frame) with image `shape`.
    xy_key, depth_key, bbox_key: names of the outputs added to the batch (None = not computed).
    corners_per_object: group size for `bbox_key` (8 for box corners).
    """

    def __init__(self, key, camera=None, xy_key="xy", depth_key=None, bbox_key=None, corners_per_object=8,
                 origin="upper-left", view_key=None, proj_key=None, shape=None):
        if camera is None and not (view_key and proj_key and shape):
            raise ValueError("give a Camera, or view_key + proj_key + shape")
        assert origin in ["upper-left", "lower-left"]
        self.key, self.camera = key, camera
        self.xy_key, self.depth_key, self.bbox_key = xy_key, depth_key, bbox_key
        self.corners_per_object, self.origin = corners_per_object, origin
        self.view_key, self.proj_key, self.shape = view_key, proj_key, shape

    def apply(self, batch, device):
        pts = batch[self.key]
        if pts.dtype != torch.float64:
            pts = pts.to(torch.float64)
        pts = pts.contiguous()
        cam = self.camera
        if cam is None:
            cam = Camera.__new__(Camera)
            cam.device = device
            cam.view_matrix = batch[self.view_key].to(torch.float64).contiguous()
            cam.proj_matrix = batch[self.proj_key].to(torch.float64).contiguous()
            cam.shape = tuple(self.shape)
        out = cam._launch(pts, False, True, self.depth_key is not None, self.origin,
                          bbox_corners=self.corners_per_object if self.bbox_key else 0)
        made = []
        for name, k in ((self.xy_key, "xy"), (self.depth_key, "z"), (self.bbox_key, "bbox")):
            if name and k in out:
                batch[name] = out[k]
                made.append(out[k])
        return made
