"""Synthetic Cube-scene frames on the GPU: the host face of kernel K3 (csrc/raster.cu).

Takes the place, in benchmarks and self-checks, of the reference's draw + read-back
(pkg_blender/blendtorch/btb/offscreen.py:76-92), which needs Blender and OpenGL.  No graphics API is
involved: the frames are computed by a CUDA rasteriser and stay in device memory, ready for K1.
"""
import ctypes as C

import torch

from . import native as N


class RasterParams(C.Structure):
    _fields_ = [("corners", C.c_void_p), ("view", C.c_void_p), ("proj", C.c_void_p),
                ("batch", C.c_int32), ("ncubes", C.c_int32),
                ("view_per_frame", C.c_int32), ("proj_per_frame", C.c_int32),
                ("height", C.c_int32), ("width", C.c_int32), ("bottom_up", C.c_int32),
                ("far_depth", C.c_float), ("background", C.c_uint8 * 4), ("light", C.c_float * 3),
                ("reserved", C.c_int32),
                ("colors", C.c_void_p), ("image_out", C.c_void_p), ("depth_out", C.c_void_p)]


N.lib.bt_raster_launch.restype = C.c_int
N.lib.bt_raster_launch.argtypes = [C.POINTER(RasterParams), C.c_void_p]


def render_cubes(corners, view, proj, shape, bottom_up=True, with_depth=True, background=(0, 0, 0, 255),
                 light=(0.3, 0.5, 1.0), colors=None, far_depth=float("inf"), device="cuda"):
    """Rasterise `corners` (B, M, 8, 3) world-space box corners (default-cube corner order) seen
    through `view` / `proj` ((4,4) or (B,4,4)) into uint8 RGBA frames (B, H, W, 4) and, optionally,
    linear depth (B, H, W).  `bottom_up=True` writes rows in GL read-back order (what the reference
    flips at offscreen.py:95-96)."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("render_cubes runs on the GPU; there is no CPU implementation here")
    pts = torch.as_tensor(corners, dtype=torch.float64, device=device).contiguous()
    if pts.dim() == 3:
        pts = pts.unsqueeze(0)
    if pts.dim() != 4 or pts.shape[2:] != (8, 3):
        raise ValueError("corners must be (B, M, 8, 3) or (M, 8, 3)")
    V = torch.as_tensor(view, dtype=torch.float64, device=device).contiguous()
    P = torch.as_tensor(proj, dtype=torch.float64, device=device).contiguous()
    B, M = pts.shape[0], pts.shape[1]
    H, W = shape
    img = torch.empty((B, H, W, 4), dtype=torch.uint8, device=device)
    depth = torch.empty((B, H, W), dtype=torch.float32, device=device) if with_depth else None
    p = RasterParams()
    p.corners, p.view, p.proj = pts.data_ptr(), V.data_ptr(), P.data_ptr()
    p.batch, p.ncubes = B, M
    p.view_per_frame, p.proj_per_frame = int(V.dim() == 3), int(P.dim() == 3)
    p.height, p.width, p.bottom_up = H, W, int(bool(bottom_up))
    p.far_depth = far_depth
    for i in range(4):
        p.background[i] = int(background[i])
    for i in range(3):
        p.light[i] = float(light[i])
    col = None
    if colors is not None:
        col = torch.as_tensor(colors, dtype=torch.uint8, device=device).contiguous()
        if col.shape != (M, 4):
            raise ValueError("colors must be (M, 4) uint8")
        p.colors = col.data_ptr()
    p.image_out = img.data_ptr()
    p.depth_out = depth.data_ptr() if depth is not None else None
    with torch.cuda.device(device):
        N.check(N.lib.bt_raster_launch(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return (img, depth) if with_depth else img
