"""Oracle: the synthetic Cube-scene rasteriser (TEST INFRASTRUCTURE).

There is no reference code for this: the reference draws through Blender/OpenGL
(pkg_blender/blendtorch/btb/offscreen.py:76-92), which is unavailable on this box.  This is the
scalar/numpy statement of the algorithm csrc/raster.cu implements -- same operations, same order,
float64 for vertex projection (the arithmetic of camera.py:84-136) and float32, one rounding per
operation, for everything in screen space -- so that the CUDA kernel can be checked bit for bit.
"""
import numpy as np

F = np.float32

TRIANGLES = np.array([[0, 1, 3], [0, 3, 2], [4, 6, 7], [4, 7, 5], [0, 4, 5], [0, 5, 1],
                      [2, 3, 7], [2, 7, 6], [0, 2, 6], [0, 6, 4], [1, 5, 7], [1, 7, 3]])
PALETTE = np.array([[230, 60, 60], [60, 200, 80], [70, 110, 240], [240, 200, 60],
                    [200, 80, 220], [60, 210, 210], [250, 140, 50], [180, 180, 180]], dtype=np.float32)


def _dot4(m, x, y, z, w):
    acc = x * m[0]
    acc = acc + y * m[1]
    acc = acc + z * m[2]
    acc = acc + w * m[3]
    return acc


def _edge(ax, ay, bx, by, px, py):
    return F(F(F(bx - ax) * F(py - ay)) - F(F(by - ay) * F(px - ax)))


def _edge_grid(ax, ay, bx, by, px, py):
    return ((bx - ax) * (py - ay)).astype(F) - ((by - ay) * (px - ax)).astype(F)


def render(corners, view, proj, height, width, bottom_up=True, background=(0, 0, 0, 255),
           light=(0.3, 0.5, 1.0), colors=None, far_depth=np.inf):
    """corners: (ncubes, 8, 3) float64 world coordinates.  Returns (image uint8 (H,W,4), depth f32 (H,W))."""
    corners = np.asarray(corners, dtype=np.float64).reshape(-1, 3)
    V, P = np.asarray(view, dtype=np.float64), np.asarray(proj, dtype=np.float64)
    perspective = P[3, 2] != 0.0
    n = len(corners)
    sx, sy, sd = np.zeros(n, F), np.zeros(n, F), np.zeros(n, F)
    cam = np.zeros((n, 3), F)
    ok = np.zeros(n, bool)
    for i, (x, y, z) in enumerate(corners):
        vx, vy, vz, vw = (_dot4(V[r], x, y, z, 1.0) for r in range(4))
        qx, qy, qw = _dot4(P[0], vx, vy, vz, vw), _dot4(P[1], vx, vy, vz, vw), _dot4(P[3], vx, vy, vz, vw)
        depth = -vz
        ok[i] = qw > 1e-9 and depth > 1e-9
        px = py = 0.0
        if ok[i]:
            nx, ny = qx / qw, qy / qw
            px = ((nx + 1.0) * 0.5) * float(width)
            py = (1.0 - (ny + 1.0) * 0.5) * float(height)
        sx[i], sy[i], sd[i] = F(px), F(py), F(depth)
        cam[i] = (F(vx), F(vy), F(vz))
    lx, ly, lz = (F(v) for v in light)
    ll = np.sqrt(F(F(F(lx * lx) + F(ly * ly)) + F(lz * lz)))

    ys, xs = np.meshgrid(np.arange(height, dtype=F) + F(0.5), np.arange(width, dtype=F) + F(0.5), indexing="ij")
    best = np.full((height, width), np.finfo(F).max, dtype=F)
    img = np.empty((height, width, 4), np.uint8)
    img[...] = np.asarray(background, np.uint8)
    ncubes = n // 8
    for cube in range(ncubes):
        for k in range(12):
            i0, i1, i2 = (cube * 8 + TRIANGLES[k][j] for j in range(3))
            if not (ok[i0] and ok[i1] and ok[i2]):
                continue
            x0, y0, x1, y1, x2, y2 = sx[i0], sy[i0], sx[i1], sy[i1], sx[i2], sy[i2]
            area = _edge(x0, y0, x1, y1, x2, y2)
            if area == 0:
                continue
            if perspective:
                q0, q1, q2 = F(F(1) / sd[i0]), F(F(1) / sd[i1]), F(F(1) / sd[i2])
            else:
                q0, q1, q2 = sd[i0], sd[i1], sd[i2]
            e1, e2 = cam[i1] - cam[i0], cam[i2] - cam[i0]
            nx = F(F(e1[1] * e2[2]) - F(e1[2] * e2[1]))
            ny = F(F(e1[2] * e2[0]) - F(e1[0] * e2[2]))
            nz = F(F(e1[0] * e2[1]) - F(e1[1] * e2[0]))
            nl = np.sqrt(F(F(F(nx * nx) + F(ny * ny)) + F(nz * nz)))
            s = F(0)
            if nl > 0 and ll > 0:
                d = F(F(F(nx * lx) + F(ny * ly)) + F(nz * lz))
                s = np.abs(F(d / F(nl * ll)))
            shade = F(F(0.25) + F(F(0.75) * s))
            base = PALETTE[cube & 7] if colors is None else np.asarray(colors, np.float32)[cube, :3]
            rgb = np.minimum((base * shade).astype(F), F(255)).astype(np.uint8)   # truncating cast
            w0 = _edge_grid(x1, y1, x2, y2, xs, ys)
            w1 = _edge_grid(x2, y2, x0, y0, xs, ys)
            w2 = _edge_grid(x0, y0, x1, y1, xs, ys)
            if area < 0:
                w0, w1, w2 = -w0, -w1, -w2
            inside = (w0 >= 0) & (w1 >= 0) & (w2 >= 0)
            if not inside.any():
                continue
            aa = np.abs(area)
            with np.errstate(divide="ignore", invalid="ignore"):
                b0, b1, b2 = (w0 / aa).astype(F), (w1 / aa).astype(F), (w2 / aa).astype(F)
                q = ((b0 * q0).astype(F) + (b1 * q1).astype(F)).astype(F) + (b2 * q2).astype(F)
                d = (F(1) / q).astype(F) if perspective else q
            win = inside & (d < best)
            best[win] = d[win]
            img[win, :3] = rgb
            img[win, 3] = 255
    depth = np.where(best == np.finfo(F).max, F(far_depth), best).astype(F)
    if bottom_up:
        img, depth = img[::-1], depth[::-1]
    return np.ascontiguousarray(img), np.ascontiguousarray(depth)
