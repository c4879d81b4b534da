// K3: synthetic Cube-scene frame source on the GPU.
//
// Stands in, for benchmarks and self-checks, for the one step of the reference that needs a
// graphics API: pkg_blender/blendtorch/btb/offscreen.py:76-92 (Eevee draw into a GPUOffScreen and
// glGetTexImage read-back).  Nothing here uses or imitates a graphics stack: it is a plain CUDA
// edge-function rasteriser for boxes (12 triangles each) with a per-pixel depth test and flat
// shading, writing RGBA8 frames -- optionally bottom-up, the order a GL read-back delivers and the
// reason offscreen.py:95-96 flips -- and linear camera-space depth.  Vertices go through exactly
// the arithmetic of K2 (camera.py:84-136), so projected annotation corners land on the rasterised
// silhouettes, which gives an end-to-end consistency check of frames against annotations and a
// producer that can feed K1 from HBM at rates no PCIe link allows.
//
// Arithmetic is written with explicit single-rounding intrinsics (no FMA contraction) in a fixed
// order, so the numpy restatement in oracle/raster.py reproduces every pixel bit for bit.
#include <cuda_runtime.h>

#include <cfloat>

#include "common.h"

namespace {

constexpr int kTile = 16;     // pixels per side of a tile: one pixel per thread
// kRegion (template parameter): tiles per side of the region a CTA rasterises.  The projection and triangle
// set-up below are done once per CTA, so a 4x4 region shares them between 16 tiles; small batches use 2x2
// regions to keep enough CTAs in flight.
constexpr int kMaxCubes = 32;
constexpr int kTrisPerCube = 12;

// corner order: (x, y, z) in {+1,-1}^3 with z fastest, then y, then x (oracle/camera.py
// default_cube_vertices, tests/test_camera.py:19-40).  Two triangles per face.
__constant__ unsigned char kTri[kTrisPerCube][3] = {
    {0, 1, 3}, {0, 3, 2},  // x = +1
    {4, 6, 7}, {4, 7, 5},  // x = -1
    {0, 4, 5}, {0, 5, 1},  // y = +1
    {2, 3, 7}, {2, 7, 6},  // y = -1
    {0, 2, 6}, {0, 6, 4},  // z = +1
    {1, 5, 7}, {1, 7, 3},  // z = -1
};

__constant__ unsigned char kPalette[8][3] = {{230, 60, 60},  {60, 200, 80},  {70, 110, 240}, {240, 200, 60},
                                             {200, 80, 220}, {60, 210, 210}, {250, 140, 50}, {180, 180, 180}};

struct Tri {
  float x0, y0, x1, y1, x2, y2;
  float q0, q1, q2;  // per-vertex depth (orthographic) or reciprocal depth (perspective)
  float area;
  unsigned int rgba;
};

__device__ __forceinline__ double dot4(const double* m, double x, double y, double z, double w) {
  double acc = __dmul_rn(x, m[0]);
  acc = __dadd_rn(acc, __dmul_rn(y, m[1]));
  acc = __dadd_rn(acc, __dmul_rn(z, m[2]));
  acc = __dadd_rn(acc, __dmul_rn(w, m[3]));
  return acc;
}

__device__ __forceinline__ float edge(float ax, float ay, float bx, float by, float px, float py) {
  // (bx-ax)*(py-ay) - (by-ay)*(px-ax), each operation rounded once
  return __fsub_rn(__fmul_rn(__fsub_rn(bx, ax), __fsub_rn(py, ay)), __fmul_rn(__fsub_rn(by, ay), __fsub_rn(px, ax)));
}

template <int kRegion>
__global__ void __launch_bounds__(kTile* kTile) k3_raster(bt_raster_params p) {
  __shared__ float sx[kMaxCubes * 8], sy[kMaxCubes * 8], sd[kMaxCubes * 8];
  __shared__ float cx[kMaxCubes * 8], cy[kMaxCubes * 8], cz[kMaxCubes * 8];
  __shared__ unsigned char ok[kMaxCubes * 8];
  __shared__ Tri tris[kMaxCubes * kTrisPerCube];
  __shared__ float bminx[kMaxCubes * kTrisPerCube], bmaxx[kMaxCubes * kTrisPerCube];
  __shared__ float bminy[kMaxCubes * kTrisPerCube], bmaxy[kMaxCubes * kTrisPerCube];
  __shared__ unsigned char live[kMaxCubes * kTrisPerCube];
  __shared__ unsigned short region_order[kMaxCubes * kTrisPerCube];   // triangles that can touch the region
  __shared__ unsigned short order[kMaxCubes * kTrisPerCube];          // ... and the current tile
  __shared__ int nregion, nlive;

  const int b = blockIdx.z;
  const int tid = threadIdx.y * kTile + threadIdx.x;
  const int ncorners = p.ncubes * 8;
  const double* V = p.view + (p.view_per_frame ? (size_t)b * 16 : 0);
  const double* P = p.proj + (p.proj_per_frame ? (size_t)b * 16 : 0);
  const bool perspective = P[14] != 0.0;  // row 3, column 2: -1 for a perspective matrix

  // phase A: project the corners exactly as K2 does (float64), keep float32 copies
  for (int i = tid; i < ncorners; i += kTile * kTile) {
    const double* c = p.corners + ((size_t)b * ncorners + i) * 3;
    const double x = c[0], y = c[1], z = c[2];
    const double vx = dot4(V + 0, x, y, z, 1.0), vy = dot4(V + 4, x, y, z, 1.0);
    const double vz = dot4(V + 8, x, y, z, 1.0), vw = dot4(V + 12, x, y, z, 1.0);
    const double qx = dot4(P + 0, vx, vy, vz, vw), qy = dot4(P + 4, vx, vy, vz, vw);
    const double qw = dot4(P + 12, vx, vy, vz, vw);
    const double depth = -vz;
    const bool good = qw > 1e-9 && depth > 1e-9;
    double px = 0.0, py = 0.0;
    if (good) {
      const double nx = __ddiv_rn(qx, qw), ny = __ddiv_rn(qy, qw);
      px = __dmul_rn(__dmul_rn(__dadd_rn(nx, 1.0), 0.5), (double)p.width);
      py = __dmul_rn(__dsub_rn(1.0, __dmul_rn(__dadd_rn(ny, 1.0), 0.5)), (double)p.height);
    }
    sx[i] = (float)px; sy[i] = (float)py; sd[i] = (float)depth;
    cx[i] = (float)vx; cy[i] = (float)vy; cz[i] = (float)vz;
    ok[i] = good;
  }
  __syncthreads();

  // phase B: triangle set-up, tile culling, flat shading (one thread per triangle; live[t] says
  // whether triangle t can touch this tile at all)
  const float rx0 = (float)(blockIdx.x * kTile * kRegion), ry0 = (float)(blockIdx.y * kTile * kRegion);
  const float rx1 = rx0 + (float)(kTile * kRegion), ry1 = ry0 + (float)(kTile * kRegion);
  const float ll = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.light[0], p.light[0]), __fmul_rn(p.light[1], p.light[1])),
                                        __fmul_rn(p.light[2], p.light[2])));
  for (int t = tid; t < p.ncubes * kTrisPerCube; t += kTile * kTile) {
    live[t] = 0;
    const int cube = t / kTrisPerCube, k = t - cube * kTrisPerCube;
    const int i0 = cube * 8 + kTri[k][0], i1 = cube * 8 + kTri[k][1], i2 = cube * 8 + kTri[k][2];
    if (!(ok[i0] && ok[i1] && ok[i2])) continue;  // a corner behind the camera: the triangle is dropped
    Tri tr;
    tr.x0 = sx[i0]; tr.y0 = sy[i0]; tr.x1 = sx[i1]; tr.y1 = sy[i1]; tr.x2 = sx[i2]; tr.y2 = sy[i2];
    tr.area = edge(tr.x0, tr.y0, tr.x1, tr.y1, tr.x2, tr.y2);
    if (tr.area == 0.0f) continue;
    const float minx = fminf(tr.x0, fminf(tr.x1, tr.x2)), maxx = fmaxf(tr.x0, fmaxf(tr.x1, tr.x2));
    const float miny = fminf(tr.y0, fminf(tr.y1, tr.y2)), maxy = fmaxf(tr.y0, fmaxf(tr.y1, tr.y2));
    if (maxx < rx0 || minx > rx1 || maxy < ry0 || miny > ry1) continue;  // conservative: cannot touch this region
    bminx[t] = minx; bmaxx[t] = maxx; bminy[t] = miny; bmaxy[t] = maxy;
    if (perspective) {
      tr.q0 = __fdiv_rn(1.0f, sd[i0]); tr.q1 = __fdiv_rn(1.0f, sd[i1]); tr.q2 = __fdiv_rn(1.0f, sd[i2]);
    } else {
      tr.q0 = sd[i0]; tr.q1 = sd[i1]; tr.q2 = sd[i2];
    }
    // face normal in camera space, |n . l| lighting (two-sided)
    const float e1x = __fsub_rn(cx[i1], cx[i0]), e1y = __fsub_rn(cy[i1], cy[i0]), e1z = __fsub_rn(cz[i1], cz[i0]);
    const float e2x = __fsub_rn(cx[i2], cx[i0]), e2y = __fsub_rn(cy[i2], cy[i0]), e2z = __fsub_rn(cz[i2], cz[i0]);
    const float nx = __fsub_rn(__fmul_rn(e1y, e2z), __fmul_rn(e1z, e2y));
    const float ny = __fsub_rn(__fmul_rn(e1z, e2x), __fmul_rn(e1x, e2z));
    const float nz = __fsub_rn(__fmul_rn(e1x, e2y), __fmul_rn(e1y, e2x));
    const float nl = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nx, nx), __fmul_rn(ny, ny)), __fmul_rn(nz, nz)));
    float s = 0.0f;
    if (nl > 0.0f && ll > 0.0f) {
      const float d = __fadd_rn(__fadd_rn(__fmul_rn(nx, p.light[0]), __fmul_rn(ny, p.light[1])), __fmul_rn(nz, p.light[2]));
      s = fabsf(__fdiv_rn(d, __fmul_rn(nl, ll)));
    }
    const float shade = __fadd_rn(0.25f, __fmul_rn(0.75f, s));
    unsigned int rgba = 0xff000000u;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float base = p.colors ? (float)p.colors[cube * 4 + c] : (float)kPalette[cube & 7][c];
      const float v = fminf(__fmul_rn(base, shade), 255.0f);
      rgba |= (unsigned int)(unsigned char)v << (8 * c);  // truncating, like the reference's uint8 casts
    }
    tr.rgba = rgba;
    tris[t] = tr;
    live[t] = 1;
  }
  __syncthreads();
  // order-preserving compaction of the triangles that can touch this region (most regions see a few
  // of the 12*ncubes): warp 0 scans the flags 32 at a time with ballots
  if (tid < 32) {
    int base = 0;
    for (int t0 = 0; t0 < p.ncubes * kTrisPerCube; t0 += 32) {
      const int t = t0 + tid;
      const bool on = t < p.ncubes * kTrisPerCube && live[t];
      const unsigned m = __ballot_sync(0xffffffffu, on);
      if (on) region_order[base + __popc(m & ((1u << tid) - 1u))] = (unsigned short)t;
      base += __popc(m);
    }
    if (tid == 0) nregion = base;
  }
  __syncthreads();

  const unsigned int background = (unsigned int)p.background[0] | ((unsigned int)p.background[1] << 8) |
                                  ((unsigned int)p.background[2] << 16) | ((unsigned int)p.background[3] << 24);
  for (int tile = 0; tile < kRegion * kRegion; ++tile) {
    const int tile_x = blockIdx.x * kRegion + (tile % kRegion), tile_y = blockIdx.y * kRegion + (tile / kRegion);
    if (tile_x * kTile >= p.width || tile_y * kTile >= p.height) continue;   // (uniform: the whole CTA skips)
    // the region's triangles that can touch this tile, same order, same conservative test
    const float tx0 = (float)(tile_x * kTile), ty0 = (float)(tile_y * kTile);
    const float tx1 = tx0 + (float)kTile, ty1 = ty0 + (float)kTile;
    if (tid < 32) {
      int base = 0;
      for (int j0 = 0; j0 < nregion; j0 += 32) {
        const int j = j0 + tid;
        bool on = false;
        unsigned short t = 0;
        if (j < nregion) {
          t = region_order[j];
          on = !(bmaxx[t] < tx0 || bminx[t] > tx1 || bmaxy[t] < ty0 || bminy[t] > ty1);
        }
        const unsigned m = __ballot_sync(0xffffffffu, on);
        if (on) order[base + __popc(m & ((1u << tid) - 1u))] = t;
        base += __popc(m);
      }
      if (tid == 0) nlive = base;
    }
    __syncthreads();

    // phase C: one pixel per thread, nearest depth wins; ties keep the lower triangle index
    const int px = tile_x * kTile + threadIdx.x, py = tile_y * kTile + threadIdx.y;
    if (px < p.width && py < p.height) {
      const float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
      float best = FLT_MAX;
      unsigned int color = background;
      for (int j = 0; j < nlive; ++j) {
        const Tri& tr = tris[order[j]];
        float w0 = edge(tr.x1, tr.y1, tr.x2, tr.y2, fx, fy);
        float w1 = edge(tr.x2, tr.y2, tr.x0, tr.y0, fx, fy);
        float w2 = edge(tr.x0, tr.y0, tr.x1, tr.y1, fx, fy);
        if (tr.area < 0.0f) { w0 = -w0; w1 = -w1; w2 = -w2; }
        if (w0 < 0.0f || w1 < 0.0f || w2 < 0.0f) continue;
        const float aa = fabsf(tr.area);
        const float b0 = __fdiv_rn(w0, aa), b1 = __fdiv_rn(w1, aa), b2 = __fdiv_rn(w2, aa);
        float q = __fadd_rn(__fadd_rn(__fmul_rn(b0, tr.q0), __fmul_rn(b1, tr.q1)), __fmul_rn(b2, tr.q2));
        const float d = perspective ? __fdiv_rn(1.0f, q) : q;
        if (d < best) { best = d; color = tr.rgba; }
      }
      const int row = p.bottom_up ? p.height - 1 - py : py;
      const size_t o = ((size_t)b * p.height + row) * p.width + px;
      reinterpret_cast<unsigned int*>(p.image_out)[o] = color;
      if (p.depth_out) p.depth_out[o] = best == FLT_MAX ? p.far_depth : best;
    }
    __syncthreads();   // the tile list is rewritten for the next tile
  }
}

}  // namespace

extern "C" int bt_raster_launch(const bt_raster_params* p, void* stream) {
  if (!p || !p->corners || !p->view || !p->proj || !p->image_out)
    return bt::fail(BT_E_INVALID, "bt_raster_launch: null pointer");
  if (p->batch < 0 || p->height < 1 || p->width < 1) return bt::fail(BT_E_INVALID, "bad extents");
  if (p->ncubes < 0 || p->ncubes > kMaxCubes) return bt::fail(BT_E_TOO_LARGE, "at most %d boxes per frame, got %d", kMaxCubes, p->ncubes);
  if (reinterpret_cast<uintptr_t>(p->image_out) % 4) return bt::fail(BT_E_INVALID, "image_out must be 4-byte aligned");
  if (p->batch == 0) return BT_OK;
  if (p->batch > 65535) return bt::fail(BT_E_TOO_LARGE, "at most 65535 frames per launch");
  auto regions = [&](int side) { return (uint64_t)((p->width + side - 1) / side) * ((p->height + side - 1) / side) * p->batch; };
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (regions(kTile * 4) >= 2048) {   // enough 64x64 regions to fill the GPU several times over
    constexpr int kSide = kTile * 4;
    dim3 grid((p->width + kSide - 1) / kSide, (p->height + kSide - 1) / kSide, p->batch);
    k3_raster<4><<<grid, dim3(kTile, kTile), 0, st>>>(*p);
  } else {
    constexpr int kSide = kTile * 2;
    dim3 grid((p->width + kSide - 1) / kSide, (p->height + kSide - 1) / kSide, p->batch);
    k3_raster<2><<<grid, dim3(kTile, kTile), 0, st>>>(*p);
  }
  BT_CUDA_TRY(cudaGetLastError());
  return BT_OK;
}
