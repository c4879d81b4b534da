// K2: batched annotation projection world -> camera -> clip -> NDC -> pixel for sm_100a.
//
// Restates, for [batch][npoints] points in one launch, the float64 numpy arithmetic of
// pkg_blender/blendtorch/btb/camera.py:84-136 and btb/utils.py:112-121 (hom / dehom):
//   with depth:    cam = [x y z 1] @ V^T ; d = -cam.z ; clip = cam @ P^T ; ndc = clip.xyz / clip.w
//   without depth: clip = [x y z 1] @ (P @ V)^T ; ndc = clip.xyz / clip.w
//   pixel:         xy = (ndc.xy + 1) * 0.5 ; if upper-left: y = 1 - y ; xy *= [w, h]
// The path is tiny (64 points per frame in BASELINE config 3) and latency-bound: one thread per
// point, matrices read through the read-only path (every thread of a frame reads the same 32
// doubles, L1-resident), one launch for the whole batch.  Products are accumulated left to right
// with separate multiply and add (no FMA contraction) so results track numpy's; numpy itself may
// use FMA inside BLAS, hence the tests allow 1e-12 relative, far inside the 1e-6 budget.
// The per-object pixel bbox is a min/max over `corners_per_object` consecutive lanes by shuffle.
#include <cuda_runtime.h>

#include "common.h"

namespace {

struct K2Args {
  bt_project_params p;
};

__device__ __forceinline__ double dot4(const double* __restrict__ m, double x, double y, double z, double w) {
  // (x*m0 + y*m1) + z*m2 + w*m3, left to right like a scalar loop over k
  double acc = __dmul_rn(x, __ldg(m + 0));
  acc = __dadd_rn(acc, __dmul_rn(y, __ldg(m + 1)));
  acc = __dadd_rn(acc, __dmul_rn(z, __ldg(m + 2)));
  acc = __dadd_rn(acc, __dmul_rn(w, __ldg(m + 3)));
  return acc;
}

__global__ void __launch_bounds__(256) k2_project(K2Args args) {
  const bt_project_params& p = args.p;
  const long long total = (long long)p.batch * p.npoints;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < total;
  double px = 0.0, py = 0.0;
  if (live) {
    const int b = (int)(i / p.npoints);
    const double* pt = p.points + i * 3;
    const double x = __ldg(pt), y = __ldg(pt + 1), z = __ldg(pt + 2);
    const double* V = p.view + (p.view_per_frame ? (size_t)b * 16 : 0);
    const double* P = p.proj + (p.proj_per_frame ? (size_t)b * 16 : 0);
    double cx, cy, cz, cw;
    if (p.input_is_ndc) {
      cx = x; cy = y; cz = z; cw = 1.0;
    } else if (p.with_depth) {
      const double vx = dot4(V + 0, x, y, z, 1.0), vy = dot4(V + 4, x, y, z, 1.0);
      const double vz = dot4(V + 8, x, y, z, 1.0), vw = dot4(V + 12, x, y, z, 1.0);
      if (p.depth_out) p.depth_out[i] = -vz;
      cx = dot4(P + 0, vx, vy, vz, vw);
      cy = dot4(P + 4, vx, vy, vz, vw);
      cz = dot4(P + 8, vx, vy, vz, vw);
      cw = dot4(P + 12, vx, vy, vz, vw);
    } else {
      // m = P @ V, then one product (camera.py:111-112)
      double m[16];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          double acc = __dmul_rn(__ldg(P + r * 4 + 0), __ldg(V + 0 * 4 + c));
          acc = __dadd_rn(acc, __dmul_rn(__ldg(P + r * 4 + 1), __ldg(V + 1 * 4 + c)));
          acc = __dadd_rn(acc, __dmul_rn(__ldg(P + r * 4 + 2), __ldg(V + 2 * 4 + c)));
          acc = __dadd_rn(acc, __dmul_rn(__ldg(P + r * 4 + 3), __ldg(V + 3 * 4 + c)));
          m[r * 4 + c] = acc;
        }
      auto row = [&](int r) {
        double acc = __dmul_rn(x, m[r * 4 + 0]);
        acc = __dadd_rn(acc, __dmul_rn(y, m[r * 4 + 1]));
        acc = __dadd_rn(acc, __dmul_rn(z, m[r * 4 + 2]));
        acc = __dadd_rn(acc, m[r * 4 + 3]);
        return acc;
      };
      cx = row(0); cy = row(1); cz = row(2); cw = row(3);
    }
    const double nx = p.input_is_ndc ? cx : __ddiv_rn(cx, cw);
    const double ny = p.input_is_ndc ? cy : __ddiv_rn(cy, cw);
    const double nz = p.input_is_ndc ? cz : __ddiv_rn(cz, cw);
    if (p.ndc_out) {
      p.ndc_out[i * 3 + 0] = nx;
      p.ndc_out[i * 3 + 1] = ny;
      p.ndc_out[i * 3 + 2] = nz;
    }
    px = __dmul_rn(__dadd_rn(nx, 1.0), 0.5);
    py = __dmul_rn(__dadd_rn(ny, 1.0), 0.5);
    if (p.origin_upper_left) py = __dsub_rn(1.0, py);
    px = __dmul_rn(px, (double)p.width);
    py = __dmul_rn(py, (double)p.height);
    if (p.xy_out) {
      p.xy_out[i * 2 + 0] = px;
      p.xy_out[i * 2 + 1] = py;
    }
  }
  if (p.bbox_out) {
    // groups of corners_per_object consecutive points never straddle a warp: the launcher
    // requires a power of two <= 32, and npoints % corners_per_object == 0.
    double lox = px, loy = py, hix = px, hiy = py;
    const unsigned full = 0xffffffffu;
    for (int d = p.corners_per_object >> 1; d > 0; d >>= 1) {
      auto sh = [&](double v) {
        int lo = __shfl_xor_sync(full, __double2loint(v), d);
        int hi = __shfl_xor_sync(full, __double2hiint(v), d);
        return __hiloint2double(hi, lo);
      };
      lox = fmin(lox, sh(lox)); loy = fmin(loy, sh(loy));
      hix = fmax(hix, sh(hix)); hiy = fmax(hiy, sh(hiy));
    }
    if (live && (i % p.corners_per_object) == 0) {
      double* o = p.bbox_out + (i / p.corners_per_object) * 4;
      o[0] = lox; o[1] = loy; o[2] = hix; o[3] = hiy;
    }
  }
}

}  // namespace

extern "C" int bt_project_launch(const bt_project_params* p, void* stream) {
  if (!p || !p->points || (!p->input_is_ndc && (!p->view || !p->proj)))
    return bt::fail(BT_E_INVALID, "bt_project_launch: null pointer");
  if (p->batch < 0 || p->npoints < 0) return bt::fail(BT_E_INVALID, "negative extent");
  if (p->depth_out && !p->with_depth) return bt::fail(BT_E_INVALID, "depth_out needs with_depth");
  if (p->bbox_out) {
    int k = p->corners_per_object;
    if (k < 1 || k > 32 || (k & (k - 1)) || p->npoints % k)
      return bt::fail(BT_E_INVALID, "corners_per_object must be a power of two <= 32 dividing npoints");
  }
  long long total = (long long)p->batch * p->npoints;
  if (total == 0) return BT_OK;
  K2Args a;
  a.p = *p;
  int grid = (int)((total + 255) / 256);
  k2_project<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  BT_CUDA_TRY(cudaGetLastError());
  return BT_OK;
}
