// K1: fused uint8 frame preprocessing for sm_100a (B200).
//
// One launch per batch does what the reference spreads over numpy and torch on the CPU
// (see include/bt_b200.h for the cited lines): vertical flip, truncating gamma on RGB, float
// conversion, (x-mean)/std with IEEE fp32 subtract and divide, HWC->CHW, and the stack into a
// batch.  The kernel is a pure HBM stream (no reuse): 1 byte read and sizeof(T) bytes written per
// channel value, so everything here is about wide coalesced accesses, enough bytes in flight and
// as few instructions per byte as possible.
//
// Because the input is uint8, every output value is one of 256 per channel.  Each CTA first
// builds that table in shared memory *with the reference's own expression* (gamma table entry,
// then __fsub_rn/__fdiv_rn, then the output cast), so the per-pixel work is one shared-memory
// lookup per byte and the results are exact by construction -- no FMA contraction, no reciprocal
// multiply (which is 1.5e-5 relative off).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>

#include "common.h"

namespace {

constexpr int kThreads = 256;
// Grid cap in CTAs per SM.  Measured on B200 (profiles/r1_k1_tuning.txt): a grid that covers the
// batch in one pass per thread beats a persistent grid-stride loop (0.97 vs 0.85 of the measured
// HBM peak at 64 x 1024^2 RGBA), so the cap only guards against absurd grids.
constexpr int kDefaultCtasPerSm = 128;
constexpr int kMaxC = 4;

struct K1Args {
  const uint8_t* src;
  void* dst;
  uint32_t B, H, W, C, Cout;
  uint32_t flip, use_gamma, use_norm;
  uint32_t flags;             // bit 0: k1_planar issues its first loads before building the table
  uint64_t src_image_stride;  // bytes
  const long long* src_index; // optional [B]: which source image feeds batch item b (nullptr: b itself)
  float mean[4];
  float stdev[4];
  uint8_t gamma_lut[256];
  double* partials;       // [gridDim.x][kMaxC][2] when stats are on
  double* stats_out;      // [Cout][2]
  // Optional (tuning bit 19, an A/B that lost: see launch_checked): the 4 x 256 possible outputs made on
  // the host with the reference's expression and carried as kernel parameters, for the CTAs to copy
  // instead of computing them.
  uint32_t table[kMaxC * 256];   // kMaxC*256 values of the output type, packed
  uint32_t has_table;
};

// Byte offset of the source image that feeds batch item `b`: dense batches read image b, gathered
// batches (frames resident in HBM, picked by a sampler) read image src_index[b].  The test is on a
// kernel parameter, i.e. uniform; the index array is a few hundred bytes and stays in L1.
__device__ __forceinline__ uint64_t src_image_off(const K1Args& a, uint32_t b) {
  const uint64_t img = a.src_index ? (uint64_t)__ldg(a.src_index + b) : (uint64_t)b;
  return img * a.src_image_stride;
}

// ---- output-type helpers -----------------------------------------------------------------
template <typename T> struct OutT;
template <> struct OutT<float> {
  static __device__ __forceinline__ float make(float o, uint8_t) { return o; }
  static __device__ __forceinline__ float as_float(float v) { return v; }
};
template <> struct OutT<__nv_bfloat16> {
  static __device__ __forceinline__ __nv_bfloat16 make(float o, uint8_t) { return __float2bfloat16_rn(o); }
  static __device__ __forceinline__ float as_float(__nv_bfloat16 v) { return __bfloat162float(v); }
};
template <> struct OutT<uint8_t> {
  static __device__ __forceinline__ uint8_t make(float, uint8_t g) { return g; }
  static __device__ __forceinline__ float as_float(uint8_t v) { return (float)v; }
};

// Build the per-channel value table with the reference's expression.
template <typename T>
__device__ __forceinline__ void build_lut(const K1Args& a, T* lut /* [kMaxC][256] */) {
  static_assert(kThreads == 256, "one thread per table index");
  if (a.has_table) {
    // host-built table: thread t copies words t, t+256, ... (consecutive lanes, consecutive banks)
    constexpr int kWords = kMaxC * 256 * (int)sizeof(T) / 4;
    uint32_t* dst = reinterpret_cast<uint32_t*>(lut);
#pragma unroll
    for (int k = 0; k < kWords / kThreads; ++k) dst[threadIdx.x + k * kThreads] = a.table[threadIdx.x + k * kThreads];
    __syncthreads();
    return;
  }
  // Thread v makes entry v of every channel.  The gamma byte is read once, as the 32-bit word that
  // holds it (a warp touches 8 distinct words of the parameter bank instead of 32 distinct bytes),
  // and the channel loop is unrolled so that mean/stdev are plain constant operands.
  const uint32_t v = threadIdx.x;
  uint32_t g = v;
  if (a.use_gamma) {
    const uint32_t word = reinterpret_cast<const uint32_t*>(a.gamma_lut)[v >> 2];
    g = (word >> (8 * (v & 3))) & 0xffu;
  }
#pragma unroll
  for (int c = 0; c < kMaxC; ++c) {
    const uint32_t gc = c < 3 ? g : v;   // alpha is never gamma-corrected (offscreen.py:109-111)
    float o = (float)gc;
    if (a.use_norm) o = __fdiv_rn(__fsub_rn(o, a.mean[c]), a.stdev[c]);
    lut[c * 256 + v] = OutT<T>::make(o, (uint8_t)gc);
  }
  __syncthreads();
}

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// Store N consecutive T (N*sizeof(T) in {4, 8, 16, 32, 64}) with streaming (evict-first) hints.
template <typename T, int N>
__device__ __forceinline__ void st_stream(T* dst, const T (&v)[N]) {
  constexpr int bytes = N * (int)sizeof(T);
  static_assert(bytes == 4 || bytes == 8 || bytes % 16 == 0, "unsupported store width");
  const uint32_t* w = reinterpret_cast<const uint32_t*>(v);
  if constexpr (bytes == 4) {
    __stcs(reinterpret_cast<unsigned int*>(dst), w[0]);
  } else if constexpr (bytes == 8) {
    __stcs(reinterpret_cast<uint2*>(dst), make_uint2(w[0], w[1]));
  } else {
#pragma unroll
    for (int k = 0; k < bytes / 16; ++k)
      __stcs(reinterpret_cast<uint4*>(dst) + k, make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]));
  }
}

__device__ __forceinline__ uint32_t byte_of(const uint32_t* words, int k) {
  return (words[k >> 2] >> (8 * (k & 3))) & 0xffu;
}

// ---- deterministic per-batch statistics ----------------------------------------------------
struct StatAcc {
  double s[kMaxC], q[kMaxC];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int c = 0; c < kMaxC; ++c) s[c] = q[c] = 0.0;
  }
};

__device__ __forceinline__ double shfl_xor_f64(double v, int d) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  lo = __shfl_xor_sync(0xffffffffu, lo, d);
  hi = __shfl_xor_sync(0xffffffffu, hi, d);
  return __hiloint2double(hi, lo);
}

// ---- the fold of the per-CTA rows: its own small kernel -------------------------------------------
// The producing kernel only writes one row of partial sums per CTA; no fence, no arrival counter, no
// "last CTA" tail.  (Measured: with the fold inside the producer every CTA paid a __threadfence -- a
// wait for all of its streamed output to be acknowledged -- plus an atomic round trip before it
// could retire, which cost a one-pass grid more than the whole reduction is worth.)  The kernel
// boundary orders the rows before their readers, and programmatic dependent launch hides the launch.
constexpr int kFoldThreads = 1024;
constexpr int kFoldRowsPerCta = 1024;   // rows a CTA folds when the fold is spread over several CTAs
constexpr int kFoldDirectMax = 2048;    // up to this many rows a single CTA folds them all

// out[cta][pair] = sum over the CTA's rows, fixed order: 256 threads per channel pair take rows
// grp, grp+256, ... (16-byte loads, four in flight), a shuffle tree adds the lanes of each warp and
// one thread adds the pair's eight warps in order.
__global__ void __launch_bounds__(kFoldThreads) k1_fold(const double* rows, unsigned n, unsigned rows_per_cta,
                                                        double* out, unsigned out_pairs) {
  __shared__ double warp_sum[kMaxC][kFoldThreads / 32 / kMaxC][2];
  asm volatile("griddepcontrol.wait;" ::: "memory");   // the rows of the producer are complete and visible
  asm volatile("griddepcontrol.launch_dependents;");
  const unsigned first = blockIdx.x * rows_per_cta;
  const unsigned last = first + rows_per_cta < n ? first + rows_per_cta : n;
  const int pair = threadIdx.x >> 8, grp = threadIdx.x & 255, lane = threadIdx.x & 31, w = (threadIdx.x >> 5) & 7;
  constexpr int kInFlight = 4;
  double ts = 0.0, tq = 0.0;
  for (unsigned b0 = first + grp; b0 < last; b0 += 256 * kInFlight) {
    double2 r[kInFlight];
#pragma unroll
    for (int j = 0; j < kInFlight; ++j) {
      const unsigned b = b0 + 256 * j;
      r[j] = b < last ? __ldcg(reinterpret_cast<const double2*>(rows + ((size_t)b * kMaxC + pair) * 2)) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int j = 0; j < kInFlight; ++j) { ts += r[j].x; tq += r[j].y; }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    ts += shfl_xor_f64(ts, d);
    tq += shfl_xor_f64(tq, d);
  }
  if (lane == 0) { warp_sum[pair][w][0] = ts; warp_sum[pair][w][1] = tq; }
  __syncthreads();
  if (grp < 2 && (unsigned)pair < out_pairs) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kFoldThreads / 32 / kMaxC; ++k) t += warp_sum[pair][k][grp];
    out[((size_t)blockIdx.x * out_pairs + pair) * 2 + grp] = t;
  }
}

// warp shuffle -> shared -> one partial row per CTA (k1_fold adds the rows in index order).
__device__ void finish_stats(const K1Args& a, StatAcc& acc) {
  __shared__ double red[kThreads / 32][kMaxC][2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // Warp reduction of the 8 values (s0,q0,..,s3,q3) with 9 double shuffles instead of 40: every
  // round halves both the lanes that share a value and the values a lane still carries
  // (exchange with lane^16: keep 4, lane^8: keep 2, lane^4: keep 1), then two plain rounds.
  // Lane l ends up owning value ((l>>4)&1)*4 + ((l>>3)&1)*2 + ((l>>2)&1).  Fixed order: deterministic.
  {
    double v[8], w4[4], w2[2];
#pragma unroll
    for (int c = 0; c < kMaxC; ++c) { v[2 * c] = acc.s[c]; v[2 * c + 1] = acc.q[c]; }
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const double keep = h16 ? v[i + 4] : v[i], send = h16 ? v[i] : v[i + 4];
      w4[i] = keep + shfl_xor_f64(send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const double keep = h8 ? w4[i + 2] : w4[i], send = h8 ? w4[i] : w4[i + 2];
      w2[i] = keep + shfl_xor_f64(send, 8);
    }
    const double keep = h4 ? w2[1] : w2[0], send = h4 ? w2[0] : w2[1];
    double t = keep + shfl_xor_f64(send, 4);
    t += shfl_xor_f64(t, 2);
    t += shfl_xor_f64(t, 1);
    if ((lane & 3) == 0) {
      const int idx = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
      red[warp][idx >> 1][idx & 1] = t;
    }
  }
  __syncthreads();
  if (threadIdx.x < kMaxC * 2) {
    int c = threadIdx.x >> 1, k = threadIdx.x & 1;
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w][c][k];
    a.partials[((size_t)blockIdx.x * kMaxC + c) * 2 + k] = t;
  }
}

// ---- variant 1: generic scalar (any shape) -----------------------------------------------
template <typename T, bool CHW, bool STATS>
__global__ void __launch_bounds__(kThreads) k1_generic(const __grid_constant__ K1Args a) {
  __shared__ T lut[kMaxC * 256];
  build_lut<T>(a, lut);
  StatAcc acc;
  if (STATS) acc.clear();
  const uint64_t per_image = (uint64_t)a.H * a.W * a.Cout;
  const uint64_t total = per_image * a.B;
  T* dst = static_cast<T*>(a.dst);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t b = (uint32_t)(i / per_image);
    uint32_t r = (uint32_t)(i - (uint64_t)b * per_image);
    uint32_t c, y, x;
    if (CHW) { c = r / (a.H * a.W); uint32_t p = r - c * a.H * a.W; y = p / a.W; x = p - y * a.W; }
    else { uint32_t p = r / a.Cout; c = r - p * a.Cout; y = p / a.W; x = p - y * a.W; }
    uint32_t sy = a.flip ? a.H - 1 - y : y;
    uint8_t v = a.src[src_image_off(a, b) + ((uint64_t)sy * a.W + x) * a.C + c];
    T o = lut[c * 256 + v];
    dst[i] = o;
    if (STATS) {
      float f = OutT<T>::as_float(o);
#pragma unroll
      for (int cc = 0; cc < kMaxC; ++cc)
        if (cc == (int)c) { acc.s[cc] += f; acc.q[cc] += (double)f * f; }
    }
  }
  if (STATS) finish_stats(a, acc);
}

// ---- variant 2: register-vectorised, planar (CHW) output ---------------------------------------
// One "unit" is PPT consecutive pixels of one row: PPT*C bytes in (the widest aligned loads that
// divide it: 16, 8 or 4 bytes), PPT values out per plane (16-byte stores where PPT*sizeof(T)
// allows).  A warp therefore reads 32*PPT*C contiguous bytes and writes 32*PPT*sizeof(T)
// contiguous bytes per plane.  UNROLL units per thread are loaded before any is consumed, to keep
// UNROLL*PPT*C bytes in flight per thread.  Index arithmetic is 32-bit with host-computed
// multiply-shift division (the launcher keeps units per launch below 2^31).
struct FastDiv {
  uint32_t mul, shift, div;
};
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv& d) {
  return (__umulhi(n, d.mul) + n) >> d.shift;
}

struct PlanarGeom {
  FastDiv upr;   // units per row
  FastDiv rows;  // rows per image (H)
  uint32_t total_units;
};

template <int LW>
__device__ __forceinline__ void ld_words(const uint8_t* p, uint32_t* w) {
  if constexpr (LW == 16) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
  } else if constexpr (LW == 8) {
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p));
  } else {
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w[0]) : "l"(p));
  }
}

template <typename T, int C, int PPT, int UNROLL, bool STATS>
__global__ void __launch_bounds__(kThreads, (STATS || sizeof(T) == 2) ? 4 : 0) k1_planar(const __grid_constant__ K1Args a, PlanarGeom g) {
  constexpr int UB = PPT * C;                              // unit bytes
  constexpr int LW = (UB % 16 == 0) ? 16 : (UB % 8 == 0) ? 8 : 4;
  static_assert(UB % 4 == 0, "unit must be whole 32-bit words");
  constexpr int WORDS = UB / 4;
  __shared__ T lut[kMaxC * 256];
  // Programmatic dependent launch: let the next launch in the stream begin its own prologue now and
  // wait for the producer of `src` only when it is needed (both instructions are no-ops when the
  // launch carries no programmatic dependency).  By default the first batch of global loads is
  // issued *before* the table is built, so that the table build (4 IEEE divisions per thread and a
  // barrier) hides under the DRAM latency of those loads instead of preceding it: measured
  // 0.88 vs 0.83 of the HBM peak at batch 8 on four streams, 0.91 vs 0.82 for bf16 at 64 x 1024^2
  // (profiles/r1_k1_loads_first.txt).
  // (with statistics the dependent is the small fold kernel: it is released at the end instead, so that
  // its CTA does not sit on an SM for the whole run)
  if (!STATS) asm volatile("griddepcontrol.launch_dependents;");
  const bool loads_first = (a.flags & 1u) != 0;   // tuning: issue the first loads before the table build
  if (!loads_first) build_lut<T>(a, lut);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  StatAcc acc;
  if (STATS) acc.clear();

  const uint32_t plane = a.H * a.W;                        // < 2^30 (checked by the launcher)
  T* dst = static_cast<T*>(a.dst);
  const uint32_t stride = gridDim.x * blockDim.x;
  bool need_lut = loads_first;

  // (the first pass runs for every thread, live or not, so that the table build inside it is
  // reached by the whole CTA)
  for (uint32_t u0 = blockIdx.x * blockDim.x + threadIdx.x; need_lut || u0 < g.total_units; u0 += stride * UNROLL) {
    uint32_t in[UNROLL][WORDS];
    uint64_t out_off[UNROLL];
    bool live[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t u = u0 + j * stride;
      live[j] = u < g.total_units && u >= u0;              // (second test: wrap-around guard)
      if (live[j]) {
        const uint32_t row = fdiv(u, g.upr);
        const uint32_t ux = u - row * g.upr.div;
        const uint32_t b = fdiv(row, g.rows);
        const uint32_t y = row - b * g.rows.div;
        const uint32_t sy = a.flip ? a.H - 1 - y : y;
        const uint8_t* sp = a.src + src_image_off(a, b) + ((uint64_t)sy * a.W + (uint64_t)ux * PPT) * C;
#pragma unroll
        for (int l = 0; l < UB / LW; ++l) ld_words<LW>(sp + l * LW, &in[j][l * (LW / 4)]);
        out_off[j] = ((uint64_t)b * a.Cout) * plane + (uint64_t)y * a.W + (uint64_t)ux * PPT;
      }
    }
    if (need_lut) {  // the loads above are in flight while the table is built
      build_lut<T>(a, lut);
      need_lut = false;
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      if (!live[j]) continue;
      const uint32_t* words = in[j];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        if (c < (int)a.Cout) {
          T vals[PPT];
          float fs = 0.f, fq = 0.f;
#pragma unroll
          for (int p = 0; p < PPT; ++p) {
            const uint32_t v = byte_of(words, p * C + c);
            // uint8 output has no normalisation and gamma leaves alpha alone: channel 3 passes through
            if (sizeof(T) == 1 && c == 3) vals[p] = OutT<T>::make((float)v, (uint8_t)v);
            else vals[p] = lut[c * 256 + v];
            if (STATS) { float f = OutT<T>::as_float(vals[p]); fs += f; fq = fmaf(f, f, fq); }
          }
          st_stream<T, PPT>(dst + out_off[j] + (uint64_t)c * plane, vals);
          if (STATS) { acc.s[c] += (double)fs; acc.q[c] += (double)fq; }
        }
      }
    }
  }
  if (STATS) {
    asm volatile("griddepcontrol.launch_dependents;");
    finish_stats(a, acc);
  }
}

// ---- variant 3: bulk-async (TMA engine) staged input, planar output -----------------------------
// The A/B partner of k1_planar that BASELINE.json's north_star asks to be weighed: input rows are
// brought into shared memory by the bulk-copy engine (`cp.async.bulk` -> SASS UBLKCP) in 4 KiB
// segments through a kStages-deep mbarrier ring, issued by one elected thread of a persistent CTA;
// all threads then take 16 bytes from shared memory, look the values up and store planes exactly
// as k1_planar does.  Loads cost no registers and no LSU issue slots, and up to
// (kStages-1) * 4 KiB per CTA are in flight independent of occupancy.  RGBA, planar output, W % 4 == 0.
constexpr int kBulkStages = 4;
constexpr int kBulkSegBytes = kThreads * 16;   // 4096 B = 1024 RGBA pixels per segment

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct BulkGeom {
  FastDiv spr;   // segments per row
  FastDiv rows;  // rows per image (H)
  uint32_t total_segments;
};

template <typename T>
__global__ void __launch_bounds__(kThreads) k1_bulk(const __grid_constant__ K1Args a, BulkGeom g) {
  __shared__ __align__(128) uint8_t stage[kBulkStages][kBulkSegBytes];
  __shared__ __align__(8) uint64_t full[kBulkStages];
  __shared__ T lut[kMaxC * 256];
  const uint32_t row_bytes = a.W * 4;
  const uint32_t plane = a.H * a.W;
  T* dst = static_cast<T*>(a.dst);

  // where segment `seg` lives: source address, byte count, output offset of its first pixel
  auto locate = [&](uint32_t seg, const uint8_t*& src, uint32_t& bytes, uint64_t& out_off) {
    const uint32_t row = fdiv(seg, g.spr);
    const uint32_t sx = seg - row * g.spr.div;
    const uint32_t b = fdiv(row, g.rows);
    const uint32_t y = row - b * g.rows.div;
    const uint32_t sy = a.flip ? a.H - 1 - y : y;
    const uint32_t off = sx * kBulkSegBytes;
    bytes = row_bytes - off < (uint32_t)kBulkSegBytes ? row_bytes - off : (uint32_t)kBulkSegBytes;
    src = a.src + src_image_off(a, b) + (uint64_t)sy * row_bytes + off;
    out_off = ((uint64_t)b * a.Cout) * plane + (uint64_t)y * a.W + (off >> 2);
  };

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kBulkStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("griddepcontrol.launch_dependents;");
  build_lut<T>(a, lut);  // ends with __syncthreads(): the barriers are initialised for everybody
  asm volatile("griddepcontrol.wait;" ::: "memory");

  const uint32_t first = blockIdx.x, step = gridDim.x;
  // prologue: fill the ring
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kBulkStages; ++s) {
      const uint32_t seg = first + s * step;
      if (seg < g.total_segments) {
        const uint8_t* src; uint32_t bytes; uint64_t out_off;
        locate(seg, src, bytes, out_off);
        mbar_expect_tx(&full[s], bytes);
        bulk_load(stage[s], src, bytes, &full[s]);
      }
    }
  }
  uint32_t it = 0;
  for (uint32_t seg = first; seg < g.total_segments; seg += step, ++it) {
    const int s = it % kBulkStages;
    const uint32_t parity = (it / kBulkStages) & 1;
    const uint8_t* src; uint32_t bytes; uint64_t out_off;
    locate(seg, src, bytes, out_off);
    mbar_wait(&full[s], parity);
    const bool mine = threadIdx.x * 16u < bytes;
    uint4 in = make_uint4(0, 0, 0, 0);
    if (mine) in = *reinterpret_cast<const uint4*>(&stage[s][threadIdx.x * 16]);
    __syncthreads();  // everybody has taken its 16 bytes: the stage may be refilled
    if (threadIdx.x == 0) {
      const uint32_t next = seg + kBulkStages * step;
      if (next < g.total_segments) {
        const uint8_t* nsrc; uint32_t nbytes; uint64_t noff;
        locate(next, nsrc, nbytes, noff);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic reads before the async write
        mbar_expect_tx(&full[s], nbytes);
        bulk_load(stage[s], nsrc, nbytes, &full[s]);
      }
    }
    if (mine) {
      const uint32_t* words = reinterpret_cast<const uint32_t*>(&in);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c < (int)a.Cout) {
          T vals[4];
#pragma unroll
          for (int p = 0; p < 4; ++p) vals[p] = lut[c * 256 + byte_of(words, p * 4 + c)];
          st_stream<T, 4>(dst + out_off + threadIdx.x * 4 + (uint64_t)c * plane, vals);
        }
      }
    }
  }
}

// ---- variant 2b: register-vectorised, interleaved (HWC) output ---------------------------------
// Rows are streams of W*C bytes.  A unit is IB = 16/sizeof(T) input bytes of one row, so that every
// thread's result is exactly one 16-byte store and a warp writes 512 contiguous bytes per store
// instruction (fp32: 4 bytes in -> float4 out; bf16: 8 -> 16; u8: 16 -> 16).  Needs Cout == C.
// IDENT (u8 without gamma) skips the table: the kernel is then a flipping copy.
template <typename T, int UNROLL, bool STATS, bool IDENT, int CFIX>
__global__ void __launch_bounds__(kThreads) k1_interleaved(const __grid_constant__ K1Args a, PlanarGeom g) {
  constexpr int IB = 16 / (int)sizeof(T);
  constexpr int LW = IB;  // one load per unit
  __shared__ T lut[kMaxC * 256];
  // Loads before the table build, as in k1_planar, pay for 8- and 16-bit outputs (uint8 at 64 x 1024^2:
  // 0.96 vs 0.92 of peak, bf16 batch 8: 7.6 vs 8.7 us); fp32 measures the same or slightly worse, so it
  // keeps the table-first structure (and its code) unchanged.
  constexpr bool kCanLoadFirst = !IDENT && sizeof(T) < 4;
  const bool loads_first = kCanLoadFirst && (a.flags & 1u) != 0;
  if (!IDENT && !loads_first) build_lut<T>(a, lut);
  StatAcc acc;
  if (STATS) acc.clear();
  const uint32_t row_bytes = a.W * a.C;
  T* dst = static_cast<T*>(a.dst);
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t C = a.C;
  bool need_lut = loads_first;

  for (uint32_t u0 = blockIdx.x * blockDim.x + threadIdx.x; (kCanLoadFirst && need_lut) || u0 < g.total_units;
       u0 += stride * UNROLL) {
    uint32_t in[UNROLL][IB / 4];
    uint64_t out_off[UNROLL];
    uint32_t phase[UNROLL];
    bool live[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t u = u0 + j * stride;
      live[j] = u < g.total_units && u >= u0;
      if (live[j]) {
        const uint32_t row = fdiv(u, g.upr);
        const uint32_t ux = u - row * g.upr.div;
        const uint32_t b = fdiv(row, g.rows);
        const uint32_t y = row - b * g.rows.div;
        const uint32_t sy = a.flip ? a.H - 1 - y : y;
        ld_words<LW>(a.src + src_image_off(a, b) + (uint64_t)sy * row_bytes + (uint64_t)ux * IB, in[j]);
        out_off[j] = (uint64_t)row * row_bytes + (uint64_t)ux * IB;
        phase[j] = (ux * (uint32_t)IB) % C;
      }
    }
    if constexpr (kCanLoadFirst) {
      if (need_lut) {  // the loads above are in flight while the table is built
        build_lut<T>(a, lut);
        need_lut = false;
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      if (!live[j]) continue;
      T vals[IB];
      uint32_t c = CFIX ? 0u : phase[j];   // RGBA units start on a pixel boundary (IB is a multiple of 4)
#pragma unroll
      for (int k = 0; k < IB; ++k) {
        if (CFIX == 4) c = k & 3;
        const uint32_t v = byte_of(in[j], k);
        // uint8 output has no normalisation and gamma leaves alpha alone: channel 3 passes through
        if (IDENT || (sizeof(T) == 1 && CFIX == 4 && (k & 3) == 3)) vals[k] = OutT<T>::make((float)v, (uint8_t)v);
        else vals[k] = lut[c * 256 + v];
        if (STATS) {
          float f = OutT<T>::as_float(vals[k]);
#pragma unroll
          for (int cc = 0; cc < kMaxC; ++cc)
            if (cc == (int)c) { acc.s[cc] += f; acc.q[cc] += (double)f * f; }
        }
        if (CFIX == 0) c = (c + 1 == C) ? 0 : c + 1;
      }
      st_stream<T, IB>(dst + out_off[j], vals);
    }
  }
  if (STATS) finish_stats(a, acc);
}

// ---- variant 2c: one pixel per thread (any width) -----------------------------------------------
// For rows that are not a whole number of 16-byte units (odd widths): every thread takes one pixel
// (a 32-bit load for C == 4, byte loads otherwise) and writes C_out scalars; lanes are consecutive
// pixels, so planar stores are 128 contiguous bytes per warp instruction and loads are dense.
template <typename T, int C, bool CHW, int UNROLL, bool STATS>
__global__ void __launch_bounds__(kThreads) k1_pixel(const __grid_constant__ K1Args a, PlanarGeom g) {
  __shared__ T lut[kMaxC * 256];
  const bool loads_first = (a.flags & 1u) != 0;   // as in k1_planar: the table build hides under the first loads
  if (!loads_first) build_lut<T>(a, lut);
  StatAcc acc;
  if (STATS) acc.clear();
  const uint32_t plane = a.H * a.W;
  T* dst = static_cast<T*>(a.dst);
  const uint32_t stride = gridDim.x * blockDim.x;
  bool need_lut = loads_first;
  for (uint32_t u0 = blockIdx.x * blockDim.x + threadIdx.x; need_lut || u0 < g.total_units; u0 += stride * UNROLL) {
    uint32_t px[UNROLL];
    uint32_t bb[UNROLL], yy[UNROLL], xx[UNROLL];
    bool live[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t u = u0 + j * stride;
      live[j] = u < g.total_units && u >= u0;
      if (live[j]) {
        const uint32_t row = fdiv(u, g.upr);   // upr = W here: units are pixels
        const uint32_t x = u - row * g.upr.div;
        const uint32_t b = fdiv(row, g.rows);
        const uint32_t y = row - b * g.rows.div;
        const uint32_t sy = a.flip ? a.H - 1 - y : y;
        const uint8_t* sp = a.src + src_image_off(a, b) + ((uint64_t)sy * a.W + x) * C;
        if constexpr (C == 4) {
          ld_words<4>(sp, &px[j]);
        } else {
          uint32_t w = 0;
#pragma unroll
          for (int c = 0; c < C; ++c) w |= (uint32_t)__ldg(sp + c) << (8 * c);
          px[j] = w;
        }
        bb[j] = b; yy[j] = y; xx[j] = x;
      }
    }
    if (need_lut) {
      build_lut<T>(a, lut);
      need_lut = false;
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      if (!live[j]) continue;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        if (c < (int)a.Cout) {
          const T o = lut[c * 256 + ((px[j] >> (8 * c)) & 0xffu)];
          uint64_t off;
          if (CHW) off = ((uint64_t)bb[j] * a.Cout + c) * plane + (uint64_t)yy[j] * a.W + xx[j];
          else off = (((uint64_t)bb[j] * a.H + yy[j]) * a.W + xx[j]) * a.Cout + c;
          const T one[1] = {o};
          if constexpr (sizeof(T) == 4) st_stream<T, 1>(dst + off, one);   // written once, never re-read here
          else dst[off] = o;
          if (STATS) { float f = OutT<T>::as_float(o); acc.s[c] += f; acc.q[c] += (double)f * f; }
        }
      }
    }
  }
  if (STATS) finish_stats(a, acc);
}

// ---- variant 2d: RGBA of any width, planar output, aligned vector stores ---------------------------
// Widths that are not a multiple of 4 leave neither the input rows nor the output plane rows on
// 16-byte boundaries, and with a flip the two are out of step with each other.  k1_pixel answers with
// one pixel and four scalar stores per thread (instruction-bound: 0.5-0.6 of the HBM peak).  Here the
// work is cut along the *output* instead: the image of item b is H*W pixels back to back and so is each
// of its planes, so a unit is one aligned group of four plane elements -- one vector store per channel
// -- whatever row they belong to.  Plane c of item b starts ((b*Cout + c)*H*W) elements into the
// output, i.e. d_c = that mod 4 elements past a boundary: thread k of an image handles, for channel c,
// the elements [4k - d_c, 4k - d_c + 4), which lie among the seven pixels 4k-3 .. 4k+3.  Those are read
// with seven 32-bit loads (through L1: neighbouring threads share them), each from wherever its row
// is in the -- possibly flipped -- source.  Every element is stored exactly once; groups that stick out
// of the plane at either end store their valid elements one by one.
// One aligned group of four elements of plane c, taken from window slots OFF .. OFF+3 (OFF = 3 - d_c).
template <typename T, bool STATS, int OFF>
__device__ __forceinline__ void window_emit(const uint32_t (&px)[7], const T* lut, int c, T* out, int e0, uint32_t plane,
                                            double& s, double& q) {
  T vals[4];
  float fs = 0.f, fq = 0.f;
  const bool whole = e0 >= 0 && (uint32_t)e0 + 4 <= plane;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const uint32_t v = (px[OFF + t] >> (8 * c)) & 0xffu;
    if (sizeof(T) == 1 && c == 3) vals[t] = OutT<T>::make((float)v, (uint8_t)v);   // alpha passes through
    else vals[t] = lut[c * 256 + v];
    if (STATS && (whole || (e0 + t >= 0 && (uint32_t)(e0 + t) < plane))) {
      const float f = OutT<T>::as_float(vals[t]);
      fs += f;
      fq = fmaf(f, f, fq);
    }
  }
  if (whole) {
    st_stream<T, 4>(out, vals);
  } else {
#pragma unroll
    for (int t = 0; t < 4; ++t)
      if (e0 + t >= 0 && (uint32_t)(e0 + t) < plane) out[t] = vals[t];
  }
  if (STATS) { s += (double)fs; q += (double)fq; }
}

template <typename T, bool STATS>
__global__ void __launch_bounds__(kThreads) k1_window(const __grid_constant__ K1Args a, PlanarGeom g) {
  __shared__ T lut[kMaxC * 256];
  build_lut<T>(a, lut);
  StatAcc acc;
  if (STATS) acc.clear();
  const uint32_t plane = a.H * a.W;                        // < 2^30 (checked by the launcher)
  T* dst = static_cast<T*>(a.dst);
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t u = blockIdx.x * blockDim.x + threadIdx.x; u < g.total_units; u += stride) {
    const uint32_t b = fdiv(u, g.rows);                    // g.rows: groups per image
    const uint32_t k = u - b * g.rows.div;
    const uint8_t* img = a.src + src_image_off(a, b);
    // the seven pixels 4k-3 .. 4k+3 (those inside the image)
    uint32_t px[7];
    const int first = (int)(4 * k) - 3;
    const int i0 = first < 0 ? 0 : first;
    uint32_t y = fdiv((uint32_t)i0, g.upr), x = (uint32_t)i0 - y * g.upr.div;   // g.upr: W
    if (first >= 0 && x + 7 <= a.W && (uint32_t)first + 7 <= plane) {
      // the usual case: all seven in one row, i.e. 28 consecutive bytes of the source
      const uint32_t sy = a.flip ? a.H - 1 - y : y;
      const unsigned int* sp = reinterpret_cast<const unsigned int*>(img + ((uint64_t)sy * a.W + x) * 4);
#pragma unroll
      for (int j = 0; j < 7; ++j) px[j] = __ldg(sp + j);
    } else {
#pragma unroll
      for (int j = 0; j < 7; ++j) {
        const int i = first + j;
        px[j] = 0;
        if (i >= i0 && (uint32_t)i < plane) {
          const uint32_t sy = a.flip ? a.H - 1 - y : y;
          px[j] = __ldg(reinterpret_cast<const unsigned int*>(img + ((uint64_t)sy * a.W + x) * 4));
          if (++x == a.W) { x = 0; ++y; }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c < (int)a.Cout) {
        const uint64_t plane_off = ((uint64_t)b * a.Cout + c) * plane;
        const int d = (int)(plane_off & 3);                // the same for every thread on this image: no divergence
        const int e0 = (int)(4 * k) - d;                   // first element of this thread's group, plane-local
        T* out = dst + plane_off + e0;
        switch (d) {
          case 0: window_emit<T, STATS, 3>(px, lut, c, out, e0, plane, acc.s[c], acc.q[c]); break;
          case 1: window_emit<T, STATS, 2>(px, lut, c, out, e0, plane, acc.s[c], acc.q[c]); break;
          case 2: window_emit<T, STATS, 1>(px, lut, c, out, e0, plane, acc.s[c], acc.q[c]); break;
          default: window_emit<T, STATS, 0>(px, lut, c, out, e0, plane, acc.s[c], acc.q[c]); break;
        }
      }
    }
  }
  if (STATS) finish_stats(a, acc);
}

// ---- row flip for frames of any element type (depth maps) ------------------------------------------
// dst[b][y] = src[index ? index[b] : b][H-1-y], rows of `upr` units of UB bytes.  A copy: HBM-bound,
// one 16-byte load and one 16-byte streaming store per unit when rows are whole 16-byte units.
struct FlipArgs {
  const uint8_t* src;
  uint8_t* dst;
  const long long* src_index;
  uint64_t row_bytes, image_bytes;
  uint32_t H;
};

template <int UB, int UNROLL>
__global__ void __launch_bounds__(kThreads) k1_flip_rows(FlipArgs a, PlanarGeom g) {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t u0 = blockIdx.x * blockDim.x + threadIdx.x; u0 < g.total_units; u0 += stride * UNROLL) {
    uint32_t w[UNROLL][UB >= 4 ? UB / 4 : 1];
    uint64_t out_off[UNROLL];
    bool live[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t u = u0 + j * stride;
      live[j] = u < g.total_units && u >= u0;
      if (live[j]) {
        const uint32_t row = fdiv(u, g.upr);
        const uint32_t ux = u - row * g.upr.div;
        const uint32_t b = fdiv(row, g.rows);
        const uint32_t y = row - b * g.rows.div;
        const uint64_t img = a.src_index ? (uint64_t)__ldg(a.src_index + b) : (uint64_t)b;
        const uint8_t* sp = a.src + img * a.image_bytes + (uint64_t)(a.H - 1 - y) * a.row_bytes + (uint64_t)ux * UB;
        if constexpr (UB >= 4) ld_words<UB>(sp, w[j]);
        else w[j][0] = __ldg(sp);
        out_off[j] = (uint64_t)row * a.row_bytes + (uint64_t)ux * UB;
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      if (!live[j]) continue;
      if constexpr (UB == 16) __stcs(reinterpret_cast<uint4*>(a.dst + out_off[j]), make_uint4(w[j][0], w[j][1], w[j][2], w[j][3]));
      else if constexpr (UB == 4) __stcs(reinterpret_cast<unsigned int*>(a.dst + out_off[j]), w[j][0]);
      else a.dst[out_off[j]] = (uint8_t)w[j][0];
    }
  }
}

thread_local int g_last_variant = 0;
thread_local int g_last_kernels = 0;

int grid_for(uint64_t units_per_thread_total, int device_sms, int ctas_per_sm) {
  uint64_t need = (units_per_thread_total + kThreads - 1) / kThreads;
  uint64_t cap = (uint64_t)device_sms * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

FastDiv make_fastdiv(uint32_t d) {
  // round-up method: q = (umulhi(n, m) + n) >> s, exact for n < 2^31
  FastDiv f;
  f.div = d;
  uint32_t s = 0;
  while ((1ull << s) < d) ++s;
  f.shift = s;
  f.mul = (uint32_t)((((1ull << s) - d) << 32) / d + 1);
  return f;
}

bool planar_geom(const K1Args& a, int ppt, PlanarGeom& g) {
  const uint64_t upr = a.W / ppt;
  const uint64_t total = (uint64_t)a.B * a.H * upr;
  if (total >= (1ull << 31) || (uint64_t)a.H * a.W >= (1ull << 30)) return false;
  g.upr = make_fastdiv((uint32_t)upr);
  g.rows = make_fastdiv(a.H);
  g.total_units = (uint32_t)total;
  return true;
}

// Launch with or without the programmatic-stream-serialization attribute (PDL).
template <typename K, typename G>
void launch_planar(K kernel, int grid, cudaStream_t st, const K1Args& a, const G& geo, int pdl) {
  if (!pdl) {
    kernel<<<grid, kThreads, 0, st>>>(a, geo);
    return;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kernel, a, geo);
}

// CTAs of `kernel` that fit on one SM (asked once per kernel).
template <auto Kernel>
int resident_ctas() {
  static int cached = 0;   // one instance per kernel; a race writes the same value twice
  if (cached == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, Kernel, kThreads, 0) != cudaSuccess || n < 1) n = 4;
    cached = n;
  }
  return cached;
}

template <typename T, bool STATS>
int launch_typed(const bt_preprocess_params* p, K1Args& a, int variant, cudaStream_t st, int grid_cap) {
  const int sms = sm_count();
  const uint64_t rows = (uint64_t)a.B * a.H;
  const bool aligned_src = (reinterpret_cast<uintptr_t>(a.src) % 16 == 0) && (a.src_image_stride % 16 == 0);
  const bool aligned_dst = reinterpret_cast<uintptr_t>(a.dst) % 16 == 0;
  const bool can_vec = aligned_src && aligned_dst && variant != 1 && variant != 4;
  const int tune_ctas = (p->reserved >> 8) & 0xff;
  auto grid_of = [&](uint64_t units, int unroll, int resident = 0) {
    int g = grid_for((units + unroll - 1) / unroll, sms, tune_ctas ? tune_ctas : kDefaultCtasPerSm);
    if (grid_cap > 0 && g > grid_cap) g = grid_cap;
    if (!tune_ctas && !STATS && resident > 0) {
      // `resident` CTAs per SM make one wave.  A grid that exceeds it by a few CTAs would run a
      // second, nearly empty wave: let a few threads take a second unit instead (an isolated batch-8
      // launch of 600 CTAs on 592 slots: 11.7 -> 10.3 us).  16-bit outputs move fewer bytes per
      // CTA and prefer the single looping wave up to the size where the 4-unit kernel takes over
      // (profiles/r1_k1_wave.txt).
      const int wave = sms * resident;
      if (g > wave && (g <= wave + wave / 10 || (sizeof(T) == 2 && units < (1u << 20)))) g = wave;
    }
    return g;
  };
#define LAUNCH_PLANAR(C_, PPT_, UNR_)                                                       \
  do {                                                                                      \
    PlanarGeom geo;                                                                         \
    if (!planar_geom(a, (PPT_), geo)) break; /* too many units for 32-bit indices */        \
    int g = grid_of(geo.total_units, (UNR_), resident_ctas<k1_planar<T, C_, PPT_, UNR_, STATS>>()); \
    launch_planar(k1_planar<T, C_, PPT_, UNR_, STATS>, g, st, a, geo, !((p->reserved >> 16) & 1)); \
    g_last_variant = 2;                                                                     \
    return g;                                                                               \
  } while (0)
  const int tune_unroll = p->reserved & 0xff;
  if (variant == 3) {
    // bulk-async staged input: RGBA, planar output, whole 4-pixel units per row, no statistics
    if (STATS || !can_vec || !p->chw || a.C != 4 || a.W % 4 != 0)
      return bt::fail(BT_E_INVALID, "the bulk-async variant takes RGBA frames with W %% 4 == 0, planar output, no statistics");
    BulkGeom bg;
    const uint64_t spr = ((uint64_t)a.W * 4 + kBulkSegBytes - 1) / kBulkSegBytes;
    const uint64_t total = rows * spr;
    if (total >= (1ull << 31) || (uint64_t)a.H * a.W >= (1ull << 30))
      return bt::fail(BT_E_INVALID, "batch too large for the bulk-async variant");
    bg.spr = make_fastdiv((uint32_t)spr);
    bg.rows = make_fastdiv(a.H);
    bg.total_segments = (uint32_t)total;
    const int per_sm = tune_ctas ? tune_ctas : 4;   // persistent CTAs, 20.5 KiB of shared memory each
    int g = (int)(total < (uint64_t)sms * per_sm ? total : (uint64_t)sms * per_sm);
    launch_planar(k1_bulk<T>, g, st, a, bg, 1);
    g_last_variant = 3;
    return g;
  }
  if (can_vec && p->chw && a.C == 4) {
    // widest unit whose stores are 16 bytes and that divides the row
    if (sizeof(T) == 1 && a.W % 16 == 0) LAUNCH_PLANAR(4, 16, 2);
    if (sizeof(T) == 2 && a.W % 8 == 0) {
      if (!STATS && tune_unroll == 4) LAUNCH_PLANAR(4, 8, 4);
      if (!STATS && tune_unroll == 1) LAUNCH_PLANAR(4, 8, 1);
      if (!STATS && tune_unroll == 44) LAUNCH_PLANAR(4, 4, 4);
      // large batches: more bytes in flight per thread (0.905 vs 0.871 of peak at 64 x 1024^2);
      // small ones keep the shorter critical path
      if (!STATS && tune_unroll == 0 && rows * (a.W / 8) >= (1u << 20)) LAUNCH_PLANAR(4, 8, 4);
      LAUNCH_PLANAR(4, 8, 2);
    }
    if (a.W % 4 == 0) {
      if (STATS && sizeof(T) == 4) {   // A/B: fewer units in flight per thread, fewer registers, more CTAs per SM
        if (tune_unroll == 1) LAUNCH_PLANAR(4, 4, 1);
        if (tune_unroll == 2) LAUNCH_PLANAR(4, 4, 2);
      }
      if (!STATS && sizeof(T) == 4) {
        if (tune_unroll == 1) LAUNCH_PLANAR(4, 4, 1);
        if (tune_unroll == 2) LAUNCH_PLANAR(4, 4, 2);
        if (tune_unroll == 8) LAUNCH_PLANAR(4, 4, 8);
        // One unit per thread (more, shorter CTAs) is 2-5 % faster up to batches of about 8 M units --
        // batch 8..64 of 640x480, 1024 x 128^2 (0.97 vs 0.92 of peak) -- and 1 % slower beyond, where
        // four units in flight per thread win (profiles/r1_k1_unroll.txt).
        if (tune_unroll == 0 && rows * (a.W / 4) < (8u << 20)) LAUNCH_PLANAR(4, 4, 1);
      }
      LAUNCH_PLANAR(4, 4, 4);
    }
  }
  if (can_vec && p->chw && a.C == 3) {
    if (a.W % 16 == 0 && tune_unroll == 16) LAUNCH_PLANAR(3, 16, 2);
    if (a.W % 8 == 0 && sizeof(T) <= 2) LAUNCH_PLANAR(3, 8, 2);
    if (a.W % 4 == 0) LAUNCH_PLANAR(3, 4, 4);
  }
  if (can_vec && p->chw && a.C == 2 && a.W % 8 == 0) LAUNCH_PLANAR(2, 8, 4);
#undef LAUNCH_PLANAR
  {
    // interleaved output (or single-channel, where planar == interleaved)
    constexpr uint32_t IB = 16 / sizeof(T);
    const uint64_t row_bytes = (uint64_t)a.W * a.C;
    if (can_vec && (!p->chw || a.C == 1) && a.Cout == a.C && row_bytes % IB == 0) {
      PlanarGeom geo;
      const uint64_t total = rows * (row_bytes / IB);
      if (total < (1ull << 31)) {
        geo.upr = make_fastdiv((uint32_t)(row_bytes / IB));
        geo.rows = make_fastdiv(a.H);
        geo.total_units = (uint32_t)total;
        int g = grid_of(total, 4);
        if (sizeof(T) == 1 && !a.use_gamma)
          k1_interleaved<T, 4, STATS, true, 0><<<g, kThreads, 0, st>>>(a, geo);
        else if (a.C == 4)
          k1_interleaved<T, 4, STATS, false, 4><<<g, kThreads, 0, st>>>(a, geo);
        else if (a.C == 1 || (sizeof(T) == 1 && !STATS))
          // one table serves every byte: a single channel, or uint8 output of 2-3 colour channels
          // (no normalisation, and gamma treats channels 0..2 alike)
          k1_interleaved<T, 4, STATS, false, 1><<<g, kThreads, 0, st>>>(a, geo);
        else
          k1_interleaved<T, 4, STATS, false, 0><<<g, kThreads, 0, st>>>(a, geo);
        g_last_variant = 2;
        return g;
      }
    }
  }
  if (variant == 2 || variant == 3)
    return bt::fail(BT_E_INVALID, "requested vectorised variant needs rows made of whole 16-byte units (W=%u C=%u)", a.W, a.C);
  if (variant != 1 && variant != 4 && p->chw && a.C == 4 && aligned_dst &&
      reinterpret_cast<uintptr_t>(a.src) % 4 == 0 && a.src_image_stride % 4 == 0) {
    // RGBA rows of any width, planar output: aligned groups of four plane elements (k1_window)
    const uint64_t plane = (uint64_t)a.H * a.W;
    const uint64_t groups = (plane + 3) / 4 + 1;           // d_c <= 3 can push one more group into the plane
    const uint64_t total = groups * a.B;
    if (plane < (1ull << 30) && total < (1ull << 31)) {
      PlanarGeom geo;
      geo.upr = make_fastdiv(a.W);
      geo.rows = make_fastdiv((uint32_t)groups);
      geo.total_units = (uint32_t)total;
      int g = grid_of(total, 1);
      k1_window<T, STATS><<<g, kThreads, 0, st>>>(a, geo);
      g_last_variant = 5;
      return g;
    }
  }
  if (variant != 1) {
    // any width: one pixel per thread
    PlanarGeom geo;
    const uint64_t total = rows * a.W;
    if (total < (1ull << 31) && (uint64_t)a.H * a.W < (1ull << 30)) {
      geo.upr = make_fastdiv(a.W);
      geo.rows = make_fastdiv(a.H);
      geo.total_units = (uint32_t)total;
      int g = grid_of(total, 4);
#define LAUNCH_PIXEL(C_)                                                                          \
  do {                                                                                            \
    if (p->chw) k1_pixel<T, C_, true, 4, STATS><<<g, kThreads, 0, st>>>(a, geo);                  \
    else k1_pixel<T, C_, false, 4, STATS><<<g, kThreads, 0, st>>>(a, geo);                        \
  } while (0)
      const bool px4 = a.C == 4 && reinterpret_cast<uintptr_t>(a.src) % 4 == 0 && a.src_image_stride % 4 == 0;
      if (px4) LAUNCH_PIXEL(4);
      else if (a.C == 3) LAUNCH_PIXEL(3);
      else if (a.C == 2) LAUNCH_PIXEL(2);
      else if (a.C == 1) LAUNCH_PIXEL(1);
#undef LAUNCH_PIXEL
      if (px4 || a.C < 4) {
        g_last_variant = 4;
        return g;
      }
    }
  }
  {
    uint64_t total = rows * a.W * a.Cout;
    int g = grid_of(total, 1);
    if (p->chw) k1_generic<T, true, STATS><<<g, kThreads, 0, st>>>(a);
    else k1_generic<T, false, STATS><<<g, kThreads, 0, st>>>(a);
    g_last_variant = 1;
    return g;
  }
}

// Scratch of the statistics reduction for a grid of at most `rows` CTAs: one row of partial sums per
// CTA and, for grids that are folded in two steps, one row per kFoldRowsPerCta of them.  Nothing in it
// has to be initialised: every row is written before the fold that reads it.
struct StatsLayout {
  uint64_t group_rows_off, bytes;
};
constexpr StatsLayout stats_layout(uint64_t rows) {
  const uint64_t row = kMaxC * 2 * sizeof(double);
  const uint64_t groups = (rows + kFoldRowsPerCta - 1) / kFoldRowsPerCta;
  return StatsLayout{rows * row, (rows * row + groups * row + 255) / 256 * 256};
}

// Enqueue the fold of `n` rows behind the kernel that writes them.
int launch_fold(const double* rows, double* group_rows, unsigned n, double* stats_out, unsigned cout, cudaStream_t st) {
  auto launch = [&](unsigned grid, const double* in, unsigned count, unsigned per_cta, double* out, unsigned pairs) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kFoldThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k1_fold, in, count, per_cta, out, pairs);
  };
  if (n <= (unsigned)kFoldDirectMax) {
    BT_CUDA_TRY(launch(1, rows, n, n, stats_out, cout));
    g_last_kernels += 1;
    return BT_OK;
  }
  const unsigned groups = (n + kFoldRowsPerCta - 1) / kFoldRowsPerCta;
  BT_CUDA_TRY(launch(groups, rows, n, kFoldRowsPerCta, group_rows, kMaxC));
  BT_CUDA_TRY(launch(1, group_rows, groups, groups, stats_out, cout));
  g_last_kernels += 2;
  return BT_OK;
}

// Caller-provided scratch covers the largest grid a launch can have (one pass per thread, capped at
// kDefaultCtasPerSm CTAs per SM of a 148-SM part); the built-in array of
// bt_preprocess_stats_static is smaller (see there).
constexpr int kStatsMaxRows = 148 * kDefaultCtasPerSm;
constexpr int kStaticRows = 148 * 8;

int launch_checked(const bt_preprocess_params* p, void* stream, int stats_rows);

}  // namespace

extern "C" {

int bt_build_gamma_lut(double coeff, uint8_t lut[256]) {
  if (!lut || !(coeff > 0.0)) return bt::fail(BT_E_INVALID, "gamma coefficient must be positive");
  // offscreen.py:106-108 in float32: x.astype(f32)/255, ** (1/coeff) with the exponent rounded to
  // float32 (numpy weak-scalar promotion), times 255.0 in float32, truncating uint8 cast.
  const float e = (float)(1.0 / coeff);
  for (int v = 0; v < 256; ++v) {
    float x = (float)v / 255.0f;
    float y = 255.0f * powf(x, e);
    lut[v] = (uint8_t)y;
  }
  return BT_OK;
}

uint64_t bt_preprocess_scratch_bytes(const bt_preprocess_params* p) {
  if (!p || !p->stats_out) return 0;
  return stats_layout(kStatsMaxRows).bytes;
}

int bt_preprocess_last_variant(void) { return g_last_variant; }
int bt_preprocess_last_kernels(void) { return g_last_kernels; }

int bt_flip_rows_launch(const void* src, void* dst, int batch, int height, uint64_t row_bytes, const int64_t* src_index,
                        void* stream) {
  if (!src || !dst) return bt::fail(BT_E_INVALID, "bt_flip_rows_launch: null pointer");
  if (batch < 0 || height < 0) return bt::fail(BT_E_INVALID, "negative extent");
  if ((uint64_t)batch * height * row_bytes == 0) return BT_OK;
  FlipArgs a;
  a.src = static_cast<const uint8_t*>(src);
  a.dst = static_cast<uint8_t*>(dst);
  a.src_index = reinterpret_cast<const long long*>(src_index);
  a.row_bytes = row_bytes;
  a.image_bytes = row_bytes * (uint64_t)height;
  a.H = (uint32_t)height;
  const bool al16 = row_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(src) % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0;
  const bool al4 = row_bytes % 4 == 0 && reinterpret_cast<uintptr_t>(src) % 4 == 0 && reinterpret_cast<uintptr_t>(dst) % 4 == 0;
  const uint32_t ub = al16 ? 16 : al4 ? 4 : 1;
  const uint64_t upr = row_bytes / ub, total = (uint64_t)batch * height * upr;
  if (total >= (1ull << 31)) return bt::fail(BT_E_TOO_LARGE, "bt_flip_rows_launch: more than 2^31 units in one launch; split the batch");
  PlanarGeom geo;
  geo.upr = make_fastdiv((uint32_t)upr);
  geo.rows = make_fastdiv((uint32_t)height);
  geo.total_units = (uint32_t)total;
  const int g = grid_for((total + 3) / 4, sm_count(), kDefaultCtasPerSm);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (ub == 16) k1_flip_rows<16, 4><<<g, kThreads, 0, st>>>(a, geo);
  else if (ub == 4) k1_flip_rows<4, 4><<<g, kThreads, 0, st>>>(a, geo);
  else k1_flip_rows<1, 4><<<g, kThreads, 0, st>>>(a, geo);
  BT_CUDA_TRY(cudaGetLastError());
  return BT_OK;
}

int bt_preprocess_launch(const bt_preprocess_params* p, void* stream) { return launch_checked(p, stream, kStatsMaxRows); }

}  // extern "C"

namespace {

// The table of all possible outputs (K1Args::table), computed with the reference's expression in IEEE
// float arithmetic -- the host's subtract and divide round exactly like __fsub_rn / __fdiv_rn -- and
// remembered per thread: consecutive launches of a pipeline use the same transform.
void fill_table(K1Args& a, int out_dtype, bool enable) {
  a.has_table = 0;
  if (!enable) return;
  struct Key {
    int out_dtype;
    uint32_t use_gamma, use_norm;
    float mean[4], stdev[4];
    uint8_t gamma[256];
  };
  thread_local Key last_key;
  thread_local uint32_t last_table[kMaxC * 256];
  thread_local bool have = false;
  Key key;
  memset(&key, 0, sizeof key);
  key.out_dtype = out_dtype;
  key.use_gamma = a.use_gamma;
  key.use_norm = a.use_norm;
  memcpy(key.mean, a.mean, sizeof key.mean);
  memcpy(key.stdev, a.stdev, sizeof key.stdev);
  if (a.use_gamma) memcpy(key.gamma, a.gamma_lut, 256);
  if (!have || memcmp(&key, &last_key, sizeof key) != 0) {
    memset(last_table, 0, sizeof last_table);
    for (int c = 0; c < kMaxC; ++c) {
      for (int v = 0; v < 256; ++v) {
        const uint32_t g = a.use_gamma ? a.gamma_lut[v] : (uint32_t)v;
        const uint32_t gc = c < 3 ? g : (uint32_t)v;   // alpha is never gamma-corrected (offscreen.py:109-111)
        volatile float o = (float)gc;                  // volatile: each step is rounded to float, nothing is fused
        if (a.use_norm) {
          volatile float d = o - a.mean[c];
          o = d / a.stdev[c];
        }
        const int i = c * 256 + v;
        if (out_dtype == BT_F32) {
          const float f = o;
          memcpy(&last_table[i], &f, 4);
        } else if (out_dtype == BT_BF16) {
          const __nv_bfloat16 h = __float2bfloat16_rn((float)o);
          memcpy(reinterpret_cast<uint8_t*>(last_table) + 2 * i, &h, 2);
        } else {
          reinterpret_cast<uint8_t*>(last_table)[i] = (uint8_t)gc;
        }
      }
    }
    last_key = key;
    have = true;
  }
  memcpy(a.table, last_table, sizeof a.table);
  a.has_table = 1;
}

// `stats_rows`: how many partial-sum rows `p->stats_scratch` has room for (bounds the grid).
int launch_checked(const bt_preprocess_params* p, void* stream, int stats_rows) {
  if (!p || !p->src || !p->dst) return bt::fail(BT_E_INVALID, "bt_preprocess_launch: null pointer");
  if (p->batch < 0 || p->height < 0 || p->width < 0) return bt::fail(BT_E_INVALID, "negative extent");
  if (p->channels < 1 || p->channels > kMaxC) return bt::fail(BT_E_INVALID, "channels must be 1..4, got %d", p->channels);
  if (p->out_channels < 1 || p->out_channels > p->channels)
    return bt::fail(BT_E_INVALID, "out_channels must be 1..channels, got %d", p->out_channels);
  if (p->out_dtype != BT_U8 && p->out_dtype != BT_F32 && p->out_dtype != BT_BF16)
    return bt::fail(BT_E_INVALID, "out_dtype must be u8, f32 or bf16");
  if (p->stats_out && !p->stats_scratch) return bt::fail(BT_E_INVALID, "stats_out needs stats_scratch");
  if ((uint64_t)p->height * p->width * p->channels >= (1ull << 32))
    return bt::fail(BT_E_INVALID, "single frames of 4 GiB or more are not supported");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  K1Args a;
  memset(&a, 0, sizeof a);
  a.src = static_cast<const uint8_t*>(p->src);
  a.dst = p->dst;
  a.B = p->batch; a.H = p->height; a.W = p->width; a.C = p->channels; a.Cout = p->out_channels;
  a.flip = p->flip != 0;
  a.use_gamma = p->use_gamma != 0;
  a.use_norm = (p->use_norm != 0) && p->out_dtype != BT_U8;
  a.flags = ((p->reserved >> 18) & 1) ^ 1;   // loads-first is the default; tuning bit 18 turns it off
  const uint64_t dense = (uint64_t)p->height * p->width * p->channels;
  a.src_image_stride = p->src_image_stride ? p->src_image_stride : dense;
  if (a.src_image_stride < dense) return bt::fail(BT_E_INVALID, "src_image_stride smaller than one frame");
  a.src_index = reinterpret_cast<const long long*>(p->src_index);
  memcpy(a.mean, p->mean, sizeof a.mean);
  memcpy(a.stdev, p->stdev, sizeof a.stdev);
  memcpy(a.gamma_lut, p->gamma_lut, 256);
  if (a.use_norm)
    for (int c = 0; c < p->channels; ++c)
      if (a.stdev[c] == 0.0f) return bt::fail(BT_E_INVALID, "stdev[%d] is zero", c);

  // A/B kept for the record (tuning bit 19; profiles/r2_k1_table_ab.txt): handing every CTA the finished
  // table as 4 KB of kernel parameters instead of letting it compute the table.  It loses badly for
  // fp32 -- 25.7 vs 9.5 us at batch 8, 154 vs 52 us at 1024 x 128^2 -- because the copy out of the
  // parameter bank is an LDC with a different address in every lane, which the constant cache serves
  // one address at a time (four such loads per thread for fp32, two for bf16: -5 %, one for uint8: +1 %).
  // Four IEEE divisions per thread, hidden under the first loads' DRAM latency, are cheaper.
  fill_table(a, p->out_dtype, ((p->reserved >> 19) & 1) != 0);
  const bool stats = p->stats_out != nullptr;
  double* group_rows = nullptr;
  if (stats) {
    const StatsLayout lay = stats_layout((uint64_t)stats_rows);
    uint8_t* base = static_cast<uint8_t*>(p->stats_scratch);
    a.partials = reinterpret_cast<double*>(base);
    group_rows = reinterpret_cast<double*>(base + lay.group_rows_off);
    a.stats_out = p->stats_out;
  }
  if ((uint64_t)a.B * a.H * a.W == 0) {
    if (stats) BT_CUDA_TRY(cudaMemsetAsync(p->stats_out, 0, sizeof(double) * 2 * a.Cout, st));
    g_last_variant = 0;
    g_last_kernels = 0;
    return BT_OK;
  }
  int g = 0;
  int cap = 0;
  if (stats) {
    // Every CTA of a statistics launch ends with a reduction (shuffles, a barrier, one row), so the
    // grid that is best without statistics -- one pass per thread -- is not best with them.  Measured
    // on B200 (profiles/r1_k1_stats_cap.txt): 16-bit and 8-bit outputs want one resident wave
    // (4 CTAs per SM) looping over the batch, fp32 wants a single wave when the batch is small enough
    // to need less than two, and 64 CTAs per SM otherwise.
    cap = stats_rows;
    if (((p->reserved >> 8) & 0xff) == 0) {
      const int sms = sm_count();
      const uint64_t ctas = ((uint64_t)a.B * a.H * a.W + 4095) / 4096;   // one pass: 16 pixels per thread
      const int per_sm = (p->out_dtype != BT_F32 || ctas <= (uint64_t)sms * 8) ? 4 : 64;
      if (sms * per_sm < cap) cap = sms * per_sm;
    }
  }
  switch (p->out_dtype) {
    case BT_F32:
      g = stats ? launch_typed<float, true>(p, a, p->variant, st, cap) : launch_typed<float, false>(p, a, p->variant, st, cap);
      break;
    case BT_BF16:
      g = stats ? launch_typed<__nv_bfloat16, true>(p, a, p->variant, st, cap)
                : launch_typed<__nv_bfloat16, false>(p, a, p->variant, st, cap);
      break;
    default:
      g = stats ? launch_typed<uint8_t, true>(p, a, p->variant, st, cap) : launch_typed<uint8_t, false>(p, a, p->variant, st, cap);
      break;
  }
  if (g < 0) return g;
  BT_CUDA_TRY(cudaGetLastError());
  g_last_kernels = 1;
  if (stats) {
    const int rc = launch_fold(a.partials, group_rows, (unsigned)g, p->stats_out, a.Cout, st);
    if (rc != BT_OK) return rc;
  }
  return BT_OK;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// bt_preprocess_stats_static: K1 with statistics for callers that must not allocate.
//
// The statistics reduction needs a little device memory besides the caller's input, output
// and stats_out: one row of partial sums per CTA (and two rows for a two-step fold).  The regular entry
// point takes that memory from the caller (`stats_scratch`), which means somebody has to
// cudaMalloc it.  Some users of this library cannot: a render or simulation plug-in that is
// handed a fixed device-memory budget by its host application, a real-time ingest loop whose
// rules forbid allocator calls after start-up, a process running next to a framework that
// already owns all free memory of the GPU.  For them the scratch lives in the array below,
// which is part of the library image: every GPU whose context loads libbt_b200.so holds these
// 76 032 bytes until the process exits, whether or not this entry point is ever called.
//
// `g_static_stats_scratch` as written in host code is only a host-side handle for the symbol,
// not a device address; the address a kernel may use is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached for the life of the process: after a
// cudaDeviceReset the cached address is stale and the library has to be reloaded (bt_b200.h says so).
// ---------------------------------------------------------------------------------------------
#include <mutex>

namespace {
constexpr uint64_t kStaticScratchBytes = stats_layout(kStaticRows).bytes;   // 76 032: 1184 rows + 2 group rows, rounded to 256
constexpr uint64_t kStaticBytesPerRow = 16384;                   // input bytes covered by one CTA's row
constexpr uint64_t kStaticMaxInputBytes = (uint64_t)kStaticRows * kStaticBytesPerRow;  // 19 398 656
constexpr int kMaxDevices = 64;
}  // namespace

static __device__ __align__(256) unsigned char g_static_stats_scratch[kStaticScratchBytes];

namespace {
std::mutex g_static_mu;
void* g_static_addr[kMaxDevices] = {nullptr};

// Device address of the static scratch on the current device; first use per device resolves it.
int static_scratch_address(void** out) {
  int dev = -1;
  BT_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return bt::fail(BT_E_INVALID, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_stats_scratch);
    if (e != cudaSuccess || !ptr)
      return bt::fail(BT_E_CUDA, "cudaGetSymbolAddress(static scratch) on device %d failed: %s", dev,
                      cudaGetErrorString(e));
    g_static_addr[dev] = ptr;
  }
  *out = g_static_addr[dev];
  return BT_OK;
}
}  // namespace

extern "C" int bt_preprocess_stats_static(const bt_preprocess_params* p, void* stream) {
  if (!p) return bt::fail(BT_E_INVALID, "bt_preprocess_stats_static: null pointer");
  if (!p->stats_out)
    return bt::fail(BT_E_INVALID, "bt_preprocess_stats_static computes statistics: stats_out must be set "
                                  "(without statistics bt_preprocess_launch needs no scratch at all)");
  if (p->batch < 0 || p->height < 0 || p->width < 0 || p->channels < 1 || p->channels > kMaxC)
    return bt::fail(BT_E_INVALID, "bt_preprocess_stats_static: bad extents");
  const uint64_t in_bytes = (uint64_t)p->batch * p->height * p->width * p->channels;
  if (in_bytes > kStaticMaxInputBytes)
    return bt::fail(BT_E_TOO_LARGE,
                    "batch of %llu input bytes exceeds the %llu bytes the built-in scratch covers (%d rows of "
                    "partial sums, one per %llu input bytes); split the batch or use bt_preprocess_launch "
                    "with caller-provided stats_scratch",
                    (unsigned long long)in_bytes, (unsigned long long)kStaticMaxInputBytes, kStaticRows,
                    (unsigned long long)kStaticBytesPerRow);
  void* scratch = nullptr;
  int rc = static_scratch_address(&scratch);
  if (rc != BT_OK) return rc;
  bt_preprocess_params q = *p;
  q.stats_scratch = scratch;  // same kernel, same launch path; only the scratch differs
  // Nothing in the array has to start at zero: whatever an earlier call left there does not matter,
  // because every row a fold reads is written by this call's kernel first (stream order).
  return launch_checked(&q, stream, kStaticRows);
}
