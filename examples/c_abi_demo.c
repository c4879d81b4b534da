/* The C ABI on its own: no Python, no torch -- device memory from the CUDA runtime, one K1 launch through
 * bt_preprocess_launch, the result checked on the host with the reference's expression
 * (offscreen.py:95-112, densityopt.py:117-119).  This is what a cgo / JNI / ctypes binding would call.
 *
 *   gcc -O2 -I include -I /usr/local/cuda/include examples/c_abi_demo.c \
 *       -L pytorch-blender_b200/blendtorch/btt/_native -lbt_b200 -L /usr/local/cuda/lib64 -lcudart -lm \
 *       -Wl,-rpath,$PWD/pytorch-blender_b200/blendtorch/btt/_native -o /tmp/c_abi_demo && /tmp/c_abi_demo
 */
#include <cuda_runtime_api.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "bt_b200.h"

#define CHECK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } } while (0)
#define CHECK_BT(x) do { int rc_ = (x); if (rc_ < 0) { fprintf(stderr, "%s: %s\n", #x, bt_last_error()); return 3; } } while (0)

int main(void) {
  enum { B = 3, H = 48, W = 64, C = 4 };
  const size_t n_in = (size_t)B * H * W * C, n_out = n_in;
  uint8_t* frames = (uint8_t*)malloc(n_in);
  float* out = (float*)malloc(n_out * sizeof(float));
  uint32_t seed = 12345u;
  for (size_t i = 0; i < n_in; ++i) { seed = seed * 1664525u + 1013904223u; frames[i] = (uint8_t)(seed >> 24); }

  if (bt_abi_version() != BT_ABI_VERSION) { fprintf(stderr, "header and library disagree\n"); return 1; }
  void *d_in = NULL, *d_out = NULL;
  CHECK_CUDA(cudaMalloc(&d_in, n_in));
  CHECK_CUDA(cudaMalloc(&d_out, n_out * sizeof(float)));
  CHECK_CUDA(cudaMemcpy(d_in, frames, n_in, cudaMemcpyHostToDevice));

  bt_preprocess_params p;
  memset(&p, 0, sizeof p);
  p.src = d_in; p.dst = d_out;
  p.batch = B; p.height = H; p.width = W; p.channels = C; p.out_channels = C;
  p.out_dtype = BT_F32; p.flip = 1; p.chw = 1; p.use_gamma = 1; p.use_norm = 1;
  for (int c = 0; c < 4; ++c) { p.mean[c] = 127.5f; p.stdev[c] = 127.5f; }
  CHECK_BT(bt_build_gamma_lut(2.2, p.gamma_lut));
  CHECK_BT(bt_preprocess_launch(&p, NULL /* default stream */));
  CHECK_CUDA(cudaMemcpy(out, d_out, n_out * sizeof(float), cudaMemcpyDeviceToHost));

  /* the reference's arithmetic, per value, in float32 */
  size_t bad = 0;
  const float e = (float)(1.0 / 2.2);
  for (int b = 0; b < B; ++b)
    for (int c = 0; c < C; ++c)
      for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
          uint8_t v = frames[(((size_t)b * H + (H - 1 - y)) * W + x) * C + c];          /* np.flipud */
          if (c < 3) v = (uint8_t)(255.0f * powf((float)v / 255.0f, e));                /* truncating gamma */
          const float want = ((float)v - 127.5f) / 127.5f;                              /* normalise, CHW */
          const float got = out[(((size_t)b * C + c) * H + y) * W + x];
          if (memcmp(&want, &got, sizeof want) != 0) ++bad;
        }
  printf("c_abi_demo: variant %d, %d kernel(s), %zu of %zu values differ\n", bt_preprocess_last_variant(),
         bt_preprocess_last_kernels(), bad, n_out);
  cudaFree(d_in); cudaFree(d_out); free(frames); free(out);
  return bad ? 4 : 0;
}
