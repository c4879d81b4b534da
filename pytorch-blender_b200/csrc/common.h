// Shared helpers for libbt_b200 (host side).
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

#include "bt_b200.h"

namespace bt {

std::string& last_error_ref();

inline int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return code;
}

inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

inline int dtype_size(int dt) {
  switch (dt) {
    case BT_U8: case BT_I8: case BT_BOOL: return 1;
    case BT_U16: case BT_I16: case BT_F16: case BT_BF16: return 2;
    case BT_U32: case BT_I32: case BT_F32: return 4;
    case BT_U64: case BT_I64: case BT_F64: return 8;
  }
  return 0;
}

}  // namespace bt

#define BT_CUDA_TRY(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess)                                                             \
      return bt::fail(BT_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                      __FILE__, __LINE__);                                             \
  } while (0)
