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

// ---- internals shared between scan.cpp, ring.cu and shm.cu --------------------------------------
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

struct bt_schema {
  std::vector<bt_column> cols;
  int batch = 0;
  uint64_t slot_bytes = 0;
};

namespace bt {

// One payload that was announced by reference (shared-memory wire) instead of carried in-line.
struct PendingRef {
  bt_shm* seg;
  int shm_slot;
  int column;
  int item;
  uint64_t nbytes;
};

int scan_fields(const uint8_t* msg, uint64_t len, std::vector<bt_field>& out);
// bt_field.payload_off of a BT_KIND_RAWREF field whose frame the receiver has already put in place
constexpr uint64_t kLandedOff = ~0ull;
// Raw-parts messages held as one buffer (frame 0 = pickled header, frames 1.. = array bytes, back to
// back): point every BT_KIND_RAWREF field of the scanned header at its frame.
int resolve_raw_parts(std::vector<bt_field>& f, const uint64_t* sizes, int nparts);
uint64_t field_bytes(const bt_field& f);
std::vector<bt_field>& tls_fields();

// Copy the scanned fields of one message into item `item` of the slot at `base`.  Fields that
// arrive as shared-memory references are appended to `refs` when it is non-null (the caller will
// DMA them straight from the segment), otherwise they are copied and released here.
struct CopyTask {
  uint8_t* dst;
  const uint8_t* src;
  uint64_t n;
};

int place_fields(const bt_schema* s, uint8_t* base, int item, const uint8_t* msg,
                 const std::vector<bt_field>& f, std::vector<PendingRef>* refs, int device,
                 std::vector<CopyTask>* tasks = nullptr);

class WorkPool {
 public:
  static WorkPool& get();
  // fn(k) for k in [0, ntasks) on up to `nthreads` threads (the caller is one of them).
  void run(int nthreads, size_t ntasks, const std::function<void(size_t)>& fn);

 private:
  void worker(int index);
  std::mutex mu_, job_mu_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> workers_;
  const std::function<void(size_t)>* fn_ = nullptr;
  size_t ntasks_ = 0;
  std::atomic<size_t> next_{0};
  int limit_ = 0, active_ = 0;
  uint64_t generation_ = 0;
};

// Streaming scan of a message that arrives in pieces (zmtp.cpp).  stream_run returns BT_OK when the
// pickle's STOP was reached, BT_E_TRUNCATED when more bytes are needed (call again with the grown
// buffer), 1 when the VM stands on a large payload that is still on the wire: stream_big_target
// says which schema column it belongs to (after checking dtype/shape/size), the receiver lands
// stream_big_len() bytes there, cuts the buffer back to stream_pos() and calls stream_run again.
void stream_begin();
int stream_run(const uint8_t* buf, uint64_t n);
uint64_t stream_pos();
uint64_t stream_big_len();
int stream_big_target(const bt_schema* schema, int* column);
int stream_fields(const uint8_t* buf, std::vector<bt_field>& out);

// Copy into staging memory that the CPU will not read again (a pinned slot or a shared-memory slot that
// only a DMA engine reads next): non-temporal stores, so that the destination lines are not first read
// into the cache for ownership -- a plain memcpy moves three bytes over the memory bus per byte copied
// (read source, read destination, write destination), this moves two.  Every path that stages by CPU copy
// is bound by the host's memory bandwidth, so that is a third off its cost.  Falls back to
// memcpy for small sizes, CPUs without AVX2, and BT_STREAM_COPY=0 (the A/B switch).
void stream_copy(void* dst, const void* src, size_t n);

// Process-wide registry of attached segments (shm.cu).
int shm_attach_cached(const char* name, size_t name_len, int device, bt_shm** out);

}  // namespace bt
