// Pinned, batch-contiguous staging ring + H2D submission.
//
// Replaces the reference's host hand-off chain for a batch (DataLoader worker -> shared-memory
// tensor -> main process -> pageable `.to(dev)`; torch collate.py:269-275 and
// examples/densityopt/densityopt.py:261,268) with: payload memcpy into a page-locked slot
// (scan.cpp) -> one cudaMemcpyAsync per batch on the caller's copy stream -> event.
#include <cuda_runtime.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "common.h"

struct bt_ring {
  int device = -1;
  uint64_t slot_bytes = 0;
  int depth = 0;
  uint8_t* base = nullptr;
  bool pinned = false;
  std::vector<cudaEvent_t> done;
  std::vector<uint8_t> in_flight;
  int next = 0;
  // shared-memory wire: payloads announced by reference
  std::vector<std::vector<bt::PendingRef>> pending;   // ingested, copy not yet issued
  std::vector<std::vector<bt::PendingRef>> copying;   // DMA issued, released when the event completes
};

namespace {
// Hand the shared-memory slots of `v` back to their producers.  The release is a compare-and-swap on
// the slot's generation (shm.cu): if the producer took a slot back while we were still reading it
// (held past its lease), the bytes we copied may belong to a newer frame -- that is reported
// (BT_E_IO, message kept in bt_last_error), never passed over.
int release_refs(std::vector<bt::PendingRef>& v) {
  int rc = BT_OK;
  std::string first;
  for (const bt::PendingRef& r : v) {
    if (bt_shm_release(r.seg, r.shm_slot) != BT_OK && rc == BT_OK) {
      rc = BT_E_IO;
      first = bt::last_error_ref();
    }
  }
  v.clear();
  if (rc != BT_OK) bt::last_error_ref() = first;
  return rc;
}
}  // namespace

namespace bt {

uint8_t* ring_slot_base(bt_ring* r, int slot) { return r->base + r->slot_bytes * (uint64_t)slot; }

int ring_place(bt_ring* r, int slot, const bt_schema* s, int item, const uint8_t* msg, const std::vector<bt_field>& f) {
  const size_t before = r->pending[slot].size();
  int rc = place_fields(s, ring_slot_base(r, slot), item, msg, f, r->pinned ? &r->pending[slot] : nullptr, r->device);
  if (rc != BT_OK) r->pending[slot].resize(before);  // the caller's compatibility decode fetches and releases them
  return rc;
}

}  // namespace bt

extern "C" {

int bt_ring_create(int device, uint64_t slot_bytes, int depth, bt_ring** out) {
  if (!out || depth < 1 || slot_bytes == 0) return bt::fail(BT_E_INVALID, "bt_ring_create: bad arguments");
  bt_ring* r = new (std::nothrow) bt_ring();
  if (!r) return bt::fail(BT_E_NOMEM, "out of memory");
  r->device = device;
  r->slot_bytes = bt::align_up(slot_bytes, 4096);
  r->depth = depth;
  r->in_flight.assign(depth, 0);
  r->pending.resize(depth);
  r->copying.resize(depth);
  if (device >= 0) {
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess)
      e = cudaHostAlloc(reinterpret_cast<void**>(&r->base), r->slot_bytes * depth, cudaHostAllocPortable);
    if (e != cudaSuccess) {
      delete r;
      return bt::fail(BT_E_CUDA, "pinned ring allocation (%llu bytes) failed: %s",
                      (unsigned long long)(slot_bytes * depth), cudaGetErrorString(e));
    }
    r->pinned = true;
    r->done.resize(depth);
    for (int i = 0; i < depth; ++i) {
      e = cudaEventCreateWithFlags(&r->done[i], cudaEventDisableTiming);
      if (e != cudaSuccess) {
        for (int j = 0; j < i; ++j) cudaEventDestroy(r->done[j]);
        cudaFreeHost(r->base);
        delete r;
        return bt::fail(BT_E_CUDA, "cudaEventCreate failed: %s", cudaGetErrorString(e));
      }
    }
  } else {
    void* p = nullptr;
    if (posix_memalign(&p, 4096, r->slot_bytes * depth) != 0) {
      delete r;
      return bt::fail(BT_E_NOMEM, "ring allocation failed");
    }
    r->base = static_cast<uint8_t*>(p);
  }
  // touch every page once so first-use faults are not charged to the ingest loop
  memset(r->base, 0, r->slot_bytes * depth);
  *out = r;
  return BT_OK;
}

int bt_ring_destroy(bt_ring* r) {
  if (!r) return BT_OK;
  if (r->pinned) {
    for (int i = 0; i < r->depth; ++i) {
      if (r->in_flight[i]) cudaEventSynchronize(r->done[i]);
      cudaEventDestroy(r->done[i]);
      release_refs(r->copying[i]);
      release_refs(r->pending[i]);
    }
    cudaFreeHost(r->base);
  } else {
    free(r->base);
  }
  delete r;
  return BT_OK;
}

int bt_ring_depth(const bt_ring* r) { return r ? r->depth : bt::fail(BT_E_INVALID, "null ring"); }
uint64_t bt_ring_slot_bytes(const bt_ring* r) { return r ? r->slot_bytes : 0; }

void* bt_ring_slot_ptr(bt_ring* r, int slot) {
  if (!r || slot < 0 || slot >= r->depth) {
    bt::fail(BT_E_INVALID, "bt_ring_slot_ptr: bad slot");
    return nullptr;
  }
  return r->base + r->slot_bytes * (uint64_t)slot;
}

int bt_ring_wait(bt_ring* r, int slot) {
  if (!r || slot < 0 || slot >= r->depth) return bt::fail(BT_E_INVALID, "bt_ring_wait: bad slot");
  if (r->pinned && r->in_flight[slot]) {
    BT_CUDA_TRY(cudaEventSynchronize(r->done[slot]));
    r->in_flight[slot] = 0;
  }
  return release_refs(r->copying[slot]);
}

int bt_ring_acquire(bt_ring* r, int* slot) {
  if (!r || !slot) return bt::fail(BT_E_INVALID, "bt_ring_acquire: bad arguments");
  int s = r->next;
  r->next = (r->next + 1) % r->depth;
  int rc = bt_ring_wait(r, s);
  if (rc != BT_OK) return rc;
  *slot = s;
  return BT_OK;
}

int bt_ring_submit(bt_ring* r, int slot, void* dev_dst, uint64_t nbytes, void* stream) {
  if (!r || slot < 0 || slot >= r->depth || !dev_dst || nbytes > r->slot_bytes)
    return bt::fail(BT_E_INVALID, "bt_ring_submit: bad arguments");
  if (!r->pinned) return bt::fail(BT_E_CUDA, "ring was created without a CUDA device; nothing to copy to");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  BT_CUDA_TRY(cudaMemcpyAsync(dev_dst, r->base + r->slot_bytes * (uint64_t)slot, nbytes, cudaMemcpyHostToDevice, st));
  BT_CUDA_TRY(cudaEventRecord(r->done[slot], st));
  r->in_flight[slot] = 1;
  return BT_OK;
}

int bt_ring_ingest(bt_ring* r, int slot, const bt_schema* s, int item, const uint8_t* msg, uint64_t len) {
  if (!r || !s || slot < 0 || slot >= r->depth || item < 0 || item >= s->batch || s->slot_bytes > r->slot_bytes)
    return bt::fail(BT_E_INVALID, "bt_ring_ingest: bad arguments");
  std::vector<bt_field>& f = bt::tls_fields();
  int rc = bt::scan_fields(msg, len, f);
  if (rc != BT_OK) return rc;
  const size_t before = r->pending[slot].size();
  rc = bt::place_fields(s, r->base + r->slot_bytes * (uint64_t)slot, item, msg, f,
                        r->pinned ? &r->pending[slot] : nullptr, r->device);
  if (rc != BT_OK) {
    // Forget (but do not release) what a rejected message referenced: the caller's compatibility
    // decode fetches those payloads itself and releases them.
    r->pending[slot].resize(before);
  }
  return rc;
}

int bt_ring_spill(bt_ring* r, int slot, const bt_schema* s) {
  if (!r || !s || slot < 0 || slot >= r->depth) return bt::fail(BT_E_INVALID, "bt_ring_spill: bad arguments");
  uint8_t* base = r->base + r->slot_bytes * (uint64_t)slot;
  for (const bt::PendingRef& p : r->pending[slot]) {
    const bt_column& c = s->cols[p.column];
    bt::stream_copy(base + c.slot_off + c.item_bytes * (uint64_t)p.item, bt_shm_slot_ptr(p.seg, p.shm_slot), p.nbytes);
  }
  return release_refs(r->pending[slot]);   // the release doubles as the "still ours after the copy" check
}

int bt_ring_submit_batch(bt_ring* r, int slot, const bt_schema* s, int n_items, void* dev_dst, void* stream) {
  if (!r || !s || slot < 0 || slot >= r->depth || !dev_dst || n_items < 0 || n_items > s->batch ||
      s->slot_bytes > r->slot_bytes)
    return bt::fail(BT_E_INVALID, "bt_ring_submit_batch: bad arguments");
  if (!r->pinned) return bt::fail(BT_E_CUDA, "ring was created without a CUDA device; nothing to copy to");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* host = r->base + r->slot_bytes * (uint64_t)slot;
  uint8_t* dev = static_cast<uint8_t*>(dev_dst);
  std::vector<bt::PendingRef>& pend = r->pending[slot];
  // how many items of each column arrived by reference
  std::vector<int> by_ref(s->cols.size(), 0);
  for (const bt::PendingRef& p : pend) by_ref[p.column]++;
  // A failure part-way must not leave the slot looking free while copies already enqueued still read
  // it, nor leave references behind to be copied twice: whatever happens, the references move to
  // `copying`, the event is recorded and the slot is marked in flight; the first error is returned.
  int rc = BT_OK;
  std::string why;
  auto note = [&](int code) {
    if (code != BT_OK && rc == BT_OK) { rc = code; why = bt::last_error_ref(); }
  };
  auto copy = [&](void* dst, const void* src, uint64_t n) {
    if (rc != BT_OK) return;
    cudaError_t e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) note(bt::fail(BT_E_CUDA, "cudaMemcpyAsync failed: %s", cudaGetErrorString(e)));
  };
  for (const bt::PendingRef& p : pend) note(bt_shm_check(p.seg, p.shm_slot));   // still ours before we read them
  for (size_t ci = 0; ci < s->cols.size(); ++ci) {
    const bt_column& c = s->cols[ci];
    if (by_ref[ci] == n_items && n_items > 0) continue;          // whole column comes from shared memory
    if (by_ref[ci] > 0 && rc == BT_OK) {                         // mixed: bring the referenced items in first
      for (const bt::PendingRef& p : pend)
        if (p.column == (int)ci)
          bt::stream_copy(host + c.slot_off + c.item_bytes * (uint64_t)p.item, bt_shm_slot_ptr(p.seg, p.shm_slot), p.nbytes);
    }
    const uint64_t nbytes = c.item_bytes * (uint64_t)n_items;
    if (nbytes) copy(dev + c.slot_off, host + c.slot_off, nbytes);
  }
  std::vector<bt::PendingRef>& fly = r->copying[slot];
  for (const bt::PendingRef& p : pend) {
    const bt_column& c = s->cols[p.column];
    if (by_ref[p.column] == n_items)
      copy(dev + c.slot_off + c.item_bytes * (uint64_t)p.item, bt_shm_slot_ptr(p.seg, p.shm_slot), p.nbytes);
    fly.push_back(p);   // (mixed columns were copied above; their slots are released with the rest)
  }
  pend.clear();
  cudaError_t e = cudaEventRecord(r->done[slot], st);
  if (e != cudaSuccess) {
    note(bt::fail(BT_E_CUDA, "cudaEventRecord failed: %s", cudaGetErrorString(e)));
    cudaStreamSynchronize(st);   // no event to wait on later: make sure nothing still reads the slot
  }
  r->in_flight[slot] = (e == cudaSuccess);
  if (rc != BT_OK) bt::last_error_ref() = why;
  return rc;
}

int bt_ring_reap(bt_ring* r) {
  if (!r) return bt::fail(BT_E_INVALID, "null ring");
  if (!r->pinned) return BT_OK;
  for (int i = 0; i < r->depth; ++i) {
    if (!r->in_flight[i] || r->copying[i].empty()) continue;
    cudaError_t e = cudaEventQuery(r->done[i]);
    if (e == cudaSuccess) {
      r->in_flight[i] = 0;
      int rc = release_refs(r->copying[i]);
      if (rc != BT_OK) return rc;
    } else if (e != cudaErrorNotReady) {
      return bt::fail(BT_E_CUDA, "cudaEventQuery failed: %s", cudaGetErrorString(e));
    }
  }
  return BT_OK;
}

int bt_device_count(int* n) {
  if (!n) return bt::fail(BT_E_INVALID, "null argument");
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess) {
    *n = 0;
    return bt::fail(BT_E_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  }
  return BT_OK;
}

int bt_gather_h2d(void* dst_dev, const void* src_host, uint64_t row_bytes, const int64_t* index, int n, void* stream) {
  if (!dst_dev || !src_host || !index || n < 0 || row_bytes == 0) return bt::fail(BT_E_INVALID, "bt_gather_h2d: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* dst = static_cast<uint8_t*>(dst_dev);
  const uint8_t* src = static_cast<const uint8_t*>(src_host);
  int i = 0;
  while (i < n) {
    if (index[i] < 0) return bt::fail(BT_E_INVALID, "bt_gather_h2d: negative row index");
    int j = i + 1;
    while (j < n && index[j] == index[j - 1] + 1) ++j;   // rows that are neighbours in the source: one copy
    BT_CUDA_TRY(cudaMemcpyAsync(dst + (uint64_t)i * row_bytes, src + (uint64_t)index[i] * row_bytes,
                                (uint64_t)(j - i) * row_bytes, cudaMemcpyHostToDevice, st));
    i = j;
  }
  return BT_OK;
}

int bt_measure_h2d(int device, uint64_t nbytes, int iters, double* gb_per_s) {
  if (!gb_per_s || nbytes == 0 || iters < 1) return bt::fail(BT_E_INVALID, "bt_measure_h2d: bad arguments");
  BT_CUDA_TRY(cudaSetDevice(device));
  void *h = nullptr, *d = nullptr;
  cudaStream_t st = nullptr;
  cudaEvent_t a = nullptr, b = nullptr;
  int rc = BT_OK;
  float ms = 0;
  cudaError_t e;
#define STEP(x) if ((e = (x)) != cudaSuccess) { rc = bt::fail(BT_E_CUDA, "%s: %s", #x, cudaGetErrorString(e)); goto out; }
  STEP(cudaHostAlloc(&h, nbytes, cudaHostAllocDefault));
  memset(h, 1, nbytes);
  STEP(cudaMalloc(&d, nbytes));
  STEP(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  STEP(cudaEventCreate(&a));
  STEP(cudaEventCreate(&b));
  for (int i = 0; i < 2; ++i) STEP(cudaMemcpyAsync(d, h, nbytes, cudaMemcpyHostToDevice, st));
  STEP(cudaEventRecord(a, st));
  for (int i = 0; i < iters; ++i) STEP(cudaMemcpyAsync(d, h, nbytes, cudaMemcpyHostToDevice, st));
  STEP(cudaEventRecord(b, st));
  STEP(cudaEventSynchronize(b));
  STEP(cudaEventElapsedTime(&ms, a, b));
  *gb_per_s = (double)nbytes * iters / (ms * 1e-3) / 1e9;
#undef STEP
out:
  if (a) cudaEventDestroy(a);
  if (b) cudaEventDestroy(b);
  if (st) cudaStreamDestroy(st);
  if (d) cudaFree(d);
  if (h) cudaFreeHost(h);
  return rc;
}

}  // extern "C"
