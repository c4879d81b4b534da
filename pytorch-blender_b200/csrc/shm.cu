// Shared-memory wire for co-located producers.  No reference counterpart:
// the reference's only transport is the pickled ZMQ message (pkg_blender/blendtorch/btb/publisher.py:43).
//
// Segment layout (POSIX shm, name "/btshm_..."):
//   [Header 64 B][SlotState 64 B x nslots][pad to 4096][slot 0][slot 1]...      slots are 4096-aligned
// A slot is BUSY from the producer's bt_shm_acquire until the consumer's bt_shm_release.  Its state
// is one atomic word, (generation << 1) | busy: every acquire bumps the generation by
// compare-and-swap (several producers may share a segment), and what travels in the message is a
// *ticket* = slot | generation << 16.  A consumer that still holds a ticket whose slot has been taken
// back (the lease below) sees the mismatch before and after its copy or DMA and fails loudly instead of
// delivering another frame's bytes; its late release is a failed CAS, not a second hand-out.  The
// consumer page-locks the whole mapping once with cudaHostRegister, after which slots are valid DMA
// sources for cudaMemcpyAsync.
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>

#include "common.h"

namespace {

constexpr uint64_t kMagic = 0x314d485354425f5fULL;  // "__BTSHM1"

struct Header {
  uint64_t magic;
  uint32_t version, nslots;
  uint64_t slot_bytes;
  uint64_t data_off;
  std::atomic<uint64_t> lease_ns;  // 0: slots are never taken back from a consumer
  uint8_t pad[24];
};
static_assert(sizeof(Header) == 64, "header is one cache line");

struct SlotState {
  std::atomic<uint32_t> word;      // (generation << 1) | busy; generation in 1..kGenMask
  uint32_t pad0;
  std::atomic<uint64_t> since_ns;  // CLOCK_MONOTONIC time of the acquire that made it busy
  uint8_t pad[48];
};
constexpr uint32_t kSegmentVersion = 2;
constexpr uint32_t kGenMask = 0x7FFF;          // 15 bits: a ticket fits a non-negative int
constexpr int kSlotBits = 16;                  // nslots <= 65536
inline uint32_t next_gen(uint32_t g) { return g >= kGenMask ? 1 : g + 1; }
static_assert(sizeof(SlotState) == 64, "one cache line per slot");

}  // namespace

struct bt_shm {
  std::string name;
  uint8_t* base = nullptr;
  uint64_t bytes = 0;
  Header* hdr = nullptr;
  SlotState* state = nullptr;
  bool creator = false;
  bool registered = false;
  uint32_t next = 0;
};

namespace {

std::mutex g_mu;
std::map<std::string, bt_shm*> g_attached;

int map_segment(const char* name, bool create, uint64_t slot_bytes, int nslots, bt_shm** out) {
  if (!name || name[0] != '/') return bt::fail(BT_E_INVALID, "shared-memory names start with '/'");
  int fd = create ? shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600) : shm_open(name, O_RDWR, 0600);
  if (fd < 0) return bt::fail(BT_E_IO, "shm_open(%s): %s", name, strerror(errno));
  uint64_t bytes = 0, data_off = 0;
  if (create) {
    slot_bytes = bt::align_up(slot_bytes, 4096);
    data_off = bt::align_up(sizeof(Header) + sizeof(SlotState) * (uint64_t)nslots, 4096);
    bytes = data_off + slot_bytes * (uint64_t)nslots;
    if (ftruncate(fd, (off_t)bytes) != 0) {
      int e = errno;
      close(fd);
      shm_unlink(name);
      return bt::fail(BT_E_IO, "ftruncate(%s, %llu): %s", name, (unsigned long long)bytes, strerror(e));
    }
  } else {
    struct stat st;
    if (fstat(fd, &st) != 0 || (uint64_t)st.st_size < sizeof(Header)) {
      close(fd);
      return bt::fail(BT_E_IO, "segment %s is not a blendtorch segment", name);
    }
    bytes = (uint64_t)st.st_size;
  }
  void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | (create ? MAP_POPULATE : 0), fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    if (create) shm_unlink(name);
    return bt::fail(BT_E_NOMEM, "mmap(%s, %llu bytes): %s", name, (unsigned long long)bytes, strerror(errno));
  }
  bt_shm* s = new (std::nothrow) bt_shm();
  if (!s) { munmap(p, bytes); return bt::fail(BT_E_NOMEM, "out of memory"); }
  s->name = name;
  s->base = static_cast<uint8_t*>(p);
  s->bytes = bytes;
  s->hdr = reinterpret_cast<Header*>(s->base);
  s->state = reinterpret_cast<SlotState*>(s->base + sizeof(Header));
  s->creator = create;
  if (create) {
    memset(s->base, 0, data_off);
    s->hdr->version = kSegmentVersion;
    uint64_t lease_ms = 30000;   // see bt_shm_set_lease
    if (const char* e = getenv("BT_SHM_LEASE_MS")) lease_ms = strtoull(e, nullptr, 10);
    s->hdr->lease_ns.store(lease_ms * 1000000ull, std::memory_order_relaxed);
    s->hdr->nslots = (uint32_t)nslots;
    s->hdr->slot_bytes = slot_bytes;
    s->hdr->data_off = data_off;
    std::atomic_thread_fence(std::memory_order_release);
    s->hdr->magic = kMagic;
  } else {
    const Header* h = s->hdr;
    // every term is bounded before it is used: a crafted header must not make state[i] or a slot
    // point outside the mapping (nslots <= 65536 keeps the products far from wrapping)
    bool ok = h->magic == kMagic && h->version == kSegmentVersion && h->nslots != 0 && h->nslots <= 65536 &&
              h->slot_bytes != 0 && h->slot_bytes <= bytes && h->data_off <= bytes &&
              h->data_off >= sizeof(Header) + sizeof(SlotState) * (uint64_t)h->nslots;
    if (ok) ok = h->slot_bytes * (uint64_t)h->nslots <= bytes - h->data_off;
    if (!ok) {
      munmap(p, bytes);
      delete s;
      return bt::fail(BT_E_IO, "segment %s has an unknown layout", name);
    }
  }
  *out = s;
  return BT_OK;
}

}  // namespace

namespace bt {

int shm_attach_cached(const char* name, size_t name_len, int device, bt_shm** out) {
  std::string key(name, name_len);
  std::lock_guard<std::mutex> lock(g_mu);
  bt_shm* s = nullptr;
  auto it = g_attached.find(key);
  if (it != g_attached.end()) {
    s = it->second;
  } else {
    int rc = map_segment(key.c_str(), false, 0, 0, &s);
    if (rc != BT_OK) return rc;
    g_attached[key] = s;
  }
  if (device >= 0 && !s->registered) {
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaHostRegister(s->base, s->bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess)
      return fail(BT_E_CUDA, "cudaHostRegister(%s, %llu bytes) failed: %s", key.c_str(),
                  (unsigned long long)s->bytes, cudaGetErrorString(e));
    s->registered = true;
  }
  *out = s;
  return BT_OK;
}

}  // namespace bt

extern "C" {

int bt_shm_create(const char* name, uint64_t slot_bytes, int nslots, bt_shm** out) {
  if (!out || slot_bytes == 0 || nslots < 1 || nslots > 65536) return bt::fail(BT_E_INVALID, "bt_shm_create: bad arguments");
  return map_segment(name, true, slot_bytes, nslots, out);
}

int bt_shm_attach(const char* name, int device, bt_shm** out) {
  if (!name || !out) return bt::fail(BT_E_INVALID, "bt_shm_attach: bad arguments");
  return bt::shm_attach_cached(name, strlen(name), device, out);
}

int bt_shm_close(bt_shm* s) {
  if (!s) return BT_OK;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_attached.find(s->name);
    if (it != g_attached.end() && it->second == s) g_attached.erase(it);
  }
  if (s->registered) cudaHostUnregister(s->base);
  munmap(s->base, s->bytes);
  if (s->creator) shm_unlink(s->name.c_str());
  delete s;
  return BT_OK;
}

int bt_shm_nslots(const bt_shm* s) { return s ? (int)s->hdr->nslots : bt::fail(BT_E_INVALID, "null segment"); }
uint64_t bt_shm_slot_bytes(const bt_shm* s) { return s ? s->hdr->slot_bytes : 0; }

void* bt_shm_slot_ptr(bt_shm* s, int ticket) {
  const int slot = ticket & ((1 << kSlotBits) - 1);
  if (!s || ticket < 0 || slot >= (int)s->hdr->nslots) {
    bt::fail(BT_E_INVALID, "bt_shm_slot_ptr: bad slot");
    return nullptr;
  }
  return s->base + s->hdr->data_off + s->hdr->slot_bytes * (uint64_t)slot;
}

static uint64_t now_ns() {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (uint64_t)t.tv_sec * 1000000000ull + (uint64_t)t.tv_nsec;
}

int bt_shm_set_lease(bt_shm* s, int64_t lease_ms) {
  if (!s || lease_ms < 0) return bt::fail(BT_E_INVALID, "bt_shm_set_lease: bad arguments");
  s->hdr->lease_ns.store((uint64_t)lease_ms * 1000000ull, std::memory_order_relaxed);
  return BT_OK;
}

int bt_shm_acquire(bt_shm* s, int timeout_ms, int* ticket) {
  if (!s || !ticket) return bt::fail(BT_E_INVALID, "bt_shm_acquire: bad arguments");
  const uint32_t n = s->hdr->nslots;
  const uint64_t t0 = now_ns();
  unsigned spins = 0;
  while (true) {
    // first free slot at or after `next`: a slot that is still held does not block the others
    const uint64_t now = now_ns();
    const uint64_t lease = s->hdr->lease_ns.load(std::memory_order_relaxed);
    for (uint32_t k = 0; k < n; ++k) {
      const uint32_t i = (s->next + k) % n;
      SlotState& st = s->state[i];
      uint32_t w = st.word.load(std::memory_order_acquire);
      // Lossy by design and opt-out (lease 0): a slot held longer than the lease is taken back so that
      // consumers that died with messages queued cannot starve the producer; the generation bump makes
      // a consumer that was merely slow notice (bt_shm_check) instead of reading the new frame.
      const bool expired = (w & 1) && lease != 0 && now - st.since_ns.load(std::memory_order_relaxed) > lease;
      if ((w & 1) && !expired) continue;
      const uint32_t gen = next_gen(w >> 1);
      if (!st.word.compare_exchange_strong(w, (gen << 1) | 1u, std::memory_order_acq_rel)) continue;  // another producer won
      st.since_ns.store(now, std::memory_order_relaxed);
      s->next = (i + 1) % n;
      *ticket = (int)(i | (gen << kSlotBits));
      return BT_OK;
    }
    if (timeout_ms == 0) return BT_E_TIMEOUT;   // a poll: no message recorded
    if (++spins < 100) continue;
    struct timespec nap = {0, 20000};  // 20 us
    nanosleep(&nap, nullptr);
    if (timeout_ms > 0 && (now_ns() - t0) / 1000000ull > (uint64_t)timeout_ms)
      return bt::fail(BT_E_TIMEOUT, "no free shared-memory slot within %d ms (consumers stalled?)", timeout_ms);
  }
}

int bt_shm_check(const bt_shm* s, int ticket) {
  const int slot = ticket & ((1 << kSlotBits) - 1);
  if (!s || ticket < 0 || slot >= (int)s->hdr->nslots) return bt::fail(BT_E_INVALID, "bt_shm_check: bad ticket");
  const uint32_t gen = (uint32_t)ticket >> kSlotBits;
  if (gen == 0) return BT_OK;   // a bare slot number: nothing to compare
  if (s->state[slot].word.load(std::memory_order_acquire) != ((gen << 1) | 1u))
    return bt::fail(BT_E_IO, "shared-memory slot %d of %s was taken back by its producer (held past the lease) "
                             "while this consumer was still using it", slot, s->name.c_str());
  return BT_OK;
}

int bt_shm_release(bt_shm* s, int ticket) {
  const int slot = ticket & ((1 << kSlotBits) - 1);
  if (!s || ticket < 0 || slot >= (int)s->hdr->nslots) return bt::fail(BT_E_INVALID, "bt_shm_release: bad slot");
  const uint32_t gen = (uint32_t)ticket >> kSlotBits;
  std::atomic<uint32_t>& word = s->state[slot].word;
  uint32_t w = word.load(std::memory_order_relaxed);
  if (gen == 0) {   // a bare slot number (tools, drain): free whatever generation holds it
    while ((w & 1) && !word.compare_exchange_weak(w, w & ~1u, std::memory_order_release)) {}
    return BT_OK;
  }
  uint32_t expect = (gen << 1) | 1u;
  if (!word.compare_exchange_strong(expect, gen << 1, std::memory_order_release))
    return bt::fail(BT_E_IO, "stale release of shared-memory slot %d of %s (taken back by its producer)", slot, s->name.c_str());
  return BT_OK;
}

}  // extern "C"
