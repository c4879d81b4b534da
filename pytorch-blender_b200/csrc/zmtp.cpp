// Direct receiver: a minimal ZMTP 3.x PULL endpoint that lands payloads straight from the kernel
// socket in the pinned staging slot.
//
// The reference receives through libzmq (`socket.recv()` at pkg_pytorch/blendtorch/btt/dataset.py:101,105):
// kernel socket buffer -> libzmq message buffer -> Python bytes -> ndarray.  rx.cpp already removes the
// Python part; this file also removes libzmq's intermediate buffer for the one thing that matters,
// the frame payload: the message is scanned *while it arrives* (scan.cpp, streaming mode), and as
// soon as the scanner stands on the raw data of an array that matches a schema column, the rest of
// that payload is recv()'d directly into its final position in the slot.  One CPU copy per payload
// byte on the consumer side (kernel -> pinned) instead of two.
//
// Wire compatibility: the peers are unmodified ZMQ PUSH sockets (btb/publisher.py:21-27: PUSH, bind).
// Implemented: ZMTP 3.0/3.1 greeting, NULL security handshake (READY with Socket-Type=PULL),
// single-frame messages (the reference wire) and multi-frame ones (the opt-in raw-parts wire: pickled
// header frame + one frame per large array, received straight into the slot), commands are skipped,
// reconnect with back-off, fair round-robin over peers, tcp:// and ipc:// endpoints.  Other security
// mechanisms are reported and the caller falls back to the libzmq receiver (rx.cpp).
#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <new>
#include <string>
#include <vector>

#include "common.h"

namespace bt {
// ring.cu: place pre-scanned fields of one message into a ring slot (handles shared-memory refs)
int ring_place(bt_ring* r, int slot, const bt_schema* s, int item, const uint8_t* msg, const std::vector<bt_field>& f);
uint8_t* ring_slot_base(bt_ring* r, int slot);
}  // namespace bt

namespace {

uint64_t now_ms() {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (uint64_t)t.tv_sec * 1000ull + (uint64_t)t.tv_nsec / 1000000ull;
}

struct Peer {
  std::string address;
  int family = AF_INET;
  sockaddr_storage addr;
  socklen_t addrlen = 0;
  int fd = -1;
  enum State { IDLE, CONNECTING, READY } state = IDLE;
  uint64_t retry_at = 0;
  uint32_t backoff_ms = 20;
  // Read-ahead: every socket read asks for as much as fits here, so the frame header, the pickle
  // prologue and the few bytes after a payload (plus the next message's header) cost one recv()
  // instead of one each.  Only what is not array payload passes through it: payload bytes beyond
  // what a read-ahead happened to bring along are still received straight into the staging slot.
  std::vector<uint8_t> ahead;
  size_t rpos = 0, wpos = 0;
};

constexpr size_t kAheadBytes = 8192 + 64;

struct External {
  uint64_t compact_pos;
  uint64_t len;
  const uint8_t* where;
};

// diagnostic counters (printed at bt_pull_destroy when BT_RX_STATS is set)
thread_local uint64_t g_recv_calls = 0, g_poll_calls = 0;

constexpr int kIoTimeoutMs = 10000;  // a message that started arriving must complete within this

// Largest frame the endpoint accepts.  Sizes come off the wire and go into resize()/reserve(): a
// garbage or hostile peer must not be able to ask for terabytes.  4 GiB by default (a 1080p RGBA frame
// with a float32 depth map is 16.6 MB), BT_MAX_MESSAGE_BYTES overrides.
uint64_t max_frame_bytes() {
  static const uint64_t v = [] {
    const char* e = getenv("BT_MAX_MESSAGE_BYTES");
    const uint64_t n = e ? strtoull(e, nullptr, 10) : 0;
    return n ? n : (4ull << 30);
  }();
  return v;
}

void close_peer(Peer& p) {
  if (p.fd >= 0) close(p.fd);
  p.fd = -1;
  p.rpos = p.wpos = 0;   // bytes of a broken connection are not the start of anything
  p.state = Peer::IDLE;
  p.retry_at = now_ms() + p.backoff_ms;
  if (p.backoff_ms < 500) p.backoff_ms *= 2;
}

// Blocking-with-deadline I/O on a non-blocking fd.
int read_exact(int fd, uint8_t* dst, uint64_t n, int timeout_ms) {
  const uint64_t deadline = now_ms() + (uint64_t)timeout_ms;
  while (n) {
    ssize_t k = recv(fd, dst, n > (1u << 30) ? (1u << 30) : (size_t)n, 0);
    ++g_recv_calls;
    if (k > 0) { dst += k; n -= (uint64_t)k; continue; }
    if (k == 0) return bt::fail(BT_E_IO, "peer closed the connection");
    if (errno == EINTR) continue;
    if (errno != EAGAIN && errno != EWOULDBLOCK) return bt::fail(BT_E_IO, "recv: %s", strerror(errno));
    const uint64_t now = now_ms();
    if (now >= deadline) return bt::fail(BT_E_IO, "peer stalled in the middle of a message");
    pollfd pf = {fd, POLLIN, 0};
    poll(&pf, 1, (int)(deadline - now));
    ++g_poll_calls;
  }
  return BT_OK;
}

int write_all(int fd, const uint8_t* src, size_t n, int timeout_ms) {
  const uint64_t deadline = now_ms() + (uint64_t)timeout_ms;
  while (n) {
    ssize_t k = send(fd, src, n, MSG_NOSIGNAL);
    if (k > 0) { src += k; n -= (size_t)k; continue; }
    if (k < 0 && errno == EINTR) continue;
    if (k < 0 && errno != EAGAIN && errno != EWOULDBLOCK) return bt::fail(BT_E_IO, "send: %s", strerror(errno));
    const uint64_t now = now_ms();
    if (now >= deadline) return bt::fail(BT_E_IO, "peer does not accept the handshake");
    pollfd pf = {fd, POLLOUT, 0};
    poll(&pf, 1, (int)(deadline - now));
  }
  return BT_OK;
}

inline size_t ahead_avail(const Peer& p) { return p.wpos - p.rpos; }

// Diagnostic switch (A/B measurements): BT_RX_READAHEAD=0 makes every read ask for exactly the bytes
// it needs, as the first version of this endpoint did.
bool readahead_enabled() {
  static const bool on = [] {
    const char* e = getenv("BT_RX_READAHEAD");
    return !(e && e[0] == '0');
  }();
  return on;
}

// One non-blocking recv() into the free part of the read-ahead buffer.
// > 0: bytes read, 0: nothing there right now, < 0: the connection is unusable.
int fill_once(Peer& p) {
  if (p.ahead.size() != kAheadBytes) p.ahead.resize(kAheadBytes);
  if (p.rpos == p.wpos) p.rpos = p.wpos = 0;
  if (p.wpos == kAheadBytes) {
    if (p.rpos == 0) return 0;   // full
    memmove(p.ahead.data(), p.ahead.data() + p.rpos, p.wpos - p.rpos);
    p.wpos -= p.rpos;
    p.rpos = 0;
  }
  while (true) {
    ssize_t k = recv(p.fd, p.ahead.data() + p.wpos, kAheadBytes - p.wpos, 0);
    if (k > 0) { p.wpos += (size_t)k; return (int)k; }
    if (k == 0) return bt::fail(BT_E_IO, "peer closed the connection");
    if (errno == EINTR) continue;
    if (errno == EAGAIN || errno == EWOULDBLOCK) return 0;
    return bt::fail(BT_E_IO, "recv: %s", strerror(errno));
  }
}

// Make at least n (<= 64) bytes available in the read-ahead buffer.
int need(Peer& p, size_t n, int timeout_ms) {
  const uint64_t deadline = now_ms() + (uint64_t)timeout_ms;
  while (ahead_avail(p) < n) {
    if (p.wpos == kAheadBytes && p.rpos > 0) {   // make room at the end
      memmove(p.ahead.data(), p.ahead.data() + p.rpos, p.wpos - p.rpos);
      p.wpos -= p.rpos;
      p.rpos = 0;
    }
    const int k = fill_once(p);
    if (k < 0) return k;
    if (k > 0) continue;
    const uint64_t now = now_ms();
    if (now >= deadline) return bt::fail(BT_E_IO, "peer stalled in the middle of a message");
    pollfd pf = {p.fd, POLLIN, 0};
    poll(&pf, 1, (int)(deadline - now));
  }
  return BT_OK;
}

// n bytes of the stream to dst.  Small requests (headers, pickle prologue, the bytes after a payload)
// go through the read-ahead buffer: when it is empty the socket is asked for a buffer-full, because
// whatever follows is needed next anyway.  Large requests (array payloads) take what the buffer
// happens to hold and recv() the rest straight into dst.
int take(Peer& p, uint8_t* dst, uint64_t n, int timeout_ms, bool small) {
  while (n) {
    size_t have = ahead_avail(p);
    if (have == 0) {
      if (!small || !readahead_enabled()) return read_exact(p.fd, dst, n, timeout_ms);
      int rc = need(p, 1, timeout_ms);
      if (rc != BT_OK) return rc;
      have = ahead_avail(p);
    }
    const size_t k = have < n ? have : (size_t)n;
    memcpy(dst, p.ahead.data() + p.rpos, k);
    p.rpos += k;
    dst += k;
    n -= k;
  }
  return BT_OK;
}

int parse_address(const std::string& a, Peer& p) {
  memset(&p.addr, 0, sizeof p.addr);
  if (a.rfind("tcp://", 0) == 0) {
    const std::string rest = a.substr(6);
    const size_t colon = rest.rfind(':');
    if (colon == std::string::npos) return bt::fail(BT_E_INVALID, "address %s has no port", a.c_str());
    std::string host = rest.substr(0, colon), port = rest.substr(colon + 1);
    if (!host.empty() && host.front() == '[' && host.back() == ']') host = host.substr(1, host.size() - 2);
    if (host == "localhost") host = "127.0.0.1";
    addrinfo hints;
    memset(&hints, 0, sizeof hints);
    hints.ai_family = AF_UNSPEC;
    hints.ai_socktype = SOCK_STREAM;
    addrinfo* res = nullptr;
    if (getaddrinfo(host.c_str(), port.c_str(), &hints, &res) != 0 || !res)
      return bt::fail(BT_E_INVALID, "cannot resolve %s", a.c_str());
    memcpy(&p.addr, res->ai_addr, res->ai_addrlen);
    p.addrlen = res->ai_addrlen;
    p.family = res->ai_family;
    freeaddrinfo(res);
    return BT_OK;
  }
  if (a.rfind("ipc://", 0) == 0) {
    const std::string path = a.substr(6);
    sockaddr_un* un = reinterpret_cast<sockaddr_un*>(&p.addr);
    if (path.empty() || path.size() >= sizeof un->sun_path) return bt::fail(BT_E_INVALID, "bad ipc path in %s", a.c_str());
    un->sun_family = AF_UNIX;
    memcpy(un->sun_path, path.c_str(), path.size());
    p.addrlen = (socklen_t)(offsetof(sockaddr_un, sun_path) + path.size());
    if (path[0] == '@') un->sun_path[0] = '\0';  // abstract namespace, as libzmq spells it
    else p.addrlen += 1;
    p.family = AF_UNIX;
    return BT_OK;
  }
  return bt::fail(BT_E_UNSUPPORTED, "the direct receiver speaks tcp:// and ipc:// only (%s)", a.c_str());
}

// ZMTP 3.0 greeting + NULL handshake, synchronously (the peer is a live libzmq engine).
int handshake(int fd) {
  uint8_t g[64];
  memset(g, 0, sizeof g);
  g[0] = 0xFF; g[8] = 0x01; g[9] = 0x7F;   // signature
  g[10] = 3; g[11] = 0;                      // version 3.0
  memcpy(g + 12, "NULL", 4);                 // mechanism, zero padded to 20 bytes; as-server = 0; filler
  int rc = write_all(fd, g, sizeof g, 2000);
  if (rc != BT_OK) return rc;
  uint8_t peer[64];
  rc = read_exact(fd, peer, sizeof peer, 2000);
  if (rc != BT_OK) return rc;
  if (peer[0] != 0xFF || peer[9] != 0x7F || peer[10] < 3)
    return bt::fail(BT_E_UNSUPPORTED, "peer does not speak ZMTP 3");
  if (memcmp(peer + 12, "NULL\0", 5) != 0)
    return bt::fail(BT_E_UNSUPPORTED, "peer uses a security mechanism other than NULL");
  static const uint8_t ready[] = {0x04, 26, 5, 'R', 'E', 'A', 'D', 'Y', 11, 'S', 'o', 'c', 'k', 'e', 't', '-', 'T', 'y', 'p', 'e',
                                  0, 0, 0, 4, 'P', 'U', 'L', 'L'};
  rc = write_all(fd, ready, sizeof ready, 2000);
  if (rc != BT_OK) return rc;
  // the peer's READY (a command frame)
  uint8_t h[9];
  rc = read_exact(fd, h, 2, 2000);
  if (rc != BT_OK) return rc;
  if (!(h[0] & 0x04)) return bt::fail(BT_E_UNSUPPORTED, "expected a ZMTP command, got flags 0x%02x", h[0]);
  uint64_t size = h[1];
  if (h[0] & 0x02) {
    rc = read_exact(fd, h + 2, 7, 2000);
    if (rc != BT_OK) return rc;
    size = 0;
    for (int i = 1; i < 9; ++i) size = (size << 8) | h[i];
  }
  if (size < 6 || size > 4096) return bt::fail(BT_E_UNSUPPORTED, "unexpected handshake command size");
  std::vector<uint8_t> body(size);
  rc = read_exact(fd, body.data(), size, 2000);
  if (rc != BT_OK) return rc;
  if (body[0] != 5 || memcmp(&body[1], "READY", 5) != 0) return bt::fail(BT_E_UNSUPPORTED, "peer refused the connection");
  // a PULL socket takes messages from PUSH peers only (ZMTP 3.0 socket type compatibility)
  static const char want[] = "\x0bSocket-Type\0\0\0\x04PUSH";
  bool push = false;
  for (uint64_t i = 6; i + sizeof want - 1 <= size; ++i)
    if (memcmp(body.data() + i, want, sizeof want - 1) == 0) push = true;
  if (!push) return bt::fail(BT_E_UNSUPPORTED, "peer is not a PUSH socket");
  return BT_OK;
}

}  // namespace

struct bt_pull {
  std::vector<Peer> peers;
  size_t rr = 0;
  std::vector<uint8_t> cb;    // message in progress, large payloads cut out
  std::vector<uint8_t> held;  // a complete message kept for the caller (all frames back to back)
  std::vector<uint64_t> held_parts;  // sizes of its frames (one entry for ordinary single-frame messages)
  bool has_held = false;
  std::vector<External> ext;
  std::vector<uint8_t> spill;   // frames of a raw-parts message that could not go straight to the slot
  std::vector<pollfd> pfds;
  uint64_t messages = 0, bytes = 0;
};

namespace {

void start_connect(Peer& p) {
  p.fd = socket(p.family, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
  if (p.fd < 0) { p.retry_at = now_ms() + 200; return; }
  static const int rcvbuf = [] {
    const char* e = getenv("BT_PULL_RCVBUF");   // bytes; tuning knob
    return e ? atoi(e) : (1 << 20);
  }();
  if (rcvbuf > 0) setsockopt(p.fd, SOL_SOCKET, SO_RCVBUF, &rcvbuf, sizeof rcvbuf);
  int rc = connect(p.fd, reinterpret_cast<sockaddr*>(&p.addr), p.addrlen);
  if (rc == 0 || errno == EINPROGRESS) {
    p.state = Peer::CONNECTING;
  } else {
    close_peer(p);
  }
}

// A TCP socket that keeps trying a local port on which nobody listens yet can end up connected to
// *itself*: when the kernel happens to pick the destination port as the source port, the connection
// opens simultaneously with itself.  It would then pass its own greeting back to us, look like a
// publisher that never sends anything, and -- worse -- occupy the very port the publisher is about to
// bind ("Address already in use" in the publisher).  Consumers started before their producers, as
// blendtorch's launch order allows, retry often enough to meet this.
bool connected_to_itself(int fd) {
  sockaddr_storage a, b;
  socklen_t la = sizeof a, lb = sizeof b;
  if (getsockname(fd, reinterpret_cast<sockaddr*>(&a), &la) != 0 || getpeername(fd, reinterpret_cast<sockaddr*>(&b), &lb) != 0)
    return false;
  if (a.ss_family != b.ss_family) return false;
  if (a.ss_family == AF_INET) {
    const sockaddr_in* x = reinterpret_cast<const sockaddr_in*>(&a);
    const sockaddr_in* y = reinterpret_cast<const sockaddr_in*>(&b);
    return x->sin_port == y->sin_port && x->sin_addr.s_addr == y->sin_addr.s_addr;
  }
  if (a.ss_family == AF_INET6) {
    const sockaddr_in6* x = reinterpret_cast<const sockaddr_in6*>(&a);
    const sockaddr_in6* y = reinterpret_cast<const sockaddr_in6*>(&b);
    return x->sin6_port == y->sin6_port && memcmp(&x->sin6_addr, &y->sin6_addr, sizeof x->sin6_addr) == 0;
  }
  return false;
}

void finish_connect(Peer& p) {
  int err = 0;
  socklen_t len = sizeof err;
  if (getsockopt(p.fd, SOL_SOCKET, SO_ERROR, &err, &len) != 0 || err != 0 || connected_to_itself(p.fd) ||
      handshake(p.fd) != BT_OK) {
    close_peer(p);
    return;
  }
  p.state = Peer::READY;
  p.backoff_ms = 20;
}

// Drive the connections; return the index of a READY peer with bytes to read, or -1 after timeout_ms.
int wait_readable(bt_pull* q, int timeout_ms) {
  const uint64_t deadline = now_ms() + (uint64_t)(timeout_ms < 0 ? 0 : timeout_ms);
  const size_t npeers = q->peers.size();
  // Fast path while every connection is up: the peer whose turn it is usually has bytes waiting --
  // read ahead earlier, or in its socket, which one speculative recv() finds out more cheaply than
  // a poll() over all peers.  Turn order is kept, so fair queuing is too.
  bool all_ready = npeers > 0;
  for (const Peer& p : q->peers) all_ready = all_ready && p.state == Peer::READY;
  if (all_ready && readahead_enabled()) {
    int tries = 0;
    for (size_t k = 0; k < npeers && tries < 2; ++k) {
      const size_t i = (q->rr + k) % npeers;
      Peer& p = q->peers[i];
      if (ahead_avail(p) == 0) {
        ++tries;
        const int got = fill_once(p);
        if (got < 0) { close_peer(p); break; }
        if (got == 0) continue;
      }
      q->rr = (i + 1) % npeers;
      return (int)i;
    }
  }
  while (true) {
    const uint64_t now = now_ms();
    q->pfds.clear();
    uint64_t next_retry = ~0ull;
    for (Peer& p : q->peers) {
      if (p.state == Peer::IDLE && now >= p.retry_at) start_connect(p);
      pollfd pf = {-1, 0, 0};
      if (p.state == Peer::CONNECTING) pf = {p.fd, POLLOUT, 0};
      else if (p.state == Peer::READY) pf = {p.fd, POLLIN, 0};
      else if (p.retry_at < next_retry) next_retry = p.retry_at;
      q->pfds.push_back(pf);
    }
    int wait = now >= deadline ? 0 : (int)(deadline - now);
    for (const Peer& p : q->peers)
      if (p.state == Peer::READY && ahead_avail(p) > 0) wait = 0;   // something is readable already
    if (next_retry != ~0ull) {
      const int until = next_retry > now ? (int)(next_retry - now) : 0;
      if (until < wait) wait = until;
    }
    poll(q->pfds.data(), q->pfds.size(), wait);
    const size_t n = q->peers.size();
    for (size_t k = 0; k < n; ++k) {
      Peer& p = q->peers[k];
      if (p.state == Peer::CONNECTING && (q->pfds[k].revents & (POLLOUT | POLLERR | POLLHUP))) finish_connect(p);
    }
    for (size_t k = 0; k < n; ++k) {
      const size_t i = (q->rr + k) % n;
      Peer& p = q->peers[i];
      if (p.state != Peer::READY) continue;
      if ((q->pfds[i].revents & POLLIN) || ahead_avail(p) > 0) {
        q->rr = (i + 1) % n;
        return (int)i;
      }
      if (q->pfds[i].revents & (POLLERR | POLLHUP | POLLNVAL)) close_peer(p);
    }
    if (now_ms() >= deadline) return -1;
  }
}

// Read one frame header.  *is_msg = 0 for a command (skipped), 1 for a message frame of `*size` bytes;
// *more = 1 when further frames of the same message follow (raw-parts wire).
int read_frame_header(Peer& p, int* is_msg, uint64_t* size, int* more) {
  uint8_t h[9];
  int rc = take(p, h, 2, kIoTimeoutMs, true);
  if (rc != BT_OK) return rc;
  uint64_t sz = h[1];
  if (h[0] & 0x02) {
    rc = take(p, h + 2, 7, kIoTimeoutMs, true);
    if (rc != BT_OK) return rc;
    sz = 0;
    for (int i = 1; i < 9; ++i) sz = (sz << 8) | h[i];
  }
  if (h[0] & 0x04) {  // command (PING, ERROR, ...): skip
    std::vector<uint8_t> skip(sz > 65536 ? 65536 : sz);
    uint64_t left = sz;
    while (left) {
      const uint64_t k = left < skip.size() ? left : skip.size();
      rc = take(p, skip.data(), k, kIoTimeoutMs, true);
      if (rc != BT_OK) return rc;
      left -= k;
    }
    *is_msg = 0;
    return BT_OK;
  }
  if (sz > max_frame_bytes())
    return bt::fail(BT_E_IO, "frame of %llu bytes exceeds the limit of %llu (BT_MAX_MESSAGE_BYTES)",
                    (unsigned long long)sz, (unsigned long long)max_frame_bytes());
  *is_msg = 1;
  *size = sz;
  *more = h[0] & 0x01;
  return BT_OK;
}

constexpr int kMaxParts = 64;            // frames per message the endpoint accepts
constexpr uint64_t kMaxHeader = 1 << 20;  // pickled header frame of a raw-parts message

// Append the remaining frames of a multipart message to q->held (sizes to q->held_parts).
int read_more_frames(bt_pull* q, Peer& p) {
  while (true) {
    int is_msg = 0, more = 0;
    uint64_t size = 0;
    int rc = read_frame_header(p, &is_msg, &size, &more);
    if (rc != BT_OK) return rc;
    if (!is_msg) continue;
    if ((int)q->held_parts.size() >= kMaxParts) return bt::fail(BT_E_IO, "message has more than %d frames", kMaxParts);
    const size_t at = q->held.size();
    if (at + size > max_frame_bytes()) return bt::fail(BT_E_IO, "message exceeds the size limit (BT_MAX_MESSAGE_BYTES)");
    q->held.resize(at + size);
    rc = take(p, q->held.data() + at, size, kIoTimeoutMs, size <= kAheadBytes);
    if (rc != BT_OK) return rc;
    q->held_parts.push_back(size);
    if (!more) return BT_OK;
  }
}

// Whole message (every frame) into q->held.
int read_whole(bt_pull* q, Peer& p, uint64_t size, int more) {
  q->held.resize(size);
  q->held_parts.assign(1, size);
  int rc = take(p, q->held.data(), size, kIoTimeoutMs, size <= kAheadBytes);
  if (rc != BT_OK) return rc;
  if (more) {
    rc = read_more_frames(q, p);
    if (rc != BT_OK) return rc;
  }
  q->has_held = true;
  return BT_OK;
}

// Rebuild the original bytes of a partly streamed message (compact buffer + landed payloads) and
// read the rest of it, so that the caller can hand it to pickle.loads.
int rebuild_into_held(bt_pull* q, Peer& p, uint64_t msg_size, uint64_t consumed) {
  std::vector<uint8_t>& out = q->held;
  out.clear();
  out.reserve(msg_size);
  uint64_t from = 0;
  for (const External& e : q->ext) {
    out.insert(out.end(), q->cb.begin() + from, q->cb.begin() + e.compact_pos);
    out.insert(out.end(), e.where, e.where + e.len);
    from = e.compact_pos;
  }
  out.insert(out.end(), q->cb.begin() + from, q->cb.end());
  const uint64_t have = out.size();
  if (have != consumed) return bt::fail(BT_E_IO, "internal: streamed byte count mismatch");
  out.resize(msg_size);
  int rc = take(p, out.data() + have, msg_size - have, kIoTimeoutMs, msg_size - have <= kAheadBytes);
  if (rc != BT_OK) return rc;
  q->held_parts.assign(1, msg_size);
  q->has_held = true;
  return BT_OK;
}

// Receive one message from `p` and place it as item `item` of the slot, payloads going straight
// from the socket into the slot.  BT_OK: placed.  BT_E_UNSUPPORTED / BT_E_SCHEMA: the message is in
// q->held for the caller's pickle.loads path.  BT_E_IO: the connection is unusable.
int receive_streaming(bt_pull* q, Peer& p, uint64_t msg_size, bt_ring* ring, int slot, const bt_schema* schema, int item) {
  constexpr uint64_t kChunk = 4096;
  uint8_t* base = bt::ring_slot_base(ring, slot);
  q->cb.clear();
  q->ext.clear();
  bt::stream_begin();
  uint64_t consumed = 0;
  int why = BT_OK;       // != BT_OK: why this message has to go through the caller's decode
  bool complete = false;
  while (!complete && why == BT_OK) {
    const uint64_t want = msg_size - consumed < kChunk ? msg_size - consumed : kChunk;
    if (want == 0) { why = BT_E_UNSUPPORTED; break; }  // the pickle wants more bytes than the frame has
    const size_t old = q->cb.size();
    q->cb.resize(old + want);
    int rc = take(p, q->cb.data() + old, want, kIoTimeoutMs, true);
    if (rc != BT_OK) return rc;
    consumed += want;
    // run the scanner over what is buffered; it may stop on several large payloads in a row
    while (true) {
      rc = bt::stream_run(q->cb.data(), q->cb.size());
      if (rc == BT_E_TRUNCATED) break;                 // needs the next chunk
      if (rc == BT_OK) { complete = true; break; }     // STOP reached
      if (rc != 1) { why = rc; break; }                // outside the subset
      const uint64_t len = bt::stream_big_len(), pos = bt::stream_pos();
      const uint64_t have = q->cb.size() - pos;        // payload bytes that came with the last chunk
      int col = -1;
      const int trc = bt::stream_big_target(schema, &col);
      if (trc != BT_OK || len - have > msg_size - consumed) {
        // cannot land it: its first bytes stay at the end of the compact buffer, which is where
        // they belong in the original byte stream, so the rebuild below stays correct
        why = trc == BT_E_SCHEMA ? BT_E_SCHEMA : BT_E_UNSUPPORTED;
        break;
      }
      const bt_column& c = schema->cols[col];
      uint8_t* dst = base + c.slot_off + c.item_bytes * (uint64_t)item;
      memcpy(dst, q->cb.data() + pos, have);
      q->cb.resize(pos);
      rc = take(p, dst + have, len - have, kIoTimeoutMs, false);
      if (rc != BT_OK) return rc;
      consumed += len - have;
      q->ext.push_back(External{pos, len, dst});
    }
  }
  if (why == BT_OK) {
    if (consumed < msg_size) {  // bytes after the pickle's STOP: not ours to interpret
      std::vector<uint8_t> tail(msg_size - consumed);
      int rc = take(p, tail.data(), tail.size(), kIoTimeoutMs, tail.size() <= kAheadBytes);
      if (rc != BT_OK) return rc;
      q->cb.insert(q->cb.end(), tail.begin(), tail.end());
      consumed = msg_size;
    }
    std::vector<bt_field>& f = bt::tls_fields();
    int rc = bt::stream_fields(q->cb.data(), f);
    if (rc == BT_OK) rc = bt::ring_place(ring, slot, schema, item, q->cb.data(), f);
    if (rc == BT_OK) return BT_OK;
    why = rc == BT_E_SCHEMA ? BT_E_SCHEMA : BT_E_UNSUPPORTED;
  }
  const std::string reason = bt::last_error_ref();
  int rc = rebuild_into_held(q, p, msg_size, consumed);
  if (rc != BT_OK) return rc;
  bt::last_error_ref() = reason;
  return why;
}

// Raw-parts message (DataPublisher(raw_parts=True)): frame 0 is the pickled header, in which every
// large array is a ("__btraw__", k, dtype, shape) marker; frame k carries that array's bytes.  Frames
// that match a schema column are received straight into the slot -- no scan of the payload, no
// intermediate buffer on either side.  Return values as receive_streaming; on BT_E_UNSUPPORTED /
// BT_E_SCHEMA the whole message (all frames, back to back) is in q->held, sizes in q->held_parts.
int receive_raw_parts(bt_pull* q, Peer& p, uint64_t header_size, bt_ring* ring, int slot, const bt_schema* schema, int item) {
  if (header_size > kMaxHeader) {
    int rc = read_whole(q, p, header_size, 1);
    if (rc != BT_OK) return rc;
    return bt::fail(BT_E_UNSUPPORTED, "header frame of %llu bytes", (unsigned long long)header_size);
  }
  q->cb.resize(header_size);
  int rc = take(p, q->cb.data(), header_size, kIoTimeoutMs, true);
  if (rc != BT_OK) return rc;
  std::vector<bt_field>& f = bt::tls_fields();
  int why = bt::scan_fields(q->cb.data(), header_size, f);
  std::string reason = why == BT_OK ? std::string() : bt::last_error_ref();
  uint8_t* base = bt::ring_slot_base(ring, slot);
  struct Part { uint64_t size; const uint8_t* landed; uint64_t spill_off; };
  std::vector<Part> parts;
  q->spill.clear();
  int more = 1;
  while (more) {
    int is_msg = 0;
    uint64_t size = 0;
    rc = read_frame_header(p, &is_msg, &size, &more);
    if (rc != BT_OK) return rc;
    if (!is_msg) { more = 1; continue; }
    const int index = (int)parts.size() + 1;
    if (index >= kMaxParts) return bt::fail(BT_E_IO, "message has more than %d frames", kMaxParts);
    uint8_t* dst = nullptr;
    bt_field* owner = nullptr;
    if (why == BT_OK) {
      for (bt_field& fi : f)
        if (fi.kind == BT_KIND_RAWREF && fi.ival == index && fi.payload_len == 0) { owner = &fi; break; }
      if (owner) {
        for (const bt_column& c : schema->cols) {
          if (strcmp(c.key, owner->key) != 0) continue;
          if (c.kind == BT_KIND_NDARRAY && c.dtype == owner->dtype && c.ndim == owner->ndim && !owner->fortran &&
              memcmp(c.shape, owner->shape, sizeof(int64_t) * c.ndim) == 0 && c.item_bytes == size)
            dst = base + c.slot_off + c.item_bytes * (uint64_t)item;
          break;
        }
      }
      if (!dst) {
        why = BT_E_SCHEMA;
        char msg[160];
        snprintf(msg, sizeof msg, "frame %d of %llu bytes matches no column of the schema", index, (unsigned long long)size);
        reason = msg;
      }
    }
    if (dst) {
      rc = take(p, dst, size, kIoTimeoutMs, false);
      if (rc != BT_OK) return rc;
      owner->payload_off = bt::kLandedOff;
      owner->payload_len = size;
      parts.push_back(Part{size, dst, 0});
    } else {
      const uint64_t at = q->spill.size();
      if (at + size > max_frame_bytes()) return bt::fail(BT_E_IO, "message exceeds the size limit (BT_MAX_MESSAGE_BYTES)");
      q->spill.resize(at + size);
      rc = take(p, q->spill.data() + at, size, kIoTimeoutMs, size <= kAheadBytes);
      if (rc != BT_OK) return rc;
      parts.push_back(Part{size, nullptr, at});
    }
  }
  if (why == BT_OK) {
    rc = bt::ring_place(ring, slot, schema, item, q->cb.data(), f);
    if (rc == BT_OK) return BT_OK;
    why = rc == BT_E_SCHEMA ? BT_E_SCHEMA : BT_E_UNSUPPORTED;
    reason = bt::last_error_ref();
  }
  // hand the message to the caller's decode: every frame back to back
  q->held.assign(q->cb.begin(), q->cb.end());
  q->held_parts.assign(1, header_size);
  for (const Part& part : parts) {
    const uint8_t* src = part.landed ? part.landed : q->spill.data() + part.spill_off;
    q->held.insert(q->held.end(), src, src + part.size);
    q->held_parts.push_back(part.size);
  }
  q->has_held = true;
  bt::last_error_ref() = reason;
  return why == BT_E_SCHEMA ? BT_E_SCHEMA : BT_E_UNSUPPORTED;
}

// Place the held message (received before the schema existed) as item `item` of the slot.
int ingest_held(bt_pull* q, bt_ring* ring, int slot, const bt_schema* schema, int item) {
  if (q->held_parts.size() <= 1) return bt_ring_ingest(ring, slot, schema, item, q->held.data(), q->held.size());
  std::vector<bt_field>& f = bt::tls_fields();
  int rc = bt::scan_fields(q->held.data(), q->held_parts[0], f);
  if (rc != BT_OK) return rc;
  bt::resolve_raw_parts(f, q->held_parts.data(), (int)q->held_parts.size());
  return bt::ring_place(ring, slot, schema, item, q->held.data(), f);
}

}  // namespace

extern "C" {

int bt_pull_create(bt_pull** out) {
  if (!out) return bt::fail(BT_E_INVALID, "bt_pull_create: null argument");
  bt_pull* q = new (std::nothrow) bt_pull();
  if (!q) return bt::fail(BT_E_NOMEM, "out of memory");
  *out = q;
  return BT_OK;
}

int bt_pull_connect(bt_pull* q, const char* address) {
  if (!q || !address) return bt::fail(BT_E_INVALID, "bt_pull_connect: null argument");
  try {
    Peer p;
    p.address = address;
    int rc = parse_address(p.address, p);
    if (rc != BT_OK) return rc;
    q->peers.push_back(p);
  } catch (const std::bad_alloc&) {
    return bt::fail(BT_E_NOMEM, "out of memory");
  }
  return BT_OK;
}

int bt_pull_destroy(bt_pull* q) {
  if (!q) return BT_OK;
  if (getenv("BT_RX_STATS") && q->messages)
    fprintf(stderr, "bt_pull: %llu messages, %.1f MB each, %.1f recv and %.1f poll calls per message (payload reads)\n",
            (unsigned long long)q->messages, (double)q->bytes / q->messages / 1e6, (double)g_recv_calls / q->messages,
            (double)g_poll_calls / q->messages);
  for (Peer& p : q->peers)
    if (p.fd >= 0) close(p.fd);
  delete q;
  return BT_OK;
}

static int pull_recv(bt_pull* q, int timeout_ms) {
  if (q->has_held) return BT_OK;
  const uint64_t deadline = now_ms() + (uint64_t)(timeout_ms < 0 ? 0 : timeout_ms);
  while (true) {
    const uint64_t now = now_ms();
    const int i = wait_readable(q, now >= deadline ? 0 : (int)(deadline - now));
    if (i < 0) return BT_E_TIMEOUT;
    Peer& p = q->peers[i];
    int is_msg = 0, more = 0;
    uint64_t size = 0;
    int rc = read_frame_header(p, &is_msg, &size, &more);
    if (rc == BT_OK && is_msg) rc = read_whole(q, p, size, more);
    if (rc != BT_OK) { close_peer(p); continue; }
    if (is_msg) return BT_OK;
  }
}

int bt_pull_recv(bt_pull* q, int timeout_ms) {
  if (!q) return bt::fail(BT_E_INVALID, "null receiver");
  try {   // sizes off the wire end up in vector::resize: nothing may unwind through the C ABI
    return pull_recv(q, timeout_ms);
  } catch (const std::bad_alloc&) {
    q->has_held = false;
    return bt::fail(BT_E_NOMEM, "out of memory while receiving a message");
  } catch (const std::exception& e) {
    q->has_held = false;
    return bt::fail(BT_E_NOMEM, "receiving a message failed: %s", e.what());
  }
}

int bt_pull_pending(bt_pull* q, const uint8_t** data, uint64_t* len) {
  if (!q || !data || !len) return bt::fail(BT_E_INVALID, "bt_pull_pending: bad arguments");
  if (!q->has_held) {
    *data = nullptr;
    *len = 0;
    return bt::fail(BT_E_INVALID, "no message is held");
  }
  *data = q->held.data();
  *len = q->held.size();
  return BT_OK;
}

int bt_pull_pending_parts(bt_pull* q, uint64_t* sizes, int max_parts) {
  if (!q || (max_parts > 0 && !sizes)) return bt::fail(BT_E_INVALID, "bt_pull_pending_parts: bad arguments");
  if (!q->has_held) return bt::fail(BT_E_INVALID, "no message is held");
  const int n = (int)q->held_parts.size();
  for (int i = 0; i < n && i < max_parts; ++i) sizes[i] = q->held_parts[i];
  return n;
}

int bt_pull_drop(bt_pull* q) {
  if (!q) return bt::fail(BT_E_INVALID, "null receiver");
  q->has_held = false;
  return BT_OK;
}

static int pull_fill(bt_pull* q, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
                     int slice_ms, int* n_done);

int bt_pull_fill(bt_pull* q, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
                 int slice_ms, int* n_done) {
  if (!q || !ring || !schema || !n_done || first_item < 0 || max_items < 0 || first_item + max_items > schema->batch)
    return bt::fail(BT_E_INVALID, "bt_pull_fill: bad arguments");
  // payloads are received straight into the slot: it has to exist and to hold the schema's layout
  if (slot < 0 || slot >= bt_ring_depth(ring) || schema->slot_bytes > bt_ring_slot_bytes(ring))
    return bt::fail(BT_E_INVALID, "bt_pull_fill: slot %d is outside the ring, or the ring's slots are smaller than the schema", slot);
  *n_done = 0;
  try {
    return pull_fill(q, ring, slot, schema, first_item, max_items, slice_ms, n_done);
  } catch (const std::bad_alloc&) {
    return bt::fail(BT_E_NOMEM, "out of memory while receiving a message");
  } catch (const std::exception& e) {
    return bt::fail(BT_E_NOMEM, "receiving a message failed: %s", e.what());
  }
}

static int pull_fill(bt_pull* q, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
                     int slice_ms, int* n_done) {
  while (*n_done < max_items) {
    if (q->has_held) {  // the first message of the stream, received before the schema existed
      int rc = ingest_held(q, ring, slot, schema, first_item + *n_done);
      if (rc != BT_OK) return rc;
      q->has_held = false;
      ++*n_done;
      continue;
    }
    const int i = wait_readable(q, slice_ms);
    if (i < 0) return BT_E_TIMEOUT;
    Peer& p = q->peers[i];
    int is_msg = 0, more = 0;
    uint64_t size = 0;
    int rc = read_frame_header(p, &is_msg, &size, &more);
    if (rc != BT_OK) { close_peer(p); continue; }
    if (!is_msg) continue;
    rc = more ? receive_raw_parts(q, p, size, ring, slot, schema, first_item + *n_done)
              : receive_streaming(q, p, size, ring, slot, schema, first_item + *n_done);
    if (rc == BT_E_UNSUPPORTED || rc == BT_E_SCHEMA) return rc;  // message is held for the caller
    if (rc != BT_OK) { close_peer(p); continue; }
    q->messages++;
    q->bytes += size;
    ++*n_done;
    bt_ring_reap(ring);
  }
  return BT_OK;
}

}  // extern "C"
