// Native publisher: a minimal ZMTP 3.x PUSH endpoint, the producer end of the ingest wire.
//
// The reference's producers are Blender processes running `DataPublisher`
// (pkg_blender/blendtorch/btb/publisher.py:21-43): a libzmq PUSH socket that binds, SNDHWM, IMMEDIATE=1,
// `send_pyobj`.  Every byte of a frame is copied by pickle, by pyzmq into a zmq_msg and by libzmq's
// I/O thread into the kernel socket, and every producer costs an interpreter plus an I/O thread.
// Producers that already hold pickled messages -- replay of recordings, synthetic load, simulators
// written in C -- need none of that: here a message goes from the caller's memory to the kernel
// socket with one copy (`sendmsg` of frame header + body), framed exactly as libzmq frames it, and
// one thread can keep many endpoints busy.  Peers are unmodified ZMQ PULL sockets
// (pkg_pytorch/blendtorch/btt/dataset.py:73-78) or this library's direct receiver (zmtp.cpp).
//
// Kept from the reference's socket options: only completed connections get messages (IMMEDIATE=1),
// peers are served round-robin and a peer that cannot take more is skipped (PUSH load balancing), the
// sender blocks when nobody can take a message (SNDHWM back-pressure: the queue is the kernel socket
// buffer plus one message in progress per peer).
#include <arpa/inet.h>
#include <fcntl.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "common.h"

namespace {

uint64_t now_ms() {
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (uint64_t)t.tv_sec * 1000ull + (uint64_t)t.tv_nsec / 1000000ull;
}

constexpr int kMaxParts = 64;
constexpr int kHandshakeMs = 2000;

// One frame being written: 1/9-byte ZMTP header, then the body.
struct Frame {
  uint8_t hdr[9];
  int hdr_len = 0;
  const uint8_t* body = nullptr;
  uint64_t len = 0;
};

struct PushPeer {
  int fd = -1;
  // message in progress (nframes == 0: idle)
  Frame frames[kMaxParts];
  int nframes = 0, cur = 0;
  uint64_t off = 0;              // bytes of frames[cur] (header + body) already written
  std::vector<uint8_t> small;    // owns the patched template of a shared-memory message
  int ticket = -1;               // shared-memory slot announced by the message in progress
};

struct Endpoint {
  std::string address;           // as resolved (what peers connect to)
  std::string unlink_path;       // ipc:// socket file to remove
  int listen_fd = -1;
  std::vector<PushPeer> peers;
  size_t rr = 0;
  int64_t started = 0;           // messages whose sending has begun
};

int read_exact(int fd, uint8_t* dst, size_t n, int timeout_ms) {
  const uint64_t deadline = now_ms() + (uint64_t)timeout_ms;
  while (n) {
    ssize_t k = recv(fd, dst, n, 0);
    if (k > 0) { dst += k; n -= (size_t)k; continue; }
    if (k == 0) return bt::fail(BT_E_IO, "peer closed the connection");
    if (errno == EINTR) continue;
    if (errno != EAGAIN && errno != EWOULDBLOCK) return bt::fail(BT_E_IO, "recv: %s", strerror(errno));
    const uint64_t now = now_ms();
    if (now >= deadline) return bt::fail(BT_E_IO, "peer stalled during the handshake");
    pollfd pf = {fd, POLLIN, 0};
    poll(&pf, 1, (int)(deadline - now));
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
    if (now >= deadline) return bt::fail(BT_E_IO, "peer does not take the handshake");
    pollfd pf = {fd, POLLOUT, 0};
    poll(&pf, 1, (int)(deadline - now));
  }
  return BT_OK;
}

// ZMTP 3.0 greeting + NULL handshake from the binding side (READY with Socket-Type=PUSH).  The peer is
// a live ZMTP engine that sends its own greeting at once, so this is done synchronously.
int handshake(int fd) {
  uint8_t g[64];
  memset(g, 0, sizeof g);
  g[0] = 0xFF; g[8] = 0x01; g[9] = 0x7F;
  g[10] = 3; g[11] = 0;
  memcpy(g + 12, "NULL", 4);
  int rc = write_all(fd, g, sizeof g, kHandshakeMs);
  if (rc != BT_OK) return rc;
  uint8_t peer[64];
  rc = read_exact(fd, peer, sizeof peer, kHandshakeMs);
  if (rc != BT_OK) return rc;
  if (peer[0] != 0xFF || peer[9] != 0x7F || peer[10] < 3) return bt::fail(BT_E_UNSUPPORTED, "peer does not speak ZMTP 3");
  if (memcmp(peer + 12, "NULL\0", 5) != 0) return bt::fail(BT_E_UNSUPPORTED, "peer uses a security mechanism other than NULL");
  static const uint8_t ready[] = {0x04, 26, 5, 'R', 'E', 'A', 'D', 'Y', 11, 'S', 'o', 'c', 'k', 'e', 't', '-', 'T', 'y', 'p', 'e',
                                  0, 0, 0, 4, 'P', 'U', 'S', 'H'};
  rc = write_all(fd, ready, sizeof ready, kHandshakeMs);
  if (rc != BT_OK) return rc;
  uint8_t h[9];
  rc = read_exact(fd, h, 2, kHandshakeMs);
  if (rc != BT_OK) return rc;
  if (!(h[0] & 0x04)) return bt::fail(BT_E_UNSUPPORTED, "expected a ZMTP command, got flags 0x%02x", h[0]);
  uint64_t size = h[1];
  if (h[0] & 0x02) {
    rc = read_exact(fd, h + 2, 7, kHandshakeMs);
    if (rc != BT_OK) return rc;
    size = 0;
    for (int i = 1; i < 9; ++i) size = (size << 8) | h[i];
  }
  if (size < 6 || size > 4096) return bt::fail(BT_E_UNSUPPORTED, "unexpected handshake command size");
  uint8_t body[4096];
  rc = read_exact(fd, body, (size_t)size, kHandshakeMs);
  if (rc != BT_OK) return rc;
  if (body[0] != 5 || memcmp(body + 1, "READY", 5) != 0) return bt::fail(BT_E_UNSUPPORTED, "peer refused the connection");
  // a PUSH socket only talks to PULL peers (ZMTP 3.0, socket type compatibility)
  static const char want[] = "\x0bSocket-Type\0\0\0\x04PULL";
  bool pull = false;
  for (uint64_t i = 6; i + sizeof want - 1 <= size; ++i)
    if (memcmp(body + i, want, sizeof want - 1) == 0) pull = true;
  if (!pull) return bt::fail(BT_E_UNSUPPORTED, "peer is not a PULL socket");
  return BT_OK;
}

void set_frame(Frame& f, const uint8_t* body, uint64_t len, bool more) {
  f.body = body;
  f.len = len;
  if (len <= 255) {
    f.hdr[0] = more ? 0x01 : 0x00;
    f.hdr[1] = (uint8_t)len;
    f.hdr_len = 2;
  } else {
    f.hdr[0] = more ? 0x03 : 0x02;
    for (int i = 0; i < 8; ++i) f.hdr[1 + i] = (uint8_t)(len >> (8 * (7 - i)));
    f.hdr_len = 9;
  }
}

void close_peer(PushPeer& p) {
  if (p.fd >= 0) close(p.fd);
  p.fd = -1;
  p.nframes = 0;
}

// Accept whatever is waiting on the listening socket.
void accept_pending(Endpoint& e) {
  while (e.listen_fd >= 0) {
    int fd = accept4(e.listen_fd, nullptr, nullptr, SOCK_NONBLOCK | SOCK_CLOEXEC);
    if (fd < 0) return;
    int one = 1;
    static const int sndbuf = [] {
      const char* e = getenv("BT_PUSH_SNDBUF");   // bytes; tuning knob
      return e ? atoi(e) : (1 << 20);
    }();
    if (sndbuf > 0) setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &sndbuf, sizeof sndbuf);
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);   // fails harmlessly on AF_UNIX
    if (handshake(fd) != BT_OK) { close(fd); continue; }
    PushPeer* slot = nullptr;
    for (PushPeer& p : e.peers)
      if (p.fd < 0) { slot = &p; break; }
    if (!slot) {
      e.peers.emplace_back();
      slot = &e.peers.back();
    }
    slot->fd = fd;
    slot->nframes = 0;
    slot->ticket = -1;
  }
}

// Write as much of the peer's message in progress as the socket takes.
// 1: message complete, 0: socket full, <0: connection lost.
int pump(PushPeer& p) {
  while (p.cur < p.nframes) {
    Frame& f = p.frames[p.cur];
    iovec iov[2];
    int niov = 0;
    if (p.off < (uint64_t)f.hdr_len) {
      iov[niov].iov_base = f.hdr + p.off;
      iov[niov].iov_len = (size_t)(f.hdr_len - p.off);
      ++niov;
      if (f.len) {
        iov[niov].iov_base = const_cast<uint8_t*>(f.body);
        iov[niov].iov_len = (size_t)f.len;
        ++niov;
      }
    } else {
      const uint64_t done = p.off - f.hdr_len;
      iov[0].iov_base = const_cast<uint8_t*>(f.body) + done;
      iov[0].iov_len = (size_t)(f.len - done);
      niov = 1;
    }
    msghdr mh;
    memset(&mh, 0, sizeof mh);
    mh.msg_iov = iov;
    mh.msg_iovlen = niov;
    ssize_t k = sendmsg(p.fd, &mh, MSG_NOSIGNAL | MSG_DONTWAIT);
    if (k < 0) {
      if (errno == EINTR) continue;
      if (errno == EAGAIN || errno == EWOULDBLOCK) return 0;
      return -1;
    }
    p.off += (uint64_t)k;
    if (p.off == (uint64_t)f.hdr_len + f.len) {
      ++p.cur;
      p.off = 0;
    }
  }
  p.nframes = 0;
  return 1;
}

// Anything a PULL peer sends after the handshake is a command (heartbeat) or the end of the connection.
bool peer_alive(PushPeer& p) {
  uint8_t junk[256];
  while (true) {
    ssize_t k = recv(p.fd, junk, sizeof junk, MSG_DONTWAIT);
    if (k > 0) continue;
    if (k == 0) return false;
    return errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR;
  }
}

int bind_endpoint(const std::string& a, Endpoint& e) {
  if (a.rfind("tcp://", 0) == 0) {
    const std::string rest = a.substr(6);
    const size_t colon = rest.rfind(':');
    if (colon == std::string::npos) return bt::fail(BT_E_INVALID, "address %s has no port", a.c_str());
    std::string host = rest.substr(0, colon), port = rest.substr(colon + 1);
    if (host == "*") host = "0.0.0.0";
    if (host == "localhost") host = "127.0.0.1";
    if (port == "*") port = "0";
    addrinfo hints;
    memset(&hints, 0, sizeof hints);
    hints.ai_family = AF_UNSPEC;
    hints.ai_socktype = SOCK_STREAM;
    hints.ai_flags = AI_PASSIVE | AI_NUMERICSERV;
    addrinfo* res = nullptr;
    if (getaddrinfo(host.c_str(), port.c_str(), &hints, &res) != 0 || !res) return bt::fail(BT_E_INVALID, "cannot resolve %s", a.c_str());
    int fd = socket(res->ai_family, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    int one = 1;
    if (fd >= 0) setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    if (fd < 0 || bind(fd, res->ai_addr, res->ai_addrlen) != 0 || listen(fd, 64) != 0) {
      const int err = errno;
      if (fd >= 0) close(fd);
      freeaddrinfo(res);
      return bt::fail(BT_E_IO, "cannot bind %s: %s", a.c_str(), strerror(err));
    }
    sockaddr_storage ss;
    socklen_t sl = sizeof ss;
    getsockname(fd, reinterpret_cast<sockaddr*>(&ss), &sl);
    const int bound = ntohs(ss.ss_family == AF_INET6 ? reinterpret_cast<sockaddr_in6*>(&ss)->sin6_port
                                                     : reinterpret_cast<sockaddr_in*>(&ss)->sin_port);
    freeaddrinfo(res);
    e.listen_fd = fd;
    e.address = "tcp://" + (host == "0.0.0.0" ? std::string("127.0.0.1") : host) + ":" + std::to_string(bound);
    return BT_OK;
  }
  if (a.rfind("ipc://", 0) == 0) {
    const std::string path = a.substr(6);
    sockaddr_un un;
    memset(&un, 0, sizeof un);
    if (path.empty() || path.size() >= sizeof un.sun_path) return bt::fail(BT_E_INVALID, "bad ipc path in %s", a.c_str());
    un.sun_family = AF_UNIX;
    memcpy(un.sun_path, path.c_str(), path.size());
    socklen_t len = (socklen_t)(offsetof(sockaddr_un, sun_path) + path.size());
    if (path[0] == '@') un.sun_path[0] = '\0';
    else { len += 1; unlink(path.c_str()); }
    int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
    if (fd < 0 || bind(fd, reinterpret_cast<sockaddr*>(&un), len) != 0 || listen(fd, 64) != 0) {
      const int err = errno;
      if (fd >= 0) close(fd);
      return bt::fail(BT_E_IO, "cannot bind %s: %s", a.c_str(), strerror(err));
    }
    e.listen_fd = fd;
    e.address = a;
    if (path[0] != '@') e.unlink_path = path;
    return BT_OK;
  }
  return bt::fail(BT_E_UNSUPPORTED, "the native publisher binds tcp:// and ipc:// only (%s)", a.c_str());
}

}  // namespace

struct bt_push {
  std::vector<Endpoint*> eps;   // stable addresses: endpoints are driven by different threads
};

namespace {

// Wait until something can happen on the endpoints [first, first+count): a peer socket takes bytes or
// dies, or a connection arrives.
void wait_any(bt_push* p, int first, int count, int timeout_ms, std::vector<pollfd>& pfds, bool idle_writable) {
  pfds.clear();
  for (int i = first; i < first + count; ++i) {
    Endpoint& e = *p->eps[i];
    pfds.push_back(pollfd{e.listen_fd, POLLIN, 0});
    for (PushPeer& q : e.peers)
      if (q.fd >= 0) pfds.push_back(pollfd{q.fd, (short)(POLLIN | ((q.nframes || idle_writable) ? POLLOUT : 0)), 0});
  }
  poll(pfds.data(), pfds.size(), timeout_ms);
}

}  // namespace

extern "C" {

int bt_push_create(bt_push** out) {
  if (!out) return bt::fail(BT_E_INVALID, "bt_push_create: null argument");
  bt_push* p = new (std::nothrow) bt_push();
  if (!p) return bt::fail(BT_E_NOMEM, "out of memory");
  *out = p;
  return BT_OK;
}

int bt_push_bind(bt_push* p, const char* address) {
  if (!p || !address) return bt::fail(BT_E_INVALID, "bt_push_bind: null argument");
  Endpoint* e = new (std::nothrow) Endpoint();
  if (!e) return bt::fail(BT_E_NOMEM, "out of memory");
  int rc = bind_endpoint(address, *e);
  if (rc != BT_OK) { delete e; return rc; }
  try {
    e->peers.reserve(16);
    p->eps.push_back(e);
  } catch (const std::bad_alloc&) {
    close(e->listen_fd);
    delete e;
    return bt::fail(BT_E_NOMEM, "out of memory");
  }
  return (int)p->eps.size() - 1;
}

int bt_push_endpoint(bt_push* p, int endpoint, char* address_out, int capacity) {
  if (!p || endpoint < 0 || endpoint >= (int)p->eps.size() || !address_out || capacity < 1)
    return bt::fail(BT_E_INVALID, "bt_push_endpoint: bad arguments");
  const std::string& a = p->eps[endpoint]->address;
  if ((int)a.size() + 1 > capacity) return bt::fail(BT_E_TOO_LARGE, "address needs %zu bytes", a.size() + 1);
  memcpy(address_out, a.c_str(), a.size() + 1);
  return BT_OK;
}

int bt_push_destroy(bt_push* p) {
  if (!p) return BT_OK;
  for (Endpoint* e : p->eps) {
    for (PushPeer& q : e->peers) close_peer(q);
    if (e->listen_fd >= 0) close(e->listen_fd);
    if (!e->unlink_path.empty()) unlink(e->unlink_path.c_str());
    delete e;
  }
  delete p;
  return BT_OK;
}

int bt_push_peers(bt_push* p, int endpoint) {
  if (!p || endpoint < 0 || endpoint >= (int)p->eps.size()) return bt::fail(BT_E_INVALID, "bt_push_peers: bad endpoint");
  Endpoint& e = *p->eps[endpoint];
  try {
    accept_pending(e);
  } catch (const std::bad_alloc&) {
    return bt::fail(BT_E_NOMEM, "out of memory");
  }
  int n = 0;
  for (PushPeer& q : e.peers) {
    if (q.fd >= 0 && q.nframes == 0 && !peer_alive(q)) close_peer(q);
    if (q.fd >= 0) ++n;
  }
  return n;
}

int bt_push_send(bt_push* p, int endpoint, const uint8_t* const* parts, const uint64_t* lens, int nparts, int timeout_ms) {
  if (!p || endpoint < 0 || endpoint >= (int)p->eps.size() || !parts || !lens || nparts < 1 || nparts > kMaxParts)
    return bt::fail(BT_E_INVALID, "bt_push_send: bad arguments");
  for (int i = 0; i < nparts; ++i)
    if (!parts[i] && lens[i]) return bt::fail(BT_E_INVALID, "bt_push_send: null frame");
  Endpoint& e = *p->eps[endpoint];
  const uint64_t deadline = timeout_ms < 0 ? ~0ull : now_ms() + (uint64_t)timeout_ms;
  std::vector<pollfd> pfds;
  try {
    while (true) {
      accept_pending(e);
      const size_t n = e.peers.size();
      for (size_t k = 0; k < n; ++k) {
        PushPeer& q = e.peers[(e.rr + k) % n];
        if (q.fd < 0) continue;
        if (!peer_alive(q)) { close_peer(q); continue; }   // hung up while idle: not a consumer any more
        for (int i = 0; i < nparts; ++i) set_frame(q.frames[i], parts[i], lens[i], i + 1 < nparts);
        q.nframes = nparts;
        q.cur = 0;
        q.off = 0;
        int rc = pump(q);
        if (rc == 0 && q.cur == 0 && q.off == 0) { q.nframes = 0; continue; }   // took nothing: try the next peer
        // this peer has (part of) the message: finish it there
        // a message that has started must complete or the connection is dropped (10 s, like the receiver)
        const uint64_t give_up = now_ms() + 10000;
        while (rc == 0) {
          const uint64_t now = now_ms();
          if (now >= give_up) { rc = -1; break; }
          pollfd pf = {q.fd, POLLOUT, 0};
          poll(&pf, 1, (int)(give_up - now));
          rc = pump(q);
        }
        if (rc < 0) { close_peer(q); continue; }   // lost mid-message: the message goes to another peer
        e.rr = ((e.rr + k) % n + 1) % n;
        ++e.started;
        return BT_OK;
      }
      const uint64_t now = now_ms();
      if (now >= deadline) return BT_E_TIMEOUT;
      const uint64_t left = deadline == ~0ull ? 100 : deadline - now;
      wait_any(p, endpoint, 1, (int)(left < 100 ? left : 100), pfds, true);
      for (PushPeer& q : e.peers)
        if (q.fd >= 0 && !peer_alive(q)) close_peer(q);
    }
  } catch (const std::bad_alloc&) {
    return bt::fail(BT_E_NOMEM, "out of memory");
  }
}

int bt_push_run(bt_push* p, const bt_push_job* job) {
  if (!p || !job || job->first_endpoint < 0 || job->n_endpoints < 1 ||
      job->first_endpoint + job->n_endpoints > (int)p->eps.size() || job->pool_n < 1 || !job->pool || !job->pool_len)
    return bt::fail(BT_E_INVALID, "bt_push_run: bad arguments");
  if (job->seg && !job->tmpl_ticket_off)
    return bt::fail(BT_E_INVALID, "bt_push_run: the shared-memory wire needs ticket offsets");
  if (job->seg && (job->payload == nullptr) != (job->payload_len == nullptr))
    return bt::fail(BT_E_INVALID, "bt_push_run: payload and payload_len go together");
  const bool in_place = job->seg && !job->payload;   // frames are produced inside the slots (see bt_b200.h)
  if (in_place)
    for (int i = 0; i < job->n_endpoints; ++i)
      if (bt_shm_nslots(job->seg[i]) != job->pool_n)
        return bt::fail(BT_E_INVALID, "bt_push_run: in-place publishing needs one template per slot (pool_n == slots)");
  const int first = job->first_endpoint, count = job->n_endpoints;
  for (int i = first; i < first + count; ++i) p->eps[i]->started = 0;
  std::vector<pollfd> pfds;
  uint64_t last_accept = 0;
  try {
    while (true) {
      if (job->stop && *job->stop) return BT_OK;
      bool progress = false, busy = false, all_started = true;
      const uint64_t now = now_ms();
      const bool look = now - last_accept >= 20;
      if (look) last_accept = now;
      for (int ei = 0; ei < count; ++ei) {
        Endpoint& e = *p->eps[first + ei];
        if (look || e.peers.empty()) accept_pending(e);
        const bool more = job->messages < 0 || e.started < job->messages;
        if (more) all_started = false;
        const size_t n = e.peers.size();
        for (size_t k = 0; k < n; ++k) {
          PushPeer& q = e.peers[(e.rr + k) % n];
          if (q.fd < 0) continue;
          if (q.nframes == 0) {
            if (!(job->messages < 0 || e.started < job->messages)) continue;
            // a consumer that has hung up gets no new message (its FIN is waiting to be read: one
            // non-blocking recv); what was already on its way when it left is lost with it, as in libzmq
            if (!peer_alive(q)) { close_peer(q); progress = true; continue; }
            size_t m = (size_t)ei * job->pool_n + (size_t)(e.started % job->pool_n);
            if (job->seg) {
              bt_shm* seg = job->seg[ei];
              int ticket = -1;
              if (bt_shm_acquire(seg, 0, &ticket) != BT_OK) { busy = true; continue; }   // consumers hold every slot
              if (in_place) {
                // the producer wrote this frame where it stands: slot k holds the frame template k describes
                m = (size_t)ei * job->pool_n + (size_t)(ticket & 0xFFFF);
              } else {
                if (job->payload_len[m] > bt_shm_slot_bytes(seg)) {
                  bt_shm_release(seg, ticket);
                  return bt::fail(BT_E_TOO_LARGE, "payload of %llu bytes does not fit a shared-memory slot",
                                  (unsigned long long)job->payload_len[m]);
                }
                bt::stream_copy(bt_shm_slot_ptr(seg, ticket), job->payload[m], job->payload_len[m]);
              }
              q.small.assign(job->pool[m], job->pool[m] + job->pool_len[m]);
              if ((uint64_t)job->tmpl_ticket_off[m] + 4 > q.small.size()) {
                bt_shm_release(seg, ticket);
                return bt::fail(BT_E_INVALID, "bt_push_run: ticket offset outside the template");
              }
              const uint32_t t = (uint32_t)ticket;
              memcpy(q.small.data() + job->tmpl_ticket_off[m], &t, 4);   // BININT: little-endian int32
              q.ticket = ticket;
              set_frame(q.frames[0], q.small.data(), q.small.size(), false);
            } else {
              set_frame(q.frames[0], job->pool[m], job->pool_len[m], false);
            }
            q.nframes = 1;
            q.cur = 0;
            q.off = 0;
            ++e.started;
            e.rr = ((e.rr + k) % n + 1) % n;
          }
          const uint64_t before_cur = q.cur, before_off = q.off;
          const int rc = pump(q);
          if (rc < 0) {
            // Connection lost.  A message of which nothing had been written yet is simply taken back (it
            // goes to the next peer); one that was under way is gone with the pipe, like in libzmq.  A
            // shared-memory slot it announced comes back at once either way.
            if (q.cur == 0 && q.off == 0 && before_cur == 0 && before_off == 0 && e.started > 0) --e.started;
            if (job->seg && q.ticket >= 0) bt_shm_release(job->seg[ei], q.ticket);
            q.ticket = -1;
            close_peer(q);
            progress = true;
            continue;
          }
          if (rc == 1) {
            q.ticket = -1;     // the consumer owns the slot from here
            progress = true;
            if (job->sent) ++*job->sent;
          } else {
            busy = true;
            if (q.cur != before_cur || q.off != before_off) progress = true;
          }
        }
      }
      if (all_started && !busy) {
        bool pending = false;
        for (int i = first; i < first + count; ++i)
          for (PushPeer& q : p->eps[i]->peers)
            if (q.fd >= 0 && q.nframes) pending = true;
        if (!pending) return BT_OK;
      }
      if (!progress) {
        bool sending = false;
        for (int i = first; i < first + count && !sending; ++i)
          for (PushPeer& q : p->eps[i]->peers)
            if (q.fd >= 0 && q.nframes) { sending = true; break; }
        if (!sending && busy) {
          // nothing on the wire and no free shared-memory slot: consumers release slots within
          // microseconds of their copies, sockets have nothing to tell us
          struct timespec nap = {0, 20000};
          nanosleep(&nap, nullptr);
          continue;
        }
        wait_any(p, first, count, 50, pfds, false);
        for (int i = first; i < first + count; ++i)
          for (PushPeer& q : p->eps[i]->peers)
            if (q.fd >= 0 && q.nframes == 0 && !peer_alive(q)) close_peer(q);
      }
    }
  } catch (const std::bad_alloc&) {
    return bt::fail(BT_E_NOMEM, "out of memory");
  }
}

}  // extern "C"
