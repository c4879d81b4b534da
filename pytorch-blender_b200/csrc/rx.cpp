// Receive loop in C: drain a ZMQ PULL socket straight into a staging slot, many messages per call.
//
// Replaces the per-message Python of the reference's hot loop (pkg_pytorch/blendtorch/btt/dataset.py:97-106:
// poll, recv, pickle.loads, yield) with one call per *batch* that runs without the GIL: poll ->
// zmq_msg_recv -> scan -> place payloads (bt_ring_ingest) -> close message, until the batch is full.
// libzmq is not linked: pyzmq ships its own copy, so the handful of entry points needed are passed
// in as function pointers taken from that very library (bt_zmq_api), and the socket is pyzmq's own
// (`socket.underlying`).  A ZMQ socket may only be used by one thread at a time; the receiver thread
// that owns the socket is the one that calls in here.
#include <cstring>
#include <new>

#include "common.h"

namespace {

struct zmq_msg_storage {
  alignas(8) unsigned char bytes[64];  // zmq_msg_t is 64 opaque bytes in libzmq 4.x
};

struct zmq_pollitem {
  void* socket;
  int fd;
  short events;
  short revents;
};

constexpr short kPollIn = 1;

typedef int (*fn_msg_init)(void*);
typedef int (*fn_msg_recv)(void*, void*, int);
typedef void* (*fn_msg_data)(void*);
typedef size_t (*fn_msg_size)(void*);
typedef int (*fn_msg_close)(void*);
typedef int (*fn_poll)(zmq_pollitem*, int, long);
typedef int (*fn_errno)(void);

}  // namespace

struct bt_rx {
  fn_msg_init msg_init;
  fn_msg_recv msg_recv;
  fn_msg_data msg_data;
  fn_msg_size msg_size;
  fn_msg_close msg_close;
  fn_poll poll;
  fn_errno zerrno;
  void* socket = nullptr;
  zmq_msg_storage held;
  bool has_held = false;
  uint64_t messages = 0, bytes = 0;
};

namespace {

// Wait up to timeout_ms for a message and receive it into rx->held.
int recv_one(bt_rx* rx, int timeout_ms) {
  zmq_pollitem item = {rx->socket, 0, kPollIn, 0};
  int rc = rx->poll(&item, 1, timeout_ms);
  if (rc < 0) {
    const int e = rx->zerrno();
    if (e == 4 /* EINTR */) return BT_E_TIMEOUT;
    return bt::fail(BT_E_IO, "zmq_poll failed (errno %d)", e);
  }
  if (rc == 0 || !(item.revents & kPollIn)) return BT_E_TIMEOUT;
  rx->msg_init(&rx->held);
  if (rx->msg_recv(&rx->held, rx->socket, 1 /* ZMQ_DONTWAIT */) < 0) {
    const int e = rx->zerrno();
    rx->msg_close(&rx->held);
    if (e == 11 /* EAGAIN */ || e == 4) return BT_E_TIMEOUT;
    return bt::fail(BT_E_IO, "zmq_msg_recv failed (errno %d)", e);
  }
  rx->has_held = true;
  return BT_OK;
}

void drop_held(bt_rx* rx) {
  if (rx->has_held) {
    rx->msg_close(&rx->held);
    rx->has_held = false;
  }
}

}  // namespace

extern "C" {

int bt_rx_create(const bt_zmq_api* api, void* zmq_socket, bt_rx** out) {
  if (!api || !zmq_socket || !out || !api->msg_init || !api->msg_recv || !api->msg_data || !api->msg_size ||
      !api->msg_close || !api->poll || !api->zmq_errno)
    return bt::fail(BT_E_INVALID, "bt_rx_create: incomplete ZMQ API table or null socket");
  bt_rx* rx = new (std::nothrow) bt_rx();
  if (!rx) return bt::fail(BT_E_NOMEM, "out of memory");
  rx->msg_init = reinterpret_cast<fn_msg_init>(api->msg_init);
  rx->msg_recv = reinterpret_cast<fn_msg_recv>(api->msg_recv);
  rx->msg_data = reinterpret_cast<fn_msg_data>(api->msg_data);
  rx->msg_size = reinterpret_cast<fn_msg_size>(api->msg_size);
  rx->msg_close = reinterpret_cast<fn_msg_close>(api->msg_close);
  rx->poll = reinterpret_cast<fn_poll>(api->poll);
  rx->zerrno = reinterpret_cast<fn_errno>(api->zmq_errno);
  rx->socket = zmq_socket;
  *out = rx;
  return BT_OK;
}

int bt_rx_destroy(bt_rx* rx) {
  if (!rx) return BT_OK;
  drop_held(rx);
  delete rx;
  return BT_OK;
}

int bt_rx_recv(bt_rx* rx, int timeout_ms) {
  if (!rx) return bt::fail(BT_E_INVALID, "null receiver");
  if (rx->has_held) return BT_OK;
  return recv_one(rx, timeout_ms);
}

int bt_rx_pending(bt_rx* rx, const uint8_t** data, uint64_t* len) {
  if (!rx || !data || !len) return bt::fail(BT_E_INVALID, "bt_rx_pending: bad arguments");
  if (!rx->has_held) {
    *data = nullptr;
    *len = 0;
    return bt::fail(BT_E_INVALID, "no message is held");
  }
  *data = static_cast<const uint8_t*>(rx->msg_data(&rx->held));
  *len = rx->msg_size(&rx->held);
  return BT_OK;
}

int bt_rx_drop(bt_rx* rx) {
  if (!rx) return bt::fail(BT_E_INVALID, "null receiver");
  drop_held(rx);
  return BT_OK;
}

int bt_rx_release_queued(bt_rx* rx, int* n_released) {
  if (!rx) return bt::fail(BT_E_INVALID, "null receiver");
  int released = 0;
  std::vector<bt_field>& f = bt::tls_fields();
  for (int guard = 0; guard < 100000; ++guard) {
    if (!rx->has_held && recv_one(rx, 0) != BT_OK) break;
    const uint8_t* data = static_cast<const uint8_t*>(rx->msg_data(&rx->held));
    const uint64_t len = rx->msg_size(&rx->held);
    if (bt::scan_fields(data, len, f) == BT_OK) {
      for (const bt_field& fi : f) {
        if (fi.kind != BT_KIND_SHMREF) continue;
        bt_shm* seg = nullptr;
        if (bt::shm_attach_cached(reinterpret_cast<const char*>(data + fi.payload_off), fi.payload_len, -1, &seg) == BT_OK &&
            bt_shm_release(seg, (int)fi.ival) == BT_OK)
          ++released;
      }
    }
    drop_held(rx);
  }
  if (n_released) *n_released = released;
  return BT_OK;
}

int bt_rx_fill(bt_rx* rx, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
               int slice_ms, int* n_done) {
  if (!rx || !ring || !schema || !n_done || first_item < 0 || max_items < 0)
    return bt::fail(BT_E_INVALID, "bt_rx_fill: bad arguments");
  *n_done = 0;
  while (*n_done < max_items) {
    if (!rx->has_held) {
      int rc = recv_one(rx, slice_ms);
      if (rc != BT_OK) return rc;  // BT_E_TIMEOUT: nothing arrived within the slice
    }
    const uint8_t* data = static_cast<const uint8_t*>(rx->msg_data(&rx->held));
    const uint64_t len = rx->msg_size(&rx->held);
    int rc = bt_ring_ingest(ring, slot, schema, first_item + *n_done, data, len);
    if (rc != BT_OK) return rc;    // message stays held for the caller's compatibility decode
    rx->messages++;
    rx->bytes += len;
    drop_held(rx);
    ++*n_done;
    bt_ring_reap(ring);
  }
  return BT_OK;
}

}  // extern "C"
