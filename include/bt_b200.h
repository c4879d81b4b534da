/*
 * bt_b200.h -- C ABI of the B200-native blendtorch ingest path (libbt_b200.so).
 *
 * The reference (v-wewei/pytorch-blender, blendtorch 0.2.0) is pure Python and
 * has no FFI; what a binding for this path would attach to are the Python call
 * sites cited beside every entry point below (paths relative to the reference
 * root).  All signatures are plain C: pointers, sizes, POD structs.  Device
 * pointers are `void*` (e.g. torch.Tensor.data_ptr()), streams are `void*`
 * holding a cudaStream_t (NULL = legacy default stream).  Nothing here throws;
 * every function returns BT_OK (0) or a negative bt_status, and
 * bt_last_error() gives the thread-local message of the last failure.
 *
 * There is no CPU implementation of the kernels behind this header: if no CUDA
 * device is usable the launch functions fail with BT_E_CUDA.
 */
/*
 * Environment switches read by the library (all optional):
 *   BT_MAX_MESSAGE_BYTES  largest frame / message the direct receiver accepts (default 4 GiB)
 *   BT_PULL_RCVBUF, BT_PUSH_SNDBUF   socket buffer sizes in bytes (default 1 MiB each)
 *   BT_SHM_LEASE_MS       lease of shared-memory slots at segment creation (default 30000, 0 = never taken back)
 *   BT_STREAM_COPY=0      staging copies with plain memcpy instead of non-temporal stores (A/B switch)
 *   BT_RX_READAHEAD=0     direct receiver without its read-ahead buffer (A/B switch), BT_RX_STATS=1 prints its counters
 */
#ifndef BT_B200_H
#define BT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BT_ABI_VERSION 3

typedef enum bt_status {
  BT_OK = 0,
  BT_E_INVALID = -1,     /* bad argument */
  BT_E_UNSUPPORTED = -2, /* pickle uses something outside the whitelisted subset:
                            caller falls back to pickle.loads (reference behaviour) */
  BT_E_TRUNCATED = -3,   /* message ends inside an opcode or payload */
  BT_E_SCHEMA = -4,      /* message does not match the batch schema */
  BT_E_CUDA = -5,        /* CUDA runtime error, text in bt_last_error() */
  BT_E_NOMEM = -6,
  BT_E_IO = -7,
  BT_E_TIMEOUT = -8,
  BT_E_TOO_LARGE = -9    /* request exceeds a fixed capacity of the entry point */
} bt_status;

/* element types (numpy dtype strings in comments) */
typedef enum bt_dtype {
  BT_U8 = 0,  /* |u1 */
  BT_I8 = 1,  /* |i1 */
  BT_U16 = 2, /* <u2 */
  BT_I16 = 3, /* <i2 */
  BT_U32 = 4, /* <u4 */
  BT_I32 = 5, /* <i4 */
  BT_U64 = 6, /* <u8 */
  BT_I64 = 7, /* <i8 */
  BT_F16 = 8, /* <f2 */
  BT_F32 = 9, /* <f4 */
  BT_F64 = 10,/* <f8 */
  BT_BOOL = 11,/* |b1 */
  BT_BF16 = 12 /* output only (no numpy dtype) */
} bt_dtype;

typedef enum bt_kind {
  BT_KIND_NDARRAY = 0, /* numpy.ndarray or numpy scalar (ndim 0) */
  BT_KIND_INT = 1,     /* Python int fitting int64  -> collates to int64  */
  BT_KIND_FLOAT = 2,   /* Python float              -> collates to float64 */
  BT_KIND_BOOL = 3,    /* Python bool               -> collates to bool    */
  BT_KIND_NONE = 4,
  BT_KIND_STR = 5,     /* UTF-8 span in the message */
  BT_KIND_OTHER = 6,   /* parsed but not representable as a column */
  BT_KIND_SHMREF = 7,  /* ndarray announced by reference into a shared-memory segment (bt_shm_*):
                          dtype/ndim/shape describe it, payload_off/len span the segment NAME inside
                          the message, ival is the slot index */
  BT_KIND_RAWREF = 8   /* ndarray whose bytes travel as a separate frame of the same multipart message
                          (raw-parts wire, DataPublisher(raw_parts=True)): dtype/ndim/shape describe it,
                          ival is the 1-based index of the frame; payload_off/len are 0 until the
                          receiver has the frame (then they say where its bytes are) */
} bt_kind;

#define BT_MAX_DIMS 8
#define BT_MAX_KEY 64

/* One top-level entry of a message dict, located without unpickling. */
typedef struct bt_field {
  char key[BT_MAX_KEY];        /* NUL-terminated UTF-8 key (truncated keys -> BT_E_UNSUPPORTED) */
  int32_t kind;                /* bt_kind */
  int32_t dtype;               /* bt_dtype, for BT_KIND_NDARRAY */
  int32_t ndim;
  int32_t fortran;             /* 1 = payload is in Fortran order */
  int64_t shape[BT_MAX_DIMS];
  uint64_t payload_off;        /* byte offset of the raw data / string inside the message */
  uint64_t payload_len;
  int64_t ival;                /* BT_KIND_INT / BT_KIND_BOOL */
  double fval;                 /* BT_KIND_FLOAT */
} bt_field;

const char* bt_last_error(void);
int bt_abi_version(void);

/* ------------------------------------------------------------------------------------------
 * Message framing.
 * Replaces `pickle.loads(data)` / `socket.recv_pyobj()` at
 * pkg_pytorch/blendtorch/btt/dataset.py:103,105 and `Unpickler.load()` at
 * pkg_pytorch/blendtorch/btt/file.py:108 for messages of the wire schema
 * pkg_blender/blendtorch/btb/publisher.py:30-43 (a dict of str -> ndarray | int | float ...).
 * Walks the pickle opcode stream (protocols 2-5, whitelisted opcodes only, no code execution,
 * no allocation of payload-sized memory) and reports where each top-level value lives.
 * Returns the number of fields (>= 0) or a negative bt_status.
 * ------------------------------------------------------------------------------------------ */
int bt_scan_message(const uint8_t* msg, uint64_t len, bt_field* out, int max_fields);

/* Batch schema: the column layout of one batch-contiguous staging slot.
 * Replaces what torch.utils.data.default_collate derives per batch
 * (torch/utils/data/_utils/collate.py:118-313: ndarray -> stack, int -> int64, float -> float64)
 * for the dicts yielded at btt/dataset.py:106 and btt/file.py:108.
 * Every column starts at a 256-byte aligned offset and holds `batch` items back to back. */
typedef struct bt_schema bt_schema;

typedef struct bt_column {
  char key[BT_MAX_KEY];
  int32_t kind;                /* NDARRAY, INT, FLOAT or BOOL */
  int32_t dtype;               /* element type of the column (INT -> BT_I64, FLOAT -> BT_F64, BOOL -> BT_BOOL) */
  int32_t ndim;
  int32_t reserved;
  int64_t shape[BT_MAX_DIMS];  /* per-item shape */
  uint64_t item_bytes;
  uint64_t slot_off;           /* offset of item 0 inside the slot */
} bt_column;

int bt_schema_create(const uint8_t* first_msg, uint64_t len, int batch, bt_schema** out);
int bt_schema_destroy(bt_schema* s);
int bt_schema_num_columns(const bt_schema* s);
int bt_schema_column(const bt_schema* s, int i, bt_column* out);
uint64_t bt_schema_slot_bytes(const bt_schema* s);

/* Scan `msg` and copy every column's item `item` into the slot at `slot_base` (host memory,
 * normally pinned).  One memcpy per payload byte: this is the single host copy of the new
 * path (reference: recv -> bytes -> ndarray -> shm batch -> pageable staging,
 * btt/dataset.py:100-106 + collate.py:246-289 + examples/densityopt/densityopt.py:261,268). */
int bt_schema_ingest(const bt_schema* s, void* slot_base, int item, const uint8_t* msg, uint64_t len);

/* Same for `n` messages at once (replay from an mmap'ed .btr, btt/file.py:100-108), items
 * first_item .. first_item+n-1, spread over `nthreads` host threads (<=1: calling thread).
 * Returns BT_OK or the first failing message's status; `*failed_index` (optional) gets its index. */
int bt_schema_ingest_many(const bt_schema* s, void* slot_base, int first_item, int n,
                          const uint8_t* const* msgs, const uint64_t* lens, int nthreads,
                          int* failed_index);

/* ------------------------------------------------------------------------------------------
 * Pinned staging ring.
 * Replaces the pageable `.to(dev)` at examples/densityopt/densityopt.py:261,268 and the
 * shared-memory hand-off of DataLoader workers: `depth` slots of `slot_bytes` page-locked
 * host memory, each with a CUDA event that marks its last H2D copy complete.
 * device < 0 creates an unpinned ring (plain aligned host memory) for tools without a GPU;
 * bt_ring_submit on such a ring fails with BT_E_CUDA.
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_ring bt_ring;

int bt_ring_create(int device, uint64_t slot_bytes, int depth, bt_ring** out);
int bt_ring_destroy(bt_ring* r);
int bt_ring_depth(const bt_ring* r);
uint64_t bt_ring_slot_bytes(const bt_ring* r);
void* bt_ring_slot_ptr(bt_ring* r, int slot);
/* Next slot in round-robin order; blocks until its previous H2D copy has completed. */
int bt_ring_acquire(bt_ring* r, int* slot);
/* cudaMemcpyAsync(dev_dst, slot, nbytes, H2D, stream) + event record. */
int bt_ring_submit(bt_ring* r, int slot, void* dev_dst, uint64_t nbytes, void* stream);
/* Block until the slot's last submitted copy is complete. */
int bt_ring_wait(bt_ring* r, int slot);

/* Ring-level ingest: bt_schema_ingest into slot `slot` of the ring, except that payloads announced
 * by shared-memory reference (BT_KIND_SHMREF) are not copied by the CPU at all: they are remembered
 * and bt_ring_submit_batch DMAs them from the (page-locked) segment straight into the device batch.
 * The references are released back to their publisher once that copy has completed. */
int bt_ring_ingest(bt_ring* r, int slot, const bt_schema* s, int item, const uint8_t* msg, uint64_t len);
/* H2D of a filled slot holding `n_items` items: one cudaMemcpyAsync per column from the pinned slot,
 * or one per item from shared memory for columns that arrived by reference; then the slot's event. */
int bt_ring_submit_batch(bt_ring* r, int slot, const bt_schema* s, int n_items, void* dev_dst, void* stream);
/* Copy the slot's pending shared-memory payloads into the pinned slot now and release them (used
 * when a stream stalls with a partly filled batch, so that publishers never wait on a held slot). */
int bt_ring_spill(bt_ring* r, int slot, const bt_schema* s);
/* Non-blocking: release references of slots whose copies have completed. */
int bt_ring_reap(bt_ring* r);
/* Batch assembly by DMA from a recording kept in page-locked host memory (no reference counterpart:
 * file.py:100-108 re-reads and re-unpickles the file every epoch): row index[i] of the host array
 * `src_host` (rows of row_bytes) goes to row i of `dst_dev`, on `stream`; neighbouring source rows are
 * merged into one copy.  No CPU touches the payload.  src_host must be page-locked for the copies to
 * be asynchronous. */
int bt_gather_h2d(void* dst_dev, const void* src_host, uint64_t row_bytes, const int64_t* index, int n, void* stream);

/* ------------------------------------------------------------------------------------------
 * Receive loop in C.
 * Replaces the per-message Python of pkg_pytorch/blendtorch/btt/dataset.py:97-106 (poll, recv,
 * pickle.loads, yield): one call fills a whole batch slot from a ZMQ PULL socket without the GIL.
 * libzmq is not linked; the caller passes the entry points of the libzmq its binding already
 * uses (pyzmq: ctypes.CDLL of its bundled libzmq) and the raw socket (`socket.underlying`).
 * A receiver must only be used by the thread that owns the socket.
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_zmq_api {
  void* msg_init;   /* int   zmq_msg_init(zmq_msg_t*)              */
  void* msg_recv;   /* int   zmq_msg_recv(zmq_msg_t*, void*, int)  */
  void* msg_data;   /* void* zmq_msg_data(zmq_msg_t*)              */
  void* msg_size;   /* size_t zmq_msg_size(const zmq_msg_t*)       */
  void* msg_close;  /* int   zmq_msg_close(zmq_msg_t*)             */
  void* poll;       /* int   zmq_poll(zmq_pollitem_t*, int, long)  */
  void* zmq_errno;  /* int   zmq_errno(void)                       */
} bt_zmq_api;

typedef struct bt_rx bt_rx;

int bt_rx_create(const bt_zmq_api* api, void* zmq_socket, bt_rx** out);
int bt_rx_destroy(bt_rx* rx);
/* Wait up to timeout_ms for one message and hold it (BT_OK), or BT_E_TIMEOUT. */
int bt_rx_recv(bt_rx* rx, int timeout_ms);
/* The held message (valid until bt_rx_drop / the next fill). */
int bt_rx_pending(bt_rx* rx, const uint8_t** data, uint64_t* len);
int bt_rx_drop(bt_rx* rx);
/* Receive and ingest (bt_ring_ingest) up to max_items messages into items first_item.. of `slot`.
 * Returns BT_OK when max_items were placed; BT_E_TIMEOUT when no message arrived for slice_ms
 * (call again); BT_E_UNSUPPORTED / BT_E_SCHEMA when a message needs the caller's pickle.loads
 * path -- that message stays held (bt_rx_pending), the caller places it and calls bt_rx_drop.
 * *n_done = messages placed by this call in every case. */
int bt_rx_fill(bt_rx* rx, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
               int slice_ms, int* n_done);
/* Before closing the socket: take every message already queued on it (without waiting) and hand
 * the shared-memory slots they reference back to their publishers, so that a consumer that stops
 * early does not strand slots.  (Slots stranded anyway are reclaimed by the publisher after 30 s.) */
int bt_rx_release_queued(bt_rx* rx, int* n_released);

/* ------------------------------------------------------------------------------------------
 * Direct receiver: a minimal ZMTP 3 PULL endpoint (tcp:// and ipc://) that needs no libzmq and lands
 * array payloads straight from the kernel socket in the staging slot -- one CPU copy per payload byte
 * on the consumer side.  Replaces `socket.recv()` + `pickle.loads` of
 * pkg_pytorch/blendtorch/btt/dataset.py:98-106 against unmodified PUSH publishers
 * (pkg_blender/blendtorch/btb/publisher.py:21-27).  Same calling contract as bt_rx_*.
 * Like the reference's socket it connects to every address given (dataset.py:77-78), takes messages
 * fairly from whichever publisher has one, and retries publishers that are not up yet.
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_pull bt_pull;

int bt_pull_create(bt_pull** out);
int bt_pull_connect(bt_pull* q, const char* address);   /* "tcp://host:port" or "ipc://path" */
int bt_pull_destroy(bt_pull* q);
int bt_pull_recv(bt_pull* q, int timeout_ms);            /* hold one whole message (BT_OK) or BT_E_TIMEOUT */
int bt_pull_pending(bt_pull* q, const uint8_t** data, uint64_t* len);
/* Frames of the held message: returns their number and writes up to max_parts sizes.  1 for the
 * reference wire.  More for the opt-in raw-parts wire (no reference counterpart; publisher.py:43 sends
 * one pickled frame): frame 0 is the pickled dict in which every large array is the tuple
 * ("__btraw__", k, dtype_str, shape) and frame k holds that array's bytes; bt_pull_pending then spans
 * all frames back to back.  bt_pull_fill receives such frames straight into their schema columns. */
int bt_pull_pending_parts(bt_pull* q, uint64_t* sizes, int max_parts);
int bt_pull_drop(bt_pull* q);
int bt_pull_fill(bt_pull* q, bt_ring* ring, int slot, const bt_schema* schema, int first_item, int max_items,
                 int slice_ms, int* n_done);

/* ------------------------------------------------------------------------------------------
 * Shared-memory wire (opt-in, co-located producers).
 * No counterpart in the reference, whose only transport is the pickled ZMQ message
 * (btb/publisher.py:43).  A producer writes each large array into a slot of a POSIX shared-memory
 * segment and sends, over the same ZMQ socket, the usual pickled dict with the array replaced by the
 * tuple ("__btshm__", segment_name, ticket, dtype_str, shape).  The consumer maps the segment once,
 * page-locks it with cudaHostRegister, and copies slots to the device by DMA: zero CPU copies on the
 * consumer side.  A slot is busy from bt_shm_acquire (producer) until bt_shm_release (consumer).
 *
 * ticket = slot | generation << 16.  Every hand-out of a slot bumps its generation with a
 * compare-and-swap (several producers may share a segment), the ticket travels in the message, and
 * the consumer compares generations before and after it reads the slot (bt_shm_check; bt_shm_release
 * is a compare-and-swap on the same word).  So a slot that its producer took back while a slow
 * consumer still held a reference is *detected* (BT_E_IO from the ingest / ring calls), not read as if
 * nothing had happened, and the late release cannot free the slot under its new owner.
 * Taking slots back: bt_shm_acquire reuses a busy slot once it has been held longer than the
 * segment's lease (default 30 s, environment BT_SHM_LEASE_MS at creation, bt_shm_set_lease; 0 =
 * never).  That protects producers from consumers that died with messages queued; it is lossy for
 * a live consumer that is slower than the lease, which then fails loudly as described.
 * Functions that take `ticket` also accept a bare slot number (generation 0 = compare nothing).
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_shm bt_shm;

int bt_shm_create(const char* name, uint64_t slot_bytes, int nslots, bt_shm** out); /* producer; nslots <= 65536 */
int bt_shm_attach(const char* name, int device, bt_shm** out); /* consumer; device >= 0 page-locks it */
int bt_shm_close(bt_shm* s);                                    /* creator also unlinks the name */
int bt_shm_nslots(const bt_shm* s);
uint64_t bt_shm_slot_bytes(const bt_shm* s);
void* bt_shm_slot_ptr(bt_shm* s, int ticket);
/* Producer: ticket of the first free slot in round-robin order; BT_E_TIMEOUT after timeout_ms
 * (<0: wait forever, 0: poll once). */
int bt_shm_acquire(bt_shm* s, int timeout_ms, int* ticket);
int bt_shm_set_lease(bt_shm* s, int64_t lease_ms);
/* Consumer: BT_OK while the slot still belongs to this ticket, BT_E_IO once it was taken back. */
int bt_shm_check(const bt_shm* s, int ticket);
/* Consumer: free the slot; BT_E_IO (and no effect) when the ticket is stale. */
int bt_shm_release(bt_shm* s, int ticket);

/* ------------------------------------------------------------------------------------------
 * Native publisher: a minimal ZMTP 3 PUSH endpoint (tcp:// and ipc://, NULL security), the producer
 * end of the ingest wire without libzmq or an interpreter per producer.
 * Replaces pkg_blender/blendtorch/btb/publisher.py:21-43 (`DataPublisher`: PUSH socket, bind,
 * SNDHWM, IMMEDIATE=1, `send_pyobj`) for producers that already hold pickled messages (replaying
 * recordings, synthetic load, simulators written in C): messages go from the caller's memory to
 * the kernel socket with one copy, framed exactly as libzmq frames them, so unmodified
 * `btt.RemoteIterableDataset` consumers (pkg_pytorch/blendtorch/btt/dataset.py:73-78,98-105) and the
 * direct receiver above take them alike.  Semantics kept from the reference's socket options:
 * messages only go to peers whose connection is complete (IMMEDIATE=1, publisher.py:26), peers are
 * served round-robin skipping those that cannot take more (PUSH load balancing), and the sender
 * blocks when no peer can take a message (SNDHWM back-pressure, publisher.py:24; the queue here is
 * the kernel socket buffer plus one message in progress per peer).
 * One bt_push holds any number of bound endpoints ("publishers").  Threading: bind every endpoint before
 * any thread sends; after that an endpoint belongs to one thread at a time -- bt_push_send, bt_push_run
 * over a range that contains it, and bt_push_peers all touch its connection list -- while different
 * endpoints may be driven by different threads concurrently.
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_push bt_push;

int bt_push_create(bt_push** out);
/* Bind one more endpoint ("tcp://host:port", port 0 or "*" = any free port; "ipc://path").  Returns
 * its index (>= 0).  bt_push_endpoint gives the address peers connect to (resolved port). */
int bt_push_bind(bt_push* p, const char* address);
int bt_push_endpoint(bt_push* p, int endpoint, char* address_out, int capacity);
int bt_push_destroy(bt_push* p);
/* Peers of `endpoint` whose handshake is complete (accepts pending connections first). */
int bt_push_peers(bt_push* p, int endpoint);
/* Send one message of `nparts` frames (1 = the reference wire; more = the raw-parts wire, frame 0 the
 * pickled header) to the next peer that can take it; blocks up to timeout_ms (<0: forever) while no
 * peer is connected or every peer's socket is full, then BT_E_TIMEOUT. */
int bt_push_send(bt_push* p, int endpoint, const uint8_t* const* parts, const uint64_t* lens, int nparts, int timeout_ms);
/* Synthetic load / replay loop, to be called from one thread per disjoint endpoint range: endpoint e in
 * [first, first + count) sends messages pool[e - first][0..pool_n) in a cycle, `messages` of them (<0:
 * until *stop becomes non-zero), keeping every connected peer's socket busy.  pool is an array of
 * count * pool_n (pointer, length) pairs, endpoint-major.
 * With `seg` non-null the payload goes over the shared-memory wire instead: for message m,
 * payload[m] (payload_len[m] bytes) is copied into a free slot and tmpl[m] -- the pickled message with
 * the array replaced by ("__btshm__", name, ticket, dtype, shape), whose ticket is a 4-byte BININT at
 * byte tmpl_ticket_off[m] -- is sent with that slot's ticket patched in.
 * With seg non-null and payload NULL frames are published *in place*, the way a renderer that draws
 * straight into shared memory does: slot k already holds the frame that tmpl[k] describes (pool_n must
 * equal the slot count), publishing is acquiring a free slot and sending its template -- no producer copy.
 * *sent (optional) is updated as messages go out.  Returns BT_OK when all were sent or stop was seen. */
typedef struct bt_push_job {
  int first_endpoint, n_endpoints, pool_n, reserved;
  int64_t messages;                 /* per endpoint; < 0: until stopped */
  const uint8_t* const* pool;       /* [n_endpoints * pool_n] messages (or templates when seg != NULL) */
  const uint64_t* pool_len;
  bt_shm* const* seg;               /* NULL, or one segment per endpoint */
  const uint8_t* const* payload;    /* [n_endpoints * pool_n] array bytes for the shared-memory wire */
  const uint64_t* payload_len;
  const uint32_t* tmpl_ticket_off;  /* [n_endpoints * pool_n] */
  volatile int* stop;               /* may be NULL */
  volatile int64_t* sent;           /* may be NULL; total over the range */
} bt_push_job;
int bt_push_run(bt_push* p, const bt_push_job* job);

/* ------------------------------------------------------------------------------------------
 * K1: fused frame preprocessing on the device.
 * Replaces, for a whole batch in one launch:
 *   np.flipud                    pkg_blender/blendtorch/btb/offscreen.py:95-96
 *   gamma (_color_correct)       pkg_blender/blendtorch/btb/offscreen.py:105-112
 *                                == examples/datagen/generate.py:10-14
 *   astype(f32), (x-m)/s, CHW    examples/densityopt/densityopt.py:117-119
 *   stack into a batch           torch default_collate (collate.py:246-289)
 * Semantics (identical to composing the cited lines):
 *   v  = src[b, flip ? H-1-y : y, x, c]                         (uint8)
 *   g  = (use_gamma && c < 3) ? gamma_lut[v] : v                (uint8, truncating gamma)
 *   o  = use_norm ? (float(g) - mean[c]) / std[c] : float(g)    (IEEE fp32 sub and div)
 *   dst[b, c, y, x] (chw) or dst[b, y, x, c] = cast(o)          (f32 / bf16 RNE / u8 = g)
 * Only the first `out_channels` channels are written (out_channels <= channels).
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_preprocess_params {
  const void* src;        /* device, uint8 [batch][height][width][channels] */
  void* dst;              /* device, out_dtype */
  int32_t batch, height, width, channels; /* channels in 1..4 */
  int32_t out_channels;   /* 1..channels */
  int32_t out_dtype;      /* BT_U8, BT_F32 or BT_BF16 */
  int32_t flip;           /* 1 = vertical flip */
  int32_t chw;            /* 1 = planar output [b][c][y][x], 0 = interleaved [b][y][x][c] */
  int32_t use_gamma;
  int32_t use_norm;       /* ignored for BT_U8 output */
  int32_t variant;        /* 0 = auto, 1 = generic scalar (one value per thread), 2 = register-vectorised
                             (fails if rows are not whole vector units), 3 = bulk-async (cp.async.bulk + mbarrier ring)
                             staged input for RGBA planar output, kept as the measured A/B of 2; 4 = one pixel per
                             thread (any width); auto may also pick, for RGBA rows of any width and planar output, the
                             kernel that stores aligned groups of four plane elements, reported as 5 by
                             bt_preprocess_last_variant() */
  int32_t reserved;       /* 0; developer tuning knob (units per thread | CTAs-per-SM cap << 8) */
  uint64_t src_image_stride; /* bytes between consecutive images in src; 0 = dense */
  float mean[4];
  float stdev[4];
  uint8_t gamma_lut[256];
  /* Optional per-batch statistics of the values written (north_star "optional per-batch
   * stats"; no reference counterpart): stats_out[c*2+0] = sum, [c*2+1] = sum of squares, as
   * float64, deterministic (fixed reduction order).  Needs `stats_scratch` of
   * bt_preprocess_scratch_bytes() bytes of device memory (uninitialised is fine); the sums are
   * added up by one or two small kernels enqueued behind K1 on the same stream.  NULL = off. */
  double* stats_out;
  void* stats_scratch;
  /* Optional gather (no reference counterpart: the reference re-reads and re-decodes the file every
   * epoch, file.py:100-108): device array of `batch` int64; batch item b is read from image
   * src_index[b] of `src` (stride src_image_stride, or one dense frame).  Lets a sampler draw
   * batches from frames that are resident in HBM (a recording cached on the device) with no copy
   * pass.  The caller guarantees every index is a valid image of `src`.  NULL = item b reads image b. */
  const int64_t* src_index;
} bt_preprocess_params;

/* The 256-entry table of offscreen.py:105-112 for one coefficient:
 * lut[v] = uint8(255.0f * powf(v / 255.0f, (float)(1.0 / coeff))), truncating. */
int bt_build_gamma_lut(double coeff, uint8_t lut[256]);
uint64_t bt_preprocess_scratch_bytes(const bt_preprocess_params* p);
int bt_preprocess_launch(const bt_preprocess_params* p, void* stream);
/* Which variant the last bt_preprocess_launch on this thread ran, and how many kernels it enqueued
 * (K1, plus one or two small fold kernels when statistics are on). */
int bt_preprocess_last_variant(void);
int bt_preprocess_last_kernels(void);

/* Vertical flip of frames of any element type: dst[b][y][:] = src[src_index ? src_index[b] : b][height-1-y][:],
 * rows of row_bytes bytes, dense images.  For fields K1 does not take (a float32 depth map read back
 * bottom-up has the same order as the colour buffer that pkg_blender/blendtorch/btb/offscreen.py:95-96
 * flips with np.flipud; BASELINE config 3 carries such a map).  A copy at HBM speed. */
int bt_flip_rows_launch(const void* src, void* dst, int batch, int height, uint64_t row_bytes, const int64_t* src_index,
                        void* stream);

/* K1 with per-batch statistics for callers that cannot allocate device memory.
 * (No counterpart in the reference: it never touches the device.)
 *
 * Same semantics and same kernel as bt_preprocess_launch with `stats_out` set, but the scratch
 * the statistics need (one row of partial sums per CTA) is a fixed array compiled into
 * the library instead of caller memory, and the call performs no cudaMalloc / cudaFree or any
 * other allocating routine.  Meant for hosts that give a plug-in a fixed device-memory budget,
 * for ingest loops that may not call an allocator after start-up, and for processes sharing a
 * GPU whose memory another framework has already claimed.  The caller still passes src, dst
 * and stats_out; `stats_scratch` in the params is ignored.
 *
 * Limits and rules:
 *  - batch*height*width*channels must not exceed 19 398 656 bytes (1184 rows of partial sums,
 *    one per 16 KiB of input); larger batches fail with BT_E_TOO_LARGE and a message saying so.
 *  - The array costs 76 032 bytes on every GPU that loads the library, until the process exits.
 *  - The first call on each device resolves the array's device address (and, like any first
 *    launch, loads the kernels); make that first call before starting any stream capture.
 *    The address is remembered per device ordinal for the life of the process: a host that resets
 *    the device (cudaDeviceReset) or destroys its primary context must reload the library before
 *    calling this again, because the remembered address belongs to the context that is gone.
 *  - All calls on a device share the one array.  Nothing in it has to start at zero: the partial
 *    sums are written by each call's kernel before its fold reads them, so back-to-back calls on
 *    one stream are fine; calls whose kernels can overlap on the GPU (different streams or
 *    threads without ordering between them) race on it and give wrong statistics. */
int bt_preprocess_stats_static(const bt_preprocess_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * K2: batched annotation projection world -> NDC -> pixel (+ linear depth, + pixel bboxes).
 * Replaces Camera.world_to_ndc  pkg_blender/blendtorch/btb/camera.py:84-112,
 *          Camera.ndc_to_pixel  pkg_blender/blendtorch/btb/camera.py:115-136,
 *          hom / dehom          pkg_blender/blendtorch/btb/utils.py:112-121
 * for `batch` frames of `npoints` points each, float64 arithmetic as in the reference.
 * The pixel-bbox reduction (min/max over each object's `corners_per_object` projected corners)
 * has no reference counterpart (camera.py:165-189 stops at the corners); BASELINE config 3 asks
 * for it.
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_project_params {
  const double* points;   /* device [batch][npoints][3] world coordinates */
  const double* view;     /* device [batch or 1][4][4] row-major */
  const double* proj;     /* device [batch or 1][4][4] row-major */
  int32_t batch, npoints;
  int32_t view_per_frame; /* 1: one matrix per frame, 0: shared */
  int32_t proj_per_frame;
  int32_t height, width;  /* image shape for ndc_to_pixel */
  int32_t origin_upper_left; /* 1 = 'upper-left' (default of the reference), 0 = 'lower-left' */
  int32_t with_depth;     /* 1 = two products + depth (camera.py:106-110), 0 = fused P@V (:111-112) */
  double* ndc_out;        /* optional device [batch][npoints][3] */
  double* xy_out;         /* optional device [batch][npoints][2] pixels */
  double* depth_out;      /* optional device [batch][npoints], needs with_depth */
  int32_t corners_per_object; /* for bbox_out; npoints % corners_per_object == 0, <= 32 and a power of two */
  int32_t input_is_ndc;   /* 1: `points` already holds NDC (ndc_to_pixel alone, camera.py:115-136);
                             view/proj are ignored and may be NULL */
  double* bbox_out;       /* optional device [batch][npoints/corners_per_object][4] = xmin,ymin,xmax,ymax */
} bt_project_params;

int bt_project_launch(const bt_project_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * K3: synthetic Cube-scene frame source on the GPU.
 * Stands in for the draw + read-back of pkg_blender/blendtorch/btb/offscreen.py:76-92 in benchmarks
 * and self-checks, without any graphics API: an edge-function rasteriser for boxes (12 triangles
 * each), nearest-depth test, flat two-sided shading.  Corners are projected with K2's arithmetic
 * (camera.py:84-136), so annotation corners projected by K2 lie on the rasterised silhouettes.
 * Triangles with a corner at or behind the camera plane are dropped (no clipping).
 * ------------------------------------------------------------------------------------------ */
typedef struct bt_raster_params {
  const double* corners;  /* device [batch][ncubes][8][3] world coordinates; corner order (x,y,z) in
                             {+1,-1}^3 with z fastest, then y, then x (the default-cube order) */
  const double* view;     /* device [batch or 1][4][4] row-major */
  const double* proj;     /* device [batch or 1][4][4] row-major */
  int32_t batch, ncubes;  /* ncubes <= 32 */
  int32_t view_per_frame, proj_per_frame;
  int32_t height, width;
  int32_t bottom_up;      /* 1: row 0 of the output is the bottom row (GL read-back order, the input
                             K1's flip expects); 0: top row first */
  float far_depth;        /* depth written where nothing was hit */
  uint8_t background[4];  /* RGBA */
  float light[3];         /* light direction in camera space (any length) */
  int32_t reserved;
  const uint8_t* colors;  /* optional device [ncubes][4] base colours; NULL = built-in palette */
  uint8_t* image_out;     /* device [batch][height][width][4] uint8 RGBA */
  float* depth_out;       /* optional device [batch][height][width] linear camera-space depth */
} bt_raster_params;

int bt_raster_launch(const bt_raster_params* p, void* stream);

/* ------------------------------------------------------------------------------------------
 * Small CUDA helpers so that callers without a CUDA binding can drive the path
 * (reference counterpart: none; the reference never touches the device inside the library).
 * ------------------------------------------------------------------------------------------ */
int bt_device_count(int* n);
int bt_measure_h2d(int device, uint64_t nbytes, int iters, double* gb_per_s); /* pinned H2D roofline */

#ifdef __cplusplus
}
#endif
#endif /* BT_B200_H */
