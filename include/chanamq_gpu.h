/* libchanagpu.so — C ABI of the B200 routing path for ChanaMQ.
 *
 * The reference (Scala/Akka, no FFI of any kind) routes one message at a time through
 *   trait QueueMatcher { subscribe; unsubscribe; lookup }          QueueMatcher.scala:11-27
 * owned by one ExchangeEntity actor (ExchangeEntity.scala:126, chosen by exchange type at
 * ExchangeEntity.scala:149-154 / 211-216).  This library sits behind that seam: a Scala
 * `GpuMatcher extends QueueMatcher` binds the entry points below through JNA and
 * hands a whole `Publishs` batch (ExchangeEntity.scala:293-308) to one chq_route call.
 *
 * Conventions: every function returns CHQ_OK (0) or a negative status; nothing throws or aborts
 * across the boundary; caller buffers are only read during the call; results point into
 * library-owned pinned memory that stays valid until the next route call on the same table.
 * Calls on ONE table must be serialised by the caller (one exchange = one actor = one thread in
 * the reference, ExchangeEntity.scala:107); bind / unbind / commit may come from another thread (the
 * reference binds from a Future callback, ExchangeEntity.scala:337-338) and are locked against each other.
 * Different tables are independent.  The table-less device-level calls (chq_decode_device, chq_fanout_device,
 * chq_merge_slices_device*) keep ONE scratch per GPU: any thread may call them, the library queues their
 * launches one call after the other, but two such calls of the same kind whose kernels overlap on one GPU
 * (different streams) would share that scratch — order them on the device (same stream, or an event).
 * chq_decode (host buffers) holds the lock for its whole call and is safe from any number of threads.
 *
 * Paths relative to /root/reference/chana-mq-server/src/main/scala/chana/mq/amqp/server/.
 */
#ifndef CHANAMQ_GPU_H
#define CHANAMQ_GPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct chq_table chq_table; /* one per exchange; opaque */

/* AMQP.ExchangeType (chana-mq-base/.../model/AMQP.scala:33-46) */
enum { CHQ_DIRECT = 0, CHQ_FANOUT = 1, CHQ_TOPIC = 2, CHQ_HEADERS = 3 };
/* COMPAT = what the reference computes: '*' is one word, '#' is an ordinary word
 *          (TrieMatcher.WILDCARD, engine/QueueMatcher.scala:70), headers routed as topic
 *          (entity/ExchangeEntity.scala:153,215).
 * SPEC   = AMQP 0-9-1 text: '#' is zero or more words. */
enum { CHQ_MODE_COMPAT = 0, CHQ_MODE_SPEC = 1 };

enum {
  CHQ_OK = 0,
  CHQ_E_ZERO_WORDS = -1,   /* "." / ".." : the reference throws ArrayIndexOutOfBounds (QueueMatcher.scala:420,528) */
  CHQ_E_KEY_TOO_LONG = -2, /* more than 255 bytes: not an AMQP shortstr */
  CHQ_E_OVERFLOW = -3,     /* result larger than the ids capacity given / than 2^32-1 ids */
  CHQ_E_CUDA = -4,         /* CUDA failure; chq_last_error() has the text.  There is no CPU fallback. */
  CHQ_E_ARG = -5,
  CHQ_E_NOT_COMMITTED = -6,
  CHQ_E_SCRATCH = -7       /* device scratch arena exhausted (device-level entry points only) */
};

const char* chq_last_error(void); /* thread-local text of the last failure */

/* ---- table life cycle (ExchangeEntity Declare / load, entity/ExchangeEntity.scala:137-165, 211-216) ---- */
/* device = CUDA ordinal.  device = -1 makes a host-only table: bind/commit/stats/queue names work (the
 * compiler can be checked without a GPU) but every route call fails with CHQ_E_CUDA. */
int32_t chq_table_create(chq_table** out, int32_t exchange_type, int32_t mode, int32_t device);
int32_t chq_table_destroy(chq_table* t);

/* QueueMatcher.subscribe as called from queueBind (entity/ExchangeEntity.scala:338):
 * subscriber = TopicSubscriber(queue, isDurable) (entity/ExchangeEntity.scala:97).
 * Binding an identical (key, queue, durable) triple again is a no-op (engine/QueueMatcher.scala:443-446). */
int32_t chq_bind(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                 uint8_t durable);
/* The same with the binding's arguments (Queue.Bind arguments -> ExchangeEntity.QueueBind.arguments,
 * entity/ExchangeEntity.scala:90): args_table = the CONTENTS of the AMQP field table (after its length word),
 * typed as on the wire.  Only a CHQ_HEADERS table in CHQ_MODE_SPEC looks at them: x-match = "all" (default)
 * or "any"; other "x-" names are ignored; a void argument asks for the header's presence, any other for the
 * same type tag and the same bytes.  A malformed table is CHQ_E_ARG. */
int32_t chq_bind_args(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                      uint8_t durable, const uint8_t* args_table, uint32_t args_len);
/* bulk form of the same call, for replaying the stored binds at start-up (entity/ExchangeEntity.scala:155-157) */
int32_t chq_bind_many(chq_table* t, const uint8_t* keys_blob, const uint32_t* key_offsets,
                      const uint8_t* queues_blob, const uint32_t* queue_offsets, const uint8_t* durable,
                      uint32_t n);
/* QueueMatcher.unsubscribe as called from queueUnbind(queue, routingKey) (entity/ExchangeEntity.scala:380).
 * Clean set removal: the reference's tombstone defect is not reproduced. */
int32_t chq_unbind(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                   uint8_t durable);
/* queueUnbind(queue): drop every binding of a deleted queue (entity/ExchangeEntity.scala:344-351) */
int32_t chq_unbind_queue(chq_table* t, const uint8_t* queue, uint32_t queue_len);
/* Make the binds / unbinds since the last commit visible to routing.
 * The first commit (and chq_commit_full always) compiles the whole binding set and uploads it; queue ids are assigned
 * there: the rank of the queue name among names that hold a binding, ascending by bytes, so ascending id = canonical
 * order.  Later commits apply the changes IN PLACE where they can (the reference's subscribe is O(depth) and visible
 * at once, engine/QueueMatcher.scala:385-392, 415-460): a new queue on a known or new pattern, a new pattern word, an
 * unbind — microseconds of host work and a few small copies behind the batches in flight.  A queue name the table has
 * not seen takes the next free id, so until the next full commit ascending id is still the order inside a result but
 * no longer name order for those queues; a queue that lost its last binding keeps its id (and stays in
 * chq_queue_name) until then.  Changes that cannot be made in place (a pattern shape the table has not
 * seen, an index more than half full, a '#' that widens the word rows, fan-out and headers tables, more than 4096
 * changes at once) fall back to the full compile by themselves.  chq_stats counts both kinds. */
int32_t chq_commit(chq_table* t);
int32_t chq_commit_full(chq_table* t);
int32_t chq_n_queues(const chq_table* t, uint32_t* n);
int32_t chq_queue_name(const chq_table* t, uint32_t queue_id, const uint8_t** name, uint32_t* len);

/* ---- batched lookup: the per-message body of ExchangeEntity.publish (entity/ExchangeEntity.scala:293-302)
 *      for a whole Publishs batch.  Message i gets queue_ids[offsets[i] .. offsets[i+1]) — the SET of
 *      queue names over all matching bindings (entity/ExchangeEntity.scala:300), ascending;
 *      any_durable[i] = OR of the matched subscribers' isDurable (entity/ExchangeEntity.scala:302);
 *      err[i] = 1 for a zero-word key, 2 for a key longer than 255 bytes (no queues then);
 *      headers exchanges: 3 for a malformed headers table, 4 for more than 64 header fields whose names
 *      some binding mentions. ---- */
typedef struct {
  uint32_t n;
  const uint32_t* offsets;   /* n+1 */
  const uint32_t* queue_ids; /* offsets[n]; with CHQ_IDS_U24 the same ids as 3-byte values, see chq_set_ids_format */
  const uint8_t* any_durable; /* n */
  const uint8_t* err;        /* n */
  int32_t ids_format;        /* CHQ_IDS_U32 or CHQ_IDS_U24: how queue_ids is laid out */
} chq_result;

/* ---- a smaller host result (no counterpart in the reference: the JVM path never moves a result between memories).
 *      End to end, chq_route is the device->host copy of the queue ids and nothing else.  A table with
 *      fewer than 2^24 queues may ask for them as 3 bytes each, little-endian, back to back — a quarter less to
 *      copy: with b = (const uint8_t*)result.queue_ids, id j of the batch is b[3j] | b[3j+1] << 8 | b[3j+2] << 16,
 *      and message i's ids are j in [offsets[i], offsets[i+1]) as before.  The array is padded with zero bytes to a
 *      multiple of 4.  Applies to chq_route / chq_route_submit + chq_route_wait (host results) only; the device
 *      entry points always write uint32.  A batch submitted while the table has 2^24 queues or more is refused
 *      (CHQ_E_ARG).  chq_unpack_u24 expands n_ids packed ids into uint32 on the host (plain C loop; a consumer
 *      that walks the ids once reads them in place instead). ---- */
#define CHQ_IDS_U32 0
#define CHQ_IDS_U24 1
int32_t chq_set_ids_format(chq_table* t, int32_t format);
void chq_unpack_u24(const uint8_t* packed, uint64_t n_ids, uint32_t* out);

/* Host buffers in, host result out (H2D, kernels, D2H inside).  keys_blob / key_offsets may be
 * ordinary or pinned host memory; pinned (chq_host_alloc) skips one staging copy.
 * headers_blob / headers_offsets: raw AMQP field tables per message for CHQ_HEADERS in SPEC mode, else NULL. */
int32_t chq_route(chq_table* t, const uint8_t* keys_blob, const uint32_t* key_offsets, uint32_t n,
                  const uint8_t* headers_blob, const uint32_t* headers_offsets, chq_result* out);

/* Pipelined form: up to CHQ_SLOTS batches in flight on one table (H2D of batch i+1 overlaps the
 * kernels and D2H of batch i).  submit() returns at once; wait() blocks until that slot's result is
 * on the host.  A slot's result stays valid until the slot is submitted again. */
#define CHQ_SLOTS 4
int32_t chq_route_submit(chq_table* t, int32_t slot, const uint8_t* keys_blob, const uint32_t* key_offsets,
                         uint32_t n);
int32_t chq_route_wait(chq_table* t, int32_t slot, chq_result* out);

/* Device buffers in, device result out, asynchronous on `stream` (a cudaStream_t; NULL = default
 * stream).  For a CHQ_HEADERS table in CHQ_MODE_SPEC the "keys" are the messages' headers tables (field
 * table contents), here and in chq_route_keys_device (which takes the hdr_tbl_off / hdr_tbl_len of decode records).  d_status receives a chq status word once the stream reaches it; d_total the id count.
 * d_ids holds ids_cap entries; a larger result sets CHQ_E_OVERFLOW and writes nothing past it. */
int32_t chq_route_device(chq_table* t, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                         uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                         uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream);

/* No counterpart in the reference (it has no device code at all); added for embedders that are not
 * allowed to allocate.
 *
 * Same contract and same kernels as chq_route_device, but the call allocates NOTHING: no cudaMalloc,
 * no cudaFree, no allocating library routine.  Who needs that: a broker thread that routes inside a
 * latency budget and cannot afford a cudaMalloc (it synchronises the whole device and can stall for
 * milliseconds while other streams are busy); a process that has handed all device memory to a pool
 * and would get cudaErrorMemoryAllocation from a fresh request; a deployment whose rule is that
 * nothing is allocated after start-up, so that steady-state routing can never fail for lack of memory.
 *
 * The scratch the launches need (prefix-sum state, the interned word ids and list descriptors of the batch, the arena
 * of the rare merge slow path) sits in one fixed array inside the library image: about 11 MiB of
 * device memory on every GPU that loads libchanagpu.so, held until the process exits.  Hence the limits:
 *   n                        <= CHQ_STATIC_MAX_KEYS  (65 536 keys per call)
 *   n x the table's word stride <= CHQ_STATIC_MAX_WORDS (524 288; the stride is 8 for ordinary tables,
 *                               so the first limit is the one that bites; it is 256 when a pattern has
 *                               '#' before its end or more than 16 words, which leaves 2 048 keys per call)
 * A larger request is refused with CHQ_E_ARG and a message saying which limit was hit; nothing is
 * launched.  The slow-path arena holds 1 Mi queue ids; a batch that needs more reports CHQ_E_SCRATCH
 * in *d_status, exactly like chq_route_device.  A CHQ_HEADERS table in CHQ_MODE_SPEC is refused with
 * CHQ_E_ARG: x-match routing needs a per-batch descriptor arena that this fixed array does not carry
 * (such a table routed through here would silently match nothing); use chq_route_device for it.
 *
 * The first call on a device looks up the array's device address (cudaGetSymbolAddress) and sizes
 * the launch; make that first call during start-up, before any stream capture begins.
 * All calls share the one array, and it keeps the previous call's contents (the function clears the
 * part that has to start at zero).  Two calls whose work overlaps on the GPU — different streams,
 * different tables, different threads — race on it and corrupt each other: serialise them. */
#define CHQ_STATIC_MAX_KEYS 65536u
#define CHQ_STATIC_MAX_WORDS 524288u
int32_t chq_route_device_static(chq_table* t, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                                uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                                uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream);

/* ---- binding-slice sharding (BASELINE.json configs[3]: binding i lives on GPU i mod G; every GPU routes ALL
 *      keys against its slice; the per-slice lists are all-gathered and merged).  No counterpart in the
 *      reference (one exchange = one actor there, entity/ExchangeEntity.scala:107).
 *      chq_declare_queues makes queue names count for the id numbering of a table although nothing is bound
 *      to them on this GPU, so that every slice numbers the whole exchange's queues identically. ---- */
int32_t chq_declare_queues(chq_table* t, const uint8_t* queues_blob, const uint32_t* queue_offsets, uint32_t n);
/* Union, per message, of n_slices (<= 8) CSR results of the same n keys, laid out as the all-gather leaves
 * them: offsets [n_slices][n+1], ids [n_slices][slice_cap], any_durable [n_slices][n].  Asynchronous on
 * `stream`; result and status exactly as chq_route_device. */
int32_t chq_merge_slices_device(int32_t device, const uint32_t* d_slice_offsets, const uint32_t* d_slice_ids, uint64_t slice_cap,
                                const uint8_t* d_slice_durable, uint32_t n_slices, uint32_t n, uint32_t* d_offsets,
                                uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable, uint64_t* d_total,
                                int32_t* d_status, void* stream);
/* The same with flags: CHQ_MERGE_DISJOINT = the caller knows that no queue is bound on two slices (true when
 * every queue has a single binding), so the per-message unions need no de-duplication — about twice as fast. */
enum { CHQ_MERGE_DISJOINT = 1 };
int32_t chq_merge_slices_device2(int32_t device, const uint32_t* d_slice_offsets, const uint32_t* d_slice_ids, uint64_t slice_cap,
                                const uint8_t* d_slice_durable, uint32_t n_slices, uint32_t n, uint32_t* d_offsets,
                                uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable, uint64_t* d_total,
                                int32_t* d_status, uint32_t flags, void* stream);

/* ---- the same layout as ONE call per batch, with the exchange between the GPUs inside the library (NCCL over NVLink
 *      plus direct loads from the peers' result buffers).  One process per GPU; one chq_shard per (exchange, rank).
 *      No counterpart in the reference (entity/ExchangeEntity.scala:107).  Per batch every rank is given the SAME n
 *      keys; rank g comes back with the finished result of ITS range of messages, [g n / G, (g+1) n / G): all ranks
 *      together hold the batch's result, every queue id crossed NVLink at most once.
 *        chq_shard_unique_id   on one rank; the 128 bytes reach the others by any means (the broker's own cluster
 *                              channel, a file, MPI, torch.distributed)
 *        chq_shard_create      collective (ncclCommInitRank): rank, world (<= 8), CUDA device of this process
 *        chq_shard_bind_many   every rank is given the same bindings, numbered first_index, first_index + 1, ...;
 *                              binding i is kept by rank i mod world, every queue name is declared on every rank
 *                              (single bindings can also go through chq_shard_table + chq_bind / chq_declare_queues)
 *        chq_shard_commit      chq_commit of the slice
 *        chq_shard_reserve     collective: buffers for batches of up to max_keys keys; slice_cap = ids one slice may
 *                              produce for a whole batch, pull_cap = ids of ONE slice that may fall into one rank's
 *                              message range.  The ranks exchange cudaIpc handles of their slice buffers here.
 *        chq_shard_route_device  collective, asynchronous on `stream`.  d_offsets[(end - first) + 1], d_queue_ids,
 *                              d_any_durable[end - first], d_err[end - first] describe the messages [*first_message,
 *                              *end_message) of the batch (offsets start at 0); *d_status = CHQ_OK, or the first failure
 *                              of ANY rank's part (CHQ_E_OVERFLOW: slice_cap / pull_cap / ids_cap too small;
 *                              CHQ_E_SCRATCH), the same on every rank — no rank is left waiting in a collective.
 *      NCCL is loaded on first use (dlopen of libnccl.so.2, the copy already in the process if there is one); without
 *      it these calls return CHQ_E_CUDA.  chq_shard_last_error() has the text of this file's failures. ---- */
typedef struct chq_shard chq_shard;
#define CHQ_SHARD_ID_BYTES 128
int32_t chq_shard_unique_id(uint8_t* id128);
int32_t chq_shard_create(chq_shard** out, const uint8_t* id128, int32_t rank, int32_t world, int32_t device,
                         int32_t exchange_type, int32_t mode);
int32_t chq_shard_destroy(chq_shard* s);
chq_table* chq_shard_table(chq_shard* s);
int32_t chq_shard_bind_many(chq_shard* s, const uint8_t* keys_blob, const uint32_t* key_offsets, const uint8_t* queues_blob,
                            const uint32_t* queue_offsets, const uint8_t* durable, uint32_t n, uint32_t first_index);
int32_t chq_shard_commit(chq_shard* s);
/* 1 when no queue name was bound on two slices so far (the unions then skip de-duplication) */
int32_t chq_shard_is_disjoint(const chq_shard* s);
int32_t chq_shard_reserve(chq_shard* s, uint32_t max_keys, uint64_t slice_cap, uint64_t pull_cap);
int32_t chq_shard_route_device(chq_shard* s, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                               uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                               uint8_t* d_err, uint64_t* d_total, int32_t* d_status, uint32_t* first_message,
                               uint32_t* end_message, void* stream);
/* Device time of the four stages of chq_shard_route_device (local route, all-gather, pull, merge), from CUDA events on
 * the caller's stream: synchronises the device, returns the sums (ms) over the batches since the last call (at most 64
 * are kept), then switches recording on or off. */
int32_t chq_shard_stage_ms(chq_shard* s, int32_t enable, double* stage_ms, uint32_t* batches);
const char* chq_shard_last_error(void);

/* ---- decode seam: FrameParser.onReceive (chana-mq-base/.../engine/FrameParser.scala:78-158) +
 *      CommandAssembler.onReceive (.../engine/CommandAssembler.scala:56-130) + Basic.Publish arguments
 *      (.../method/Basic.scala:45) + BasicProperties.readFrom (.../model/BasicProperties.scala:42-96) + the
 *      publish fields FrameStage derives (engine/FrameStage.scala:483-508), for MANY connection segments at once.
 *
 *      bytes[seg_offsets[s] .. seg_offsets[s+1]) is what connection s has received and not yet consumed; it
 *      must start on a frame boundary.  The call is stateless: seg_consumed[s] = bytes up to the end of the
 *      last COMPLETE command (or idle heartbeat); the caller keeps the rest and prepends it next time (the
 *      reference keeps that tail inside FrameParser / CommandAssembler instead).  One record per assembled
 *      command, segment by segment in stream order.  seg_err[s]: 0, 501 = FRAME_ERROR (bad end marker,
 *      wrong frame type for the assembler state, body overshoot; chana-mq-base/.../model/ErrorCodes.scala:17)
 *      or 1 = the reference throws here (unknown class id, truncated payload, continuation flag, negative
 *      size); the segment stops at its first error, seg_err_off[s] = the offending frame, and the records
 *      before it are still returned (FrameStage discards them and closes: engine/FrameStage.scala:320-335).
 *      Heartbeat frames are skipped (engine/FrameStage.scala:301-303). ---- */
typedef struct {
  uint64_t method_off;  /* offset of the METHOD frame in `bytes` */
  uint64_t body_off;    /* payload of the first non-empty BODY frame (0: none) */
  uint64_t body_size;   /* bodySize of the content header (what the BODY frames carry in total) */
  int64_t ttl;          /* expiration as Long.parseLong reads it; valid if CHQ_PUB_HAS_TTL */
  uint32_t exch_off, exch_len; /* kind 1: exchange name.   kind 0: the method frame's payload */
  uint32_t key_off, key_len;   /* kind 1: routing key.     kind 0: class id, method id */
  uint32_t hdr_tbl_off, hdr_tbl_len; /* contents of the `headers` field table (after its length word) */
  uint32_t seg;
  uint16_t channel;
  uint8_t kind;         /* 1 = Basic.Publish, decoded;  0 = any other method, handed back undecoded */
  uint8_t flags;        /* CHQ_PUB_* */
} chq_publish_rec;      /* 64 bytes */
enum {
  CHQ_PUB_MANDATORY = 1, CHQ_PUB_IMMEDIATE = 2,
  CHQ_PUB_PERSIST = 4,      /* deliveryMode == 2 */
  CHQ_PUB_HAS_TTL = 8, CHQ_PUB_HAS_HEADERS = 16,
  CHQ_PUB_HAS_PROPS = 32    /* a content header was decoded */
};
/* host buffers (H2D, kernels, D2H inside); at most 4 GiB per call */
int32_t chq_decode(const uint8_t* bytes, uint64_t len, const uint64_t* seg_offsets, uint32_t n_seg, chq_publish_rec* recs,
                   uint64_t recs_cap, uint64_t* n_recs, uint64_t* seg_consumed, int32_t* seg_err, uint64_t* seg_err_off);
/* device buffers, asynchronous on `stream`; *d_status = CHQ_OK or CHQ_E_OVERFLOW (more commands than recs_cap) */
int32_t chq_decode_device(const uint8_t* d_bytes, uint64_t len, const uint64_t* d_seg_offsets, uint32_t n_seg,
                          chq_publish_rec* d_recs, uint64_t recs_cap, uint64_t* d_n_recs, uint64_t* d_seg_consumed,
                          int32_t* d_seg_err, uint64_t* d_seg_err_off, int32_t* d_status, void* stream);
/* routing keys of the Basic.Publish records, compacted on the device (order not preserved: d_rec_index[j]
 * names the record key j came from) — the input of chq_route_keys_device */
int32_t chq_publish_keys_device(const chq_publish_rec* d_recs, const uint64_t* d_n_recs, uint64_t recs_cap,
                                uint32_t* d_key_off, uint32_t* d_key_len, uint32_t* d_rec_index, uint32_t* d_n_pub,
                                void* stream);
/* chq_route_device for keys scattered inside a byte batch: key i = d_bytes[d_key_off[i] .. + d_key_len[i]) */
int32_t chq_route_keys_device(chq_table* t, const uint8_t* d_bytes, const uint32_t* d_key_off, const uint32_t* d_key_len,
                              uint32_t n, uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap,
                              uint8_t* d_any_durable, uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream);

/* ---- the step after the match: per-queue message lists and per-message flags ----
 * Replaces the tail of ExchangeEntity.publish (chana-mq-server/.../entity/ExchangeEntity.scala:300-307): the
 * reference appends Msg(msgId, ...) to queueToMsgs(q) for every matched queue, message by message, so each
 * queue's vector is in publish order.  Here the CSR (d_offsets, d_queue_ids) of a routed batch of n messages is
 * transposed on the device into d_q_offsets[n_queues + 1] / d_q_msgs[total]: queue q receives the messages
 * d_q_msgs[d_q_offsets[q] .. d_q_offsets[q + 1]) — indices into the batch, ascending = publish order.
 * n_queues = chq_n_queues of the table that routed the batch.  Asynchronous on `stream`; *d_status = CHQ_OK,
 * CHQ_E_OVERFLOW (more ids than q_cap; nothing is written past it) or CHQ_E_ARG (an id >= n_queues). */
int32_t chq_fanout_device(const uint32_t* d_offsets, const uint32_t* d_queue_ids, uint32_t n, uint32_t n_queues,
                          uint32_t* d_q_offsets, uint32_t* d_q_msgs, uint64_t q_cap, int32_t* d_status, void* stream);
/* Per-message results of the publish loop: d_refer_count[i] = number of queues (MessageEntity.Refer count,
 * ExchangeEntity.scala:303) and d_out_flags[i] = CHQ_FAN_* bits:
 *   PERSISTENT       exchange durable && message persistent && any matched subscriber durable (ExchangeEntity.scala:302)
 *   NON_ROUTED       mandatory && no queue                                                      (ExchangeEntity.scala:319-323)
 *   NON_DELIVERABLE  immediate && no matched queue has a consumer                               (ExchangeEntity.scala:324-330)
 * d_pub_flags[i] = CHQ_PUB_MANDATORY | CHQ_PUB_IMMEDIATE | CHQ_PUB_PERSIST (the low bits of chq_publish_rec.flags;
 * NULL = none set); d_queue_has_consumer = one byte per queue id (NULL = no queue has a consumer: the broker keeps
 * that knowledge in its QueueEntity actors and uploads it when it changes). */
enum { CHQ_FAN_PERSISTENT = 1, CHQ_FAN_NON_ROUTED = 2, CHQ_FAN_NON_DELIVERABLE = 4 };
int32_t chq_publish_flags_device(const uint32_t* d_offsets, const uint32_t* d_queue_ids, const uint8_t* d_any_durable,
                                 const uint8_t* d_pub_flags, uint32_t n, int32_t exchange_durable,
                                 const uint8_t* d_queue_has_consumer, uint32_t n_queues, uint32_t* d_refer_count,
                                 uint8_t* d_out_flags, void* stream);
const char* chq_fanout_last_error(void);

/* ---- delivery-side encode: Basic.Deliver + content header + body frames ----
 * Replaces AMQCommand.render for deliveries (chana-mq-base/.../model/AMQCommand.scala:30-65; the method
 * arguments of .../method/Basic.scala:189-205, AMQContentHeader.toFrame .../model/AMQContentHeader.scala:50-57,
 * Frame.writeTo .../model/Frame.scala:220-226).  Delivery d sends the Basic.Publish that record `rec` of a decoded
 * batch describes (chq_decode_device left d_bytes / d_recs on the device) to one consumer: METHOD frame
 * (consumer tag, delivery tag, redelivered, exchange, routing key), HEADER frame (class 60, weight 0, bodySize, the
 * publish's property bytes as they arrived) and the body re-cut into frames of at most frame_max - 8 payload bytes
 * (frame_max 0: one frame).  Output d of the batch is d_out[d_out_offsets[d] .. d_out_offsets[d + 1]), in delivery
 * order, ready to be written to the consumers' sockets.  Asynchronous on `stream`; *d_status = CHQ_OK,
 * CHQ_E_OVERFLOW (out_cap too small: whole deliveries that do not fit are left out, d_out_offsets is complete) or
 * CHQ_E_ARG (a delivery names a record that is not a decoded Basic.Publish; it gets zero bytes).
 * The reference re-encodes the properties from its BasicProperties object; the copy made here is identical whenever
 * that re-encoding is the identity (no repeated names and at most four entries per field table — from five entries
 * on the reference emits its Map's hash order, which nothing pins). */
typedef struct {
  uint64_t delivery_tag;
  uint32_t rec;          /* index into d_recs */
  uint32_t ctag_off;     /* consumer tag = d_ctags[ctag_off .. ctag_off + ctag_len) */
  uint16_t channel;
  uint8_t ctag_len;
  uint8_t redelivered;
  uint32_t pad;
} chq_delivery;          /* 24 bytes */
int32_t chq_deliver_encode_device(const uint8_t* d_bytes, const chq_publish_rec* d_recs, uint64_t n_recs, const uint8_t* d_ctags,
                                  const chq_delivery* d_deliveries, uint32_t n_del, uint32_t frame_max, uint8_t* d_out,
                                  uint64_t out_cap, uint64_t* d_out_offsets, int32_t* d_status, void* stream);
const char* chq_encode_last_error(void);

/* No counterpart in the reference.  Puts the CALLING thread on the NUMA node of `device`: CPU affinity = that node's
 * CPUs (within what the thread may use), memory policy = prefer that node — for everything the thread allocates from
 * then on, the library's pinned staging and result buffers included.  Call it once per routing thread / process before
 * chq_table_create when a box holds several GPUs: result buffers that all sit on one socket make the other socket's
 * GPUs copy across the socket link.  *node = the node (-1: unknown or a single-node machine, nothing changed),
 * *n_cpus = CPUs in the new affinity mask.  Linux only. */
int32_t chq_numa_bind_to_device(int32_t device, int32_t* node, int32_t* n_cpus);

/* pinned host memory for callers that want zero-copy staging (JNA: new Pointer(addr)) */
int32_t chq_host_alloc(void** p, uint64_t bytes);
int32_t chq_host_free(void* p);

/* counters as a JSON object: bindings, queues, groups, shapes, word-id stride, batches, messages,
 * ids, device bytes of the table */
int32_t chq_stats(const chq_table* t, char* json, uint32_t cap);
/* Per-kernel device time, measured with CUDA events recorded on the launching stream around k_tokenize
 * and the route kernels (k_probe + k_offsets + k_write + k_write_rest).  The call synchronises the device, returns the
 * sums (ms) and the number of batches since the last call (at most 256 are kept), then switches recording
 * on or off. */
int32_t chq_profile(chq_table* t, int32_t enable, double* tokenize_ms, double* route_ms, uint32_t* launches);
/* The same, one sum per kernel: kernel_ms[0..4] = k_tokenize (or k_headers_match), k_probe, k_offsets, k_write,
 * k_write_rest. */
int32_t chq_profile_kernels(chq_table* t, int32_t enable, double* kernel_ms, uint32_t* launches);
/* how many of this library's kernels have been launched by this process (all tables) */
uint64_t chq_kernel_launches(void);
/* measurement hook: cycle totals of k_write_rest by kind of message (32 counters: cycles, messages, ids per
 * bin, see kernels.cu); all zero unless the library was built with -DCHQ_CLOCKS.  Reading resets them. */
int32_t chq_debug_clocks(uint64_t* out32);
/* test hook: structural invariants of the compiled table on the host (lists ascending and inside the table, ids in
 * range, every group that shares a queue with another one flagged for de-duplication); CHQ_E_ARG + text on a breach */
int32_t chq_debug_check(chq_table* t);
/* test hook: lower the per-message limits that select the slower merge paths (0 = default) */
int32_t chq_debug_limits(chq_table* t, uint32_t max_slots, uint32_t max_small);

#ifdef __cplusplus
}
#endif
#endif
