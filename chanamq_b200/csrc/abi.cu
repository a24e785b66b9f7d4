// C ABI of libchanagpu.so (declarations and reference citations: include/chanamq_gpu.h).
// Host logic only: table life cycle, staging, stream plumbing.  All matching happens in
// kernels.cu; there is no CPU fallback — a CUDA failure is returned as CHQ_E_CUDA.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include "../../include/chanamq_gpu.h"
#include "host_table.hpp"
#include "kernels.h"

using namespace chq;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CU(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e__ = (call);                                                                     \
    if (e__ != cudaSuccess)                                                                       \
      return fail(CHQ_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));               \
  } while (0)

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;  // elements
  int ensure(size_t need) {
    if (need <= cap) return CHQ_OK;
    if (p) CU(cudaFree(p));
    p = nullptr;
    cap = 0;
    size_t want = need + need / 4 + 64;
    CU(cudaMalloc(&p, want * sizeof(T)));
    cap = want;
    return CHQ_OK;
  }
  int upload(const T* src, size_t n, cudaStream_t st) {
    int rc = ensure(n ? n : 1);
    if (rc) return rc;
    if (n) CU(cudaMemcpyAsync(p, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
    return CHQ_OK;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};
template <typename T>
struct PinBuf {
  T* p = nullptr;
  size_t cap = 0;
  int ensure(size_t need) {
    if (need <= cap) return CHQ_OK;
    if (p) CU(cudaFreeHost(p));
    p = nullptr;
    cap = 0;
    size_t want = need + need / 4 + 64;
    CU(cudaMallocHost(&p, want * sizeof(T)));
    cap = want;
    return CHQ_OK;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
};

// scratch a route launch needs beyond its inputs and outputs
struct Workspace {
  DevBuf<uint32_t> wid;
  DevBuf<uint8_t> nwords;
  DevBuf<uint64_t> state;   // [0..1] = control words, [2..] = per-tile scan state
  DevBuf<uint32_t> scratch;
  DevBuf<uint32_t> cnt, meta;   // per message: id count, list count | class | flags (k_probe -> k_offsets, k_write)
  DevBuf<uint2> mlist;          // per message: its lists (offset, length), kRouteSlots each
  DevBuf<uint32_t> rest;        // messages left to k_write_rest
  DevBuf<uint32_t> desc_k;      // headers exchange: matched lists per message (k_headers_match -> k_route)
  DevBuf<uint32_t> desc_start;
  DevBuf<uint8_t> desc_flags;
  DevBuf<uint2> desc_lists;
  size_t scratch_want = 16u << 20;
  size_t desc_want = 1u << 20;  // descriptor arena (grown after CHQ_E_SCRATCH)
  // optional per-kernel timing (chq_profile): events on the launching stream around each kernel
  static constexpr int kProfRing = 256;
  cudaEvent_t ev[kProfRing][6] = {};  // start, after tokenize, k_probe, k_offsets, k_write, k_write_rest
  int prof_n = 0;
  void release() {
    wid.release(); nwords.release(); state.release(); scratch.release(); cnt.release(); meta.release(); mlist.release(); rest.release();
    desc_k.release(); desc_start.release(); desc_flags.release(); desc_lists.release();
    for (auto& e : ev) for (auto& x : e) if (x) { cudaEventDestroy(x); x = nullptr; }
    prof_n = 0;
  }
};

struct RouteSlot {
  cudaStream_t st = nullptr;
  cudaEvent_t meta_done = nullptr;
  Workspace ws;
  PinBuf<uint8_t> h_blob;
  PinBuf<uint32_t> h_off;
  DevBuf<uint8_t> d_blob;
  DevBuf<uint32_t> d_off;
  DevBuf<uint32_t> d_offsets, d_ids;
  DevBuf<uint32_t> d_packed;  // CHQ_IDS_U24: the ids as 3-byte values (pack.cu), what the host gets instead of d_ids
  int32_t ids_format = CHQ_IDS_U32;  // of the batch in flight
  DevBuf<uint8_t> d_durable, d_err;
  DevBuf<uint64_t> d_total;   // [0] total, [1] status (as int32 in low word)
  PinBuf<uint32_t> r_offsets, r_ids;
  PinBuf<uint8_t> r_durable, r_err;
  PinBuf<uint64_t> r_total;
  uint32_t n = 0;
  uint64_t last_total = 0;   // ids of the previous batch on this slot: sizes d_ids for the next one
  bool busy = false;
  void release() {
    ws.release(); h_blob.release(); h_off.release(); d_blob.release(); d_off.release(); d_offsets.release();
    d_ids.release(); d_packed.release(); d_durable.release(); d_err.release(); d_total.release(); r_offsets.release();
    r_ids.release(); r_durable.release(); r_err.release(); r_total.release();
    if (meta_done) cudaEventDestroy(meta_done);
    if (st) cudaStreamDestroy(st);
  }
};

}  // namespace

struct chq_table {
  HostTable host;
  CompiledTable ct;
  // written under mu by bind / commit (possibly on a Future's thread), read by the routing thread
  std::atomic<bool> committed{false};   // device copy is current: routing allowed
  std::atomic<bool> compiled{false};    // ct is current: queue names / stats valid
  bool host_only = false;   // device == -1: compile-only table for host-side checks, cannot route
  bool uploaded = false;    // a full compile has been uploaded: later commits may patch it in place
  uint64_t n_full = 0, n_incremental = 0;
  int device = 0, n_sms = 148;
  // device copy of the compiled table
  DevBuf<DictSlot> d_dict;
  DevBuf<uint8_t> d_dict_bytes;
  DevBuf<WordSlot> d_wdict;
  DevBuf<uint32_t> d_wdict_off;
  DevBuf<Shape> d_shapes;
  DevBuf<uint16_t> d_exact;
  DevBuf<uint16_t> d_tags;
  DevBuf<uint32_t> d_slot_group;
  DevBuf<Group> d_groups;
  DevBuf<uint32_t> d_lits, d_rems, d_qlist;
  DevBuf<DictSlot> d_vdict;
  DevBuf<uint8_t> d_vdict_bytes;
  DevBuf<uint32_t> d_presence, d_all_atoms;
  DevBuf<AtomRec> d_atoms;
  DevBuf<AllGroup> d_all_groups;
  TableView tv{};
  RouteSlot slots[CHQ_SLOTS];
  Workspace dev_ws;  // for chq_route_device
  uint32_t max_slots = kRouteSlots, max_small = kRouteSmall;
  std::atomic<int32_t> ids_format{CHQ_IDS_U32};  // host results of chq_route / chq_route_wait (chq_set_ids_format)
  bool profile = false;
  uint64_t n_batches = 0, n_messages = 0, n_ids = 0;
  std::mutex mu;  // bind may arrive from a Future callback off the actor thread (ExchangeEntity.scala:337-338)
  chq_table(int type, int mode) : host(type, mode) {}
};

namespace {

int set_device(chq_table* t) {
  CU(cudaSetDevice(t->device));
  return CHQ_OK;
}

int upload_table(chq_table* t) {
  CompiledTable& c = t->ct;
  cudaStream_t st = nullptr;
  int rc;
  if ((rc = t->d_dict.upload(c.dict.data(), c.dict.size(), st))) return rc;
  if ((rc = t->d_dict_bytes.upload(c.dict_bytes.data(), c.dict_bytes.size(), st))) return rc;
  if ((rc = t->d_wdict.upload(c.wdict.data(), c.wdict.size(), st))) return rc;
  if ((rc = t->d_wdict_off.upload(c.wdict_off.data(), c.wdict_off.size(), st))) return rc;
  if ((rc = t->d_shapes.upload(c.shapes.data(), c.shapes.size(), st))) return rc;
  if ((rc = t->d_exact.upload(c.exact_rng.data(), c.exact_rng.size(), st))) return rc;
  if ((rc = t->d_tags.upload(c.tags.data(), c.tags.size(), st))) return rc;
  if ((rc = t->d_slot_group.upload(c.slot_group.data(), c.slot_group.size(), st))) return rc;
  if ((rc = t->d_groups.upload(c.groups.data(), c.groups.size(), st))) return rc;
  if ((rc = t->d_lits.upload(c.lits.data(), c.lits.size(), st))) return rc;
  if ((rc = t->d_rems.upload(c.rems.data(), c.rems.size(), st))) return rc;
  if ((rc = t->d_qlist.upload(c.qlist.data(), c.qlist.size(), st))) return rc;
  if (c.headers_mode) {
    if ((rc = t->d_vdict.upload(c.vdict.data(), c.vdict.size(), st))) return rc;
    if ((rc = t->d_vdict_bytes.upload(c.vdict_bytes.data(), c.vdict_bytes.size(), st))) return rc;
    if ((rc = t->d_presence.upload(c.presence_atom.data(), c.presence_atom.size(), st))) return rc;
    if ((rc = t->d_all_atoms.upload(c.all_atoms.data(), c.all_atoms.size(), st))) return rc;
    if ((rc = t->d_atoms.upload(c.atoms.data(), c.atoms.size(), st))) return rc;
    if ((rc = t->d_all_groups.upload(c.all_groups.data(), c.all_groups.size(), st))) return rc;
  }
  CU(cudaStreamSynchronize(st));
  TableView& v = t->tv;
  v = TableView{};
  if (c.headers_mode) {
    v.vdict = t->d_vdict.p;
    v.vdict_bytes = t->d_vdict_bytes.p;
    v.vdict_mask = (uint32_t)c.vdict.size() - 1;
    v.presence_atom = t->d_presence.p;
    v.atoms = t->d_atoms.p;
    v.all_groups = t->d_all_groups.p;
    v.all_atoms = t->d_all_atoms.p;
    v.always_off = c.always_off;
    v.always_len = c.always_len;
    v.always_flags = c.always_flags;
  }
  v.dict = t->d_dict.p;
  v.dict_bytes = t->d_dict_bytes.p;
  v.dict_mask = (uint32_t)c.dict.size() - 1;
  v.wdict = t->d_wdict.p;
  v.wdict_off = t->d_wdict_off.p;
  v.wdict_mask = (uint32_t)c.wdict.size() - 1;
  v.shapes = t->d_shapes.p;
  v.exact_rng = t->d_exact.p;
  v.hashy_begin = c.hashy_begin;
  v.hashy_end = c.hashy_end;
  v.tags = t->d_tags.p;
  v.slot_group = t->d_slot_group.p;
  v.bucket_mask = (uint32_t)(c.tags.size() / kBucketSlots) - 1;
  v.groups = t->d_groups.p;
  v.lits = t->d_lits.p;
  v.rems = t->d_rems.p;
  v.qlist = t->d_qlist.p;
  v.wstride = c.wstride;
  v.n_groups = (uint32_t)c.groups.size();
  v.exchange_type = (uint32_t)t->host.type();
  v.fanout_group = c.fanout_group;
  return CHQ_OK;
}

// Enqueue tokenise + route for n keys already on the device.  Everything is asynchronous on st.
int enqueue_route(chq_table* t, Workspace& ws, const uint8_t* d_blob, const uint32_t* d_off, const uint32_t* d_klen, uint32_t n,
                  uint32_t* d_offsets, uint32_t* d_ids, uint64_t ids_cap, uint8_t* d_durable, uint8_t* d_err,
                  uint64_t* d_total, int32_t* d_status, cudaStream_t st) {
  NvtxRange nvtx_range("chq: tokenize + route launches");
  const uint32_t n_tiles = route_scan_tiles(n);
  const bool headers = t->ct.headers_mode;  // d_blob/d_off/d_klen then describe the messages' headers tables
  int rc;
  if (headers) {
    if ((rc = ws.desc_k.ensure((size_t)n + 1))) return rc;
    if ((rc = ws.desc_start.ensure((size_t)n + 1))) return rc;
    if ((rc = ws.desc_flags.ensure((size_t)n + 1))) return rc;
    if ((rc = ws.desc_lists.ensure(std::max<size_t>((size_t)n * kHeaderListsPerMessage, ws.desc_want)))) return rc;
  } else {
    if ((rc = ws.wid.ensure((size_t)n * t->tv.wstride + 4))) return rc;
    if ((rc = ws.nwords.ensure((size_t)n + 1))) return rc;
  }
  if ((rc = ws.state.ensure((size_t)n_tiles + kStateHead))) return rc;
  if ((rc = ws.scratch.ensure(ws.scratch_want))) return rc;
  if ((rc = ws.cnt.ensure((size_t)n + 1))) return rc;
  if ((rc = ws.meta.ensure((size_t)n + 1))) return rc;
  if ((rc = ws.mlist.ensure((size_t)n * kRouteSlots + 1))) return rc;
  if ((rc = ws.rest.ensure((size_t)n + 1))) return rc;
  CU(cudaMemsetAsync(ws.state.p, 0, ((size_t)n_tiles + kStateHead) * sizeof(uint64_t), st));
  if (n == 0) {
    CU(cudaMemsetAsync(d_offsets, 0, sizeof(uint32_t), st));
    CU(cudaMemsetAsync(d_total, 0, sizeof(uint64_t), st));
    CU(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    return CHQ_OK;
  }
  const int split = (t->host.type() == CHQ_TOPIC || t->host.type() == CHQ_HEADERS) ? 1 : 0;
  cudaEvent_t* pe = nullptr;
  if (t->profile && ws.prof_n < Workspace::kProfRing) {
    pe = ws.ev[ws.prof_n++];
    for (int i = 0; i < 6; ++i)
      if (!pe[i]) CU(cudaEventCreate(&pe[i]));
    CU(cudaEventRecord(pe[0], st));
  }
  if (headers)
    CU(launch_headers_match(d_blob, d_off, d_klen, n, t->tv, ws.desc_k.p, ws.desc_start.p, ws.desc_flags.p, ws.desc_lists.p,
                            (uint32_t)std::min<size_t>(ws.desc_lists.cap, 0xFFFFFFFFu),
                            reinterpret_cast<uint32_t*>(ws.state.p) + kCtrlDescBump, reinterpret_cast<uint32_t*>(ws.state.p) + kCtrlStatus,
                            d_err, st));
  else if (d_klen) CU(launch_tokenize_keys(d_blob, d_off, d_klen, n, t->tv, split, ws.wid.p, ws.nwords.p, d_err, st));
  else CU(launch_tokenize(d_blob, d_off, n, t->tv, split, ws.wid.p, ws.nwords.p, d_err, st));
  if (pe) CU(cudaEventRecord(pe[1], st));
  RouteParams p{};
  p.tv = t->tv;
  p.wid = ws.wid.p;
  p.nwords = ws.nwords.p;
  p.n = n;
  p.n_tiles = n_tiles;
  p.offsets = d_offsets;
  p.ids = d_ids;
  p.ids_cap = ids_cap;
  p.durable = d_durable;
  p.total = d_total;
  p.ctrl = reinterpret_cast<uint32_t*>(ws.state.p);
  p.tile_state = ws.state.p + kStateHead;
  p.cnt = ws.cnt.p;
  p.meta = ws.meta.p;
  p.mlist = ws.mlist.p;
  p.rest = ws.rest.p;
  p.scratch = ws.scratch.p;
  p.scratch_cap = ws.scratch.cap;
  p.max_slots = t->max_slots;
  p.max_small = t->max_small;
  if (headers) { p.desc_k = ws.desc_k.p; p.desc_start = ws.desc_start.p; p.desc_flags = ws.desc_flags.p; p.desc_lists = ws.desc_lists.p; }
#ifdef CHQ_DEBUG_SWITCHES  // experiment builds only: 1 = k_write writes nothing, 2 = k_probe probes nothing
  { const char* dbg = getenv("CHQ_DEBUG"); p.debug = dbg ? (uint32_t)atoi(dbg) : 0u; }
#endif
  CU(launch_route(p, t->n_sms, st, pe ? pe + 2 : nullptr));
  if (pe) CU(cudaEventRecord(pe[5], st));
  CU(launch_finish(p.ctrl, d_status, st));
  t->n_batches++;
  t->n_messages += n;
  return CHQ_OK;
}

bool is_pinned(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

int slot_init(RouteSlot& s) {
  if (!s.st) CU(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
  if (!s.meta_done) CU(cudaEventCreateWithFlags(&s.meta_done, cudaEventDisableTiming));
  return CHQ_OK;
}

int slot_launch(chq_table* t, RouteSlot& s) {
  int rc = enqueue_route(t, s.ws, s.d_blob.p, s.d_off.p, nullptr, s.n, s.d_offsets.p, s.d_ids.p, s.d_ids.cap, s.d_durable.p,
                         s.d_err.p, s.d_total.p, reinterpret_cast<int32_t*>(s.d_total.p + 1), s.st);
  if (rc) return rc;
  if (s.ids_format == CHQ_IDS_U24) {
    if ((rc = s.d_packed.ensure(3 * (s.d_ids.cap / 4 + 1)))) return rc;
    CU(launch_pack24(s.d_ids.p, s.d_total.p, s.d_ids.cap, s.d_packed.p, t->n_sms, s.st));
  }
  CU(cudaMemcpyAsync(s.r_total.p, s.d_total.p, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.st));
  CU(cudaMemcpyAsync(s.r_offsets.p, s.d_offsets.p, ((size_t)s.n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.st));
  if (s.n) {
    CU(cudaMemcpyAsync(s.r_durable.p, s.d_durable.p, s.n, cudaMemcpyDeviceToHost, s.st));
    CU(cudaMemcpyAsync(s.r_err.p, s.d_err.p, s.n, cudaMemcpyDeviceToHost, s.st));
  }
  CU(cudaEventRecord(s.meta_done, s.st));
  return CHQ_OK;
}

}  // namespace

extern "C" {

const char* chq_last_error(void) { return g_err.c_str(); }

int32_t chq_table_create(chq_table** out, int32_t exchange_type, int32_t mode, int32_t device) {
  if (!out || exchange_type < CHQ_DIRECT || exchange_type > CHQ_HEADERS || (mode != CHQ_MODE_COMPAT && mode != CHQ_MODE_SPEC))
    return fail(CHQ_E_ARG, "chq_table_create: bad argument");
  if (device < -1) return fail(CHQ_E_ARG, "chq_table_create: bad device");
  if (device >= 0) {
    int count = 0;
    CU(cudaGetDeviceCount(&count));  // fails loudly when there is no GPU: no CPU path exists
    if (device >= count) return fail(CHQ_E_ARG, "chq_table_create: no such CUDA device");
  }
  chq_table* t = new (std::nothrow) chq_table(exchange_type, mode);
  if (!t) return fail(CHQ_E_ARG, "out of host memory");
  t->device = device;
  t->host_only = device < 0;  // compiles and reports, but every route call fails: there is no CPU matcher
  if (const char* s = getenv("CHQ_SCRATCH_IDS")) {  // test hook: a small slow-path arena, to exercise the grow-and-retry path
    const size_t want = (size_t)strtoull(s, nullptr, 10);
    if (want >= 64) {
      for (auto& sl : t->slots) sl.ws.scratch_want = want;
      t->dev_ws.scratch_want = want;
    }
  }
  if (!t->host_only) {
    cudaDeviceProp prop{};
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
      delete t;
      return fail(CHQ_E_CUDA, "cudaGetDeviceProperties failed");
    }
    t->n_sms = prop.multiProcessorCount;
  }
  *out = t;
  return CHQ_OK;
}

int32_t chq_table_destroy(chq_table* t) {
  if (!t) return CHQ_OK;
  if (t->host_only) { delete t; return CHQ_OK; }
  cudaSetDevice(t->device);
  cudaDeviceSynchronize();
  for (auto& s : t->slots) s.release();
  t->dev_ws.release();
  t->d_dict.release(); t->d_dict_bytes.release(); t->d_wdict.release(); t->d_wdict_off.release(); t->d_shapes.release(); t->d_exact.release();
  t->d_vdict.release(); t->d_vdict_bytes.release(); t->d_presence.release(); t->d_all_atoms.release();
  t->d_atoms.release(); t->d_all_groups.release();
  t->d_tags.release(); t->d_slot_group.release(); t->d_groups.release(); t->d_lits.release(); t->d_rems.release(); t->d_qlist.release();
  delete t;
  return CHQ_OK;
}

int32_t chq_bind(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                 uint8_t durable) {
  if (!t || (!key && key_len) || (!queue && queue_len)) return fail(CHQ_E_ARG, "chq_bind: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  return t->host.bind(std::string_view((const char*)key, key_len), std::string_view((const char*)queue, queue_len), durable != 0);
}

int32_t chq_bind_args(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                      uint8_t durable, const uint8_t* args_table, uint32_t args_len) {
  if (!t || (!key && key_len) || (!queue && queue_len) || (!args_table && args_len)) return fail(CHQ_E_ARG, "chq_bind_args: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  int rc = t->host.bind(std::string_view((const char*)key, key_len), std::string_view((const char*)queue, queue_len), durable != 0,
                        std::string_view((const char*)args_table, args_len));
  if (rc == CHQ_E_ARG) return fail(rc, "chq_bind_args: malformed arguments table");
  return rc;
}

int32_t chq_bind_many(chq_table* t, const uint8_t* keys_blob, const uint32_t* key_offsets, const uint8_t* queues_blob,
                      const uint32_t* queue_offsets, const uint8_t* durable, uint32_t n) {
  if (!t || !key_offsets || !queue_offsets) return fail(CHQ_E_ARG, "chq_bind_many: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  int first_err = CHQ_OK;
  for (uint32_t i = 0; i < n; ++i) {
    int rc = t->host.bind(std::string_view((const char*)keys_blob + key_offsets[i], key_offsets[i + 1] - key_offsets[i]),
                          std::string_view((const char*)queues_blob + queue_offsets[i], queue_offsets[i + 1] - queue_offsets[i]),
                          durable ? durable[i] != 0 : false);
    if (rc && !first_err) first_err = rc;
  }
  return first_err;
}

int32_t chq_unbind(chq_table* t, const uint8_t* key, uint32_t key_len, const uint8_t* queue, uint32_t queue_len,
                   uint8_t durable) {
  if (!t) return fail(CHQ_E_ARG, "chq_unbind: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  return t->host.unbind(std::string_view((const char*)key, key_len), std::string_view((const char*)queue, queue_len), durable != 0);
}

int32_t chq_unbind_queue(chq_table* t, const uint8_t* queue, uint32_t queue_len) {
  if (!t) return fail(CHQ_E_ARG, "chq_unbind_queue: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  return t->host.unbind_queue(std::string_view((const char*)queue, queue_len));
}

int32_t chq_declare_queues(chq_table* t, const uint8_t* queues_blob, const uint32_t* queue_offsets, uint32_t n) {
  if (!t || (n && (!queues_blob || !queue_offsets))) return fail(CHQ_E_ARG, "chq_declare_queues: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  t->committed = t->compiled = false;
  for (uint32_t i = 0; i < n; ++i)
    t->host.declare_queue(std::string_view((const char*)queues_blob + queue_offsets[i], queue_offsets[i + 1] - queue_offsets[i]));
  return CHQ_OK;
}

namespace {

// The changes an incremental commit made on the host, copied over the device copy (a handful of small copies per
// bind).  false = an array outgrew its device buffer: the caller uploads the whole table again.
int patch_table(chq_table* t, const std::vector<Patch>& patches, bool* fits) {
  CompiledTable& c = t->ct;
  *fits = c.wdict.size() <= t->d_wdict.cap && c.wdict_off.size() <= t->d_wdict_off.cap && c.dict_bytes.size() <= t->d_dict_bytes.cap &&
          c.tags.size() <= t->d_tags.cap && c.slot_group.size() <= t->d_slot_group.cap && c.groups.size() <= t->d_groups.cap &&
          c.lits.size() <= t->d_lits.cap && c.rems.size() <= t->d_rems.cap && c.qlist.size() <= t->d_qlist.cap;
  if (!*fits) return CHQ_OK;
  for (const Patch& p : patches) {
    if (p.n == 0) continue;
    switch (p.array) {
      case kPaWdict: CU(cudaMemcpy(t->d_wdict.p + p.off, c.wdict.data() + p.off, p.n * sizeof(WordSlot), cudaMemcpyHostToDevice)); break;
      case kPaWdictOff: CU(cudaMemcpy(t->d_wdict_off.p + p.off, c.wdict_off.data() + p.off, p.n * 4, cudaMemcpyHostToDevice)); break;
      case kPaDictBytes: CU(cudaMemcpy(t->d_dict_bytes.p + p.off, c.dict_bytes.data() + p.off, p.n, cudaMemcpyHostToDevice)); break;
      case kPaTags: CU(cudaMemcpy(t->d_tags.p + p.off, c.tags.data() + p.off, p.n * 2, cudaMemcpyHostToDevice)); break;
      case kPaSlotGroup: CU(cudaMemcpy(t->d_slot_group.p + p.off, c.slot_group.data() + p.off, p.n * 4, cudaMemcpyHostToDevice)); break;
      case kPaGroups: CU(cudaMemcpy(t->d_groups.p + p.off, c.groups.data() + p.off, p.n * sizeof(Group), cudaMemcpyHostToDevice)); break;
      case kPaLits: CU(cudaMemcpy(t->d_lits.p + p.off, c.lits.data() + p.off, p.n * 4, cudaMemcpyHostToDevice)); break;
      case kPaRems: CU(cudaMemcpy(t->d_rems.p + p.off, c.rems.data() + p.off, p.n * 4, cudaMemcpyHostToDevice)); break;
      case kPaQlist: CU(cudaMemcpy(t->d_qlist.p + p.off, c.qlist.data() + p.off, p.n * 4, cudaMemcpyHostToDevice)); break;
      default: break;
    }
  }
  // (a copy from pageable memory returns once the bytes are staged, not once they have landed: the batches that
  // follow run on non-blocking streams and must not overtake them)
  CU(cudaStreamSynchronize(nullptr));
  t->tv.n_groups = (uint32_t)c.groups.size();
  return CHQ_OK;
}

int commit_impl(chq_table* t, bool allow_incremental) {
  NvtxRange nvtx_range("chq_commit");
  std::lock_guard<std::mutex> g(t->mu);
  if (t->host_only) {  // no device copy: the same decision, host side only
    std::vector<Patch> patches;
    if (allow_incremental && t->uploaded && t->host.incremental_ready() && t->host.apply_pending(t->ct, patches)) {
      t->n_incremental++;
    } else {
      t->host.compile(t->ct);
      t->uploaded = true;
      t->n_full++;
    }
    t->compiled = true;
    return CHQ_OK;
  }
  int rc = set_device(t);
  if (rc) return rc;
  // In place when the table is on the device already and the changes allow it (host_inc.cpp): the reference's bind is
  // O(depth) and visible at once (QueueMatcher.scala:385-392); here it is a few small copies behind the batches in
  // flight.  Anything else: compile the whole binding set again.
  if (allow_incremental && t->uploaded && t->host.incremental_ready()) {
    std::vector<Patch> patches;
    if (t->host.apply_pending(t->ct, patches)) {
      t->compiled = true;
      if (!patches.empty()) {
        CU(cudaDeviceSynchronize());  // nothing in flight may read a half-patched table
        bool fits = false;
        if ((rc = patch_table(t, patches, &fits))) return rc;
        if (!fits && (rc = upload_table(t))) return rc;
      }
      t->n_incremental++;
      t->committed = true;
      return CHQ_OK;
    }
  }
  CU(cudaDeviceSynchronize());  // nothing in flight may still read the old table
  t->host.compile(t->ct);
  t->compiled = true;
  if ((rc = upload_table(t))) return rc;
  t->uploaded = true;
  t->n_full++;
  t->committed = true;
  return CHQ_OK;
}

}  // namespace

int32_t chq_commit(chq_table* t) {
  if (!t) return fail(CHQ_E_ARG, "chq_commit: bad argument");
  return commit_impl(t, true);
}

int32_t chq_commit_full(chq_table* t) {
  if (!t) return fail(CHQ_E_ARG, "chq_commit_full: bad argument");
  return commit_impl(t, false);
}

int32_t chq_n_queues(const chq_table* t, uint32_t* n) {
  if (!t || !n) return fail(CHQ_E_ARG, "chq_n_queues: bad argument");
  if (!t->compiled) return fail(CHQ_E_NOT_COMMITTED, "table not committed");
  *n = (uint32_t)t->ct.queue_names.size();
  return CHQ_OK;
}

int32_t chq_queue_name(const chq_table* t, uint32_t queue_id, const uint8_t** name, uint32_t* len) {
  if (!t || !name || !len) return fail(CHQ_E_ARG, "chq_queue_name: bad argument");
  if (!t->compiled) return fail(CHQ_E_NOT_COMMITTED, "table not committed");
  if (queue_id >= t->ct.queue_names.size()) return fail(CHQ_E_ARG, "chq_queue_name: id out of range");
  *name = (const uint8_t*)t->ct.queue_names[queue_id].data();
  *len = (uint32_t)t->ct.queue_names[queue_id].size();
  return CHQ_OK;
}

int32_t chq_route_submit(chq_table* t, int32_t slot, const uint8_t* keys_blob, const uint32_t* key_offsets, uint32_t n) {
  if (!t || slot < 0 || slot >= CHQ_SLOTS || (n && (!key_offsets))) return fail(CHQ_E_ARG, "chq_route_submit: bad argument");
  if (t->host_only) return fail(CHQ_E_CUDA, "chq_route: host-only table (device -1) cannot route; there is no CPU matcher");
  if (!t->committed) return fail(CHQ_E_NOT_COMMITTED, "chq_route: table has uncommitted binds (call chq_commit)");
  int rc = set_device(t);
  if (rc) return rc;
  NvtxRange nvtx_range("chq_route_submit (H2D + launches)");
  RouteSlot& s = t->slots[slot];
  if ((rc = slot_init(s))) return rc;
  if (s.busy) CU(cudaStreamSynchronize(s.st));
  s.n = n;
  s.ids_format = t->ids_format.load();
  if (s.ids_format == CHQ_IDS_U24 && t->ct.queue_names.size() >= (1u << 24))
    return fail(CHQ_E_ARG, "chq_route: CHQ_IDS_U24 needs a table of fewer than 2^24 queues");
  const size_t blob_len = n ? key_offsets[n] : 0;
  if ((rc = s.d_blob.ensure(blob_len + 16))) return rc;
  if ((rc = s.d_off.ensure((size_t)n + 1))) return rc;
  if ((rc = s.d_offsets.ensure((size_t)n + 1))) return rc;
  if ((rc = s.d_durable.ensure((size_t)n + 1))) return rc;
  if ((rc = s.d_err.ensure((size_t)n + 1))) return rc;
  if ((rc = s.d_total.ensure(2))) return rc;
  // the id buffer follows what the exchange produces: the last batch's total with headroom (first batch: a guess;
  // a batch that overflows is run again by chq_route_wait with the exact size)
  if ((rc = s.d_ids.ensure(std::max<size_t>((size_t)n * 16, (size_t)(s.last_total + s.last_total / 4)) + 1024))) return rc;
  if ((rc = s.r_total.ensure(2))) return rc;
  if ((rc = s.r_offsets.ensure((size_t)n + 1))) return rc;
  if ((rc = s.r_durable.ensure((size_t)n + 1))) return rc;
  if ((rc = s.r_err.ensure((size_t)n + 1))) return rc;
  if (n) {
    const uint8_t* src_blob = keys_blob;
    const uint32_t* src_off = key_offsets;
    if (!is_pinned(keys_blob)) {
      if ((rc = s.h_blob.ensure(blob_len + 1))) return rc;
      memcpy(s.h_blob.p, keys_blob, blob_len);
      src_blob = s.h_blob.p;
    }
    if (!is_pinned(key_offsets)) {
      if ((rc = s.h_off.ensure((size_t)n + 1))) return rc;
      memcpy(s.h_off.p, key_offsets, ((size_t)n + 1) * sizeof(uint32_t));
      src_off = s.h_off.p;
    }
    if (blob_len) CU(cudaMemcpyAsync(s.d_blob.p, src_blob, blob_len, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(s.d_off.p, src_off, ((size_t)n + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s.st));
  }
  if ((rc = slot_launch(t, s))) return rc;
  s.busy = true;
  return CHQ_OK;
}

int32_t chq_route_wait(chq_table* t, int32_t slot, chq_result* out) {
  if (!t || slot < 0 || slot >= CHQ_SLOTS || !out) return fail(CHQ_E_ARG, "chq_route_wait: bad argument");
  NvtxRange nvtx_range("chq_route_wait (D2H of the result)");
  RouteSlot& s = t->slots[slot];
  if (!s.busy) return fail(CHQ_E_ARG, "chq_route_wait: slot was not submitted");
  int rc = set_device(t);
  if (rc) return rc;
  for (int attempt = 0;; ++attempt) {
    CU(cudaEventSynchronize(s.meta_done));
    const uint64_t total = s.r_total.p[0];
    const int32_t status = (int32_t)(uint32_t)s.r_total.p[1];
    if (status == CHQ_OK) break;
    if (attempt >= 3 || total > 0xFFFFFFFFull) {
      s.busy = false;
      return fail(status, status == CHQ_E_SCRATCH ? "device scratch arena exhausted" : "result does not fit 2^32-1 ids");
    }
    // grow what was too small and run the batch again (inputs are still on the device)
    CU(cudaStreamSynchronize(s.st));
    if (status == CHQ_E_OVERFLOW) { if ((rc = s.d_ids.ensure(total + 1024))) return rc; }
    else {
      // the bump pointers kept counting past the end, so they say how much the batch really needs
      uint64_t need_scratch = 0, need_desc = 0;
      uint32_t desc_used = 0;
      CU(cudaMemcpy(&need_scratch, s.ws.state.p + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost));
      CU(cudaMemcpy(&desc_used, reinterpret_cast<uint32_t*>(s.ws.state.p) + kCtrlDescBump, sizeof(uint32_t), cudaMemcpyDeviceToHost));
      need_desc = desc_used;
      if (need_scratch > s.ws.scratch.cap) s.ws.scratch_want = (size_t)(need_scratch + need_scratch / 8 + 1024);
      if (need_desc > s.ws.desc_lists.cap) s.ws.desc_want = (size_t)(need_desc + need_desc / 8 + 1024);
    }
    if ((rc = slot_launch(t, s))) return rc;
  }
  const uint64_t total = s.r_total.p[0];
  // (CHQ_IDS_U24: 3 bytes per id, the copy rounded up to whole words — the packed array is padded with zeros)
  const bool u24 = s.ids_format == CHQ_IDS_U24;
  const uint64_t id_words = u24 ? (3 * total + 3) / 4 : total;
  if ((rc = s.r_ids.ensure(id_words + 1))) return rc;
  if (total) CU(cudaMemcpyAsync(s.r_ids.p, u24 ? s.d_packed.p : s.d_ids.p, id_words * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.st));
  CU(cudaStreamSynchronize(s.st));
  s.busy = false;
  s.last_total = total;
  t->n_ids += total;
  out->n = s.n;
  out->offsets = s.r_offsets.p;
  out->queue_ids = s.r_ids.p;
  out->ids_format = s.ids_format;
  out->any_durable = s.r_durable.p;
  out->err = s.r_err.p;
  return CHQ_OK;
}

int32_t chq_set_ids_format(chq_table* t, int32_t format) {
  if (!t || (format != CHQ_IDS_U32 && format != CHQ_IDS_U24)) return fail(CHQ_E_ARG, "chq_set_ids_format: bad argument");
  t->ids_format.store(format);  // batches already submitted keep the format they were submitted with
  return CHQ_OK;
}

void chq_unpack_u24(const uint8_t* packed, uint64_t n_ids, uint32_t* out) {
  for (uint64_t j = 0; j < n_ids; ++j, packed += 3) out[j] = (uint32_t)packed[0] | ((uint32_t)packed[1] << 8) | ((uint32_t)packed[2] << 16);
}

int32_t chq_route(chq_table* t, const uint8_t* keys_blob, const uint32_t* key_offsets, uint32_t n,
                  const uint8_t* headers_blob, const uint32_t* headers_offsets, chq_result* out) {
  if (t && t->host.type() == CHQ_HEADERS && t->host.mode() == CHQ_MODE_SPEC) {
    // x-match routing: the messages' headers tables are the input, the routing keys play no part
    if (n && (!headers_blob || !headers_offsets)) return fail(CHQ_E_ARG, "chq_route: a headers exchange needs headers_blob/headers_offsets");
    keys_blob = headers_blob;
    key_offsets = headers_offsets;
  }
  int rc = chq_route_submit(t, 0, keys_blob, key_offsets, n);
  if (rc) return rc;
  return chq_route_wait(t, 0, out);
}

int32_t chq_route_device(chq_table* t, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                         uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                         uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream) {
  if (!t || !d_offsets || !d_total || !d_status || (n && (!d_key_offsets || !d_any_durable || !d_err)))
    return fail(CHQ_E_ARG, "chq_route_device: bad argument");
  if (t->host_only) return fail(CHQ_E_CUDA, "chq_route_device: host-only table (device -1) cannot route");
  if (!t->committed) return fail(CHQ_E_NOT_COMMITTED, "chq_route_device: table has uncommitted binds");
  int rc = set_device(t);
  if (rc) return rc;
  return enqueue_route(t, t->dev_ws, d_keys_blob, d_key_offsets, nullptr, n, d_offsets, d_queue_ids, ids_cap, d_any_durable,
                       d_err, d_total, d_status, (cudaStream_t)stream);
}

int32_t chq_route_keys_device(chq_table* t, const uint8_t* d_bytes, const uint32_t* d_key_off, const uint32_t* d_key_len,
                              uint32_t n, uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap,
                              uint8_t* d_any_durable, uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream) {
  if (!t || !d_offsets || !d_total || !d_status || (n && (!d_key_off || !d_key_len || !d_any_durable || !d_err)))
    return fail(CHQ_E_ARG, "chq_route_keys_device: bad argument");
  if (t->host_only) return fail(CHQ_E_CUDA, "chq_route_keys_device: host-only table (device -1) cannot route");
  if (!t->committed) return fail(CHQ_E_NOT_COMMITTED, "chq_route_keys_device: table has uncommitted binds");
  int rc = set_device(t);
  if (rc) return rc;
  return enqueue_route(t, t->dev_ws, d_bytes, d_key_off, d_key_len, n, d_offsets, d_queue_ids, ids_cap, d_any_durable,
                       d_err, d_total, d_status, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// chq_route_device_static: the same tokenise + route launches as chq_route_device, for callers
// that must not allocate.  Everything the launches need beyond the caller's inputs and outputs
// (scan state, the interned-word rows, the slow-path merge arena) lives in ONE fixed array that is
// part of the library image on every GPU that loads it.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr uint32_t kStaticMaxKeys = CHQ_STATIC_MAX_KEYS;     // keys per call
constexpr uint32_t kStaticMaxWords = CHQ_STATIC_MAX_WORDS;   // keys x word-id stride of the table
constexpr uint32_t kStaticArenaIds = 1u << 20;               // slow-path merge arena, in queue ids
constexpr size_t kStaticTiles = (kStaticMaxKeys + kRouteScanTile - 1) / kRouteScanTile;
constexpr size_t kStaticOffState = 0;                                                            // u64[head + tiles]
constexpr size_t kStaticOffNwords = ((kStateHead + kStaticTiles) * 8 + 255) & ~size_t(255);      // u8[keys]
constexpr size_t kStaticOffCnt = (kStaticOffNwords + kStaticMaxKeys + 255) & ~size_t(255);       // u32[keys]
constexpr size_t kStaticOffMeta = kStaticOffCnt + (size_t)kStaticMaxKeys * 4;                    // u32[keys]
constexpr size_t kStaticOffRest = kStaticOffMeta + (size_t)kStaticMaxKeys * 4;                   // u32[keys]
constexpr size_t kStaticOffMlist = kStaticOffRest + (size_t)kStaticMaxKeys * 4;                  // uint2[keys][kRouteSlots]
constexpr size_t kStaticOffWid = kStaticOffMlist + (size_t)kStaticMaxKeys * kRouteSlots * 8;     // u32[words]
constexpr size_t kStaticOffArena = kStaticOffWid + (size_t)kStaticMaxWords * 4;                  // u32[arena]
constexpr size_t kStaticBytes = kStaticOffArena + (size_t)kStaticArenaIds * 4;

// About 11 MiB of device memory, held by every GPU that loads this library until the process exits.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// In host code the name g_static_scratch is only a handle for the runtime, not a device address;
// the address a kernel can use comes from cudaGetSymbolAddress, asked once per device.
unsigned char* g_static_ptr[64] = {};

int static_scratch_ptr(int device, unsigned char** out) {
  if (device < 0 || device >= 64) return fail(CHQ_E_ARG, "chq_route_device_static: device ordinal out of range");
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    CU(cudaGetSymbolAddress(&p, g_static_scratch));  // the table's device is current here
    g_static_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_static_ptr[device];
  return CHQ_OK;
}

}  // namespace

int32_t chq_route_device_static(chq_table* t, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                                uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                                uint8_t* d_err, uint64_t* d_total, int32_t* d_status, void* stream) {
  if (!t || !d_offsets || !d_total || !d_status || (n && (!d_key_offsets || !d_any_durable || !d_err)))
    return fail(CHQ_E_ARG, "chq_route_device_static: bad argument");
  if (t->host_only) return fail(CHQ_E_CUDA, "chq_route_device_static: host-only table (device -1) cannot route");
  if (!t->committed) return fail(CHQ_E_NOT_COMMITTED, "chq_route_device_static: table has uncommitted binds");
  if (t->ct.headers_mode)
    return fail(CHQ_E_ARG, "chq_route_device_static: a headers exchange in CHQ_MODE_SPEC matches on the messages' header "
                           "tables and needs a descriptor arena this entry point does not carry; use chq_route_device");
  if (n > kStaticMaxKeys)
    return fail(CHQ_E_ARG, "chq_route_device_static: more than CHQ_STATIC_MAX_KEYS keys in one call (split the batch, or "
                           "use chq_route_device, which allocates)");
  if ((uint64_t)n * t->tv.wstride > kStaticMaxWords)
    return fail(CHQ_E_ARG, "chq_route_device_static: keys x the table's word stride exceeds CHQ_STATIC_MAX_WORDS (tables "
                           "with '#' inside a pattern keep 256 words per key: send fewer keys per call)");
  int rc = set_device(t);
  if (rc) return rc;
  unsigned char* base = nullptr;
  if ((rc = static_scratch_ptr(t->device, &base))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t n_tiles = route_scan_tiles(n);
  uint64_t* state = reinterpret_cast<uint64_t*>(base + kStaticOffState);
  // the array keeps whatever the previous call left: the control words and the per-tile scan state
  // must start at zero; the rows, counts and arena are fully rewritten before they are read
  CU(cudaMemsetAsync(state, 0, ((size_t)n_tiles + kStateHead) * sizeof(uint64_t), st));
  if (n == 0) {
    CU(cudaMemsetAsync(d_offsets, 0, sizeof(uint32_t), st));
    CU(cudaMemsetAsync(d_total, 0, sizeof(uint64_t), st));
    CU(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    return CHQ_OK;
  }
  uint8_t* nwords = base + kStaticOffNwords;
  uint32_t* wid = reinterpret_cast<uint32_t*>(base + kStaticOffWid);
  const int split = (t->host.type() == CHQ_TOPIC || t->host.type() == CHQ_HEADERS) ? 1 : 0;
  CU(launch_tokenize(d_keys_blob, d_key_offsets, n, t->tv, split, wid, nwords, d_err, st));
  RouteParams p{};
  p.tv = t->tv;
  p.wid = wid;
  p.nwords = nwords;
  p.n = n;
  p.n_tiles = n_tiles;
  p.offsets = d_offsets;
  p.ids = d_queue_ids;
  p.ids_cap = ids_cap;
  p.durable = d_any_durable;
  p.total = d_total;
  p.ctrl = reinterpret_cast<uint32_t*>(state);
  p.tile_state = state + kStateHead;
  p.cnt = reinterpret_cast<uint32_t*>(base + kStaticOffCnt);
  p.meta = reinterpret_cast<uint32_t*>(base + kStaticOffMeta);
  p.mlist = reinterpret_cast<uint2*>(base + kStaticOffMlist);
  p.rest = reinterpret_cast<uint32_t*>(base + kStaticOffRest);
  p.scratch = reinterpret_cast<uint32_t*>(base + kStaticOffArena);
  p.scratch_cap = kStaticArenaIds;
  p.max_slots = t->max_slots;
  p.max_small = t->max_small;
  p.debug = 0;
  CU(launch_route(p, t->n_sms, st));
  CU(launch_finish(p.ctrl, d_status, st));
  t->n_batches++;
  t->n_messages += n;
  return CHQ_OK;
}

// ---------------------------------------------------------------------------------------------
// binding-slice sharding: union of the per-slice results of one batch (after the all-gather)
// ---------------------------------------------------------------------------------------------
namespace {
Workspace g_merge_ws[64];
std::mutex g_merge_mu[64];  // host side of the per-GPU workspace; calls on one GPU must not overlap on the device (header)
}

int32_t chq_merge_slices_device(int32_t device, const uint32_t* d_slice_offsets, const uint32_t* d_slice_ids, uint64_t slice_cap,
                                const uint8_t* d_slice_durable, uint32_t n_slices, uint32_t n, uint32_t* d_offsets,
                                uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable, uint64_t* d_total,
                                int32_t* d_status, void* stream) {
  return chq_merge_slices_device2(device, d_slice_offsets, d_slice_ids, slice_cap, d_slice_durable, n_slices, n, d_offsets,
                                  d_queue_ids, ids_cap, d_any_durable, d_total, d_status, 0u, stream);
}

int32_t chq_merge_slices_device2(int32_t device, const uint32_t* d_slice_offsets, const uint32_t* d_slice_ids, uint64_t slice_cap,
                                 const uint8_t* d_slice_durable, uint32_t n_slices, uint32_t n, uint32_t* d_offsets,
                                 uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable, uint64_t* d_total,
                                 int32_t* d_status, uint32_t flags, void* stream) {
  if (device < 0 || device >= 64 || !d_slice_offsets || !d_slice_ids || !d_slice_durable || !d_offsets || !d_any_durable ||
      !d_total || !d_status || n_slices == 0 || n_slices > (uint32_t)kRouteSlots)
    return fail(CHQ_E_ARG, "chq_merge_slices_device: bad argument (1..8 slices)");
  if ((uint64_t)n_slices * slice_cap > 0xFFFFFFFFull) return fail(CHQ_E_ARG, "chq_merge_slices_device: gathered ids exceed 2^32-1");
  CU(cudaSetDevice(device));
  std::lock_guard<std::mutex> lock(g_merge_mu[device]);
  Workspace& ws = g_merge_ws[device];
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t n_tiles = route_scan_tiles(n);
  int rc;
  if ((rc = ws.state.ensure((size_t)n_tiles + kStateHead))) return rc;
  if ((rc = ws.scratch.ensure(ws.scratch_want))) return rc;
  if ((rc = ws.cnt.ensure((size_t)n + 1))) return rc;
  if ((rc = ws.meta.ensure((size_t)n + 1))) return rc;
  if ((rc = ws.mlist.ensure((size_t)n * kRouteSlots + 1))) return rc;
  if ((rc = ws.rest.ensure((size_t)n + 1))) return rc;
  CU(cudaMemsetAsync(ws.state.p, 0, ((size_t)n_tiles + kStateHead) * sizeof(uint64_t), st));
  if (n == 0) {
    CU(cudaMemsetAsync(d_offsets, 0, sizeof(uint32_t), st));
    CU(cudaMemsetAsync(d_total, 0, sizeof(uint64_t), st));
    CU(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    return CHQ_OK;
  }
  cudaDeviceProp prop{};
  static int sms[64] = {};
  if (!sms[device]) { CU(cudaGetDeviceProperties(&prop, device)); sms[device] = prop.multiProcessorCount; }
  RouteParams p{};
  p.tv.qlist = d_slice_ids;
  p.tv.wstride = 4;
  p.n = n;
  p.n_tiles = n_tiles;
  p.offsets = d_offsets;
  p.ids = d_queue_ids;
  p.ids_cap = ids_cap;
  p.durable = d_any_durable;
  p.total = d_total;
  p.ctrl = reinterpret_cast<uint32_t*>(ws.state.p);
  p.tile_state = ws.state.p + kStateHead;
  p.cnt = ws.cnt.p;
  p.meta = ws.meta.p;
  p.mlist = ws.mlist.p;
  p.rest = ws.rest.p;
  p.scratch = ws.scratch.p;
  p.scratch_cap = ws.scratch.cap;
  p.max_slots = kRouteSlots;
  p.max_small = kRouteSmall;
  p.slice_offsets = d_slice_offsets;
  p.slice_durable = d_slice_durable;
  p.n_slices = n_slices;
  p.slice_cap = slice_cap;
  p.slice_disjoint = (flags & CHQ_MERGE_DISJOINT) ? 1u : 0u;
  CU(launch_route(p, sms[device], st));
  CU(launch_finish(p.ctrl, d_status, st));
  return CHQ_OK;
}

int32_t chq_host_alloc(void** p, uint64_t bytes) {
  if (!p) return fail(CHQ_E_ARG, "chq_host_alloc: bad argument");
  CU(cudaMallocHost(p, bytes ? bytes : 1));
  return CHQ_OK;
}
int32_t chq_host_free(void* p) {
  if (p) CU(cudaFreeHost(p));
  return CHQ_OK;
}

int32_t chq_stats(const chq_table* t, char* json, uint32_t cap) {
  if (!t || !json || !cap) return fail(CHQ_E_ARG, "chq_stats: bad argument");
  const CompiledTable& c = t->ct;
  const uint64_t table_bytes = c.dict.size() * sizeof(DictSlot) + c.dict_bytes.size() + c.wdict.size() * (sizeof(WordSlot) + 4) + c.shapes.size() * sizeof(Shape) +
                               c.tags.size() * 2 + c.slot_group.size() * 4 + c.groups.size() * sizeof(Group) +
                               (c.lits.size() + c.rems.size() + c.qlist.size()) * 4;
  int w = snprintf(json, cap,
                   "{\"bindings\": %llu, \"queues\": %zu, \"groups\": %zu, \"heads\": %llu, \"shapes\": %zu, "
                   "\"words\": %zu, \"pattern_words\": %llu, \"wstride\": %u, \"table_bytes\": %llu, "
                   "\"batches\": %llu, \"messages\": %llu, \"ids\": %llu, \"full_commits\": %llu, "
                   "\"incremental_commits\": %llu, \"committed\": %s}",
                   (unsigned long long)c.n_bindings, c.queue_names.size(), c.groups.size(), (unsigned long long)c.n_heads,
                   c.shapes.size(), (size_t)c.n_words, (unsigned long long)c.pattern_words, c.wstride,
                   (unsigned long long)table_bytes, (unsigned long long)t->n_batches, (unsigned long long)t->n_messages,
                   (unsigned long long)t->n_ids, (unsigned long long)t->n_full, (unsigned long long)t->n_incremental,
                   t->compiled ? "true" : "false");
  return (w < 0 || (uint32_t)w >= cap) ? fail(CHQ_E_ARG, "chq_stats: buffer too small") : CHQ_OK;
}

static int32_t profile_collect(chq_table* t, int32_t enable, double ms[5], uint32_t* launches, const char* who) {
  if (!t) return fail(CHQ_E_ARG, std::string(who) + ": bad argument");
  if (t->host_only) return fail(CHQ_E_CUDA, std::string(who) + ": host-only table");
  int rc = set_device(t);
  if (rc) return rc;
  CU(cudaDeviceSynchronize());
  uint32_t cnt = 0;
  for (int k = 0; k < 5; ++k) ms[k] = 0;
  Workspace* all[CHQ_SLOTS + 1];
  for (int i = 0; i < CHQ_SLOTS; ++i) all[i] = &t->slots[i].ws;
  all[CHQ_SLOTS] = &t->dev_ws;
  for (Workspace* ws : all) {
    for (int i = 0; i < ws->prof_n; ++i) {
      for (int k = 0; k < 5; ++k) {
        float a = 0;
        CU(cudaEventElapsedTime(&a, ws->ev[i][k], ws->ev[i][k + 1]));
        ms[k] += a;
      }
      ++cnt;
    }
    ws->prof_n = 0;
  }
  if (launches) *launches = cnt;
  t->profile = enable != 0;
  return CHQ_OK;
}

int32_t chq_profile(chq_table* t, int32_t enable, double* tokenize_ms, double* route_ms, uint32_t* launches) {
  double ms[5];
  int rc = profile_collect(t, enable, ms, launches, "chq_profile");
  if (rc) return rc;
  if (tokenize_ms) *tokenize_ms = ms[0];
  if (route_ms) *route_ms = ms[1] + ms[2] + ms[3] + ms[4];
  return CHQ_OK;
}

int32_t chq_profile_kernels(chq_table* t, int32_t enable, double* kernel_ms, uint32_t* launches) {
  if (!kernel_ms) return fail(CHQ_E_ARG, "chq_profile_kernels: bad argument");
  return profile_collect(t, enable, kernel_ms, launches, "chq_profile_kernels");
}

uint64_t chq_kernel_launches(void) { return kernel_launches() ; }

int32_t chq_debug_clocks(uint64_t* out32) {
  if (!out32) return fail(CHQ_E_ARG, "chq_debug_clocks: bad argument");
  CU(cudaDeviceSynchronize());
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "");
  CU(read_clocks(reinterpret_cast<unsigned long long*>(out32)));
  return CHQ_OK;
}

int32_t chq_debug_check(chq_table* t) {
  // test hook: structural invariants of the compiled table on the host (after in-place commits above all)
  if (!t) return fail(CHQ_E_ARG, "chq_debug_check: bad argument");
  std::lock_guard<std::mutex> g(t->mu);
  const CompiledTable& c = t->ct;
  if (c.headers_mode) return CHQ_OK;
  std::vector<uint32_t> uses(c.queue_names.size(), 0);
  for (size_t gi = 0; gi < c.groups.size(); ++gi) {
    const Group& gr = c.groups[gi];
    if ((size_t)gr.list_off + gr.list_len > c.qlist.size()) return fail(CHQ_E_ARG, "group " + std::to_string(gi) + ": list outside qlist");
    for (uint32_t i = 0; i < gr.list_len; ++i) {
      const uint32_t q = c.qlist[gr.list_off + i];
      if (q >= c.queue_names.size()) return fail(CHQ_E_ARG, "group " + std::to_string(gi) + ": id out of range");
      if (i && c.qlist[gr.list_off + i - 1] >= q) return fail(CHQ_E_ARG, "group " + std::to_string(gi) + ": list not ascending");
      uses[q]++;
    }
    if (gr.shape >= c.shapes.size()) return fail(CHQ_E_ARG, "group " + std::to_string(gi) + ": shape out of range");
  }
  for (size_t gi = 0; gi < c.groups.size(); ++gi) {
    const Group& gr = c.groups[gi];
    for (uint32_t i = 0; i < gr.list_len; ++i)
      if (uses[c.qlist[gr.list_off + i]] > 1 && !(gr.flags & kShared))
        return fail(CHQ_E_ARG, "group " + std::to_string(gi) + ": shares a queue with another group but is not flagged");
  }
  return CHQ_OK;
}

int32_t chq_debug_limits(chq_table* t, uint32_t max_slots, uint32_t max_small) {
  if (!t) return fail(CHQ_E_ARG, "chq_debug_limits: bad argument");
  t->max_slots = (max_slots == 0 || max_slots > (uint32_t)kRouteSlots) ? kRouteSlots : max_slots;
  t->max_small = (max_small == 0 || max_small > (uint32_t)kRouteSmall) ? kRouteSmall : max_small;
  return CHQ_OK;
}

}  // extern "C"
