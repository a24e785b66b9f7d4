// Binding-slice sharding of ONE exchange over the GPUs of a box, behind the C ABI (include/chanamq_gpu.h,
// chq_shard_*): BASELINE.json configs[3], SURVEY.md §8e.  The reference has no counterpart (one exchange is one
// actor on one thread, chana-mq-server/.../entity/ExchangeEntity.scala:107).
//
// One process per GPU.  Binding i lives on rank i mod G; every rank numbers the queues of the WHOLE exchange
// (chq_declare_queues), so ids agree across slices.  Per batch, on every rank, all on the caller's stream:
//   1. the local route: ALL keys against this rank's slice (chq_route_device) -> slice CSR in a buffer the other
//      ranks can read (cudaIpc);
//   2. the small arrays over NCCL / NVLink, in one group: every rank sends every owner the slice offsets and durable
//      flags of that owner's message range (ncclSend / ncclRecv: 1/G of what an all-gather of the whole arrays would
//      move to every rank — at 8 GPUs that was 42 MB per rank and step, 0.2 ms), and the status words are
//      all-gathered; the collective is also what orders step 3 behind every rank's step 1.  CHQ_SHARD_ALLGATHER=1
//      selects the plain ncclAllGather of the whole arrays instead;
//   3. k_pull: rank g owns the messages [g n / G, (g+1) n / G).  It fetches, from every slice, just the ids of its
//      own range: plain coalesced loads from the peers' buffers over NVLink into local memory (each id crosses NVLink
//      once, nothing is padded, no size goes through the host), and rebases the offsets;
//   4. the per-message union of the G slices of its range (the route kernels in slice mode).
// The slice buffers are double-buffered: batch b + 2 overwrites what batch b left only after the all-gather of batch
// b + 1 has completed, which every rank enters after its merge of batch b.
//
// NCCL is reached through dlopen (the copy already in the process if there is one, e.g. torch's), so the library
// itself loads on machines without NCCL; chq_shard_* then fail with CHQ_E_CUDA and say why.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/chanamq_gpu.h"
#include "kernels.h"

namespace {

thread_local std::string g_shard_err;
int sfail(int code, const std::string& m) { g_shard_err = m; return code; }
#define SCU(call)                                                                                              \
  do {                                                                                                         \
    cudaError_t e__ = (call);                                                                                  \
    if (e__ != cudaSuccess) return sfail(CHQ_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));     \
  } while (0)

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string why;
};
NcclApi g_nccl;
std::once_flag g_nccl_once;

const NcclApi& nccl() {
  std::call_once(g_nccl_once, [] {
    NcclApi& a = g_nccl;
    a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // the copy the process already uses (torch's), if any
    if (!a.lib) a.lib = dlopen("libnccl.so.2", RTLD_NOW);
    if (!a.lib) { a.why = std::string("libnccl.so.2 not found: ") + dlerror(); return; }
    auto sym = [&](const char* n) { void* p = dlsym(a.lib, n); if (!p && a.why.empty()) a.why = std::string("NCCL lacks ") + n; return p; };
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
    a.Send = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
    a.Recv = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
  });
  return g_nccl;
}
#define SNC(call)                                                                                                  \
  do {                                                                                                             \
    ncclResult_t r__ = (call);                                                                                     \
    if (r__ != ncclSuccess) return sfail(CHQ_E_CUDA, std::string(#call) + ": " + nccl().GetErrorString(r__));      \
  } while (0)

constexpr int kMaxWorld = chq::kRouteSlots;   // the merge takes up to eight lists per message

}  // namespace

struct chq_shard {
  int rank = 0, world = 1, device = 0;
  chq_table* table = nullptr;
  ncclComm_t comm = nullptr;
  // which slice holds each queue name's bindings (kept while binding): no name on two slices = the unions need no
  // de-duplication
  std::unordered_map<std::string, int> queue_slice;
  bool disjoint = true;
  // reserved by chq_shard_reserve
  uint32_t max_keys = 0;
  uint64_t slice_cap = 0, pull_cap = 0;
  unsigned char* loc[2] = {nullptr, nullptr};        // slice result of a batch: offsets[n+2] | durable | err | ids
  const uint32_t* peer_ids[2][kMaxWorld] = {};        // the ids part of every rank's loc[b] (own: local pointer)
  void* peer_base[2][kMaxWorld] = {};                 // what cudaIpcOpenMemHandle returned (to close)
  uint32_t* all_off = nullptr;                        // [G][max_keys + 2] gathered
  uint8_t* all_dur = nullptr;                         // [G][max_keys]
  uint32_t* sub_off = nullptr;                        // [G][max range + 1] rebased
  uint8_t* sub_dur = nullptr;                         // [G][max range]
  uint32_t* got = nullptr;                            // [G][pull_cap] the ids of this rank's range, slice by slice
  uint64_t* local_total = nullptr;                    // [2]: total, status of the local route
  int32_t* pull_status = nullptr;
  int32_t* statuses = nullptr;                        // [G] gathered status words of the local routes
  bool whole_arrays = false;                          // CHQ_SHARD_ALLGATHER=1: all-gather the whole small arrays
  const uint32_t** d_peer_ids = nullptr;              // device copy of peer_ids [2][kMaxWorld]
  unsigned batch = 0;
  // stage timing (chq_shard_stage_ms): events around the four stages of the last kStageRing batches
  static constexpr int kStageRing = 64;
  cudaEvent_t ev[kStageRing][5] = {};
  int ev_n = 0;
  bool timing = false;
  size_t off_bytes() const { return ((size_t)(max_keys + 2) * 4 + 255) & ~size_t(255); }
  size_t dur_bytes() const { return ((size_t)max_keys + 255) & ~size_t(255); }
  uint32_t* loc_off(int b) const { return reinterpret_cast<uint32_t*>(loc[b]); }
  uint8_t* loc_dur(int b) const { return loc[b] + off_bytes(); }
  uint8_t* loc_err(int b) const { return loc[b] + off_bytes() + dur_bytes(); }
  uint32_t* loc_ids(int b) const { return reinterpret_cast<uint32_t*>(loc[b] + off_bytes() + 2 * dur_bytes()); }
};

namespace {

// Rank g's part of step 3 (see the head of the file), one grid row per slice r: the ids of the messages [m0, m1)
// from slice r's result (a peer's memory for r != g: plain loads over NVLink) into got[r][...], that slice's offsets
// rebased to the range, its durable flags, and the status words of all local routes folded into *status.
// Slice r's offsets of the range are all_off[r * off_stride + first + 0 .. nn], its durable flags all_dur[r * dur_stride +
// first + 0 .. nn) (first = 0 when only the ranges were exchanged, = the range's first message after a whole all-gather).
__global__ void __launch_bounds__(256) k_pull(const uint32_t* __restrict__ all_off, size_t off_stride,
                                              const uint8_t* __restrict__ all_dur, size_t dur_stride, uint32_t first, uint32_t nn,
                                              const int32_t* __restrict__ statuses, const uint32_t* const* __restrict__ peer_ids,
                                              uint32_t* __restrict__ got, uint64_t pull_cap, uint32_t* __restrict__ sub_off,
                                              uint8_t* __restrict__ sub_dur, int32_t* __restrict__ status) {
  const uint32_t r = blockIdx.y, m0 = first;
  const uint32_t* so = all_off + (size_t)r * off_stride;
  const uint32_t a = so[m0], b = so[m0 + nn];
  const uint64_t len = b - a, take = len < pull_cap ? len : pull_cap;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
  if (tid == 0) {
    const int32_t st = statuses[r];                    // slice r's local route: CHQ_OK, CHQ_E_OVERFLOW, CHQ_E_SCRATCH
    if (st != CHQ_OK) atomicMin(status, st);
    if (len > pull_cap) atomicMin(status, (int32_t)CHQ_E_OVERFLOW);
  }
  const uint32_t* src = peer_ids[r] + a;
  uint32_t* dst = got + (size_t)r * pull_cap;
  // 16-byte loads on the (possibly remote) source, four per thread in flight: a trip over NVLink takes microseconds,
  // the link is only kept busy by many wide requests.  The source decides the alignment; the local side takes what comes.
  const size_t head = min((size_t)((4u - (a & 3u)) & 3u), (size_t)take);
  if (tid < head) dst[tid] = src[tid];
  const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
  uint32_t* d = dst + head;
  const size_t n4 = (take - head) / 4;
  size_t i = tid;
  for (; i + 3 * nthr < n4; i += 4 * nthr) {
    const uint4 v0 = __ldcs(s4 + i), v1 = __ldcs(s4 + i + nthr), v2 = __ldcs(s4 + i + 2 * nthr), v3 = __ldcs(s4 + i + 3 * nthr);
    uint32_t* o0 = d + 4 * i;
    uint32_t* o1 = d + 4 * (i + nthr);
    uint32_t* o2 = d + 4 * (i + 2 * nthr);
    uint32_t* o3 = d + 4 * (i + 3 * nthr);
    o0[0] = v0.x; o0[1] = v0.y; o0[2] = v0.z; o0[3] = v0.w;
    o1[0] = v1.x; o1[1] = v1.y; o1[2] = v1.z; o1[3] = v1.w;
    o2[0] = v2.x; o2[1] = v2.y; o2[2] = v2.z; o2[3] = v2.w;
    o3[0] = v3.x; o3[1] = v3.y; o3[2] = v3.z; o3[3] = v3.w;
  }
  for (; i < n4; i += nthr) {
    const uint4 v = __ldcs(s4 + i);
    uint32_t* o = d + 4 * i;
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
  for (size_t j = head + 4 * n4 + tid; j < take; j += nthr) dst[j] = src[j];
  uint32_t* o = sub_off + (size_t)r * (nn + 1);
  for (size_t j = tid; j <= nn; j += nthr) {
    const uint64_t v = so[m0 + j] - a;
    o[j] = (uint32_t)(v < pull_cap ? v : pull_cap);    // (after an overflow the merge still stays inside got[r])
  }
  const uint8_t* du = all_dur + (size_t)r * dur_stride + m0;
  for (size_t j = tid; j < nn; j += nthr) sub_dur[(size_t)r * nn + j] = du[j];
}

// the pull's verdict overrides the merge's
__global__ void k_shard_status(const int32_t* pull_status, int32_t* status) {
  if (*pull_status != CHQ_OK) *status = *pull_status;
}

void shard_release(chq_shard* s) {
  for (int b = 0; b < 2; ++b) {
    for (int r = 0; r < kMaxWorld; ++r)
      if (s->peer_base[b][r]) { cudaIpcCloseMemHandle(s->peer_base[b][r]); s->peer_base[b][r] = nullptr; }
    if (s->loc[b]) { cudaFree(s->loc[b]); s->loc[b] = nullptr; }
  }
  cudaFree(s->all_off); cudaFree(s->all_dur); cudaFree(s->sub_off); cudaFree(s->sub_dur); cudaFree(s->got);
  cudaFree(s->local_total); cudaFree(s->pull_status); cudaFree(s->d_peer_ids); cudaFree(s->statuses);
  s->statuses = nullptr;
  for (auto& e : s->ev) for (auto& x : e) if (x) { cudaEventDestroy(x); x = nullptr; }
  s->ev_n = 0;
  s->all_off = s->sub_off = s->got = nullptr;
  s->all_dur = s->sub_dur = nullptr;
  s->local_total = nullptr; s->pull_status = nullptr; s->d_peer_ids = nullptr;
  s->max_keys = 0;
}

}  // namespace

extern "C" {

const char* chq_shard_last_error(void) { return g_shard_err.c_str(); }

int32_t chq_shard_unique_id(uint8_t* id128) {
  if (!id128) return sfail(CHQ_E_ARG, "chq_shard_unique_id: bad argument");
  const NcclApi& nc = nccl();
  if (!nc.why.empty()) return sfail(CHQ_E_CUDA, "chq_shard_unique_id: " + nc.why);
  static_assert(sizeof(ncclUniqueId) == CHQ_SHARD_ID_BYTES, "CHQ_SHARD_ID_BYTES is NCCL's unique id");
  ncclUniqueId id;
  SNC(nc.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return CHQ_OK;
}

int32_t chq_shard_create(chq_shard** out, const uint8_t* id128, int32_t rank, int32_t world, int32_t device,
                         int32_t exchange_type, int32_t mode) {
  if (!out || !id128 || world < 1 || world > kMaxWorld || rank < 0 || rank >= world)
    return sfail(CHQ_E_ARG, "chq_shard_create: bad argument (1..8 ranks)");
  const NcclApi& nc = nccl();
  if (!nc.why.empty()) return sfail(CHQ_E_CUDA, "chq_shard_create: " + nc.why);
  chq_shard* s = new (std::nothrow) chq_shard();
  if (!s) return sfail(CHQ_E_ARG, "out of host memory");
  s->rank = rank; s->world = world; s->device = device;
  int rc = chq_table_create(&s->table, exchange_type, mode, device);
  if (rc) { g_shard_err = chq_last_error(); delete s; return rc; }
  if (cudaSetDevice(device) != cudaSuccess) { chq_table_destroy(s->table); delete s; return sfail(CHQ_E_CUDA, "cudaSetDevice failed"); }
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclResult_t r = nc.CommInitRank(&s->comm, world, id, rank);   // collective: every rank of the exchange calls it
  if (r != ncclSuccess) {
    chq_table_destroy(s->table);
    delete s;
    return sfail(CHQ_E_CUDA, std::string("ncclCommInitRank: ") + nc.GetErrorString(r));
  }
  *out = s;
  return CHQ_OK;
}

int32_t chq_shard_destroy(chq_shard* s) {
  if (!s) return CHQ_OK;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();
  shard_release(s);
  if (s->comm) nccl().CommDestroy(s->comm);
  chq_table_destroy(s->table);
  delete s;
  return CHQ_OK;
}

chq_table* chq_shard_table(chq_shard* s) { return s ? s->table : nullptr; }

int32_t chq_shard_bind_many(chq_shard* s, const uint8_t* keys_blob, const uint32_t* key_offsets, const uint8_t* queues_blob,
                            const uint32_t* queue_offsets, const uint8_t* durable, uint32_t n, uint32_t first_index) {
  if (!s || !key_offsets || !queue_offsets) return sfail(CHQ_E_ARG, "chq_shard_bind_many: bad argument");
  int first_err = CHQ_OK;
  for (uint32_t i = 0; i < n; ++i) {
    const int owner = (int)(((uint64_t)first_index + i) % (uint32_t)s->world);
    const uint8_t* q = queues_blob + queue_offsets[i];
    const uint32_t ql = queue_offsets[i + 1] - queue_offsets[i];
    auto ins = s->queue_slice.emplace(std::string((const char*)q, ql), owner);
    if (!ins.second && ins.first->second != owner) s->disjoint = false;
    if (owner != s->rank) continue;
    int rc = chq_bind(s->table, keys_blob + key_offsets[i], key_offsets[i + 1] - key_offsets[i], q, ql, durable ? durable[i] : 0);
    if (rc && !first_err) { first_err = rc; g_shard_err = chq_last_error(); }
  }
  // every slice numbers the queues of the whole exchange
  int rc = chq_declare_queues(s->table, queues_blob, queue_offsets, n);
  if (rc && !first_err) { first_err = rc; g_shard_err = chq_last_error(); }
  return first_err;
}

int32_t chq_shard_commit(chq_shard* s) {
  if (!s) return sfail(CHQ_E_ARG, "chq_shard_commit: bad argument");
  int rc = chq_commit_full(s->table);   // every slice must number the queues alike: by name rank
  if (rc) g_shard_err = chq_last_error();
  return rc;
}

int32_t chq_shard_is_disjoint(const chq_shard* s) { return s && s->disjoint ? 1 : 0; }

int32_t chq_shard_reserve(chq_shard* s, uint32_t max_keys, uint64_t slice_cap, uint64_t pull_cap) {
  if (!s || max_keys == 0 || slice_cap == 0 || pull_cap == 0) return sfail(CHQ_E_ARG, "chq_shard_reserve: bad argument");
  if ((uint64_t)s->world * pull_cap > 0xFFFFFFFFull || slice_cap > 0xFFFFFFFFull)
    return sfail(CHQ_E_ARG, "chq_shard_reserve: capacities exceed 2^32-1 ids");
  const NcclApi& nc = nccl();
  SCU(cudaSetDevice(s->device));
  SCU(cudaDeviceSynchronize());
  shard_release(s);
  s->max_keys = max_keys; s->slice_cap = slice_cap; s->pull_cap = pull_cap;
  const int G = s->world;
  const size_t loc_bytes = s->off_bytes() + 2 * s->dur_bytes() + (size_t)slice_cap * 4 + 256;
  const uint32_t max_range = max_keys / G + 2;
  for (int b = 0; b < 2; ++b) SCU(cudaMalloc(&s->loc[b], loc_bytes));
  SCU(cudaMalloc(&s->all_off, (size_t)G * (max_keys + 2) * 4));
  SCU(cudaMalloc(&s->all_dur, (size_t)G * max_keys + 16));
  SCU(cudaMalloc(&s->sub_off, (size_t)G * (max_range + 1) * 4));
  SCU(cudaMalloc(&s->sub_dur, (size_t)G * max_range + 16));
  SCU(cudaMalloc(&s->got, (size_t)G * pull_cap * 4 + 256));
  SCU(cudaMalloc(&s->local_total, 2 * sizeof(uint64_t)));
  SCU(cudaMalloc(&s->pull_status, sizeof(int32_t)));
  SCU(cudaMalloc(&s->statuses, sizeof(int32_t) * kMaxWorld));
  { const char* e = getenv("CHQ_SHARD_ALLGATHER"); s->whole_arrays = e && atoi(e) != 0; }
  SCU(cudaMalloc(&s->d_peer_ids, sizeof(uint32_t*) * 2 * kMaxWorld));
  // every rank learns where the others keep their slice results: the cudaIpc handles of both buffers, and the sizes
  // they were reserved with (which must agree), travel by one all-gather
  struct Hello { cudaIpcMemHandle_t h[2]; uint64_t max_keys, slice_cap, pull_cap, pad; };
  Hello mine{};
  for (int b = 0; b < 2; ++b) SCU(cudaIpcGetMemHandle(&mine.h[b], s->loc[b]));
  mine.max_keys = max_keys; mine.slice_cap = slice_cap; mine.pull_cap = pull_cap;
  Hello *d_mine = nullptr, *d_all = nullptr;
  SCU(cudaMalloc(&d_mine, sizeof(Hello)));
  SCU(cudaMalloc(&d_all, sizeof(Hello) * G));
  SCU(cudaMemcpy(d_mine, &mine, sizeof(Hello), cudaMemcpyHostToDevice));
  SNC(nc.AllGather(d_mine, d_all, sizeof(Hello), ncclUint8, s->comm, nullptr));
  SCU(cudaStreamSynchronize(nullptr));
  std::vector<Hello> all(G);
  SCU(cudaMemcpy(all.data(), d_all, sizeof(Hello) * G, cudaMemcpyDeviceToHost));
  cudaFree(d_mine);
  cudaFree(d_all);
  const size_t ids_at = s->off_bytes() + 2 * s->dur_bytes();
  for (int r = 0; r < G; ++r) {
    if (all[r].max_keys != max_keys || all[r].slice_cap != slice_cap || all[r].pull_cap != pull_cap)
      return sfail(CHQ_E_ARG, "chq_shard_reserve: the ranks of one exchange must reserve the same sizes");
    for (int b = 0; b < 2; ++b) {
      if (r == s->rank) { s->peer_ids[b][r] = s->loc_ids(b); continue; }
      void* base = nullptr;
      SCU(cudaIpcOpenMemHandle(&base, all[r].h[b], cudaIpcMemLazyEnablePeerAccess));   // (one process per GPU)
      s->peer_base[b][r] = base;
      s->peer_ids[b][r] = reinterpret_cast<const uint32_t*>(static_cast<unsigned char*>(base) + ids_at);
    }
  }
  SCU(cudaMemcpy(s->d_peer_ids, s->peer_ids, sizeof(uint32_t*) * 2 * kMaxWorld, cudaMemcpyHostToDevice));
  s->batch = 0;
  return CHQ_OK;
}

int32_t chq_shard_route_device(chq_shard* s, const uint8_t* d_keys_blob, const uint32_t* d_key_offsets, uint32_t n,
                               uint32_t* d_offsets, uint32_t* d_queue_ids, uint64_t ids_cap, uint8_t* d_any_durable,
                               uint8_t* d_err, uint64_t* d_total, int32_t* d_status, uint32_t* first_message,
                               uint32_t* end_message, void* stream) {
  if (!s || !d_offsets || !d_total || !d_status || !first_message || !end_message || (n && (!d_key_offsets || !d_any_durable || !d_err)))
    return sfail(CHQ_E_ARG, "chq_shard_route_device: bad argument");
  if (!s->max_keys) return sfail(CHQ_E_ARG, "chq_shard_route_device: call chq_shard_reserve first");
  if (n == 0 || n > s->max_keys) return sfail(CHQ_E_ARG, "chq_shard_route_device: 1 .. max_keys keys per batch (chq_shard_reserve)");
  chq::NvtxRange nvtx_range("chq_shard_route_device (route, all-gather, pull, merge)");
  const NcclApi& nc = nccl();
  SCU(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int G = s->world, b = (int)(s->batch++ & 1u);
  const uint32_t m0 = (uint32_t)(((uint64_t)s->rank * n) / G), m1 = (uint32_t)(((uint64_t)(s->rank + 1) * n) / G), nn = m1 - m0;
  *first_message = m0;
  *end_message = m1;
  cudaEvent_t* pe = nullptr;
  if (s->timing && s->ev_n < chq_shard::kStageRing) {
    pe = s->ev[s->ev_n++];
    for (int i = 0; i < 5; ++i)
      if (!pe[i]) SCU(cudaEventCreate(&pe[i]));
    SCU(cudaEventRecord(pe[0], st));
  }
  // 1. all keys against this rank's slice; the status word goes behind the offsets so that it is gathered with them
  uint32_t* lo = s->loc_off(b);
  int rc = chq_route_device(s->table, d_keys_blob, d_key_offsets, n, lo, s->loc_ids(b), s->slice_cap, s->loc_dur(b), s->loc_err(b),
                            s->local_total, reinterpret_cast<int32_t*>(lo + n + 1), st);
  if (rc) { g_shard_err = chq_last_error(); return rc; }
  if (pe) SCU(cudaEventRecord(pe[1], st));
  // 2. the small arrays: every owner gets its range of every slice's offsets and durable flags (or, on request,
  //    everybody gets everything), and everybody gets every status word
  size_t off_stride, dur_stride;
  uint32_t first;
  SNC(nc.GroupStart());
  SNC(nc.AllGather(lo + n + 1, s->statuses, 1, ncclInt32, s->comm, st));
  if (s->whole_arrays) {
    SNC(nc.AllGather(lo, s->all_off, (size_t)n + 2, ncclUint32, s->comm, st));
    SNC(nc.AllGather(s->loc_dur(b), s->all_dur, (size_t)n, ncclUint8, s->comm, st));
    off_stride = (size_t)n + 2; dur_stride = n; first = m0;
  } else {
    off_stride = (size_t)nn + 1; dur_stride = nn; first = 0;
    for (int g = 0; g < G; ++g) {
      if (g == s->rank) continue;
      const uint32_t g0 = (uint32_t)(((uint64_t)g * n) / G), g1 = (uint32_t)(((uint64_t)(g + 1) * n) / G);
      SNC(nc.Send(lo + g0, (size_t)(g1 - g0) + 1, ncclUint32, g, s->comm, st));
      SNC(nc.Recv(s->all_off + (size_t)g * off_stride, (size_t)nn + 1, ncclUint32, g, s->comm, st));
      if (g1 > g0) SNC(nc.Send(s->loc_dur(b) + g0, (size_t)(g1 - g0), ncclUint8, g, s->comm, st));
      if (nn) SNC(nc.Recv(s->all_dur + (size_t)g * dur_stride, (size_t)nn, ncclUint8, g, s->comm, st));
    }
  }
  SNC(nc.GroupEnd());
  if (!s->whole_arrays) {  // this rank's own slice of its own range
    SCU(cudaMemcpyAsync(s->all_off + (size_t)s->rank * off_stride, lo + m0, ((size_t)nn + 1) * 4, cudaMemcpyDeviceToDevice, st));
    if (nn) SCU(cudaMemcpyAsync(s->all_dur + (size_t)s->rank * dur_stride, s->loc_dur(b) + m0, nn, cudaMemcpyDeviceToDevice, st));
  }
  if (pe) SCU(cudaEventRecord(pe[2], st));
  // 3. the ids of this rank's message range from every slice (peers: over NVLink), offsets rebased
  SCU(cudaMemsetAsync(s->pull_status, 0, sizeof(int32_t), st));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
  const unsigned rows = (unsigned)G, cols = (unsigned)((sms * 8 + G - 1) / G);
  k_pull<<<dim3(cols, rows), 256, 0, st>>>(s->all_off, off_stride, s->all_dur, dur_stride, first, nn, s->statuses,
                                           s->d_peer_ids + (size_t)b * kMaxWorld, s->got, s->pull_cap, s->sub_off, s->sub_dur,
                                           s->pull_status);
  chq::note_launches(1);
  SCU(cudaGetLastError());
  if (nn) SCU(cudaMemcpyAsync(d_err, s->loc_err(b) + m0, nn, cudaMemcpyDeviceToDevice, st));
  if (pe) SCU(cudaEventRecord(pe[3], st));
  // 4. per-message union of the G slices over the range
  rc = chq_merge_slices_device2(s->device, s->sub_off, s->got, s->pull_cap, s->sub_dur, (uint32_t)G, nn, d_offsets, d_queue_ids,
                                ids_cap, d_any_durable, d_total, d_status, s->disjoint ? CHQ_MERGE_DISJOINT : 0u, st);
  if (rc) { g_shard_err = chq_last_error(); return rc; }
  k_shard_status<<<1, 1, 0, st>>>(s->pull_status, d_status);
  chq::note_launches(1);
  SCU(cudaGetLastError());
  if (pe) SCU(cudaEventRecord(pe[4], st));
  return CHQ_OK;
}

int32_t chq_shard_stage_ms(chq_shard* s, int32_t enable, double* stage_ms, uint32_t* batches) {
  if (!s || !stage_ms) return sfail(CHQ_E_ARG, "chq_shard_stage_ms: bad argument");
  SCU(cudaSetDevice(s->device));
  SCU(cudaDeviceSynchronize());
  for (int k = 0; k < 4; ++k) stage_ms[k] = 0;
  for (int i = 0; i < s->ev_n; ++i)
    for (int k = 0; k < 4; ++k) {
      float a = 0;
      SCU(cudaEventElapsedTime(&a, s->ev[i][k], s->ev[i][k + 1]));
      stage_ms[k] += a;
    }
  if (batches) *batches = (uint32_t)s->ev_n;
  s->ev_n = 0;
  s->timing = enable != 0;
  return CHQ_OK;
}

}  // extern "C"
