// The step after the match: ExchangeEntity.publish turns the per-message queue sets into
// per-QUEUE message lists in publish order and a few per-message flags
// (chana-mq-server/.../entity/ExchangeEntity.scala:300-307, 312-330):
//
//   queues foreach { q => queueToMsgs += (q -> (queueToMsgs(q) :+ Msg(msgId, ...))) }        EX:305-307
//   isPersistent = exchange.isDurable && msg.isPersist && OR(subscriber.isDurable)            EX:302
//   Refer(msgId, queues.size, isPersistent)                                                   EX:303
//   isNonRouted = mandatory && queues.isEmpty ; isNonDeliverable = immediate && no consumer   EX:319-330
//
// Here: the CSR message -> queue ids that the route kernels wrote is transposed into a CSR
// queue -> message indices (ascending = publish order), entirely on the device:
//   k_fan_count    histogram of the queue ids                       (one atomic per id)
//   k_fan_scan_*   exclusive prefix sum over the queues             (tile sums, their scan, tile scans)
//   k_fan_scatter  every id drops its message index into its queue's range (order by arrival)
//   k_fan_sort_*   every queue's range is put in ascending order (registers / shared memory / global), which
//                  makes the result independent of the arrival order: identical from run to run
//   k_fan_flags    the per-message flags
// Integer work; bounded by atomic throughput into L2 and by the sorts, not by HBM.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "../../include/chanamq_gpu.h"
#include "kernels.h"

namespace {

constexpr uint32_t kFull32 = 0xFFFFFFFFu;
constexpr int kScanTileFan = 4096;        // queues per scan tile (1024 threads x 4)
constexpr uint32_t kSortSmem = 1024;      // ranges up to this many entries are sorted in shared memory by one warp
constexpr uint32_t kNoMsg = 0xFFFFFFFFu;

__global__ void k_fan_count(const uint32_t* __restrict__ ids, const uint32_t* __restrict__ offsets, uint32_t n, uint32_t n_queues,
                            uint32_t* __restrict__ qcount, uint32_t* __restrict__ status) {
  const uint64_t total = offsets[n];
  bool bad = false;
  for (uint64_t e = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; e < total; e += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t q = __ldg(ids + e);
    if (q < n_queues) atomicAdd(qcount + q, 1u);
    else bad = true;
  }
  if (bad) atomicOr(status, 1u);  // a queue id outside [0, n_queues): the caller passed the wrong table size
}

// ---- exclusive scan of qcount[0..n_queues) -> q_offsets[0..n_queues], three small kernels ----
__global__ void __launch_bounds__(1024) k_fan_scan_tiles(const uint32_t* __restrict__ qcount, uint32_t n_queues,
                                                         uint32_t* __restrict__ tile_sum) {
  __shared__ uint32_t wsum[32];
  const uint32_t base = blockIdx.x * kScanTileFan + threadIdx.x * 4;
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) s += base + i < n_queues ? qcount[base + i] : 0u;
#pragma unroll
  for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(kFull32, s, d);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t v = wsum[threadIdx.x];
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(kFull32, v, d);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024) k_fan_scan_sums(uint32_t* __restrict__ tile_sum, uint32_t n_tiles) {  // one block
  __shared__ uint32_t wtot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t b = 0; b < n_tiles; b += 1024) {
    const uint32_t i = b + threadIdx.x;
    const uint32_t v = i < n_tiles ? tile_sum[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(kFull32, incl, d);
      if ((threadIdx.x & 31) >= d) incl += t;
    }
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t wbase = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += wtot[w];
    if (i < n_tiles) tile_sum[i] = carry + wbase + incl - v;  // exclusive
    __syncthreads();
    if (threadIdx.x == 1023) carry += wbase + incl;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(1024) k_fan_scan_apply(const uint32_t* __restrict__ qcount, uint32_t n_queues,
                                                         const uint32_t* __restrict__ tile_base, uint32_t* __restrict__ q_offsets,
                                                         uint32_t* __restrict__ cursor) {
  __shared__ uint32_t wtot[32];
  const uint32_t base = blockIdx.x * kScanTileFan + threadIdx.x * 4;
  uint32_t v[4], s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[i] = base + i < n_queues ? qcount[base + i] : 0u; s += v[i]; }
  uint32_t incl = s;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(kFull32, incl, d);
    if ((threadIdx.x & 31) >= d) incl += t;
  }
  if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = incl;
  __syncthreads();
  uint32_t wbase = 0;
  for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += wtot[w];
  uint32_t run = tile_base[blockIdx.x] + wbase + incl - s;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (base + i < n_queues) { q_offsets[base + i] = run; cursor[base + i] = run; }
    run += v[i];
  }
  if (base < n_queues && base + 4 >= n_queues) q_offsets[n_queues] = run;  // the thread that holds the last queue
}

// Every id drops its message index into its queue's range.  One warp per 32 consecutive messages; the lanes
// run over a message's ids, so the id reads are coalesced.
__global__ void k_fan_scatter(const uint32_t* __restrict__ ids, const uint32_t* __restrict__ offsets, uint32_t n, uint32_t n_queues,
                              uint32_t* __restrict__ cursor, uint32_t* __restrict__ q_msgs, uint64_t q_cap) {
  const int lane = threadIdx.x & 31;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t m0 = warp * 32; m0 < n; m0 += n_warps * 32) {
    const uint32_t mine = m0 + lane;
    const uint32_t a = mine < n ? offsets[mine] : 0u, b = mine < n ? offsets[mine + 1] : 0u;
    for (int j = 0; j < 32; ++j) {
      const uint32_t lo = __shfl_sync(kFull32, a, j), hi = __shfl_sync(kFull32, b, j);
      for (uint32_t e = lo + lane; e < hi; e += 32) {
        const uint32_t q = __ldg(ids + e);
        if (q < n_queues) {
          const uint32_t slot = atomicAdd(cursor + q, 1u);
          if (slot < q_cap) q_msgs[slot] = m0 + j;
        }
      }
    }
  }
}

// Ascending order inside every queue's range.  One warp per queue: up to 32 entries in registers, up to
// kSortSmem in shared memory; longer ranges are left to k_fan_sort_big (their queue ids are appended to big[]).
__global__ void __launch_bounds__(128) k_fan_sort_small(const uint32_t* __restrict__ q_offsets, uint32_t n_queues,
                                                        uint32_t* __restrict__ q_msgs, uint64_t q_cap,
                                                        uint32_t* __restrict__ big, uint32_t* __restrict__ n_big) {
  __shared__ uint32_t buf[4][kSortSmem];
  if ((uint64_t)q_offsets[n_queues] > q_cap) return;  // CHQ_E_OVERFLOW: the ranges reach past q_msgs, nothing there is touched
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t q = warp; q < n_queues; q += n_warps) {
    const uint32_t a = q_offsets[q], len = q_offsets[q + 1] - a;
    if (len <= 1) continue;
    if (len <= 32) {
      uint32_t v = (uint32_t)lane < len ? q_msgs[a + lane] : kNoMsg;
#pragma unroll
      for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int d = size >> 1; d; d >>= 1) {
          const uint32_t other = __shfl_xor_sync(kFull32, v, d);
          const bool up = (lane & size) == 0, lower = (lane & d) == 0;
          v = (lower == up) ? min(v, other) : max(v, other);
        }
      }
      if ((uint32_t)lane < len) q_msgs[a + lane] = v;
    } else if (len <= kSortSmem) {
      uint32_t np = 64;
      while (np < len) np <<= 1;
      uint32_t* s = buf[w];
      for (uint32_t i = lane; i < np; i += 32) s[i] = i < len ? q_msgs[a + i] : kNoMsg;
      __syncwarp();
      for (uint32_t size = 2; size <= np; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
          for (uint32_t t = lane; t < (np >> 1); t += 32) {
            const uint32_t i = ((t / stride) * (stride << 1)) + (t % stride), j = i + stride;
            const bool up = (i & size) == 0;
            const uint32_t x = s[i], y = s[j];
            if ((x > y) == up) { s[i] = y; s[j] = x; }
          }
          __syncwarp();
        }
      }
      for (uint32_t i = lane; i < len; i += 32) q_msgs[a + i] = s[i];
      __syncwarp();
    } else if (lane == 0) {
      big[atomicAdd(n_big, 1u)] = q;
    }
  }
}

// Long ranges: the entries of a queue are distinct message indices below n, so they are sorted by marking
// them in a bitmap of n bits and reading the bitmap back in order.  The bitmap lives in shared memory when the
// batch is small enough for it (bitmaps == null; up to ~1.6 M messages), else in global memory, one per CTA.
__global__ void __launch_bounds__(1024) k_fan_sort_big(const uint32_t* __restrict__ q_offsets, uint32_t* __restrict__ q_msgs,
                                                       const uint32_t* __restrict__ big, const uint32_t* __restrict__ n_big,
                                                       uint32_t n, uint32_t* __restrict__ bitmaps) {
  // (on CHQ_E_OVERFLOW k_fan_sort_small returns before it queues any range, so *n_big == 0 and nothing happens here)
  extern __shared__ uint32_t smem_bm[];
  __shared__ uint32_t wtot[32];
  const uint32_t words = (n + 31) / 32;
  // word w of the bitmap sits at w + w / 32: every thread reads a run of consecutive words below, and the skew
  // keeps the 32 runs of a warp in different shared-memory banks
  const uint32_t phys_words = words + words / 32 + 1;
  uint32_t* bm = bitmaps ? bitmaps + (size_t)blockIdx.x * phys_words : smem_bm;
  auto at = [](uint32_t w) { return w + (w >> 5); };
  const uint32_t per = (words + blockDim.x - 1) / blockDim.x;  // consecutive words per thread
  for (uint32_t k = blockIdx.x; k < *n_big; k += gridDim.x) {
    const uint32_t q = big[k], a = q_offsets[q], len = q_offsets[q + 1] - a;
    for (uint32_t i = threadIdx.x; i < phys_words; i += blockDim.x) bm[i] = 0;
    __syncthreads();
    // (combining the bits of neighbouring lanes before the atomic — match_any or a segmented shuffle scan — was
    // measured and is slower than the plain shared-memory atomics)
    for (uint32_t i = threadIdx.x; i < len; i += blockDim.x) {
      const uint32_t m = q_msgs[a + i];
      atomicOr(bm + at(m >> 5), 1u << (m & 31));
    }
    __syncthreads();
    const uint32_t w0 = threadIdx.x * per, w1 = min(w0 + per, words);
    uint32_t cnt = 0;
    for (uint32_t i = w0; i < w1; ++i) cnt += __popc(bm[at(i)]);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(kFull32, incl, d);
      if ((threadIdx.x & 31) >= d) incl += t;
    }
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t wbase = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += wtot[w];
    uint32_t out = a + wbase + incl - cnt;
    for (uint32_t i = w0; i < w1; ++i)
      for (uint32_t bits = bm[at(i)]; bits; bits &= bits - 1) q_msgs[out++] = i * 32 + (__ffs(bits) - 1);
    __syncthreads();
  }
}

// Per-message outputs of the publish loop (EX:302-303, EX:319-330).  pub_flags[i]: CHQ_PUB_MANDATORY /
// CHQ_PUB_IMMEDIATE / CHQ_PUB_PERSIST as in chq_publish_rec.flags; queue_has_consumer (optional) one byte per queue.
__global__ void k_fan_flags(const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ ids,
                            const uint8_t* __restrict__ any_durable, const uint8_t* __restrict__ pub_flags, uint32_t n,
                            int exchange_durable, const uint8_t* __restrict__ queue_has_consumer, uint32_t n_queues,
                            uint32_t* __restrict__ refer_count, uint8_t* __restrict__ out_flags) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t a = offsets[i], b = offsets[i + 1], cnt = b - a;
  const uint32_t pf = pub_flags ? pub_flags[i] : 0u;
  uint32_t f = 0;
  if (exchange_durable && (pf & CHQ_PUB_PERSIST) && any_durable[i]) f |= CHQ_FAN_PERSISTENT;
  if ((pf & CHQ_PUB_MANDATORY) && cnt == 0) f |= CHQ_FAN_NON_ROUTED;
  if (pf & CHQ_PUB_IMMEDIATE) {
    bool consumer = false;
    if (queue_has_consumer)
      for (uint32_t e = a; e < b && !consumer; ++e) {
        const uint32_t q = __ldg(ids + e);
        consumer = q < n_queues && queue_has_consumer[q] != 0;
      }
    if (!consumer) f |= CHQ_FAN_NON_DELIVERABLE;
  }
  refer_count[i] = cnt;
  out_flags[i] = (uint8_t)f;
}

__global__ void k_fan_status(const uint32_t* __restrict__ offsets, uint32_t n, uint64_t q_cap, const uint32_t* __restrict__ bad,
                             int32_t* __restrict__ status) {
  *status = *bad ? CHQ_E_ARG : ((uint64_t)offsets[n] > q_cap ? CHQ_E_OVERFLOW : CHQ_OK);
}

thread_local std::string g_fan_err;
int ffail(int code, const std::string& m) { g_fan_err = m; return code; }
#define FCU(call)                                                                                              \
  do {                                                                                                         \
    cudaError_t e__ = (call);                                                                                  \
    if (e__ != cudaSuccess) return ffail(CHQ_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));     \
  } while (0)

struct FanScratch {  // per device; grown on demand
  uint32_t* qcount = nullptr; size_t qcount_cap = 0;   // [n_queues] counts, then reused as nothing else
  uint32_t* cursor = nullptr; size_t cursor_cap = 0;   // [n_queues]
  uint32_t* tiles = nullptr; size_t tiles_cap = 0;     // [tiles + 2]: tile sums; last two words: n_big, status
  uint32_t* big = nullptr; size_t big_cap = 0;         // [n_queues]
  uint32_t* bitmaps = nullptr; size_t bitmaps_cap = 0; // [CTAs x ceil(n / 32)]
};
FanScratch g_fan[64];
std::mutex g_fan_mu[64];  // host side of the scratch (growing it, queueing the launches); see chq_fanout_device

template <typename T>
int fgrow(T** p, size_t* cap, size_t need) {
  if (need <= *cap) return CHQ_OK;
  if (*p) FCU(cudaFree(*p));
  *p = nullptr; *cap = 0;
  const size_t want = need + need / 8 + 64;
  FCU(cudaMalloc(p, want * sizeof(T)));
  *cap = want;
  return CHQ_OK;
}

}  // namespace

extern "C" {

const char* chq_fanout_last_error(void) { return g_fan_err.c_str(); }

int32_t chq_fanout_device(const uint32_t* d_offsets, const uint32_t* d_queue_ids, uint32_t n, uint32_t n_queues,
                          uint32_t* d_q_offsets, uint32_t* d_q_msgs, uint64_t q_cap, int32_t* d_status, void* stream) {
  if (!d_offsets || !d_q_offsets || !d_status || (q_cap && (!d_queue_ids || !d_q_msgs)))
    return ffail(CHQ_E_ARG, "chq_fanout_device: bad argument");
  int dev = 0, sms = 148;
  FCU(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  FanScratch& sc = g_fan[dev & 63];
  // One scratch per GPU: concurrent callers queue their launches one after the other here, but the kernels of two
  // calls on different streams would still share it on the device — calls on one GPU must not overlap (header).
  std::lock_guard<std::mutex> lock(g_fan_mu[dev & 63]);
  cudaStream_t st = (cudaStream_t)stream;
  const uint32_t n_tiles = (n_queues + kScanTileFan - 1) / kScanTileFan;
  // bitmap of the long-range sort: shared memory if the batch fits, else global
  const size_t bm_words = (size_t)((n + 31) / 32) + ((n + 31) / 32) / 32 + 1;  // skewed layout, see k_fan_sort_big
  const size_t bm_bytes = bm_words * 4;
  static int smem_ok[64] = {};  // 0 unknown, 1 opted in, -1 not available
  if (smem_ok[dev & 63] == 0)
    smem_ok[dev & 63] = cudaFuncSetAttribute(k_fan_sort_big, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) == cudaSuccess ? 1 : -1;
  const bool bm_smem = smem_ok[dev & 63] == 1 && bm_bytes <= 200 * 1024;
  const uint32_t big_ctas = bm_smem ? (uint32_t)sms : 64u;
  int rc;
  if ((rc = fgrow(&sc.qcount, &sc.qcount_cap, (size_t)n_queues + 1))) return rc;
  if ((rc = fgrow(&sc.cursor, &sc.cursor_cap, (size_t)n_queues + 1))) return rc;
  if ((rc = fgrow(&sc.tiles, &sc.tiles_cap, (size_t)n_tiles + 2))) return rc;
  if ((rc = fgrow(&sc.big, &sc.big_cap, (size_t)n_queues + 1))) return rc;
  if (!bm_smem && (rc = fgrow(&sc.bitmaps, &sc.bitmaps_cap, (size_t)big_ctas * bm_words + 1))) return rc;
  uint32_t* n_big = sc.tiles + n_tiles;
  uint32_t* bad = sc.tiles + n_tiles + 1;
  FCU(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  if (n_queues == 0 || n == 0) {
    FCU(cudaMemsetAsync(d_q_offsets, 0, ((size_t)n_queues + 1) * sizeof(uint32_t), st));
    return CHQ_OK;
  }
  FCU(cudaMemsetAsync(sc.qcount, 0, (size_t)n_queues * sizeof(uint32_t), st));
  FCU(cudaMemsetAsync(n_big, 0, 2 * sizeof(uint32_t), st));
  k_fan_count<<<sms * 16, 256, 0, st>>>(d_queue_ids, d_offsets, n, n_queues, sc.qcount, bad);
  k_fan_scan_tiles<<<n_tiles, 1024, 0, st>>>(sc.qcount, n_queues, sc.tiles);
  k_fan_scan_sums<<<1, 1024, 0, st>>>(sc.tiles, n_tiles);
  k_fan_scan_apply<<<n_tiles, 1024, 0, st>>>(sc.qcount, n_queues, sc.tiles, d_q_offsets, sc.cursor);
  k_fan_scatter<<<sms * 16, 256, 0, st>>>(d_queue_ids, d_offsets, n, n_queues, sc.cursor, d_q_msgs, q_cap);
  k_fan_sort_small<<<sms * 16, 128, 0, st>>>(d_q_offsets, n_queues, d_q_msgs, q_cap, sc.big, n_big);
  k_fan_sort_big<<<big_ctas, 1024, bm_smem ? bm_bytes : 0, st>>>(d_q_offsets, d_q_msgs, sc.big, n_big, n,
                                                                 bm_smem ? nullptr : sc.bitmaps);
  k_fan_status<<<1, 1, 0, st>>>(d_offsets, n, q_cap, bad, d_status);
  chq::note_launches(8);
  FCU(cudaGetLastError());
  return CHQ_OK;
}

int32_t chq_publish_flags_device(const uint32_t* d_offsets, const uint32_t* d_queue_ids, const uint8_t* d_any_durable,
                                 const uint8_t* d_pub_flags, uint32_t n, int32_t exchange_durable,
                                 const uint8_t* d_queue_has_consumer, uint32_t n_queues, uint32_t* d_refer_count,
                                 uint8_t* d_out_flags, void* stream) {
  if (n && (!d_offsets || !d_any_durable || !d_refer_count || !d_out_flags))
    return ffail(CHQ_E_ARG, "chq_publish_flags_device: bad argument");
  if (n == 0) return CHQ_OK;
  k_fan_flags<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_offsets, d_queue_ids, d_any_durable, d_pub_flags, n,
                                                                 exchange_durable, d_queue_has_consumer, n_queues, d_refer_count,
                                                                 d_out_flags);
  chq::note_launches(1);
  FCU(cudaGetLastError());
  return CHQ_OK;
}

}  // extern "C"
