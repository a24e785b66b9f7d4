// sm_100a kernels of the routing hot path.
//
//   k_tokenize : routing-key bytes -> interned word ids        (replaces String.split + the per-level
//                                                               String hashing of TrieMatcher.lookup,
//                                                               QueueMatcher.scala:403-413, 261-263)
//   k_probe/k_offsets/k_write : word ids x compiled table -> CSR of queue ids (replaces ilookup/bLookup,
//                                                               QueueMatcher.scala:523-585, and the
//                                                               set/dedup/flag work of
//                                                               ExchangeEntity.scala:297-302)
//
// Routing a batch is k_tokenize, then three launches (k_probe -> k_offsets -> k_write): probe every
// key against the compiled table and count its queues exactly, prefix-sum the counts into CSR
// offsets, write the merged, sorted, de-duplicated id lists straight to their final position.
// Inputs are read once and outputs written once; the table (tags, groups, queue lists) is L2-resident.  Integer/byte work throughout — no tensor
// cores (there is no dense contraction on this path).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/chanamq_gpu.h"
#include "chq_common.h"
#include "kernels.h"
#include "merge.cuh"
#include "probe.cuh"
#include "weave.cuh"
#include "flat.cuh"
#include "kway.cuh"

#ifndef CHQ_MIN_CTAS
#define CHQ_MIN_CTAS 8
#endif
#ifndef CHQ_PROBE_SORT
#define CHQ_PROBE_SORT 1
#endif
#ifndef CHQ_EXP
#define CHQ_EXP 0   // timing experiments only (results are wrong): 1 no stream, 2 no placement, 4 no tiny messages,
#endif              // 8 no flat messages at all, 16 no k_write_rest launch
#ifndef CHQ_WRITE_CTAS
#define CHQ_WRITE_CTAS 8
#endif

namespace chq {

// One thread per key; the threads of a warp take one WORD per loop trip, so the dictionary probe is
// executed by all of them together.  Java split semantics: words are cut at every '.', trailing
// empty words are dropped, a key without any '.' is one word (even when empty).
// Key i is blob[off[i] .. off[i+1]) (klen == null: keys packed back to back) or blob[off[i] .. off[i]+klen[i])
// (keys scattered inside a byte batch, as the frame decoder leaves them).
// Bytes are fetched as aligned 32-bit words and shifted into place: eight key bytes = three loads that
// mostly hit L1 (neighbouring keys share the lines).  A word that straddles the edge of the readable
// range [rlo, rhi) is assembled from byte loads, so nothing outside that range is ever touched.
constexpr int kTokThreads = 256;
constexpr int kTokStage = 8;   // rows of up to this many word ids are staged in shared memory and stored coalesced

// 0x80 in every byte of x that equals '.'
__device__ __forceinline__ uint32_t dot_bytes(uint32_t x) {
  x ^= 0x2E2E2E2Eu;
  return (x - 0x01010101u) & ~x & 0x80808080u;
}

__global__ void __launch_bounds__(kTokThreads) k_tokenize(const uint8_t* __restrict__ blob, const uint32_t* __restrict__ off,
                                                          const uint32_t* __restrict__ klen, uint32_t n, TableView tv, int split,
                                                          uint32_t* __restrict__ wid, uint8_t* __restrict__ nwords,
                                                          uint8_t* __restrict__ err) {
  __shared__ uint32_t stage[kTokThreads][kTokStage + 1];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const bool staged = tv.wstride <= (uint32_t)kTokStage;
  // byte offsets below are relative to the 4-byte aligned address at or before blob
  const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(blob) & 3u);
  const uint32_t* ab = reinterpret_cast<const uint32_t*>(blob - mis);
  uint32_t v0 = 0, v1 = 0, rlo = 0, rhi = 0;
  bool active = i < n;
  if (active) {
    const uint32_t b0 = off[i], b1 = klen ? b0 + klen[i] : off[i + 1];
    v0 = b0 + mis; v1 = b1 + mis;
    if (klen) { rlo = v0; rhi = v1; }                 // scattered keys: stay inside this key
    else { rlo = mis; rhi = __ldg(off + n) + mis; }   // packed keys: stay inside the blob
    if (b1 - b0 > (uint32_t)kMaxKeyBytes) { nwords[i] = 0; err[i] = 2; active = false; }
  }
  if (staged) {
#pragma unroll
    for (int c = 0; c < kTokStage; ++c) stage[threadIdx.x][c] = 0;
  }
  uint32_t* row = wid + (size_t)i * tv.wstride;
  const bool live = active;
  uint32_t cnt = 0, last_nonempty = 0, pos = v0;
  bool dot = false;
  while (active) {
    // one word: bytes from pos up to the next '.' or the end of the key, eight at a time
    uint32_t h = kWordBasis, wl = 0, f_lo = 0, f_hi = 0;
    bool ends_with_dot = false;
    for (uint32_t q = pos;; q += 8) {
      const uint32_t k = q >> 2, sh = (q & 3u) * 8u;
      const uint32_t w0 = key_word32(ab, k, rlo, rhi), w1 = key_word32(ab, k + 1, rlo, rhi);
      const uint32_t w2 = sh ? key_word32(ab, k + 2, rlo, rhi) : 0u;
      uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
      const uint32_t rem = v1 - q;
      uint32_t p = 8;
      if (split) {
        const uint32_t dl = dot_bytes(lo), dh = dot_bytes(hi);
        p = dl ? (uint32_t)(__ffs(dl) - 1) >> 3 : (dh ? 4u + ((uint32_t)(__ffs(dh) - 1) >> 3) : 8u);
      }
      const uint32_t lim = min(rem, 8u), take = min(p, lim);
      if (take < 8) {
        if (take >= 4) hi &= (1u << (8 * (take - 4))) - 1u;
        else { hi = 0; lo &= (1u << (8 * take)) - 1u; }
      }
      if (q == pos) { f_lo = lo; f_hi = hi; }
      if (take > 0 || q == pos) h = word_chunk(h, lo, hi);
      wl += take;
      if (take < 8) { ends_with_dot = p < lim; break; }
    }
    const uint32_t id = word_lookup(tv, word_fin(h, wl), f_lo, f_hi, wl, blob + (pos - mis));
    if (cnt < tv.wstride) {
      if (staged) stage[threadIdx.x][cnt] = id;
      else row[cnt] = id;
    }
    ++cnt;
    if (wl > 0) last_nonempty = cnt;
    if (ends_with_dot) { dot = true; pos += wl + 1; }
    else active = false;
  }
  if (live) {
    const uint32_t nw = dot ? last_nonempty : 1;
    nwords[i] = (uint8_t)nw;
    err[i] = nw == 0 ? 1 : 0;
  }
  if (staged) {  // 32 rows of a warp are contiguous in wid: store them as whole lines
    __syncwarp();
    const uint32_t ws = tv.wstride, tbase = threadIdx.x - lane;
    const uint32_t first = blockIdx.x * blockDim.x + tbase;     // first key of this warp
    for (uint32_t x = lane; x < 32 * ws; x += 32) {
      const uint32_t r = x / ws, c = x - r * ws;
      if (first + r < n) wid[(size_t)first * ws + x] = stage[tbase + r][c];
    }
  }
}

// ------------------------------------------------------------------------------------------
// the fused route kernel
// ------------------------------------------------------------------------------------------
// Optional cycle accounting of k_write by kind of message (build with CHQ_NVCC_DEFS=-DCHQ_CLOCKS; read with
// chq_debug_clocks): [8*0 + b] cycles, [8*1 + b] messages, [8*2 + b] ids of bin b
// (0 sweep for big messages, 1 tiny, 2 single list, 3 slow-path copy, 4 multi-way, 5 weave <= 32 short ids,
// 6 weave > 32 short ids, 7 per-chunk header loads).
#ifdef CHQ_CLOCKS
__device__ unsigned long long g_clk[32];
#define CHQ_CLK_T0() const long long clk_t0 = clock64()
#define CHQ_CLK_ADD(bin, msgs, ids)                                              \
  do {                                                                           \
    if ((threadIdx.x & 31) == 0) {                                               \
      atomicAdd(&g_clk[bin], (unsigned long long)(clock64() - clk_t0));          \
      atomicAdd(&g_clk[8 + (bin)], (unsigned long long)(msgs));                  \
      atomicAdd(&g_clk[16 + (bin)], (unsigned long long)(ids));                  \
    }                                                                            \
  } while (0)
#else
#define CHQ_CLK_T0() do {} while (0)
#define CHQ_CLK_ADD(bin, msgs, ids) do {} while (0)
#endif

namespace {

constexpr int kWarps = kRouteThreads / 32;

struct RouteSmem {  // k_probe
  // per message: matched lists (offset, length), slot-major.  While a key is being probed the same slots hold
  // its unresolved tag hits (slot index, shape << 16 | literal mask): hit c is read before result c is written.
  uint32_t loff[kRouteSlots][kRouteThreads];
  uint32_t llen[kRouteSlots][kRouteThreads];
  WarpScratch ws[kWarps];   // for counting unions that must de-duplicate
};

// Generic slow path for one message, by one warp: every matched list is appended to a piece of the
// global scratch arena, sorted there and de-duplicated.  Returns the count; *scr_off = where.
__device__ uint32_t warp_fallback(const RouteParams& p, const uint32_t* __restrict__ row, uint32_t nw, int lane,
                                  uint64_t* scr_off, uint32_t* any_flags) {
  const TableView& tv = p.tv;
  const bool fan = tv.exchange_type == 1;
  const uint32_t L = min(nw, (uint32_t)kLenCap);
  const uint32_t e0 = nw <= (uint32_t)kLenCap ? tv.exact_rng[3 * L] : 0;
  const uint32_t n_exact = nw <= (uint32_t)kLenCap ? tv.exact_rng[3 * L + 1] - e0 : 0;
  const uint32_t n_shapes = fan ? 0 : n_exact + (tv.exact_rng[3 * L + 2] - tv.hashy_begin);
  uint64_t base = 0;
  uint32_t np = 0, total = 0, flags = 0;
  KeyCtx kctx;
  key_ctx_init(kctx, row, nw, tv.wstride);
  for (int pass = 0; pass < 2; ++pass) {
    uint32_t wpos = 0;
    for (uint32_t s0 = 0; s0 < n_shapes; s0 += 32) {
      const uint32_t si = s0 + lane;
      // each lane takes one shape; a lane can meet several groups (bucket collisions, chained
      // remainders), so the lanes walk their matches in lock-step rounds
      uint32_t mine = 0;
      LaneMatches lm;
      lane_matches_begin(lm, tv, kctx, si < n_shapes, si < n_exact ? e0 + si : tv.hashy_begin + (si - n_exact));
      for (;;) {
        GroupRec rec{};
        const bool hit = lane_matches_next(lm, tv, kctx, rec);
        if (!__any_sync(kFull, hit)) break;
        unsigned hm = __ballot_sync(kFull, hit);
        if (pass == 0) {
          if (hit) { mine += rec.list_len; flags |= rec.flags; }
        } else {
          while (hm) {
            const int src = __ffs(hm) - 1;
            hm &= hm - 1;
            const uint32_t o = __shfl_sync(kFull, rec.list_off, src), l = __shfl_sync(kFull, rec.list_len, src);
            for (uint32_t i = lane; i < l; i += 32) p.scratch[base + wpos + i] = __ldg(tv.qlist + o + i);
            wpos += l;
          }
        }
      }
      if (pass == 0) {
#pragma unroll
        for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(kFull, mine, d);
        total += mine;
      }
    }
    if (pass == 0) {
#pragma unroll
      for (int d = 16; d; d >>= 1) flags |= __shfl_xor_sync(kFull, flags, d);
      *any_flags = flags;
      if (total == 0) { *scr_off = 0; return 0; }
      np = 32;
      while (np < total) np <<= 1;
      unsigned long long got = 0;
      if (lane == 0) got = atomicAdd(reinterpret_cast<unsigned long long*>(p.ctrl + kCtrlScratchLo), (unsigned long long)np);
      base = __shfl_sync(kFull, got, 0);
      if (base + np > p.scratch_cap) {
        if (lane == 0) atomicOr(p.ctrl + kCtrlStatus, kStatusScratch);
        *scr_off = 0;
        return 0;
      }
    }
  }
  for (uint32_t i = total + lane; i < np; i += 32) p.scratch[base + i] = kNone;
  __syncwarp();
  warp_bitonic_sort(p.scratch + base, np, lane);
  const uint32_t cnt = warp_unique(p.scratch + base, total, lane);
  *scr_off = base;
  return cnt;
}

// Slow path with the lists given as (offset, length) descriptors in global memory.
__device__ uint32_t warp_fallback_desc(const RouteParams& p, const uint2* __restrict__ desc, uint32_t k, int lane, uint64_t* scr_off) {
  uint32_t total = 0;
  for (uint32_t j = 0; j < k; ++j) total += __ldg(desc + j).y;
  *scr_off = 0;
  if (total == 0) return 0;
  uint32_t np = 32;
  while (np < total) np <<= 1;
  unsigned long long got = 0;
  if (lane == 0) got = atomicAdd(reinterpret_cast<unsigned long long*>(p.ctrl + kCtrlScratchLo), (unsigned long long)np);
  const uint64_t base = __shfl_sync(kFull, got, 0);
  if (base + np > p.scratch_cap) {
    if (lane == 0) atomicOr(p.ctrl + kCtrlStatus, kStatusScratch);
    return 0;
  }
  uint32_t wpos = 0;
  for (uint32_t j = 0; j < k; ++j) {
    const uint2 d = __ldg(desc + j);
    for (uint32_t i = lane; i < d.y; i += 32) p.scratch[base + wpos + i] = __ldg(p.tv.qlist + d.x + i);
    wpos += d.y;
  }
  for (uint32_t i = total + lane; i < np; i += 32) p.scratch[base + i] = kNone;
  __syncwarp();
  warp_bitonic_sort(p.scratch + base, np, lane);
  const uint32_t cnt = warp_unique(p.scratch + base, total, lane);
  *scr_off = base;
  return cnt;
}

// Slow path when the lists are already known (lane j < k holds list j): append, sort, de-duplicate in
// the global scratch arena.
__device__ uint32_t warp_fallback_lists(const RouteParams& p, uint32_t my_off, uint32_t my_len, int k, int lane, uint64_t* scr_off) {
  uint32_t total = lane < k ? my_len : 0;
#pragma unroll
  for (int d = 16; d; d >>= 1) total += __shfl_xor_sync(kFull, total, d);
  *scr_off = 0;
  if (total == 0) return 0;
  uint32_t np = 32;
  while (np < total) np <<= 1;
  unsigned long long got = 0;
  if (lane == 0) got = atomicAdd(reinterpret_cast<unsigned long long*>(p.ctrl + kCtrlScratchLo), (unsigned long long)np);
  const uint64_t base = __shfl_sync(kFull, got, 0);
  if (base + np > p.scratch_cap) {
    if (lane == 0) atomicOr(p.ctrl + kCtrlStatus, kStatusScratch);
    return 0;
  }
  uint32_t wpos = 0;
  for (int j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, my_len, j);
    for (uint32_t i = lane; i < l; i += 32) p.scratch[base + wpos + i] = __ldg(p.tv.qlist + o + i);
    wpos += l;
  }
  for (uint32_t i = total + lane; i < np; i += 32) p.scratch[base + i] = kNone;
  __syncwarp();
  warp_bitonic_sort(p.scratch + base, np, lane);
  const uint32_t cnt = warp_unique(p.scratch + base, total, lane);
  *scr_off = base;
  return cnt;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// the route launch: three kernels on one stream
//   k_probe   one thread per message: which queue lists apply, and how many ids the union has
//   k_offsets exclusive prefix sum of the counts -> CSR offsets (single pass, chained look-back)
//   k_write   one warp per routed message (tiny ones by a single thread): the sorted union, written once
// No barrier separates a message's probing from anyone else's, and each kernel has its own
// register budget; the per-message list descriptors (64 bytes) travel through global memory.
// ------------------------------------------------------------------------------------------
namespace {

constexpr uint32_t kMetaTiny = 1u << 13, kMetaDedup = 1u << 12;  // meta = k | class << 8 | flags
constexpr uint32_t kMetaBig = 1u << 14;   // long result: k_write starts these first so that none is left for the end
constexpr uint32_t kMetaFlat = 1u << 15;  // one long list (<= kFlatMaxLong ids) + a short side, nothing shared: flat.cuh
constexpr uint32_t kBigIds = 4096;

}  // namespace

__global__ void __launch_bounds__(kRouteThreads, CHQ_MIN_CTAS) k_probe(RouteParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RouteSmem& sm = *reinterpret_cast<RouteSmem*>(smem_raw);
  const TableView& tv = p.tv;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t m = blockIdx.x * kRouteThreads + tid;

  uint32_t k = 0, ub = 0, mxlen = 0, mxidx = 0, flags = 0, cnt = 0, nw = 0;
  if (m < p.n && !p.slice_offsets && !p.desc_k) nw = p.nwords[m];
#if CHQ_PROBE_SORT
  // The number of shapes a key is probed under grows with its word count (7 .. 22 for 3 .. 6 words), so the 128
  // messages of the CTA are dealt to the threads in order of word count: the lanes of a warp then run loops of the
  // same length.  Every thread still owns one message; only which one changes (everything below is indexed by m).
  if (!p.slice_offsets && !p.desc_k) {
    __shared__ uint32_t s_bin[10];
    __shared__ uint16_t s_perm[kRouteThreads];
    if (tid < 10) s_bin[tid] = 0;
    __syncthreads();
    const uint32_t bin = min(nw, 9u);
    const uint32_t r = atomicAdd(&s_bin[bin], 1u);
    __syncthreads();
    uint32_t base = 0;
    for (uint32_t x = 0; x < bin; ++x) base += s_bin[x];
    s_perm[base + r] = (uint16_t)tid;
    __syncthreads();
    m = blockIdx.x * kRouteThreads + s_perm[tid];
    nw = m < p.n ? p.nwords[m] : 0u;
  }
#endif
  const uint32_t* row = p.wid + (size_t)m * tv.wstride;
  bool overflow = false;  // more tag hits than candidate slots: the slow path re-probes
  if (p.slice_offsets || p.desc_k) nw = m < p.n ? 1u : 0u;
#ifdef CHQ_DEBUG_SWITCHES
  if (p.debug & 2u) nw = 0;
#endif
  if (nw > 0) {
    auto record = [&](const GroupRec& rec) {
      if (k < p.max_slots) { sm.loff[k][tid] = rec.list_off; sm.llen[k][tid] = rec.list_len; }
      if (rec.list_len > mxlen) { mxlen = rec.list_len; mxidx = k; }
      ++k;
      ub += rec.list_len;
      flags |= rec.flags;
    };
    if (p.slice_offsets) {  // merge of per-slice results: the lists are given, nothing to probe
      for (uint32_t r = 0; r < p.n_slices; ++r) {
        const uint32_t* so = p.slice_offsets + (size_t)r * (p.n + 1) + m;
        const uint32_t a = __ldg(so), e = __ldg(so + 1);
        if (e > a) {
          GroupRec rec{};
          rec.list_off = (uint32_t)(r * p.slice_cap + a);
          rec.list_len = e - a;
          rec.flags = (p.slice_disjoint ? 0u : (uint32_t)kShared) | (__ldg(p.slice_durable + (size_t)r * p.n + m) ? kAnyDurable : 0);
          record(rec);
        } else if (__ldg(p.slice_durable + (size_t)r * p.n + m)) {
          flags |= kAnyDurable;
        }
      }
    } else if (p.desc_k) {  // headers exchange: k_headers_match left the lists in global memory
      const uint32_t kd = __ldg(p.desc_k + m);
      const uint2* dl = p.desc_lists + __ldg(p.desc_start + m);
      for (uint32_t j = 0; j < kd; ++j) {
        const uint2 d = __ldg(dl + j);
        GroupRec rec{};
        rec.list_off = d.x;
        rec.list_len = d.y;
        record(rec);
      }
      flags |= __ldg(p.desc_flags + m);
    } else if (tv.exchange_type == 1 /* fanout */) {
      if (tv.fanout_group != kNone) record(load_group(tv, tv.fanout_group));
    } else {
      KeyCtx kctx;
      key_ctx_init(kctx, row, nw, tv.wstride);
      const uint32_t L = min(nw, (uint32_t)kLenCap);
      const uint32_t e0 = nw <= (uint32_t)kLenCap ? tv.exact_rng[3 * L] : 0;
      const uint32_t ne = nw <= (uint32_t)kLenCap ? tv.exact_rng[3 * L + 1] - e0 : 0;
      const uint32_t nh = tv.exact_rng[3 * L + 2] - tv.hashy_begin;
      uint32_t ncand = 0;
      probe_shapes(tv, kctx, e0, ne, tv.hashy_begin, nh, [&](uint32_t slot, uint32_t shape, uint32_t lit_mask) {
        if (ncand < (uint32_t)kRouteSlots) {
          sm.loff[ncand][tid] = slot;
          sm.llen[ncand][tid] = (shape << 16) | (lit_mask & 0xFFFFu);
        }
        ++ncand;
      });
      overflow = ncand > (uint32_t)kRouteSlots;
      for (uint32_t c = 0; c < min(ncand, (uint32_t)kRouteSlots); ++c) {
        const uint32_t slot = sm.loff[c][tid], sl = sm.llen[c][tid];
        resolve_candidate(tv, kctx, slot, sl >> 16, sl & 0xFFFFu, [&](const GroupRec& rec) {
          if (k > c) overflow = true;  // a chain gave more results than hits read so far: let the slow path redo it
          else record(rec);
        });
      }
    }
  }
  // classify: 0 = count known, 1 = needs the de-duplicating merge to count, 2 = slow path (result parked in
  // scratch), 3 = count known, several long lists (k_write merges them with warp_union_long), 4 = more lists
  // than slots but few ids, lists described in global memory (k_write: warp_gather_small)
  int cls = 0;
  const bool dedup = (flags & kShared) != 0;
  // tiny: the whole result fits the 16-input sorting network of one thread (merge.cuh)
  // (a headers exchange: up to 16 one-queue lists are the usual message; those past the slots are read where
  // k_headers_match left them, here and in k_write — class 4 says so)
  const bool tiny_desc = p.desc_k != nullptr && !overflow && k > p.max_slots && ub <= (uint32_t)kTinyIds;
  const bool tiny = !overflow && k >= 1 && (k <= p.max_slots || tiny_desc) && ub <= (uint32_t)kTinyIds;
  const uint2* dl = tiny_desc ? p.desc_lists + __ldg(p.desc_start + m) : nullptr;
  auto my_off = [&](uint32_t j) { return dl ? __ldg(dl + j).x : sm.loff[j][tid]; };
  auto my_len = [&](uint32_t j) { return dl ? __ldg(dl + j).y : sm.llen[j][tid]; };
  if (tiny) {
    cnt = (k == 1 || !dedup) ? ub : tiny_union(tv.qlist, k, my_off, my_len, nullptr);
    if (tiny_desc) cls = 4;
  }
  else if (overflow) cls = 2;
  else if (p.desc_k && k > p.max_slots && k <= (uint32_t)kRouteSmall && ub <= (uint32_t)kRouteSmall) { cls = 4; cnt = ub; }
  else if (k <= 1) cnt = ub;
  else if (k > p.max_slots) cls = 2;
  // one long list + at most max_small other ids: k_write's staged weave (weave.cuh); more: warp_union_long
  // lists of similar length that share nothing (the slices of a sharded exchange, several '#' patterns): staged and
  // merged pairwise in shared memory (kway.cuh)
  // (in a table: from four lists on, none of which dominates — otherwise streaming the longest and placing the others
  // was measured faster; the slices of a sharded exchange: always — placing means k - 1 searches in global memory per
  // id there, and set the run time of k_write; two slices: 0.41 -> 0.26 ms per half batch against warp_merge2)
  else if (!dedup && k <= p.max_slots && ub <= (uint32_t)kKwayMax && p.max_small == (uint32_t)kRouteSmall &&
           (p.slice_offsets ? k >= 2 : (k >= 4 && ub - mxlen >= 2u * mxlen && ub > 64u))) { cls = 5; cnt = ub; }
  // two lists of similar length that share nothing, too long for the above: a symmetric merge through registers (warp_merge2)
  // (slices of a sharded exchange only: in a table such pairs are rare, and the few long ones, merged by one warp in
  // dependent trips, set the run time of k_write_rest — measured: +0.005 ms on the headline batch)
  else if (p.slice_offsets && !dedup && k == 2 && k <= p.max_slots && 2u * (ub - mxlen) >= mxlen && ub > 64u && ub < (1u << 30) &&
           p.max_small == (uint32_t)kRouteSmall) { cls = 6; cnt = ub; }
  // (nothing shared between the lists: the flat path of k_write takes a much longer short side)
  else if (ub - mxlen > ((dedup || mxlen > (uint32_t)kFlatMaxLong || p.max_small != (uint32_t)kRouteSmall) ? p.max_small
                                                                                                             : (uint32_t)kFlatMaxShort) ||
           ub >= (1u << 30)) { cls = 3; cnt = ub; }  // dedup: counted below
  else if (dedup) cls = 1;
  else cnt = ub;

  uint64_t scr = 0;
  __syncwarp();
  for (unsigned todo = __ballot_sync(kFull, cls == 2); todo; todo &= todo - 1) {
    const int src = __ffs(todo) - 1;
    const uint32_t mm = __shfl_sync(kFull, m, src);
    uint64_t so;
    uint32_t fl;
    uint32_t c;
    if (p.desc_k) {  // more lists than slots: read the descriptors again
      fl = __shfl_sync(kFull, flags, src);
      c = warp_fallback_desc(p, p.desc_lists + __ldg(p.desc_start + mm), __shfl_sync(kFull, k, src), lane, &so);
    } else if (p.slice_offsets) {  // lists are in shared memory (at most n_slices <= kRouteSlots of them)
      const int kk = __shfl_sync(kFull, k, src);
      uint32_t o = 0, l = 0;
      if (lane < kk) { o = sm.loff[lane][warp * 32 + src]; l = sm.llen[lane][warp * 32 + src]; }
      fl = __shfl_sync(kFull, flags, src);
      c = warp_fallback_lists(p, o, l, kk, lane, &so);
    } else {
      c = warp_fallback(p, p.wid + (size_t)mm * tv.wstride, __shfl_sync(kFull, nw, src), lane, &so, &fl);
    }
    if (lane == src) { cnt = c; flags = fl; scr = so; }
  }
  for (unsigned todo = __ballot_sync(kFull, cls == 4 && dedup && !tiny); todo; todo &= todo - 1) {
    const int src = __ffs(todo) - 1;
    const uint32_t mm = __shfl_sync(kFull, m, src);
    const uint32_t c = warp_gather_small<false>(tv.qlist, p.desc_lists + __ldg(p.desc_start + mm), __shfl_sync(kFull, k, src),
                                                true, sm.ws[warp], nullptr, 0, lane);
    if (lane == src) cnt = c;
  }
  for (unsigned todo = __ballot_sync(kFull, cls == 1 || (cls == 3 && dedup)); todo; todo &= todo - 1) {
    const int src = __ffs(todo) - 1;
    const int kk = __shfl_sync(kFull, k, src);
    const bool lng = __shfl_sync(kFull, cls, src) == 3;
    uint32_t o = 0, l = 0;
    if (lane < kk) { o = sm.loff[lane][warp * 32 + src]; l = sm.llen[lane][warp * 32 + src]; }
    const uint32_t c = lng ? warp_union_long<false>(tv.qlist, o, l, kk, true, sm.ws[warp], nullptr, 0, lane)
                           : warp_union<false>(tv.qlist, o, l, kk, true, sm.ws[warp], nullptr, 0, 0, lane);
    if (lane == src) cnt = c;
  }
  if (m >= p.n) return;
  p.durable[m] = (flags & kAnyDurable) ? 1 : 0;
  p.cnt[m] = cnt;
  // (a single list is the same thing with an empty short side)
  const bool flat = cls == 0 && !tiny && (!dedup || k == 1) && k >= 1 && k <= p.max_slots && mxlen <= (uint32_t)kFlatMaxLong;
  p.meta[m] = min(k, 255u) | ((uint32_t)cls << 8) | (dedup ? kMetaDedup : 0u) | (tiny ? kMetaTiny : 0u) |
              ((cls == 3 || cnt >= kBigIds) ? kMetaBig : 0u) | (flat ? kMetaFlat : 0u);
  uint2* lists = p.mlist + (size_t)m * kRouteSlots;
  if (cls == 2) {
    lists[0] = make_uint2((uint32_t)scr, (uint32_t)(scr >> 32));
  } else if (cls == 4) {
    lists[0] = make_uint2(__ldg(p.desc_start + m), 0u);  // where its descriptors start in the arena
  } else if (cnt > 0) {
    // the longest list goes first: k_write streams it and weaves the others in
    for (uint32_t j = 0; j < k; ++j) {
      const uint32_t src = j == 0 ? mxidx : (j == mxidx ? 0u : j);
      lists[j] = make_uint2(sm.loff[src][tid], sm.llen[src][tid]);
    }
  }
}

// Exclusive prefix sum of cnt[0..n) -> offsets[0..n], one pass: tiles of kScanTile counts, each tile adds up
// its own, publishes the aggregate and looks back over its predecessors' (tiles finish fast, so the look-back
// is short).  The last tile writes offsets[n], the total and the overflow status.
constexpr int kScanThreads = 256, kScanPer = 4, kScanTile = kScanThreads * kScanPer;
constexpr uint64_t kFlagAgg = 1ull << 62, kFlagIncl = 2ull << 62, kValMask = (1ull << 62) - 1;
static_assert(kScanTile == kRouteScanTile, "abi.cu sizes the tile state with kRouteScanTile");

__global__ void __launch_bounds__(kScanThreads) k_offsets(RouteParams p) {
  __shared__ uint64_t warp_sum[kScanThreads / 32];
  __shared__ uint64_t tile_base;
  __shared__ uint32_t tile_id;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  volatile uint64_t* tstate = p.tile_state;
  if (tid == 0) tile_id = atomicAdd(p.ctrl + kCtrlTile, 1u);  // tiles in launch order: predecessors always run
  __syncthreads();
  const uint32_t tile = tile_id;
  const uint32_t base = tile * kScanTile + tid * kScanPer;
  uint32_t v[kScanPer];
  uint64_t mine = 0;
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    v[i] = base + i < p.n ? p.cnt[base + i] : 0;
    mine += v[i];
  }
  uint64_t incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint64_t t = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sum[warp] = incl;
  __syncthreads();
  uint64_t warp_base = 0, tile_total = 0;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; ++w) {
    const uint64_t s = warp_sum[w];
    if (w < warp) warp_base += s;
    tile_total += s;
  }
  if (warp == 0) {
    if (lane == 0) {
      tstate[tile] = (tile == 0 ? kFlagIncl : kFlagAgg) | tile_total;
      __threadfence();
    }
    uint64_t excl = 0;
    for (int64_t j = (int64_t)tile - 1; j >= 0; j -= 32) {
      const int64_t idx = j - lane;
      uint64_t s;
      do {
        s = kFlagIncl;
        if (idx >= 0) s = tstate[idx];
      } while (__any_sync(kFull, (s >> 62) == 0));
      const unsigned im = __ballot_sync(kFull, (s >> 62) == 2);
      const int stop = im ? __ffs(im) - 1 : 31;
      uint64_t c = lane <= stop ? (s & kValMask) : 0;
#pragma unroll
      for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(kFull, c, d);
      excl += c;
      if (im) break;
    }
    if (lane == 0) {
      if (tile > 0) {
        tstate[tile] = kFlagIncl | (excl + tile_total);
        __threadfence();
      }
      tile_base = excl;
      if (tile == p.n_tiles - 1) {
        const uint64_t total = excl + tile_total;
        *p.total = total;
        p.offsets[p.n] = (uint32_t)min(total, (uint64_t)0xFFFFFFFFull);
        if (total > 0xFFFFFFFFull || total > p.ids_cap) atomicOr(p.ctrl + kCtrlStatus, kStatusOverflow);
      }
    }
  }
  __syncthreads();
  uint64_t pos = tile_base + warp_base + (incl - mine);
#pragma unroll
  for (int i = 0; i < kScanPer; ++i) {
    if (base + i < p.n) p.offsets[base + i] = (uint32_t)min(pos, (uint64_t)0xFFFFFFFFull);
    pos += v[i];
  }
}

// One message, by the whole warp: c ids to ids[o ...), described by its meta word mt and its row of lists.
// k_write has ONE call site for it (both passes share the loop) so that the kernel stays inside the
// instruction cache.
struct WriteCtx {  // the few fields of RouteParams it needs, by value (kernel parameters have no address)
  const uint2* desc_lists;
  const uint32_t* qlist;
  const uint2* mlist;
  const uint32_t* scratch;
  uint32_t* ids;
  uint64_t ids_cap;
  bool tma_ok;   // qlist is the table's own array (16-byte aligned, lists of 8+ ids start on 16-byte boundaries)
  KwayBuf* kway; // this warp's buffers for warp_kway (they share their shared memory with ws / wb: one path at a time)
  uint64_t* mbar; // this warp's mbarrier for the bulk copies of warp_weave
};
// (lo, ll) = this lane's entry of the message's list row (lane j < k: list j), fetched by the caller ahead of time.
__device__ __forceinline__ void write_message(const WriteCtx& p, WarpScratch& ws, WeaveBuf& wb, uint32_t& phase, uint32_t lo,
                                              uint32_t ll, uint32_t c, uint32_t mt, uint64_t o, int lane) {
  const int kk = mt & 0xFF, cl = (mt >> 8) & 0xF;
  CHQ_CLK_T0();
  if (cl == 2) {
    const uint64_t so = ((uint64_t)__shfl_sync(kFull, ll, 0) << 32) | __shfl_sync(kFull, lo, 0);
    for (uint32_t i = lane; i < c; i += 32) p.ids[o + i] = p.scratch[so + i];
    CHQ_CLK_ADD(3, 1, c);
  } else if (cl == 4) {
    warp_gather_small<true>(p.qlist, p.desc_lists + __shfl_sync(kFull, lo, 0), kk, (mt & kMetaDedup) != 0, ws, p.ids, o, lane);
    CHQ_CLK_ADD(3, 1, c);
  } else if (cl == 5) {
    if (lane >= kk) lo = ll = 0;
    warp_kway(p.qlist, lo, ll, kk, *p.kway, p.ids + o, lane);
  } else if (cl == 6) {
    warp_merge2(p.qlist, __shfl_sync(kFull, lo, 0), __shfl_sync(kFull, ll, 0), __shfl_sync(kFull, lo, 1), __shfl_sync(kFull, ll, 1),
                p.ids + o, lane);
  } else if (kk == 1) {
    const uint32_t* sp = p.qlist + __shfl_sync(kFull, lo, 0) + lane;
    uint32_t* dst = p.ids + o + lane;
    uint32_t i = lane;
    for (; i + 96 < c; i += 128, sp += 128, dst += 128) {
      const uint32_t v0 = __ldg(sp), v1 = __ldg(sp + 32), v2 = __ldg(sp + 64), v3 = __ldg(sp + 96);
      dst[0] = v0; dst[32] = v1; dst[64] = v2; dst[96] = v3;
    }
    for (; i < c; i += 32, sp += 32, dst += 32) *dst = __ldg(sp);
    CHQ_CLK_ADD(2, 1, c);
  } else {
    if (lane >= kk) lo = ll = 0;
    if (cl == 3) warp_union_long<true>(p.qlist, lo, ll, kk, (mt & kMetaDedup) != 0, ws, p.ids, o, lane);
    else warp_weave(p.qlist, lo, ll, kk, (mt & kMetaDedup) != 0, p.tma_ok, ws, wb, p.mbar, phase, p.ids + o, lane);
#ifdef CHQ_CLOCKS
    uint32_t sum = ll, mx = ll;
    for (int d = 16; d; d >>= 1) { sum += __shfl_xor_sync(kFull, sum, d); mx = max(mx, __shfl_xor_sync(kFull, mx, d)); }
    CHQ_CLK_ADD(cl == 3 ? 4 : (sum - mx <= 32 ? 5 : 6), 1, c);
#endif
  }
}

// k_write: the two kinds of message that make up almost all of a batch — tiny results (their own lane gathers, sorts
// and writes) and one long list plus a short side (flat.cuh).  Persistent warps take 32 consecutive messages per grab;
// the grab after next and the next chunk's words are fetched while the current chunk is written (a grab is an atomic
// plus three dependent loads: unhidden, that chain was a third of the kernel).  Everything else (unions of several
// long lists, de-duplicating merges, slow-path results parked in scratch, the headers exchange's many small lists)
// is put on a work list for k_write_rest, which has its own register budget.
__global__ void __launch_bounds__(kRouteThreads, CHQ_WRITE_CTAS) k_write(RouteParams p) {
  __shared__ FlatBuf fb_all[kRouteThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  FlatBuf& fb = fb_all[warp];
  if (*p.total > 0xFFFFFFFFull) return;  // offsets saturated: the launch reports CHQ_E_OVERFLOW
#ifdef CHQ_DEBUG_SWITCHES
  if (p.debug & 1u) return;
#endif
  // (through an asm move, so that the compiler keeps the pointer in one register pair and forms list addresses with
  // one multiply-add: left alone it re-derives them from the kernel parameters, five instructions per load)
  const uint32_t* ql;
  asm("mov.b64 %0, %1;" : "=l"(ql) : "l"(p.tv.qlist));
  // Work is handed out in units of eight messages: four units (one lane per message) per grab for most of the batch,
  // two and then one towards its end, where a whole 32-message chunk per warp would leave the other warps waiting for
  // the last ones (a chunk takes about a seventh of the kernel's run time).
  const uint32_t n_units = (p.n + 7) / 8;
  const uint32_t tail2 = n_units - n_units / 8, tail1 = n_units - n_units / 24;
  uint32_t u = 0, take = 4;
  auto grab = [&]() {
    uint32_t r = 0;
    if (lane == 0) r = atomicAdd(p.ctrl + kCtrlChunk, take);
    u = __shfl_sync(kFull, r, 0);
  };
  uint32_t meta, cnt, pos, m;
  auto fetch = [&]() {
    m = u * 8 + lane;
    meta = cnt = pos = 0;
    if ((uint32_t)lane < 8 * take && m < p.n) { meta = CHQ_LD_ONCE(p.meta + m); cnt = CHQ_LD_ONCE(p.cnt + m); pos = p.offsets[m]; }
  };
  for (grab(); u < n_units; grab()) {
    fetch();
    take = u >= tail1 ? 1u : (u >= tail2 ? 2u : 4u);   // size of the NEXT grab
    const bool fits = (uint64_t)pos + cnt <= p.ids_cap;  // else CHQ_E_OVERFLOW: nothing is written past ids_cap
    const bool live = cnt > 0 && fits;
    if (!(CHQ_EXP & 4) && live && (meta & kMetaTiny)) {
      const uint2* lists = p.mlist + (size_t)m * kRouteSlots;
      if (((meta >> 8) & 15u) == 4u) lists = p.desc_lists + __ldg(lists).x;   // headers exchange: the lists stayed in the descriptor arena
      auto off = [&](uint32_t j) { return __ldg(lists + j).x; };
      auto len = [&](uint32_t j) { return __ldg(lists + j).y; };
      tiny_union(ql, meta & 0xFF, off, len, p.ids + pos);
    }
    __syncwarp();
    {  // what this kernel does not do goes on k_write_rest's list: long results from the front, the others from the back
      const bool other = live && !(meta & (kMetaTiny | kMetaFlat));
      const unsigned om = __ballot_sync(kFull, other);
      if (om) {
        const bool big = (meta & kMetaBig) != 0;
        const unsigned bm = __ballot_sync(kFull, other && big), sm = om & ~bm;
        uint32_t b0 = 0, s0 = 0;
        if (lane == 0) {
          if (bm) b0 = atomicAdd(p.ctrl + kCtrlRestBig, (uint32_t)__popc(bm));
          if (sm) s0 = atomicAdd(p.ctrl + kCtrlRest, (uint32_t)__popc(sm));
        }
        b0 = __shfl_sync(kFull, b0, 0);
        s0 = __shfl_sync(kFull, s0, 0);
        const uint32_t ltm = (1u << lane) - 1u;
        if (other) p.rest[big ? b0 + __popc(bm & ltm) : p.n - 1u - (s0 + __popc(sm & ltm))] = m;
      }
    }
    // up to eight messages at a time: short ids placed by all lanes together, long lists streamed
    for (unsigned fm = (CHQ_EXP & 8) ? 0u : __ballot_sync(kFull, live && (meta & kMetaFlat)); fm;) {
      unsigned gm = fm;
      if (__popc(fm) > kFlatGroup) gm = fm & ((1u << __fns(fm, 0, kFlatGroup + 1)) - 1u);
      fm &= ~gm;
      flat_place(ql, p.mlist, p.ids, fb, gm, m, meta, cnt, pos, lane);
      const int ng = (CHQ_EXP & 1) ? 0 : __popc(gm);
#pragma unroll 1
      for (int gi = 0; gi < ng; ++gi)
        mask_stream(ql + fb.d[gi][0].x, p.ids + fb.pos[gi], fb.cnt[gi], fb.cm[gi], lane);
    }
  }
}

// The messages k_write left out, from its work list: those with long results first (so that the longest jobs start
// first instead of being met by chance near the end), eight messages per grab, one per warp trip.
__global__ void __launch_bounds__(kRouteThreads, 8) k_write_rest(RouteParams p) {
  // per warp: the scratch of the merge paths and the weave tile, or (never at the same time) the two buffers of warp_kway
  struct WarpMem { WarpScratch ws; WeaveBuf wb; };
  union __align__(16) WarpShared { WarpMem m; KwayBuf kb; };
  __shared__ WarpShared sh_all[kRouteThreads / 32];
  __shared__ uint64_t bar_all[kRouteThreads / 32];   // the bulk copies' mbarriers: outside the shared buffers
  const TableView& tv = p.tv;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_big = p.ctrl[kCtrlRestBig], n_all = n_big + p.ctrl[kCtrlRest];
  if (n_all == 0 || *p.total > 0xFFFFFFFFull) return;
#ifdef CHQ_DEBUG_SWITCHES
  if (p.debug & 1u) return;
#endif
  WarpScratch& ws = sh_all[warp].m.ws;
  WeaveBuf& wb = sh_all[warp].m.wb;
  uint32_t phase = 0;
  weave_bar_init(&bar_all[warp], lane);
  const WriteCtx wc{p.desc_lists, tv.qlist, p.mlist, p.scratch, p.ids, p.ids_cap,
                    !p.slice_offsets && (reinterpret_cast<uintptr_t>(tv.qlist) & 15u) == 0, &sh_all[warp].kb, &bar_all[warp]};
  // messages per grab: as many as keeps every warp of the grid busy (a short list handed out eight at a time would
  // leave most warps idle while a few work through their eight one after the other), at most eight
  const uint32_t n_warps = gridDim.x * (kRouteThreads / 32);
  const uint32_t kPerGrab = max(1u, min(8u, n_all / n_warps));
  for (;;) {
    CHQ_CLK_T0();
    uint32_t i0 = 0;
    if (lane == 0) i0 = atomicAdd(p.ctrl + kCtrlRestNext, kPerGrab);
    i0 = __shfl_sync(kFull, i0, 0);
    if (i0 >= n_all) break;
    // lane j < 8: item i0 + j of the list (long results sit at the front of p.rest, the others at its back)
    const uint32_t i = i0 + lane;
    uint32_t m = 0, cnt = 0, meta = 0, pos = 0;
    if ((uint32_t)lane < kPerGrab && i < n_all) {
      m = p.rest[i < n_big ? i : p.n - 1u - (i - n_big)];
      meta = p.meta[m]; cnt = p.cnt[m]; pos = p.offsets[m];
    }
    CHQ_CLK_ADD(7, 0, 0);
    // the list row of the next message is fetched while the current one is written
    unsigned todo = __ballot_sync(kFull, cnt > 0);
    auto row_of = [&](unsigned td) -> uint2 {
      if (td == 0) return make_uint2(0u, 0u);
      const uint32_t mm = __shfl_sync(kFull, m, __ffs(td) - 1);
      if (lane >= kRouteSlots) return make_uint2(0u, 0u);
      return __ldg(p.mlist + (size_t)mm * kRouteSlots + lane);
    };
    uint2 row = row_of(todo);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint2 mine = row;
      row = row_of(todo);
      write_message(wc, ws, wb, phase, mine.x, mine.y, __shfl_sync(kFull, cnt, src), __shfl_sync(kFull, meta, src),
                    __shfl_sync(kFull, pos, src), lane);
    }
  }
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
static unsigned long long g_launches = 0;
unsigned long long kernel_launches() { return g_launches; }
void note_launches(int n) { g_launches += (unsigned long long)n; }

cudaError_t launch_tokenize(const uint8_t* blob, const uint32_t* off, uint32_t n, const TableView& tv, int split,
                            uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_tokenize<<<(n + kTokThreads - 1) / kTokThreads, kTokThreads, 0, st>>>(blob, off, nullptr, n, tv, split, wid, nwords, err);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_tokenize_keys(const uint8_t* bytes, const uint32_t* key_off, const uint32_t* key_len, uint32_t n,
                                 const TableView& tv, int split, uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_tokenize<<<(n + kTokThreads - 1) / kTokThreads, kTokThreads, 0, st>>>(bytes, key_off, key_len, n, tv, split, wid, nwords, err);
  ++g_launches;
  return cudaGetLastError();
}


__global__ void k_finish(const uint32_t* ctrl, int32_t* status) {
  const uint32_t s = ctrl[kCtrlStatus];
  *status = (s & kStatusScratch) ? CHQ_E_SCRATCH : (s & kStatusOverflow) ? CHQ_E_OVERFLOW : CHQ_OK;
}

// cycle accounting (zeros unless built with -DCHQ_CLOCKS); reading resets the counters
cudaError_t read_clocks(unsigned long long* out32) {
  for (int i = 0; i < 32; ++i) out32[i] = 0;
#ifdef CHQ_CLOCKS
  cudaError_t e = cudaMemcpyFromSymbol(out32, g_clk, sizeof(unsigned long long) * 32);
  if (e != cudaSuccess) return e;
  static const unsigned long long zero[32] = {};
  return cudaMemcpyToSymbol(g_clk, zero, sizeof(zero));
#else
  return cudaSuccess;
#endif
}

cudaError_t launch_finish(const uint32_t* ctrl, int32_t* d_status, cudaStream_t st) {
  k_finish<<<1, 1, 0, st>>>(ctrl, d_status);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_route(const RouteParams& p, int n_sms, cudaStream_t st, cudaEvent_t* between) {
  if (p.n == 0) return cudaSuccess;
  static int occ_by_dev[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  static int occ_rest_by_dev[64] = {};
  int& occ = occ_by_dev[dev & 63];
  int& occ_rest = occ_rest_by_dev[dev & 63];
  if (occ == 0) {  // once per device: size the persistent grids of the write kernels
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_write, kRouteThreads, 0);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_rest, k_write_rest, kRouteThreads, 0);
    if (e != cudaSuccess) return e;
    if (occ_rest < 1) occ_rest = 1;
  }
  k_probe<<<(p.n + kRouteThreads - 1) / kRouteThreads, kRouteThreads, sizeof(RouteSmem), st>>>(p);
  if (between && (e = cudaEventRecord(between[0], st)) != cudaSuccess) return e;
  k_offsets<<<p.n_tiles, kScanThreads, 0, st>>>(p);
  if (between && (e = cudaEventRecord(between[1], st)) != cudaSuccess) return e;
  const uint64_t want = (uint64_t)n_sms * occ, chunks = (p.n + 31) / 32;
  k_write<<<(uint32_t)(chunks < want ? chunks : want), kRouteThreads, 0, st>>>(p);
  if (between && (e = cudaEventRecord(between[2], st)) != cudaSuccess) return e;
  const uint64_t want_rest = (uint64_t)n_sms * occ_rest;   // (its warps pull work from the list; usually it is short)
  if (!(CHQ_EXP & 16)) k_write_rest<<<(uint32_t)(chunks < want_rest ? chunks : want_rest), kRouteThreads, 0, st>>>(p);
  g_launches += 4;
  return cudaGetLastError();
}

}  // namespace chq
