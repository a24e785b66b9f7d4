// sm_100a kernels of the routing hot path.
//
//   k_tokenize : routing-key bytes -> interned word ids        (replaces String.split + the per-level
//                                                               String hashing of TrieMatcher.lookup,
//                                                               QueueMatcher.scala:403-413, 261-263)
//   k_route    : word ids x compiled table -> CSR of queue ids (replaces ilookup/bLookup,
//                                                               QueueMatcher.scala:523-585, and the
//                                                               set/dedup/flag work of
//                                                               ExchangeEntity.scala:297-302)
//
// k_route is ONE fused, persistent kernel per batch: probe -> exact per-message counts -> chained
// (decoupled look-back) prefix sum across tiles -> merged, sorted, de-duplicated id lists written
// straight to their final CSR position.  Inputs are read once and outputs written once; the table
// (hash buckets, groups, queue lists) is L2-resident.  Integer/byte work throughout — no tensor
// cores (there is no dense contraction on this path).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/chanamq_gpu.h"
#include "chq_common.h"
#include "kernels.h"
#include "merge.cuh"
#include "probe.cuh"

#ifndef CHQ_MIN_CTAS
#define CHQ_MIN_CTAS 8
#endif

namespace chq {

// One thread per key.  Java split semantics: words are cut at every '.', trailing
// empty words are dropped, a key without any '.' is one word (even when empty).
// Key i is blob[off[i] .. off[i+1]) (klen == null: keys packed back to back) or blob[off[i] .. off[i]+klen[i])
// (keys scattered inside a byte batch, as the frame decoder leaves them).
__global__ void __launch_bounds__(256) k_tokenize(const uint8_t* __restrict__ blob, const uint32_t* __restrict__ off,
                                                  const uint32_t* __restrict__ klen, uint32_t n, TableView tv, int split,
                                                  uint32_t* __restrict__ wid, uint8_t* __restrict__ nwords,
                                                  uint8_t* __restrict__ err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t b0 = off[i], b1 = klen ? b0 + klen[i] : off[i + 1];
  if (b1 - b0 > (uint32_t)kMaxKeyBytes) { nwords[i] = 0; err[i] = 2; return; }
  uint32_t* row = wid + (size_t)i * tv.wstride;
  uint32_t cnt = 0, last_nonempty = 0, wl = 0, wstart = b0;
  bool dot = false;
  uint64_t h = kFnvBasis, f8 = 0;
  for (uint32_t j = b0; j <= b1; ++j) {
    const bool end = (j == b1);
    const uint8_t c = end ? 0 : __ldg(blob + j);
    if (end || (split && c == '.')) {
      const uint32_t id = dict_lookup(tv, h, f8, wl, blob + wstart);
      if (cnt < tv.wstride) row[cnt] = id;
      ++cnt;
      if (wl > 0) last_nonempty = cnt;
      if (!end) dot = true;
      h = kFnvBasis; f8 = 0; wl = 0; wstart = j + 1;
    } else {
      h = fnv_step(h, c);
      if (wl < 8) f8 |= (uint64_t)c << (8 * wl);
      ++wl;
    }
  }
  const uint32_t nw = dot ? last_nonempty : 1;
  nwords[i] = (uint8_t)nw;
  err[i] = nw == 0 ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// the fused route kernel
// ------------------------------------------------------------------------------------------
namespace {

constexpr uint64_t kFlagAgg = 1ull << 62, kFlagIncl = 2ull << 62, kValMask = (1ull << 62) - 1;
constexpr int kWarps = kRouteThreads / 32;

struct RouteSmem {
  // per message: matched lists (offset, length), slot-major.  While a key is being probed the same slots hold
  // its unresolved tag hits (slot index, shape << 16 | literal mask): hit c is read before result c is written.
  uint32_t loff[kRouteSlots][kRouteThreads];
  uint32_t llen[kRouteSlots][kRouteThreads];
  uint64_t pos[kRouteThreads];                // final CSR position of the message's ids
  uint32_t cnt[kRouteThreads];
  uint16_t kcls[kRouteThreads];               // k | class << 8 | dedup << 12
  uint16_t queue[kRouteThreads];              // routed messages of this tile, any order
  WarpScratch ws[kWarps];
  uint64_t warp_sum[kWarps];
  uint64_t tile_base;
  uint32_t tile, q_n, q_next;
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

__global__ void __launch_bounds__(kRouteThreads, CHQ_MIN_CTAS) k_route(RouteParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RouteSmem& sm = *reinterpret_cast<RouteSmem*>(smem_raw);
  const TableView& tv = p.tv;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  volatile uint64_t* tstate = p.tile_state;

  for (;;) {
    __syncthreads();  // previous tile fully done with sm
    if (tid == 0) {
      sm.tile = atomicAdd(p.ctrl + kCtrlTile, 1u);
      sm.q_n = 0;
      sm.q_next = 0;
    }
    __syncthreads();
    const uint32_t tile = sm.tile;
    if (tile >= p.n_tiles) return;
    const uint32_t m = tile * kRouteThreads + tid;

    // ---- phase A: one thread per message: probe every compatible shape -------------------
    uint32_t k = 0, ub = 0, mxlen = 0, flags = 0, cnt = 0, nw = 0;
    const uint32_t* row = p.wid + (size_t)m * tv.wstride;
    if (m < p.n && !p.slice_offsets && !p.desc_k) nw = p.nwords[m];
    bool overflow = false;  // more tag hits than candidate slots: the slow path re-probes
    if (p.slice_offsets || p.desc_k) nw = m < p.n ? 1u : 0u;
    if (nw > 0 && !(p.debug & 2u)) {
      auto record = [&](const GroupRec& rec) {
        if (k < p.max_slots) { sm.loff[k][tid] = rec.list_off; sm.llen[k][tid] = rec.list_len; }
        ++k;
        ub += rec.list_len;
        mxlen = max(mxlen, rec.list_len);
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
            rec.flags = kShared | (__ldg(p.slice_durable + (size_t)r * p.n + m) ? kAnyDurable : 0);
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
    // classify: 0 = count known, 1 = needs the de-duplicating merge to count, 2 = slow path
    int cls = 0;
    const bool dedup = (flags & kShared) != 0;
    // tiny: the whole result fits the 16-input sorting network of this thread (merge.cuh)
    const bool tiny = !overflow && k >= 1 && k <= p.max_slots && ub <= (uint32_t)kTinyIds;
    auto my_off = [&](uint32_t j) { return sm.loff[j][tid]; };
    auto my_len = [&](uint32_t j) { return sm.llen[j][tid]; };
    if (tiny) cnt = (k == 1 || !dedup) ? ub : tiny_union(tv.qlist, k, my_off, my_len, nullptr);
    else if (overflow) cls = 2;
    else if (k <= 1) cnt = ub;
    else if (k > p.max_slots || ub - mxlen > p.max_small) cls = 2;
    else if (dedup) cls = 1;
    else cnt = ub;

    __syncwarp();
    for (unsigned todo = __ballot_sync(kFull, cls == 2); todo; todo &= todo - 1) {
      const int src = __ffs(todo) - 1;
      const uint32_t mm = tile * kRouteThreads + warp * 32 + src;
      uint64_t so;
      uint32_t fl;
      uint32_t c;
      if (p.desc_k) {  // more lists than shared-memory slots: read the descriptors again
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
      if (lane == src) {
        cnt = c;
        flags = fl;
        sm.loff[0][tid] = (uint32_t)so;
        sm.llen[0][tid] = (uint32_t)(so >> 32);
      }
    }
    for (unsigned todo = __ballot_sync(kFull, cls == 1); todo; todo &= todo - 1) {
      const int src = __ffs(todo) - 1;
      const int kk = __shfl_sync(kFull, k, src);
      uint32_t o = 0, l = 0;
      if (lane < kk) { o = sm.loff[lane][warp * 32 + src]; l = sm.llen[lane][warp * 32 + src]; }
      const uint32_t c = warp_union<false>(tv.qlist, o, l, kk, true, sm.ws[warp], nullptr, 0, 0, lane);
      if (lane == src) cnt = c;
    }
    if (m < p.n) p.durable[m] = (flags & kAnyDurable) ? 1 : 0;

    // ---- tile prefix sum + chained scan across tiles ---------------------------------------
    uint64_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t v = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += v;
    }
    if (lane == 31) sm.warp_sum[warp] = incl;
    // queue of routed messages for phase B (any order)
    {
      const bool queued = cnt > 0 && !tiny;  // tiny messages are written by their own thread
      const unsigned routed = __ballot_sync(kFull, queued);
      uint32_t qb = 0;
      if (lane == 0 && routed) qb = atomicAdd(&sm.q_n, (uint32_t)__popc(routed));
      qb = __shfl_sync(kFull, qb, 0);
      if (queued) sm.queue[qb + __popc(routed & ((1u << lane) - 1))] = (uint16_t)tid;
    }
    sm.cnt[tid] = cnt;
    sm.kcls[tid] = (uint16_t)(min(k, 255u) | (cls << 8) | (dedup ? 1u << 12 : 0u));
    __syncthreads();
    uint64_t warp_base = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const uint64_t v = sm.warp_sum[w];
      if (w < warp) warp_base += v;
      tile_total += v;
    }
    if (warp == 0) {
      if (lane == 0) {
        tstate[tile] = (tile == 0 ? kFlagIncl : kFlagAgg) | tile_total;
        __threadfence();
      }
      uint64_t excl = 0;
      for (int64_t j = (int64_t)tile - 1; j >= 0; j -= 32) {
        const int64_t idx = j - lane;
        uint64_t v;
        do {
          v = kFlagIncl;
          if (idx >= 0) v = tstate[idx];
        } while (__any_sync(kFull, (v >> 62) == 0));
        const unsigned im = __ballot_sync(kFull, (v >> 62) == 2);
        const int stop = im ? __ffs(im) - 1 : 31;
        uint64_t c = lane <= stop ? (v & kValMask) : 0;
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
        sm.tile_base = excl;
        if (tile == p.n_tiles - 1) {
          const uint64_t total = excl + tile_total;
          *p.total = total;
          p.offsets[p.n] = (uint32_t)min(total, (uint64_t)0xFFFFFFFFull);
          if (total > 0xFFFFFFFFull || total > p.ids_cap) atomicOr(p.ctrl + kCtrlStatus, kStatusOverflow);
        }
      }
    }
    __syncthreads();
    {
      const uint64_t pos = sm.tile_base + warp_base + (incl - cnt);
      sm.pos[tid] = pos;
      if (m < p.n) p.offsets[m] = (uint32_t)min(pos, (uint64_t)0xFFFFFFFFull);
    }
    __syncthreads();

    // ---- phase B: tiny messages by their own thread, the rest by warps off the tile's queue ----
    if (tiny && cnt > 0 && !(p.debug & 1u)) {
      const uint64_t o = sm.pos[tid];
      if (o + cnt <= p.ids_cap) tiny_union(tv.qlist, k, my_off, my_len, p.ids + o);
    }
    const uint32_t q_n = (p.debug & 1u) ? 0u : sm.q_n;
    for (;;) {
      uint32_t qi = 0;
      if (lane == 0) qi = atomicAdd(&sm.q_next, 1u);
      qi = __shfl_sync(kFull, qi, 0);
      if (qi >= q_n) break;
      const int t = sm.queue[qi];
      const uint32_t c = sm.cnt[t];
      const uint32_t kc = sm.kcls[t];
      const int kk = kc & 0xFF, cl = (kc >> 8) & 0xF;
      const uint64_t o = sm.pos[t];
      if (o + c > p.ids_cap) continue;  // does not fit: the launch reports CHQ_E_OVERFLOW; nothing is written past ids_cap
      if (cl == 2) {
        const uint64_t so = ((uint64_t)sm.llen[0][t] << 32) | sm.loff[0][t];
        for (uint32_t i = lane; i < c; i += 32) p.ids[o + i] = p.scratch[so + i];
      } else if (kk == 1) {
        const uint32_t* src = tv.qlist + sm.loff[0][t] + lane;
        uint32_t* dst = p.ids + o + lane;
        uint32_t i = lane;
        for (; i + 96 < c; i += 128, src += 128, dst += 128) {
          const uint32_t v0 = __ldg(src), v1 = __ldg(src + 32), v2 = __ldg(src + 64), v3 = __ldg(src + 96);
          dst[0] = v0; dst[32] = v1; dst[64] = v2; dst[96] = v3;
        }
        for (; i < c; i += 32, src += 32, dst += 32) *dst = __ldg(src);
      } else {
        uint32_t lo = 0, ll = 0;
        if (lane < kk) { lo = sm.loff[lane][t]; ll = sm.llen[lane][t]; }
        warp_union<true>(tv.qlist, lo, ll, kk, (kc >> 12) & 1, sm.ws[warp], p.ids, o, p.ids_cap, lane);
      }
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
  k_tokenize<<<(n + 255) / 256, 256, 0, st>>>(blob, off, nullptr, n, tv, split, wid, nwords, err);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_tokenize_keys(const uint8_t* bytes, const uint32_t* key_off, const uint32_t* key_len, uint32_t n,
                                 const TableView& tv, int split, uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_tokenize<<<(n + 255) / 256, 256, 0, st>>>(bytes, key_off, key_len, n, tv, split, wid, nwords, err);
  ++g_launches;
  return cudaGetLastError();
}

__global__ void k_finish(const uint32_t* ctrl, int32_t* status) {
  const uint32_t s = ctrl[kCtrlStatus];
  *status = (s & kStatusScratch) ? CHQ_E_SCRATCH : (s & kStatusOverflow) ? CHQ_E_OVERFLOW : CHQ_OK;
}

cudaError_t launch_finish(const uint32_t* ctrl, int32_t* d_status, cudaStream_t st) {
  k_finish<<<1, 1, 0, st>>>(ctrl, d_status);
  ++g_launches;
  return cudaGetLastError();
}

cudaError_t launch_route(const RouteParams& p, int n_sms, cudaStream_t st) {
  if (p.n_tiles == 0) return cudaSuccess;
  static int occ_by_dev[64] = {};
  const int smem = (int)sizeof(RouteSmem);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  int& occ = occ_by_dev[dev & 63];
  if (occ == 0) {  // once per device: opt in to the shared-memory size, size the persistent grid
    e = cudaFuncSetAttribute(k_route, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_route, kRouteThreads, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
  }
  const uint64_t want = (uint64_t)n_sms * occ;
  const uint32_t grid = (uint32_t)(p.n_tiles < want ? p.n_tiles : want);
  k_route<<<grid, kRouteThreads, smem, st>>>(p);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace chq
