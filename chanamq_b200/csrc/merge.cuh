// Warp-level set union of sorted queue-id lists (included by kernels.cu only).
//
// A routed message matched k >= 1 groups; each group owns a sorted, duplicate-free list of queue
// ids in the (L2-resident) table.  The result the exchange needs is the sorted union
// (ExchangeEntity.scala:300 collapses subscribers to a SET of queue names).  Typical shape: one
// long list ("a.*.*" bound by hundreds of queues) and a few short ones.  So: the short lists are
// staged in shared memory and rank-merged there; the long list is streamed from L2 in chunks of
// 32 and the short ids are woven in with shuffle-based mutual ranks — no per-element serial work,
// no atomics, coalesced stores.
#pragma once
#include "kernels.h"
#include "probe.cuh"

// Queue ids are written once and not read again by this launch group: streaming stores (evict-first), so that the
// 0.4 GB a batch writes passes through L2 without pushing the compiled table out of it.
#ifndef CHQ_STREAM_STORES
#define CHQ_STREAM_STORES 1
#endif
#if CHQ_STREAM_STORES
#define CHQ_ST_OUT(ptr, val) __stcs((ptr), (val))
#else
#define CHQ_ST_OUT(ptr, val) (*(ptr) = (val))
#endif
#ifndef CHQ_STREAM_LOADS
#define CHQ_STREAM_LOADS 0
#endif
#if CHQ_STREAM_LOADS   // per-message words k_probe left for k_write: read once
#define CHQ_LD_ONCE(ptr) __ldcs(ptr)
#else
#define CHQ_LD_ONCE(ptr) (*(ptr))
#endif

namespace chq {
namespace {

struct WarpScratch {
  uint32_t a[kRouteSmall];   // staged short lists, back to back
  uint32_t b[kRouteSmall];   // their sorted union
  uint32_t st[kRouteSlots];  // start of list j inside a[]
  uint32_t ln[kRouteSlots];  // its length (0 for the long list)
  uint32_t win[kRouteSlots][32];  // warp_union_long: the ids each short list currently shows
};

__device__ __forceinline__ uint32_t smem_lower_bound(const uint32_t* a, uint32_t n, uint32_t x) {
  uint32_t lo = 0;
  while (n) {
    const uint32_t half = n >> 1;
    if (a[lo + half] < x) { lo += half + 1; n -= half + 1; }
    else n = half;
  }
  return lo;
}
__device__ __forceinline__ uint32_t smem_upper_bound(const uint32_t* a, uint32_t n, uint32_t x) {
  uint32_t lo = 0;
  while (n) {
    const uint32_t half = n >> 1;
    if (a[lo + half] <= x) { lo += half + 1; n -= half + 1; }
    else n = half;
  }
  return lo;
}

__device__ void warp_bitonic_sort(uint32_t* a, uint32_t np, int lane) {
  for (uint32_t size = 2; size <= np; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t t = lane; t < (np >> 1); t += 32) {
        const uint32_t i = ((t / stride) * (stride << 1)) + (t % stride), j = i + stride;
        const bool up = (i & size) == 0;
        const uint32_t x = a[i], y = a[j];
        if ((x > y) == up) { a[i] = y; a[j] = x; }
      }
      __syncwarp();
    }
  }
}

// In-place removal of adjacent duplicates of a sorted array by one warp; returns the new length.
__device__ uint32_t warp_unique(uint32_t* a, uint32_t n, int lane) {
  uint32_t kept = 0, carry = 0;  // carry = the last entry of the previous 32 as it was before compaction
  for (uint32_t base = 0; base < n; base += 32) {
    const uint32_t i = base + lane;
    const uint32_t v = i < n ? a[i] : 0;
    uint32_t prev = __shfl_up_sync(kFull, v, 1);
    if (lane == 0) prev = carry;
    const bool keep = i < n && (i == 0 || prev != v);
    carry = __shfl_sync(kFull, v, 31);
    __syncwarp();
    const unsigned m = __ballot_sync(kFull, keep);
    if (keep) a[kept + __popc(m & ((1u << lane) - 1))] = v;
    kept += __popc(m);
    __syncwarp();
  }
  return kept;
}

// number of the 32 lane values `v` (ascending across lanes, padded with kNone) that are < x: 0..32
__device__ __forceinline__ uint32_t lanes_below(uint32_t v, uint32_t x) {
  uint32_t q = 0;
#pragma unroll
  for (int step = 16; step; step >>= 1) {
    const uint32_t probe = __shfl_sync(kFull, v, q + step - 1);
    if (probe < x) q += step;
  }
  const uint32_t last = __shfl_sync(kFull, v, q);
  return q + (last < x ? 1u : 0u);
}

// ---- tiny messages: at most kTinyIds ids over ALL matched lists --------------------------------------
// A third of the routed messages carry a handful of ids; giving each a whole warp costs hundreds of
// instructions of set-up for a few stores.  Instead the message's own probing thread gathers the ids,
// sorts them with a fixed 16-input bitonic network (static register indices, so all 32 lanes of the
// warp run it in lock step whatever their data) and writes the distinct ones.
constexpr int kTinyIds = 16;

__device__ __forceinline__ void sort16(uint32_t (&v)[kTinyIds]) {
#pragma unroll
  for (int size = 2; size <= kTinyIds; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
      for (int i = 0; i < kTinyIds; ++i) {
        const int j = i ^ stride;
        if (j > i) {
          const bool up = (i & size) == 0;
          const uint32_t a = v[i], b = v[j];
          const uint32_t lo = min(a, b), hi = max(a, b);
          v[i] = up ? lo : hi;
          v[j] = up ? hi : lo;
        }
      }
    }
  }
}

// Gathers list j = qlist[off(j) .. off(j)+len(j)) for j < k (at most kTinyIds ids in total), sorts, and
// returns the number of distinct ids; with out != null they are written to out[0 ...).
template <typename Off, typename Len>
__device__ __forceinline__ uint32_t tiny_union(const uint32_t* __restrict__ qlist, uint32_t k, Off off, Len len,
                                               uint32_t* __restrict__ out) {
  uint32_t tmp[kTinyIds];
#pragma unroll
  for (int s = 0; s < kTinyIds; ++s) tmp[s] = kNone;
  uint32_t e = 0;
  for (uint32_t j = 0; j < k; ++j) {
    const uint32_t o = off(j), l = len(j);
    for (uint32_t i = 0; i < l && e < (uint32_t)kTinyIds; ++i) tmp[e++] = __ldg(qlist + o + i);
  }
  uint32_t v[kTinyIds];
#pragma unroll
  for (int s = 0; s < kTinyIds; ++s) v[s] = tmp[s];
  sort16(v);
  uint32_t n = 0;
#pragma unroll
  for (int s = 0; s < kTinyIds; ++s) {
    const bool keep = v[s] != kNone && (s == 0 || v[s] != v[s - 1]);
    if (keep) {
      if (out) CHQ_ST_OUT(out + n, v[s]);
      ++n;
    }
  }
  return n;
}

// Fast path of warp_union: the short lists hold at most 32 ids together (the usual case: one long
// list plus a few one- or two-id lists).  The short ids live in one register per lane, are sorted
// with shuffles, and are woven into the streamed long list one id at a time.
//   short_len / short_st: per lane j < k, length of list j (0 for the long one) and its start
//   among the short ids.
template <bool WRITE>
__device__ uint32_t warp_union_small(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t short_len,
                                     uint32_t short_st, int k, bool dedup, uint32_t big_off, uint32_t nb, uint32_t ns,
                                     uint32_t* __restrict__ out, uint64_t pos, uint64_t cap, int lane) {
  uint32_t s = kNone;
  const unsigned short_lanes = __ballot_sync(kFull, short_len > 0);
  const int n_short = __popc(short_lanes);
  if (n_short == 2) {
    // two sorted lists: one ascending from lane 0, the other descending from lane 31 -> bitonic merge
    const int j1 = __ffs(short_lanes) - 1, j2 = 31 - __clz(short_lanes);
    const uint32_t o1 = __shfl_sync(kFull, my_off, j1), l1 = __shfl_sync(kFull, short_len, j1);
    const uint32_t o2 = __shfl_sync(kFull, my_off, j2), l2 = __shfl_sync(kFull, short_len, j2);
    if ((uint32_t)lane < l1) s = __ldg(qlist + o1 + lane);
    else if ((uint32_t)lane >= 32 - l2) s = __ldg(qlist + o2 + (31 - lane));
#pragma unroll
    for (int d = 16; d; d >>= 1) {
      const uint32_t other = __shfl_xor_sync(kFull, s, d);
      s = (lane & d) == 0 ? min(s, other) : max(s, other);
    }
  } else {
    for (unsigned todo = short_lanes; todo; todo &= todo - 1) {
      const int j = __ffs(todo) - 1;
      const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, short_len, j);
      const uint32_t st = __shfl_sync(kFull, short_st, j);
      if ((uint32_t)lane >= st && (uint32_t)lane < st + l) s = __ldg(qlist + o + (lane - st));
    }
    if (n_short > 2) {
#pragma unroll
      for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int d = size >> 1; d; d >>= 1) {
          const uint32_t other = __shfl_xor_sync(kFull, s, d);
          const bool up = (lane & size) == 0, lower = (lane & d) == 0;
          s = (lower == up) ? min(s, other) : max(s, other);
        }
      }
    }
  }
  // an id present in two short lists is emitted once
  const uint32_t prev = __shfl_up_sync(kFull, s, 1);
  const bool sdup = dedup && lane > 0 && s == prev;
  unsigned todo = __ballot_sync(kFull, (uint32_t)lane < ns && !sdup);  // short ids not yet placed, bit j = lane j
  uint32_t s_next = todo ? __shfl_sync(kFull, s, __ffs(todo) - 1) : kNone;
  uint32_t kept_total = 0;  // short ids placed so far
  uint32_t my_pos = kNone;  // where this lane's short id goes, relative to pos (kNone: dropped)
  (void)cap;

  // Weave one chunk of 32 long ids (lane value b, kNone past the end): every short id that is not
  // above the chunk is ranked with two ballots; returns how far this lane's long id moves up.
  auto weave = [&](uint32_t b, uint32_t base) -> uint32_t {
    const uint32_t bmax = __shfl_sync(kFull, b, 31);
    uint32_t shift = kept_total;
    if (todo && s_next <= bmax) {
      do {
        const int j = __ffs(todo) - 1;
        const unsigned lt = __ballot_sync(kFull, b < s_next);
        const unsigned eq = __ballot_sync(kFull, b == s_next);
        todo &= todo - 1;
        if (eq == 0) {
          const uint32_t p = __popc(lt);
          if (lane == j) my_pos = base + p + kept_total;
          if ((uint32_t)lane >= p) ++shift;
          ++kept_total;
        }
        s_next = todo ? __shfl_sync(kFull, s, __ffs(todo) - 1) : kNone;
      } while (todo && s_next <= bmax);
    }
    return shift;
  };

  // Stream the long list: full chunks without lane predicates, then the partial one.  In the partial
  // chunk the padding lanes hold kNone, so every short id still waiting is placed there, after the list.
  const uint32_t nfull = nb >> 5, rem = nb & 31;
  const uint32_t* src = qlist + big_off + lane;
  uint32_t* dst = WRITE ? out + pos + lane : nullptr;
  uint32_t base = 0;
  uint32_t c = 0;
  // four chunks per round trip to L2: the loads of a round are issued together
  for (; c + 4 <= nfull; c += 4) {
    uint32_t b4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) b4[u] = __ldg(src + 32 * u);
    src += 128;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t shift = weave(b4[u], base);
      if (WRITE) { dst[shift] = b4[u]; dst += 32; }
      base += 32;
    }
  }
  if (c < nfull) {  // one to three full chunks left
    uint32_t b4[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) b4[u] = (c + u < nfull) ? __ldg(src + 32 * u) : 0;
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      if (c + u < nfull) {
        const uint32_t shift = weave(b4[u], base);
        if (WRITE) { dst[shift] = b4[u]; dst += 32; }
        base += 32;
        src += 32;
      }
    }
  }
  if (rem) {
    const uint32_t b = (uint32_t)lane < rem ? __ldg(src) : kNone;
    const uint32_t shift = weave(b, base);
    if (WRITE && (uint32_t)lane < rem) dst[shift] = b;
  }
  // short ids above a list that ended on a chunk boundary
  if ((todo >> lane) & 1u) my_pos = nb + kept_total + __popc(todo & ((1u << lane) - 1));
  if (WRITE && my_pos != kNone) out[pos + my_pos] = s;
  return nb + kept_total + __popc(todo);
}

// Lane j < k holds list j = qlist[my_off .. my_off + my_len).  Precondition: the lists other than
// the longest hold at most kRouteSmall ids together.  `dedup` = some id may sit in two lists.
// Returns the size of the union; with WRITE the union goes to out[pos ...), never past cap.
template <bool WRITE>
__device__ uint32_t warp_union(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t my_len, int k, bool dedup,
                               WarpScratch& ws, uint32_t* __restrict__ out, uint64_t pos, uint64_t cap, int lane) {
  // the longest list is streamed; everything else is "short"
  const uint32_t len_k = lane < k ? my_len : 0;
  uint32_t nb = len_k;
#pragma unroll
  for (int d = 16; d; d >>= 1) nb = max(nb, __shfl_xor_sync(kFull, nb, d));
  const int big = __ffs(__ballot_sync(kFull, lane < k && len_k == nb)) - 1;
  const uint32_t big_off = __shfl_sync(kFull, my_off, big);
  const uint32_t short_len = (lane < k && lane != big) ? my_len : 0;
  uint32_t incl = short_len;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t v = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += v;
  }
  uint32_t ns = __shfl_sync(kFull, incl, 31);
  if (ns <= 32)  // the common case: a handful of short ids — keep them in registers
    return warp_union_small<WRITE>(qlist, my_off, short_len, incl - short_len, k, dedup, big_off, nb, ns, out, pos, cap, lane);
  if (lane < k) { ws.st[lane] = incl - short_len; ws.ln[lane] = short_len; }
  for (int j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, short_len, j);
    const uint32_t st = __shfl_sync(kFull, incl - short_len, j);
    for (uint32_t i = lane; i < l; i += 32) ws.a[st + i] = __ldg(qlist + o + i);
  }
  __syncwarp();
  const uint32_t* S = ws.a;
  if (k > 2) {
    // rank-merge: every id finds its place among the other short lists by binary search
    for (uint32_t e = lane; e < ns; e += 32) {
      int j = 0;
      while (e >= ws.st[j] + ws.ln[j]) ++j;
      const uint32_t x = ws.a[e];
      uint32_t r = e - ws.st[j];
      for (int j2 = 0; j2 < k; ++j2) {
        const uint32_t l2 = ws.ln[j2];
        if (j2 == j || l2 == 0) continue;
        r += j2 < j ? smem_upper_bound(ws.a + ws.st[j2], l2, x) : smem_lower_bound(ws.a + ws.st[j2], l2, x);
      }
      ws.b[r] = x;
    }
    __syncwarp();
    if (dedup) ns = warp_unique(ws.b, ns, lane);
    S = ws.b;
  }
  // weave the short ids into the long list, 32 x 32 at a time
  uint32_t t = 0, kept_total = 0;
  uint32_t b_next = lane < nb ? __ldg(qlist + big_off + lane) : kNone;
  for (uint32_t c0 = 0; c0 < nb; c0 += 32) {
    const uint32_t i = c0 + lane;
    const bool valid = i < nb;
    const uint32_t b = b_next;
    b_next = (i + 32 < nb) ? __ldg(qlist + big_off + i + 32) : kNone;
    const uint32_t bmax = __shfl_sync(kFull, b, min(31u, nb - c0 - 1));
    const uint32_t kept0 = kept_total;
    uint32_t shift = 0;
    while (t < ns) {
      const uint32_t s = (t + lane < ns) ? S[t + lane] : kNone;
      const bool inr = s <= bmax;  // in-range lanes form a prefix (S ascending; ids < kNone)
      const unsigned m_in = __ballot_sync(kFull, inr);
      if (m_in == 0) break;
      // p = how many ids of this chunk are < s  (s <= bmax, so p <= 31)
      uint32_t p = 0;
#pragma unroll
      for (int step = 16; step; step >>= 1) {
        const uint32_t probe = __shfl_sync(kFull, b, p + step - 1);
        if (probe < s) p += step;
      }
      const bool dup = __shfl_sync(kFull, b, p) == s;
      const bool keep = inr && !dup;
      const unsigned km = __ballot_sync(kFull, keep);
      if (WRITE && keep) {
        const uint64_t o = pos + c0 + p + kept_total + __popc(km & ((1u << lane) - 1));
        if (o < cap) out[o] = s;
      }
      // every long id moves up by the kept short ids below it
      const uint32_t q = lanes_below(inr ? s : kNone, b);
      shift += __popc(km & (q >= 32 ? kFull : ((1u << q) - 1)));
      kept_total += __popc(km);
      const uint32_t n_in = __popc(m_in);
      t += n_in;
      if (n_in < 32) break;
    }
    if (WRITE && valid) {
      const uint64_t o = pos + i + kept0 + shift;
      if (o < cap) out[o] = b;
    }
  }
  // short ids above the whole long list
  if (WRITE) {
    for (uint32_t i = t + lane; i < ns; i += 32) {
      const uint64_t o = pos + nb + kept_total + (i - t);
      if (o < cap) out[o] = S[i];
    }
  }
  __syncwarp();
  return nb + kept_total + (ns - t);
}

// Sorted union of k <= kRouteSlots sorted lists of ANY length in which no id occurs twice (lane j < k holds
// list j = qlist[my_off .. my_off + my_len)), written to out[pos ...).  Used when the lists other than the
// longest are too many ids for the staged paths above.
//
// The longest list is streamed and woven exactly as in warp_union.  The ids to weave in — the union of the
// OTHER lists — are produced a block at a time by a multi-way step: each of those lists shows its next 32
// ids (one per lane); ids up to T = the smallest last-shown id among the lists that still have 32 or more
// cannot be preceded by anything unseen, so they are final: each goes to block[lane + how many shown ids of
// the other lists are smaller] (shuffle binary searches).  The list that sets T empties its whole window, so
// a step yields at least 32 ids (or everything that is left) and at most 32 per list.
static_assert(2 * kRouteSmall >= 32 * (kRouteSlots - 1), "a block of short ids lives in WarpScratch::a..b");

// With `dedup` an id may sit in several lists: equal ids of different short lists always show up in the same
// step (both are <= T and inside their windows), get neighbouring places in the block (ties go by list index)
// and are collapsed there; a short id equal to a long one is dropped by the weave.  Returns the size of the union.
template <bool WRITE>
__device__ uint32_t warp_union_long(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t my_len, int k, bool dedup,
                                    WarpScratch& ws, uint32_t* __restrict__ out, uint64_t pos, int lane) {
  const uint32_t len_k = lane < k ? my_len : 0;
  uint32_t nb = len_k;
#pragma unroll
  for (int d = 16; d; d >>= 1) nb = max(nb, __shfl_xor_sync(kFull, nb, d));
  const int big = __ffs(__ballot_sync(kFull, lane < k && len_k == nb)) - 1;
  const uint32_t big_off = __shfl_sync(kFull, my_off, big);
  const uint32_t len_s = (lane < k && lane != big) ? my_len : 0;  // lane j: length of short list j
  uint32_t cur = 0;                                               // lane j: ids of it already handed out
  uint32_t* S = ws.a;                                             // the current block, ascending
  uint32_t blk_n = 0, blk_t = 0;
  bool more = __ballot_sync(kFull, len_s != 0) != 0;

  uint32_t kept_total = 0;
  uint32_t b_next = lane < nb ? __ldg(qlist + big_off + lane) : kNone;
  // chunks of the long list, then one more trip (c0 == nb, no long ids, every short id in range) for what is left
#pragma unroll 1
  for (uint32_t c0 = 0;; c0 = min(c0 + 32, nb)) {
    const bool tail = c0 >= nb;
    const uint32_t i = c0 + lane;
    const bool valid = i < nb;
    const uint32_t b = b_next;
    b_next = (i + 32 < nb) ? __ldg(qlist + big_off + i + 32) : kNone;
    const uint32_t bmax = tail ? kNone - 1 : __shfl_sync(kFull, b, min(31u, nb - c0 - 1));
    const uint32_t kept0 = kept_total;
    uint32_t shift = 0;
#pragma unroll 1
    for (;;) {
      if (blk_t == blk_n) {
        if (!more) break;
        // ---- next block of short ids: one multi-way step (see above) ----
        const uint32_t rem = len_s - cur;
        uint32_t w[kRouteSlots];
#pragma unroll
        for (int j = 0; j < kRouteSlots; ++j) {  // all the loads first
          const uint32_t r = __shfl_sync(kFull, rem, j), o = __shfl_sync(kFull, my_off + cur, j);
          w[j] = (uint32_t)lane < r ? __ldg(qlist + o + lane) : kNone;
        }
        uint32_t T = kNone;
        __syncwarp();  // everyone is done with the previous block
#pragma unroll
        for (int j = 0; j < kRouteSlots; ++j) {
          ws.win[j][lane] = w[j];
          const uint32_t last = __shfl_sync(kFull, w[j], 31);
          if (last != kNone) T = min(T, last);   // that list still has 32 or more
        }
        __syncwarp();
        uint32_t emitted = 0, mine = 0;
#pragma unroll 1
        for (int j = 0; j < k; ++j) {
          const uint32_t x = ws.win[j][lane];
          const bool go = x != kNone && x <= T;
          const uint32_t c = __popc(__ballot_sync(kFull, go));  // a prefix of the window
          if (c == 0) continue;
          uint32_t rank = lane;
#pragma unroll 1
          for (int j2 = 0; j2 < k; ++j2) {
            const uint32_t* W = ws.win[j2];
            if (j2 == j || W[0] == kNone) continue;
            // shown ids of list j2 below x (or equal to it, for the lists before j: equal ids keep list order);
            // at most 31 below (a full window ends at or above T >= x, a partial one in kNone), 32 only with a tie
            const uint32_t xx = (dedup && j2 < j) ? x + 1 : x;  // ids < kNone, so x + 1 does not wrap
            uint32_t q = 0;
#pragma unroll
            for (int step = 16; step; step >>= 1)
              if (W[q + step - 1] < xx) q += step;
            if (q == 31 && W[31] < xx) q = 32;
            rank += q;
          }
          if (go) S[rank] = x;
          emitted += c;
          if (lane == j) mine = c;
        }
        cur += mine;
        blk_n = emitted;
        blk_t = 0;
        more = __ballot_sync(kFull, len_s != cur) != 0;
        __syncwarp();
        if (dedup) blk_n = warp_unique(S, emitted, lane);
      }
      const uint32_t s = (blk_t + lane < blk_n) ? S[blk_t + lane] : kNone;
      const bool inr = s <= bmax;  // in-range lanes form a prefix (the block ascends; ids < kNone)
      const unsigned m_in = __ballot_sync(kFull, inr);
      if (m_in == 0) break;
      uint32_t p = 0;  // how many ids of this chunk are < s  (s <= bmax, so p <= 31; none in the last trip)
      if (!tail) {
#pragma unroll
        for (int step = 16; step; step >>= 1) {
          const uint32_t probe = __shfl_sync(kFull, b, p + step - 1);
          if (probe < s) p += step;
        }
      }
      const bool in_long = __shfl_sync(kFull, b, p) == s;  // (every lane takes part in the shuffle)
      const bool keep = inr && !(dedup && in_long);        // an id the long list has too is dropped
      const unsigned km = __ballot_sync(kFull, keep);
      if (WRITE && keep) out[pos + c0 + p + kept_total + __popc(km & ((1u << lane) - 1))] = s;
      // every long id moves up by the kept short ids below it
      const uint32_t q = lanes_below(inr ? s : kNone, b);
      shift += __popc(km & (q >= 32 ? kFull : ((1u << q) - 1)));
      const uint32_t n_in = __popc(m_in);
      kept_total += __popc(km);
      blk_t += n_in;
      if (n_in < 32 && blk_t < blk_n) break;  // the next short id lies beyond this chunk
    }
    if (WRITE && valid) out[pos + i + kept0 + shift] = b;
    if (tail) break;
  }
  __syncwarp();
  return nb + kept_total;
}

// Union of k <= 32 lists holding at most 32 ids together, given as (offset, length) descriptors in global
// memory (the headers exchange: a message typically meets a dozen one-queue lists).  One id per lane,
// shuffle bitonic sort, neighbours compared to drop duplicates.  Returns the size of the union; with WRITE
// it goes to out[pos ...).
template <bool WRITE>
__device__ uint32_t warp_gather32(const uint32_t* __restrict__ qlist, const uint2* __restrict__ desc, uint32_t k, bool dedup,
                                  uint32_t* __restrict__ out, uint64_t pos, int lane) {
  uint2 d = make_uint2(0u, 0u);
  if ((uint32_t)lane < k) d = __ldg(desc + lane);
  uint32_t incl = d.y;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const uint32_t v = __shfl_up_sync(kFull, incl, s);
    if (lane >= s) incl += v;
  }
  const uint32_t total = __shfl_sync(kFull, incl, 31);
  // lane i takes the i-th id overall: the list whose [start, start + len) holds i
  uint32_t id = kNone;
  for (uint32_t j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, d.x, j), l = __shfl_sync(kFull, d.y, j);
    const uint32_t st = __shfl_sync(kFull, incl, j) - l;
    if ((uint32_t)lane >= st && (uint32_t)lane < st + l) id = __ldg(qlist + o + (lane - st));
  }
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
    for (int dd = size >> 1; dd; dd >>= 1) {
      const uint32_t other = __shfl_xor_sync(kFull, id, dd);
      const bool up = (lane & size) == 0, lower = (lane & dd) == 0;
      id = (lower == up) ? min(id, other) : max(id, other);
    }
  }
  const uint32_t prev = __shfl_up_sync(kFull, id, 1);
  const bool keep = (uint32_t)lane < total && !(dedup && lane > 0 && id == prev);
  const unsigned km = __ballot_sync(kFull, keep);
  if (WRITE && keep) out[pos + __popc(km & ((1u << lane) - 1))] = id;
  return __popc(km);
}

// The same for up to kRouteSmall ids in any number of lists: every lane copies whole lists (they are short)
// into shared memory, then a bitonic sort there.
template <bool WRITE>
__device__ uint32_t warp_gather_small(const uint32_t* __restrict__ qlist, const uint2* __restrict__ desc, uint32_t k,
                                      bool dedup, WarpScratch& ws, uint32_t* __restrict__ out, uint64_t pos, int lane) {
  if (k <= 32) {
    uint32_t sum = (uint32_t)lane < k ? __ldg(desc + lane).y : 0u;
#pragma unroll
    for (int s = 16; s; s >>= 1) sum += __shfl_xor_sync(kFull, sum, s);
    if (sum <= 32) return warp_gather32<WRITE>(qlist, desc, k, dedup, out, pos, lane);
  }
  uint32_t wpos = 0;
  for (uint32_t base = 0; base < k; base += 32) {
    uint2 d = make_uint2(0u, 0u);
    if (base + lane < k) d = __ldg(desc + base + lane);
    uint32_t incl = d.y;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t v = __shfl_up_sync(kFull, incl, s);
      if (lane >= s) incl += v;
    }
    const uint32_t st = wpos + incl - d.y;
    for (uint32_t i = 0; i < d.y && st + i < (uint32_t)kRouteSmall; ++i) ws.a[st + i] = __ldg(qlist + d.x + i);
    wpos += __shfl_sync(kFull, incl, 31);
  }
  const uint32_t total = min(wpos, (uint32_t)kRouteSmall);
  uint32_t np = 32;
  while (np < total) np <<= 1;
  for (uint32_t i = total + lane; i < np; i += 32) ws.a[i] = kNone;
  __syncwarp();
  warp_bitonic_sort(ws.a, np, lane);
  const uint32_t cnt = dedup ? warp_unique(ws.a, total, lane) : total;
  __syncwarp();
  if (WRITE)
    for (uint32_t i = lane; i < cnt; i += 32) out[pos + i] = ws.a[i];
  __syncwarp();
  return cnt;
}

}  // namespace
}  // namespace chq
