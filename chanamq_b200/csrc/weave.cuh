// k_write's main path (included by kernels.cu only): the sorted union of ONE long queue-id list and a
// short side of at most kRouteSmall ids, which is what 98 % of the ids of a routed batch belong to
// (ExchangeEntity.scala:297-302 builds this set id by id; typical shape: "a.b.*" bound by hundreds of
// queues plus a handful of exact bindings).
//
//   1. the long list is staged in shared memory, a tile of up to kWeaveTile ids at a time, by ONE bulk
//      copy (cp.async.bulk global -> shared, completion on an mbarrier; lists of 8 ids or more start on a
//      16-byte boundary of the table's qlist, see HostTable::compile) issued before anything else, so that
//      it flies while
//   2. the short ids are gathered and sorted (registers up to 32 ids, shared memory beyond);
//   3. every short id finds its place among the staged ids by a binary search in shared memory (all of them
//      at once, one per lane, no dependent trips to L2) — place + index among the short ids = its position
//      in the OUTPUT, and each output chunk of 32 gets a bit mask of the positions short ids take;
//   4. the short ids are stored by their own lanes; the long ids are written chunk by chunk, aligned to 128
//      bytes, by one branch-free body: the lanes whose position takes a short id (or lies outside the message,
//      in the first and last chunk) sit out, the others read tile[position - bits below them in the mask].
//      No per-id loop, no shuffles, two ballots per 32 short ids.
#pragma once
#include "merge.cuh"

namespace chq {
namespace {

#ifndef CHQ_WEAVE_TILE
#define CHQ_WEAVE_TILE 576
#endif
constexpr int kWeaveTile = CHQ_WEAVE_TILE;                           // long-list ids staged at a time (multiple of 4)
constexpr int kWeaveChunks = (kWeaveTile + kRouteSmall + 31 + 31) / 32 + 1;  // output chunks one tile can touch

struct __align__(16) WeaveBuf {
  uint32_t tile[kWeaveTile + 4];   // +4: the bulk copy moves whole 16-byte units
  uint32_t cmask[kWeaveChunks < 32 ? 32 : kWeaveChunks];   // per output chunk: the positions that take a short id
};  // (the bulk copy's mbarrier is kept apart and handed in: this memory has other users between messages)

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void weave_bar_init(uint64_t* bar, int lane) {
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
}

// lane 0: bulk copy of `bytes` (multiple of 16) from src (16-byte aligned) into the tile
__device__ __forceinline__ void weave_bulk_load(WeaveBuf& wb, uint64_t* mbar, const uint32_t* src, uint32_t bytes) {
  const uint32_t bar = smem_addr(mbar), dst = smem_addr(wb.tile);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

__device__ __forceinline__ void weave_bar_wait(uint64_t* mbar, uint32_t parity) {
  const uint32_t bar = smem_addr(mbar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WEAVE_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WEAVE_DONE;\n"
      "bra WEAVE_WAIT;\n"
      "WEAVE_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}

// number of tile[0 .. tl) below x (tl >= 1, the same for all lanes: every lane runs the same number of steps)
__device__ __forceinline__ uint32_t tile_lower_bound(const uint32_t* tile, uint32_t tl, uint32_t x) {
  uint32_t base = 0, n = tl;
  while (n > 1) {
    const uint32_t half = n >> 1;
    if (tile[base + half - 1] < x) base += half;
    n -= half;
  }
  return base + (tile[base] < x ? 1u : 0u);
}

// Lane j < k holds list j = qlist[my_off .. my_off + my_len); list 0 is the longest (k_probe stores it first),
// the others hold at most kRouteSmall ids together and 2 <= k <= kRouteSlots.  `dedup` = an id may sit in two
// lists.  Writes the union to out[0 ...) (k_probe counted it: nothing is returned).  `phase` = parity of the
// warp's mbarrier, flipped by every bulk copy; tma_ok = qlist is the table's own, 16-byte aligned array.
__device__ __forceinline__ void warp_weave(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t my_len, int k, bool dedup,
                                           bool tma_ok, WarpScratch& ws, WeaveBuf& wb, uint64_t* mbar, uint32_t& phase,
                                           uint32_t* __restrict__ out, int lane) {
  const uint32_t lanebit = 1u << lane, ltmask = lanebit - 1u;
  const uint32_t out_mis = (uint32_t)(reinterpret_cast<uintptr_t>(out) >> 2) & 31u;   // where out[0] sits in its 128-byte line
  const uint32_t nb = __shfl_sync(kFull, my_len, 0), big_off = __shfl_sync(kFull, my_off, 0);
  // ---- first tile on its way before anything else ----
  uint32_t tl = min(nb, (uint32_t)kWeaveTile);
  const bool bulk = tma_ok && (big_off & 3u) == 0;
  if (bulk) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier plain stores into the tile come first
  __syncwarp();  // every lane is done with the tile of the previous message
  if (bulk) {
    if (lane == 0) weave_bulk_load(wb, mbar, qlist + big_off, ((tl + 3u) & ~3u) * 4u);
  } else {
    for (uint32_t i = lane; i < tl; i += 32) wb.tile[i] = __ldg(qlist + big_off + i);
  }
  // ---- the short side, ascending, duplicates among its own lists dropped: one id per lane (s) when there are
  //      at most 32, else ws.b[0 .. ns) ----
  uint32_t s = kNone, ns;
  bool hole = false;
  bool in_regs = true;   // the short ids sit one per lane (at most 32 of them before duplicates go), else in ws.b
  if (k == 2) {  // one short list: it is sorted already
    ns = __shfl_sync(kFull, my_len, 1);
    const uint32_t o = __shfl_sync(kFull, my_off, 1);
    in_regs = ns <= 32;
    if (in_regs) {
      if ((uint32_t)lane < ns) s = __ldg(qlist + o + lane);
    } else {
      for (uint32_t i = lane; i < ns; i += 32) ws.b[i] = __ldg(qlist + o + i);
    }
  } else {
    const uint32_t short_len = (lane > 0 && lane < k) ? my_len : 0u;
    uint32_t incl = short_len;
#pragma unroll
    for (int d = 1; d < kRouteSlots; d <<= 1) {
      const uint32_t v = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += v;
    }
    ns = __shfl_sync(kFull, incl, kRouteSlots - 1);
    const uint32_t st = incl - short_len;
    in_regs = ns <= 32;   // (decided before duplicates are removed: the ids stay where they were put)
    if (in_regs) {
      if (k == 3) {  // two sorted lists: one up from lane 0, one down from lane 31 -> a bitonic sequence
        const uint32_t o1 = __shfl_sync(kFull, my_off, 1), l1 = __shfl_sync(kFull, my_len, 1);
        const uint32_t o2 = __shfl_sync(kFull, my_off, 2), l2 = __shfl_sync(kFull, my_len, 2);
        if ((uint32_t)lane < l1) s = __ldg(qlist + o1 + lane);
        else if ((uint32_t)lane >= 32 - l2) s = __ldg(qlist + o2 + (31 - lane));
#pragma unroll
        for (int d = 16; d; d >>= 1) {
          const uint32_t other = __shfl_xor_sync(kFull, s, d);
          s = (lane & d) == 0 ? min(s, other) : max(s, other);
        }
      } else {
        for (int j = 1; j < k; ++j) {
          const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, my_len, j), sj = __shfl_sync(kFull, st, j);
          const uint32_t i = (uint32_t)lane - sj;
          if (i < l) s = __ldg(qlist + o + i);
        }
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
      if (dedup) {  // an id that sits in two short lists: neighbours after the sort; the later one is a hole,
        const uint32_t prev = __shfl_up_sync(kFull, s, 1);   // which keeps its place in the order but is never written
        hole = lane > 0 && s == prev;
      }
    } else {
      if (lane < k) { ws.st[lane] = st; ws.ln[lane] = short_len; }
      for (int j = 1; j < k; ++j) {
        const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, short_len, j), sj = __shfl_sync(kFull, st, j);
        for (uint32_t i = lane; i < l; i += 32) ws.a[sj + i] = __ldg(qlist + o + i);
      }
      __syncwarp();
      // rank-merge: every id finds its place among the other short lists by binary search
      for (uint32_t e = lane; e < ns; e += 32) {
        int j = 1;
        while (e >= ws.st[j] + ws.ln[j]) ++j;
        const uint32_t x = ws.a[e];
        uint32_t r = e - ws.st[j];
        for (int j2 = 1; j2 < k; ++j2) {
          if (j2 == j) continue;
          const uint32_t l2 = ws.ln[j2];
          r += j2 < j ? smem_upper_bound(ws.a + ws.st[j2], l2, x) : smem_lower_bound(ws.a + ws.st[j2], l2, x);
        }
        ws.b[r] = x;
      }
      __syncwarp();
      if (dedup) ns = warp_unique(ws.b, ns, lane);
    }
  }
  __syncwarp();

  // ---- tiles of the long list ----
  uint32_t kept = 0;    // short ids already written
  uint32_t sdone = 0;   // short ids (or holes) already dealt with
  for (uint32_t tb = 0;;) {
    const bool last = tb + tl >= nb;
    // output positions are relative to out[0]; chunks are cut where the ADDRESS is a multiple of 128 bytes
    const int ustart = (int)(tb + kept);
    const uint32_t a0 = (out_mis + (uint32_t)ustart) & 31u;
    const int u0 = ustart - (int)a0;   // start of the first chunk (may lie before the message)
    // positions before the message count as taken, like those of short ids: their lanes sit out
    wb.cmask[lane] = lane == 0 ? (1u << a0) - 1u : 0u;
    if (kWeaveChunks > 32 && lane < kWeaveChunks - 32) wb.cmask[32 + lane] = 0;
    if (bulk) {
      weave_bar_wait(mbar, phase);
      phase ^= 1u;
    }
    __syncwarp();
    for (uint32_t r = sdone; r < ns; r += 32) {
      const uint32_t t = r + lane;
      const uint32_t x = in_regs ? ((uint32_t)lane >= sdone ? s : kNone) : (t < ns ? ws.b[t] : kNone);
      const uint32_t lb = tile_lower_bound(wb.tile, tl, x);
      const bool live = in_regs ? ((uint32_t)lane >= sdone && (uint32_t)lane < ns) : t < ns;
      const bool in_tile = live && (lb < tl || last);   // a prefix of the live lanes: the short ids ascend
      const bool keep = in_tile && !hole && !(dedup && lb < tl && wb.tile[lb] == x);
      const unsigned im = __ballot_sync(kFull, in_tile), km = __ballot_sync(kFull, keep);
      if (keep) {
        const uint32_t q = tb + lb + kept + __popc(km & ltmask);   // its position in the output
        const uint32_t rel = q - (uint32_t)u0;
        atomicOr(&wb.cmask[rel >> 5], 1u << (rel & 31u));
        out[q] = x;
      }
      kept += __popc(km);
      const uint32_t n_in = __popc(im);
      sdone += n_in;
      if (in_regs || n_in < 32) break;
    }
    const uint32_t span = tb + tl + kept - (uint32_t)u0;   // positions from the first chunk's start to the end
    const uint32_t n_chunks = (span + 31u) >> 5;
    if (lane == 0 && (span & 31u)) atomicOr(&wb.cmask[n_chunks - 1], ~((1u << (span & 31u)) - 1u));   // past the end: taken
    __syncwarp();
    // ---- the long ids of this tile: lane at position p of a chunk reads tile[p - (taken positions before p)] ----
    const uint32_t* tp = wb.tile + lane;
    uint32_t* dp = out + (u0 + lane);
    const uint32_t* cm = wb.cmask;
#pragma unroll 4
    for (uint32_t c = 0; c < n_chunks; ++c) {
      const uint32_t m = cm[c];
      if (!(m & lanebit)) dp[32 * c] = *(tp - __popc(m & ltmask));
      tp += 32 - __popc(m);
    }
    if (last) break;
    tb += tl;
    tl = min(nb - tb, (uint32_t)kWeaveTile);
    __syncwarp();  // everyone has read the tile
    if (bulk) {
      if (lane == 0) weave_bulk_load(wb, mbar, qlist + big_off + tb, ((tl + 3u) & ~3u) * 4u);
    } else {
      for (uint32_t i = lane; i < tl; i += 32) wb.tile[i] = __ldg(qlist + big_off + tb + i);
    }
  }
}

}  // namespace
}  // namespace chq
