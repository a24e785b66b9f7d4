// k_write_rest's path for messages whose ids are spread over several lists that share nothing (included by kernels.cu
// only): the slices of a binding-sliced exchange (binding i on GPU i mod G: a message's ids are spread evenly over the
// G slice results), several '#' patterns of one publisher.  Streaming the longest list and placing every other id by
// k - 1 searches (flat.cuh) is the wrong tool there: most ids are on the short side.  Here the lists are staged back to
// back in shared memory (asynchronous 4-byte copies: the slices' lists start anywhere) and merged pairwise, level by
// level (2 lists: one level, 8 lists: three).  A level is ONE pass over the message: every lane takes an equal share of
// the level's output, finds where that share starts in its pair of runs by one binary search along the diagonal
// (merge path), and from there merges sequentially — a compare, a select and one shared-memory load per id.
// (Before: one binary search per id and level — 13 warp instructions per id at 8 slices, k_write_rest 0.36 ms for the
// 13 M ids of an eighth of a batch.)
#pragma once
#include "merge.cuh"

namespace chq {
namespace {

constexpr int kKwayMax = 1280;   // ids of one message on this path (two buffers of that size per warp: 40 KB per CTA)

struct __align__(16) KwayBuf {
  uint32_t a[kKwayMax + 4], b[kKwayMax + 4];   // (+ 4: the merge loop reads one past a run before it looks at the index)
  uint32_t start[kRouteSlots + 2];   // run r = [start[r], start[r + 1]); entries past the last run = total
};

__device__ __forceinline__ void cp_async4(uint32_t* smem_dst, const uint32_t* __restrict__ src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(src) : "memory");
}

// Lane j < k holds list j = qlist[my_off .. my_off + my_len), k <= kRouteSlots, total <= kKwayMax ids, no id in two
// lists.  Writes the sorted union to out[0 .. total).
__device__ __forceinline__ void warp_kway(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t my_len, int k, KwayBuf& kb,
                                          uint32_t* __restrict__ out, int lane) {
  const uint32_t len = lane < k ? my_len : 0u;
  uint32_t incl = len;
#pragma unroll
  for (int d = 1; d < kRouteSlots; d <<= 1) {
    const uint32_t v = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += v;
  }
  const uint32_t total = __shfl_sync(kFull, incl, kRouteSlots - 1);
  __syncwarp();   // the previous message is done with the buffers
  if (lane < kRouteSlots + 2) kb.start[lane] = lane < kRouteSlots ? incl - len : total;   // (runs past k are empty)
  for (int j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, len, j), s = __shfl_sync(kFull, incl - len, j);
    for (uint32_t i = lane; i < l; i += 32) cp_async4(&kb.a[s + i], qlist + o + i);
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
  uint32_t* A = kb.a;
  uint32_t* B = kb.b;
  // a lane's share of every level's output: an odd number of positions, so that the lanes' writes fall in different banks
  const uint32_t share = ((total + 31) >> 5) | 1u;
  const uint32_t lo = min(total, (uint32_t)lane * share), hi = min(total, lo + share);
  for (int runs = k; runs > 1; runs = (runs + 1) >> 1) {
    if (lo < hi) {
      // the pair of runs (2 p, 2 p + 1) position lo falls in: A-run [as, bs), B-run [bs, eb) -> output [as, eb)
      int pr = 0;
      while (kb.start[2 * pr + 2] <= lo) ++pr;
      uint32_t as = kb.start[2 * pr], bs = kb.start[2 * pr + 1], eb = kb.start[2 * pr + 2];
      // merge path: of the lo - as ids before position lo, how many come from the A-run?
      const uint32_t d = lo - as, la = bs - as, lb = eb - bs;
      uint32_t l = d > lb ? d - lb : 0u, h = min(d, la);
      while (l < h) {
        const uint32_t mid = (l + h) >> 1;
        if (A[as + mid] < A[bs + (d - 1u - mid)]) l = mid + 1u; else h = mid;
      }
      uint32_t ia = as + l, ib = bs + (d - l);
      for (uint32_t pos = lo;;) {
        // the part of the share inside this pair: a compare, a select and one load per id, nothing else in the loop
        // (the load is unconditional: one past a run is still inside the buffers, and the value is not used)
        const uint32_t seg = min(hi, eb);
        uint32_t va = ia < bs ? A[ia] : kNone, vb = ib < eb ? A[ib] : kNone;   // (an exhausted run shows kNone, above every id)
        for (; pos < seg; ++pos) {
          const bool ta = va < vb;
          B[pos] = ta ? va : vb;
          ia += ta ? 1u : 0u;
          ib += ta ? 0u : 1u;
          const uint32_t nx = ta ? ia : ib;
          const uint32_t got = A[nx];
          const uint32_t v = nx < (ta ? bs : eb) ? got : kNone;
          va = ta ? v : va;
          vb = ta ? vb : v;
        }
        if (pos >= hi) break;
        ++pr;   // on to the next pair, from its beginning
        ia = eb;
        bs = kb.start[2 * pr + 1];
        ib = bs;
        eb = kb.start[2 * pr + 2];
      }
    }
    __syncwarp();
    // runs 2 j and 2 j + 1 became run j
    uint32_t ns = 0;
    const int new_runs = (runs + 1) >> 1;
    if (lane < kRouteSlots + 2) ns = lane < new_runs ? kb.start[2 * lane] : total;
    __syncwarp();
    if (lane < kRouteSlots + 2) kb.start[lane] = ns;
    __syncwarp();
    uint32_t* t = A; A = B; B = t;
  }
  for (uint32_t e = lane; e < total; e += 32) CHQ_ST_OUT(out + e, A[e]);
}

// Two lists of similar length that share nothing (two slices of a sharded exchange, two '#' patterns): a symmetric
// merge through registers.  Both lists show a window of 32 ids, one per lane; everything up to T = the smaller of the
// two windows' last ids (a list that ends inside its window sets no bound) can be placed for good: an id's place =
// ids already written + its lane + the other window's ids below it (a shuffle binary search).  The list that sets T
// empties its whole window, so a trip writes at least 32 ids, about 63 when the lists interleave evenly.
__device__ __forceinline__ void warp_merge2(const uint32_t* __restrict__ qlist, uint32_t oa, uint32_t na, uint32_t ob, uint32_t nb,
                                            uint32_t* __restrict__ out, int lane) {
  uint32_t ia = 0, ib = 0;
  const uint32_t* pa = qlist + oa + lane;
  const uint32_t* pb = qlist + ob + lane;
  while (ia < na || ib < nb) {
    const uint32_t ra = na - ia, rb = nb - ib;
    const uint32_t a = (uint32_t)lane < ra ? __ldg(pa + ia) : kNone;
    const uint32_t b = (uint32_t)lane < rb ? __ldg(pb + ib) : kNone;
    const uint32_t a31 = __shfl_sync(kFull, a, 31), b31 = __shfl_sync(kFull, b, 31);
    const uint32_t T = min(ra > 32 ? a31 : kNone, rb > 32 ? b31 : kNone);
    const bool ea = a != kNone && a <= T, eb = b != kNone && b <= T;   // prefixes of the lanes: the windows ascend
    const uint32_t below_a = lanes_below(b, a), below_b = lanes_below(a, b);
    uint32_t* o = out + (ia + ib + lane);
    if (ea) CHQ_ST_OUT(o + below_a, a);
    if (eb) CHQ_ST_OUT(o + below_b, b);
    ia += __popc(__ballot_sync(kFull, ea));
    ib += __popc(__ballot_sync(kFull, eb));
  }
}

}  // namespace
}  // namespace chq
