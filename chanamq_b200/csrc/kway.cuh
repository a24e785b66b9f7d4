// k_write_rest's path for messages whose lists are of SIMILAR length and share nothing (included by kernels.cu only):
// the slices of a binding-sliced exchange (binding i on GPU i mod G: a message's ids are spread evenly over the G
// slice results), several '#' patterns of one publisher.  Streaming the longest list and placing every other id by
// k - 1 searches (flat.cuh) is the wrong tool there: most ids are on the short side.  Here the lists are staged back to
// back in shared memory and merged pairwise, level by level (2 lists: one level, 8 lists: three): every id finds its
// place in the merged pair by ONE binary search over the partner run, in shared memory, all lanes at once.
#pragma once
#include "merge.cuh"

namespace chq {
namespace {

constexpr int kKwayMax = 1024;   // ids of one message on this path (two buffers of that size per warp)

struct __align__(16) KwayBuf {
  uint32_t a[kKwayMax], b[kKwayMax];
  uint32_t start[kRouteSlots + 1];   // run r = [start[r], start[r + 1])
};

// number of a[0 .. n) below x; n is the same for all lanes, so every lane runs the same number of steps
__device__ __forceinline__ uint32_t smem_lower_bound_uniform(const uint32_t* a, uint32_t n, uint32_t x) {
  if (n == 0) return 0;
  uint32_t base = 0;
  while (n > 1) {
    const uint32_t half = n >> 1;
    if (a[base + half - 1] < x) base += half;
    n -= half;
  }
  return base + (a[base] < x ? 1u : 0u);
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
  if (lane <= kRouteSlots) kb.start[lane] = lane < kRouteSlots ? incl - len : total;   // (runs past k are empty)
  for (int j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, len, j), s = __shfl_sync(kFull, incl - len, j);
    for (uint32_t i = lane; i < l; i += 32) kb.a[s + i] = __ldg(qlist + o + i);
  }
  __syncwarp();
  uint32_t* A = kb.a;
  uint32_t* B = kb.b;
  for (int runs = k; runs > 1; runs = (runs + 1) >> 1) {
    // run by run (all lanes on the same run: its partner, and with it the number of search steps, is the same for
    // every lane): an element's place in the merged pair = its index + the partner's ids below it
    for (int r = 0; r < runs; ++r) {
      const uint32_t rs = kb.start[r], rl = kb.start[r + 1] - rs;
      const int p = r ^ 1;
      if (p < runs) {
        const uint32_t ps = kb.start[p], pl = kb.start[p + 1] - ps, base = kb.start[r & ~1];
        for (uint32_t i = lane; i < rl; i += 32) {
          const uint32_t x = A[rs + i];
          B[base + i + smem_lower_bound_uniform(A + ps, pl, x)] = x;   // (no ties: the lists share nothing)
        }
      } else {  // a run without a partner keeps its place
        for (uint32_t i = lane; i < rl; i += 32) B[rs + i] = A[rs + i];
      }
    }
    __syncwarp();
    // runs 2 j and 2 j + 1 became run j
    uint32_t ns = 0;
    const int new_runs = (runs + 1) >> 1;
    if (lane <= kRouteSlots) ns = lane < new_runs ? kb.start[2 * lane] : total;
    __syncwarp();
    if (lane <= kRouteSlots) kb.start[lane] = ns;
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
