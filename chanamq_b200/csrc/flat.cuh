// k_write's main path (included by kernels.cu only): messages whose result is ONE long queue-id list plus a short
// side of at most kRouteSmall ids from other lists, nothing shared between the lists — 98 % of the ids of a routed
// batch (ExchangeEntity.scala:297-302 builds this set id by id; typical shape: "a.b.*" bound by hundreds of
// queues plus a handful of exact bindings).  Two steps, both without any per-message serial work:
//
//   flat_place   Up to eight such messages of a 32-message chunk are taken together and their SHORT ids are dealt
//                to the lanes, one id per lane whatever message it belongs to.  A short id's place in the output
//                is (ids of the long list below it) + (ids of the message's other short lists below it) + its
//                index in its own list: a handful of independent binary searches over L2-resident lists, run by
//                32 lanes for 32 different ids at once.  The lane stores its id there and marks the position in
//                the message's bit mask.  No gathering, no sorting network, no per-message set-up by a whole warp.
//   mask_stream  The long list then only has to flow around the marked positions: 32 output positions per step,
//                cut where the ADDRESS is a multiple of 128 bytes, source index = position - marked positions
//                before it (popc over the mask), one branch-free body.  Positions before the message's start and
//                after its end are marked like short ids, so the loop has no edge cases.
//                (Measured and not kept: four consecutive positions per lane with 128-bit stores — most steps hold
//                a short id or a message edge, the lanes then diverge and both variants of the body run.)
#pragma once
#include "merge.cuh"

namespace chq {
namespace {

constexpr int kFlatGroup = 8;        // messages placed together
constexpr int kFlatMaxLong = 1024;   // longest list of a message on this path (longer ones: warp_weave)
constexpr int kFlatChunks = (31 + kFlatMaxLong + kRouteSmall + 31) / 32;  // 32-position steps a message can span

struct __align__(16) FlatBuf {
  uint32_t cm[kFlatGroup][(kFlatChunks + 4) & ~3];   // per message: marked positions, 32 per word (rows of whole uint4)
  uint2 d[kFlatGroup][kRouteSlots];                  // its lists (offset, length); [0] = the long one
  uint32_t start[kFlatGroup * kRouteSlots + 1];      // where list (g, e) starts among the group's short ids
  uint32_t msg[kFlatGroup], pos[kFlatGroup], cnt[kFlatGroup], kk[kFlatGroup];
};

// number of a[0 .. n) below x; a is in global memory (L2), every lane has its own list
__device__ __forceinline__ uint32_t ldg_lower_bound(const uint32_t* __restrict__ a, uint32_t n, uint32_t x) {
  uint32_t lo = 0;
  while (n) {
    const uint32_t half = n >> 1;
    if (__ldg(a + lo + half) < x) { lo += half + 1; n -= half + 1; }
    else n = half;
  }
  return lo;
}

// gm = the lanes of this chunk whose message belongs to the group (at most kFlatGroup); (m, meta, cnt, pos) = the
// lane's own message.  Afterwards every short id of the group's messages is in place in ids[], fb.cm[g] marks
// those places (and everything outside the message), fb.d / fb.pos / fb.cnt describe message g for quad_stream.
__device__ __forceinline__ void flat_place(const uint32_t* __restrict__ qlist, const uint2* __restrict__ mlist,
                                           uint32_t* __restrict__ ids, FlatBuf& fb, unsigned gm, uint32_t m, uint32_t meta,
                                           uint32_t cnt, uint32_t pos, int lane) {
  const uint32_t ltmask = (1u << lane) - 1u;
  const uint32_t ids_mis = (uint32_t)(reinterpret_cast<uintptr_t>(ids) >> 2) & 31u;
  const int ng = __popc(gm);
  __syncwarp();  // the previous group is done with the buffer
  if ((gm >> lane) & 1u) {
    const int g = __popc(gm & ltmask);
    fb.msg[g] = m; fb.pos[g] = pos; fb.cnt[g] = cnt; fb.kk[g] = meta & 0xFFu;
  }
  {
    static_assert(sizeof(fb.cm) % 16 == 0 && sizeof(fb.cm) / 16 <= 96, "three 16-byte stores per lane clear the masks");
    uint4* cmf = reinterpret_cast<uint4*>(&fb.cm[0][0]);
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    constexpr int kQuads = sizeof(fb.cm) / 16;
    cmf[lane] = z;
    if (lane + 32 < kQuads) cmf[lane + 32] = z;
    if (lane + 64 < kQuads) cmf[lane + 64] = z;
  }
  __syncwarp();
  // the lists: lane -> (message lane / 8 and 4 + lane / 8, list lane % 8)
  const int e = lane & 7, g0 = lane >> 3, g1 = g0 + 4;
  uint2 r0 = make_uint2(0u, 0u), r1 = make_uint2(0u, 0u);
  if (g0 < ng && (uint32_t)e < fb.kk[g0]) r0 = __ldg(mlist + (size_t)fb.msg[g0] * kRouteSlots + e);
  if (g1 < ng && (uint32_t)e < fb.kk[g1]) r1 = __ldg(mlist + (size_t)fb.msg[g1] * kRouteSlots + e);
  fb.d[g0][e] = r0;
  fb.d[g1][e] = r1;
  if (lane < ng) {  // positions outside the message count as taken: lane -> message
    const uint32_t a0 = (ids_mis + fb.pos[lane]) & 31u, span = a0 + fb.cnt[lane];   // cnt >= 1
    const uint32_t last = (span - 1u) >> 5, end = span - (last << 5);                // end in 1..32
    fb.cm[lane][0] |= (1u << a0) - 1u;
    if (end < 32u) fb.cm[lane][last] |= ~((1u << end) - 1u);
  }
  // the group's short ids, numbered list after list
  uint32_t packed = e >= 1 ? (r0.y | (r1.y << 16)) : 0u;   // two running sums in one word: a group holds < 65536 short ids
  uint32_t incl = packed;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t v = __shfl_up_sync(kFull, incl, d);
    if (lane >= d) incl += v;
  }
  const uint32_t tot = __shfl_sync(kFull, incl, 31);
  const uint32_t t0 = tot & 0xFFFFu, total = t0 + (tot >> 16);
  fb.start[lane] = (incl - packed) & 0xFFFFu;
  fb.start[32 + lane] = t0 + ((incl - packed) >> 16);
  if (lane == 0) fb.start[64] = total;
  __syncwarp();
  for (uint32_t f0 = 0; f0 < total; f0 += 32) {
    const uint32_t f = f0 + lane;
    const bool live = f < total;
    uint32_t E = 0;
#pragma unroll
    for (int s = 32; s; s >>= 1)
      if (fb.start[E + s] <= f) E += s;
    const int g = E >> 3, le = E & 7;
    const uint32_t idx = f - fb.start[E];
    uint32_t q = idx;
    if (live) {
      const uint2 de = fb.d[g][le];
      const uint32_t x = __ldg(qlist + de.x + idx);
      const uint32_t kk = fb.kk[g];
      for (uint32_t e2 = 0; e2 < kk; ++e2) {   // the ids of the message's other lists below x
        if (e2 == (uint32_t)le) continue;
        const uint2 d2 = fb.d[g][e2];
        q += ldg_lower_bound(qlist + d2.x, d2.y, x);
      }
      const uint32_t p0 = fb.pos[g];
      ids[(size_t)p0 + q] = x;
      const uint32_t rel = q + ((ids_mis + p0) & 31u);
      atomicOr(&fb.cm[g][rel >> 5], 1u << (rel & 31u));
    }
  }
  __syncwarp();
}

// The long list src[0 ...) of one message flows into out[0 .. c) around the positions marked in cm (see above).
// The two addresses of a step are formed by one multiply-add each (mad.wide): the compiler's own choice for
// pointer + 32-bit index was five instructions per step.
__device__ __forceinline__ void mask_stream(const uint32_t* __restrict__ src, uint32_t* __restrict__ out, uint32_t c,
                                            const uint32_t* cm, int lane) {
  const uint32_t a0 = (uint32_t)(reinterpret_cast<uintptr_t>(out) >> 2) & 31u;
  const uint32_t n_ch = (a0 + c + 31u) >> 5;
  const uint32_t lanebit = 1u << lane, ltmask = lanebit - 1u;
  uint64_t sa = reinterpret_cast<uint64_t>(src + lane);   // position p takes src[p - marked positions before p]
  uint32_t* dp = out - a0 + lane;                          // the lanes of a step write one 128-byte line
  auto step = [&](uint32_t m, uint32_t* d) {
    if (!(m & lanebit)) {
      const uint32_t below = __popc(m & ltmask);
      uint64_t a;
      uint32_t v;
      asm("mad.wide.s32 %0, %1, -4, %2;" : "=l"(a) : "r"(below), "l"(sa));
      asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(a));
      *d = v;
    }
    const uint32_t adv = 32u - __popc(m);
    asm("mad.wide.u32 %0, %1, 4, %0;" : "+l"(sa) : "r"(adv));
  };
  const uint4* cm4 = reinterpret_cast<const uint4*>(cm);
  uint32_t ch = 0;
  for (; ch + 4 <= n_ch; ch += 4, dp += 128) {
    const uint4 M = cm4[ch >> 2];
    step(M.x, dp); step(M.y, dp + 32); step(M.z, dp + 64); step(M.w, dp + 96);
  }
  for (; ch < n_ch; ++ch, dp += 32) step(cm[ch], dp);
}

}  // namespace
}  // namespace chq
