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
constexpr int kFlatMaxLong = 1024;   // longest list of a message on this path (longer ones: k_write_rest)
constexpr int kFlatMaxShort = 1024;  // ids in its other lists together: several long lists ('#' patterns, the slices
                                     // of a sharded exchange) are one streamed list and the others placed id by id
constexpr int kFlatChunks = (31 + kFlatMaxLong + kFlatMaxShort + 31) / 32;  // 32-position steps a message can span
constexpr int kFlatRow = (kFlatChunks + 7) & ~7;   // mask words per message: mask_stream takes eight steps per trip

struct __align__(16) FlatBuf {
  uint32_t cm[kFlatGroup][kFlatRow];                 // per message: marked positions, 32 per word
  uint2 d[kFlatGroup][kRouteSlots];                  // its lists (offset, length); [0] = the long one
  uint32_t start[kFlatGroup * kRouteSlots + 1];      // where list (g, e) starts among the group's short ids
  uint32_t msg[kFlatGroup], pos[kFlatGroup], cnt[kFlatGroup], kk[kFlatGroup];
};

// number of q[off .. off + n) below x; q is in global memory (L2), every lane has its own list
__device__ __forceinline__ uint32_t ldg_lower_bound(const uint32_t* __restrict__ q, uint32_t off, uint32_t n, uint32_t x) {
  uint32_t lo = off;
  while (n) {
    const uint32_t half = n >> 1;
    const bool below = __ldg(q + (lo + half)) < x;
    lo = below ? lo + half + 1 : lo;
    n = below ? n - half - 1 : half;
  }
  return lo - off;
}

// (Measured and not kept: galloping away from the proportional place idx * n2 / n1 when the two lists have similar
// lengths — the slices of a sharded exchange interleave, so the answer lies next to it — 0.194 -> 0.216 ms on the
// headline batch, 0.34 -> 0.43 ms with '#' patterns: the lanes' loops diverge and the plain bisection's loads are
// L1 hits after the first few.)
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
  __syncwarp();
  {  // masks: nothing marked inside the message, everything marked outside it (before its start in the first word,
     // from its end to the end of the row), so that the stream loop has no edge cases
    static_assert(kFlatRow % 4 == 0, "rows of whole uint4");
    constexpr int kMaskStores = (kFlatGroup * kFlatRow / 4 + 31) / 32;   // 16-byte stores per lane
    static_assert(kFlatGroup * kFlatMaxShort < 65536, "two 16-bit running sums in one word, below");
    uint4* cmf = reinterpret_cast<uint4*>(&fb.cm[0][0]);
#pragma unroll
    for (int r = 0; r < kMaskStores; ++r) {
      const int qd = lane + 32 * r;
      const int g = qd / (kFlatRow / 4), w0 = (qd - g * (kFlatRow / 4)) * 4;
      if (qd < kFlatGroup * kFlatRow / 4 && g < ng) {
        const uint32_t a0 = (ids_mis + fb.pos[g]) & 31u, span = a0 + fb.cnt[g];   // cnt >= 1
        auto word = [&](uint32_t w) -> uint32_t {   // positions [32 w, 32 w + 32) against [a0, span)
          const uint32_t lo = 32u * w;
          uint32_t mk = 0;
          if (lo < a0) mk = (1u << (a0 - lo)) - 1u;   // (a0 < 32: only word 0)
          if (span < lo + 32u) mk |= span <= lo ? 0xFFFFFFFFu : ~((1u << (span - lo)) - 1u);
          return mk;
        };
        cmf[qd] = make_uint4(word(w0), word(w0 + 1), word(w0 + 2), word(w0 + 3));
      }
    }
  }
  // the lists: lane -> (message lane / 8 and 4 + lane / 8, list lane % 8)
  const int e = lane & 7, g0 = lane >> 3, g1 = g0 + 4;
  uint2 r0 = make_uint2(0u, 0u), r1 = make_uint2(0u, 0u);
  if (g0 < ng && (uint32_t)e < fb.kk[g0]) r0 = __ldg(mlist + (size_t)fb.msg[g0] * kRouteSlots + e);
  if (g1 < ng && (uint32_t)e < fb.kk[g1]) r1 = __ldg(mlist + (size_t)fb.msg[g1] * kRouteSlots + e);
  fb.d[g0][e] = r0;
  fb.d[g1][e] = r1;
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
#ifdef CHQ_EXP
  if (CHQ_EXP & 2) return;
#endif
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
        q += ldg_lower_bound(qlist, d2.x, d2.y, x);
      }
      const uint32_t p0 = fb.pos[g];
      CHQ_ST_OUT(ids + ((size_t)p0 + q), x);
      const uint32_t rel = q + ((ids_mis + p0) & 31u);
      atomicOr(&fb.cm[g][rel >> 5], 1u << (rel & 31u));
    }
  }
  __syncwarp();
}

// The long list src[0 ...) of one message flows into out[0 .. c) around the positions marked in cm (see above).
// kStreamSteps steps per trip: their loads are issued together, then their stores; the row of masks is marked to
// its end, so a trip needs no bounds.
#ifndef CHQ_STREAM_STEPS
#define CHQ_STREAM_STEPS 4
#endif
#ifndef CHQ_STREAM_PIPE
#define CHQ_STREAM_PIPE 1
#endif
constexpr int kStreamSteps = CHQ_STREAM_STEPS;   // 4 or 8
static_assert(kFlatRow % kStreamSteps == 0, "a row of masks is a whole number of trips");

__device__ __forceinline__ void mask_stream(const uint32_t* __restrict__ src, uint32_t* __restrict__ out, uint32_t c,
                                            const uint32_t* cm, int lane) {
  const uint32_t a0 = (uint32_t)(reinterpret_cast<uintptr_t>(out) >> 2) & 31u;
  const uint32_t n_trips = (a0 + c + 32u * kStreamSteps - 1u) / (32u * kStreamSteps);
  const uint32_t lanebit = 1u << lane, ltmask = lanebit - 1u;
  // position p takes src[p - marked positions before p].  (The base goes through an asm move so that the compiler
  // keeps it in one register pair: left alone it re-derives it from the kernel parameters at every load, five
  // instructions instead of one multiply-add.)
  const uint32_t* base;
  asm("mov.b64 %0, %1;" : "=l"(base) : "l"(src + lane));
  int ti = 0;
  uint32_t* dp = out - a0 + lane;     // the lanes of a step write one 128-byte line
  const uint4* cm4 = reinterpret_cast<const uint4*>(cm);
  auto masks = [&](uint32_t t, uint32_t (&m)[kStreamSteps]) {
    const uint4 A = cm4[(kStreamSteps / 4) * t];
    m[0] = A.x; m[1] = A.y; m[2] = A.z; m[3] = A.w;
    if (kStreamSteps == 8) {
      const uint4 B = cm4[2 * t + 1];
      m[4] = B.x; m[5] = B.y; m[6] = B.z; m[7] = B.w;
    }
  };
  auto loads = [&](const uint32_t (&m)[kStreamSteps], uint32_t (&v)[kStreamSteps]) {
#pragma unroll
    for (int j = 0; j < kStreamSteps; ++j) {
      v[j] = 0;
      if (!(m[j] & lanebit)) v[j] = __ldg(base + (ti - (int)__popc(m[j] & ltmask)));
      ti += 32 - (int)__popc(m[j]);
    }
  };
  auto stores = [&](const uint32_t (&m)[kStreamSteps], const uint32_t (&v)[kStreamSteps]) {
#pragma unroll
    for (int j = 0; j < kStreamSteps; ++j)
      if (!(m[j] & lanebit)) CHQ_ST_OUT(dp + 32 * j, v[j]);
    dp += 32 * kStreamSteps;
  };
#if CHQ_STREAM_PIPE
  // the loads of trip t + 1 are in flight while trip t is stored
  uint32_t m0[kStreamSteps], v0[kStreamSteps];
  masks(0, m0);
  loads(m0, v0);
#pragma unroll 1
  for (uint32_t t = 1; t < n_trips; ++t) {
    uint32_t m1[kStreamSteps], v1[kStreamSteps];
    masks(t, m1);
    loads(m1, v1);
    stores(m0, v0);
#pragma unroll
    for (int j = 0; j < kStreamSteps; ++j) { m0[j] = m1[j]; v0[j] = v1[j]; }
  }
  stores(m0, v0);
#else
#pragma unroll 1
  for (uint32_t t = 0; t < n_trips; ++t) {
    uint32_t m[kStreamSteps], v[kStreamSteps];
    masks(t, m);
    loads(m, v);
    stores(m, v);
  }
#endif
}

}  // namespace
}  // namespace chq
