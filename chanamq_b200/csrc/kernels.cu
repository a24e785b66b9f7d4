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

namespace chq {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;

__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ------------------------------------------------------------------------------------------
// word dictionary probe (exact: hash, length, first 8 bytes, then the remaining bytes)
// ------------------------------------------------------------------------------------------
__device__ uint32_t dict_lookup(const TableView& tv, uint64_t fnv, uint64_t f8, uint32_t len, const uint8_t* bytes) {
  const uint64_t h = fin(fnv ^ ((uint64_t)len << 56));
  uint32_t i = (uint32_t)h & tv.dict_mask;
  for (;;) {
    const uint4 a = ldg16(&tv.dict[i]);
    const uint4 b = ldg16(reinterpret_cast<const uint8_t*>(&tv.dict[i]) + 16);
    const uint32_t id = b.y;
    if (id == 0) return kUnknownWord;
    const uint64_t sh = ((uint64_t)a.y << 32) | a.x, sf = ((uint64_t)a.w << 32) | a.z;
    if (sh == h && sf == f8 && b.x == len) {
      bool same = true;
      if (len > 8) {
        const uint8_t* d = tv.dict_bytes + b.z;
        for (uint32_t j = 8; j < len; ++j)
          if (__ldg(d + j) != __ldg(bytes + j)) { same = false; break; }
      }
      if (same) return id;
    }
    i = (i + 1) & tv.dict_mask;
  }
}

}  // namespace

// One thread per key.  Java split semantics: words are cut at every '.', trailing
// empty words are dropped, a key without any '.' is one word (even when empty).
__global__ void __launch_bounds__(256) k_tokenize(const uint8_t* __restrict__ blob, const uint32_t* __restrict__ off,
                                                  uint32_t n, TableView tv, int split, uint32_t* __restrict__ wid,
                                                  uint8_t* __restrict__ nwords, uint8_t* __restrict__ err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t b0 = off[i], b1 = off[i + 1];
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

namespace {

// ------------------------------------------------------------------------------------------
// probing
// ------------------------------------------------------------------------------------------
struct GroupRec {
  uint32_t list_off, list_len, lit_off, next, rem_off, rem_len, nlit, flags;
};
__device__ __forceinline__ GroupRec load_group(const TableView& tv, uint32_t g) {
  const uint4 a = ldg16(&tv.groups[g]);
  const uint4 b = ldg16(reinterpret_cast<const uint8_t*>(&tv.groups[g]) + 16);
  GroupRec r;
  r.list_off = a.x; r.list_len = a.y; r.lit_off = a.z; r.next = a.w;
  r.rem_off = b.x; r.rem_len = b.y & 0xFFFFu; r.nlit = (b.y >> 16) & 0xFFu; r.flags = b.y >> 24;
  return r;
}

// Head group of the chain whose (shape, literal tuple) equals the key's words at the shape's
// literal positions, or kNone.  Exact: the stored tuple is compared word by word.
__device__ __forceinline__ Shape load_shape(const TableView& tv, uint32_t s) {
  const uint4 raw = ldg16(&tv.shapes[s]);
  Shape sh;
  sh.seed = ((uint64_t)raw.y << 32) | raw.x;
  sh.lit_mask = raw.z;
  sh.kind = raw.w & 0xFFu;
  sh.len = (raw.w >> 8) & 0xFFu;
  sh.pad = 0;
  return sh;
}

__device__ uint32_t probe_head(const TableView& tv, uint32_t shape_idx, const Shape& sh,
                               const uint32_t* __restrict__ row) {
  uint64_t h = sh.seed;
  for (uint32_t m = sh.lit_mask; m; m &= m - 1) {
    const uint32_t w = __ldg(row + (__ffs(m) - 1));
    if (w == kUnknownWord) return kNone;  // a word no binding mentions only matches wildcards
    h = mix(h, w);
  }
  h = fin(h);
  const uint32_t tag = tag_of(h);
  uint32_t b = (uint32_t)h & tv.bucket_mask;
  for (;;) {
    const uint4 s01 = ldg16(&tv.slots[(size_t)b * 4]);
    const uint4 s23 = ldg16(&tv.slots[(size_t)b * 4 + 2]);
    const uint32_t tags[4] = {s01.x, s01.z, s23.x, s23.z};
    const uint32_t grps[4] = {s01.y, s01.w, s23.y, s23.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (tags[i] == 0) return kNone;
      if (tags[i] == tag) {
        const uint32_t g = grps[i];
        const Group* gp = &tv.groups[g];
        const uint32_t lit_off = __ldg(&gp->lit_off);
        if (__ldg(&gp->shape) == shape_idx) {
          bool same = true;
          uint32_t li = 0;
          for (uint32_t m = sh.lit_mask; m; m &= m - 1, ++li)
            if (__ldg(tv.lits + lit_off + li) != __ldg(row + (__ffs(m) - 1))) { same = false; break; }
          if (same) return g;
        }
      }
    }
    b = (b + 1) & tv.bucket_mask;
  }
}

// Word-level NFA for a pattern remainder (literals, '*', '#') against key words [from, nw).
// State s = remainder items consumed so far; '#' may take zero or more words.
__device__ bool remainder_matches(const TableView& tv, const GroupRec& g, const uint32_t* __restrict__ row,
                                  uint32_t from, uint32_t nw) {
  const uint32_t r = g.rem_len;
  const uint32_t* rem = tv.rems + g.rem_off;
  // reach bitset over states 0..r (r <= 128): 5 words
  uint32_t cur[5] = {1u, 0, 0, 0, 0}, nxt[5];
  auto close = [&](uint32_t* st) {  // epsilon: a '#' can be skipped
    for (uint32_t s = 0; s < r; ++s)
      if ((st[s >> 5] >> (s & 31)) & 1u)
        if (__ldg(rem + s) == kHash) st[(s + 1) >> 5] |= 1u << ((s + 1) & 31);
  };
  close(cur);
  for (uint32_t p = from; p < nw; ++p) {
    const uint32_t w = __ldg(row + p);
#pragma unroll
    for (int i = 0; i < 5; ++i) nxt[i] = 0;
    for (uint32_t s = 0; s < r; ++s) {
      if (!((cur[s >> 5] >> (s & 31)) & 1u)) continue;
      const uint32_t item = __ldg(rem + s);
      if (item == kHash) nxt[s >> 5] |= 1u << (s & 31);
      else if (item == kStar || (item == w && w != kUnknownWord)) nxt[(s + 1) >> 5] |= 1u << ((s + 1) & 31);
    }
    close(nxt);
#pragma unroll
    for (int i = 0; i < 5; ++i) cur[i] = nxt[i];
  }
  return (cur[r >> 5] >> (r & 31)) & 1u;
}

// Calls f(group index, record) for every group that matches this key.
template <typename F>
__device__ __forceinline__ void visit_shape(const TableView& tv, uint32_t s, const uint32_t* __restrict__ row,
                                            uint32_t nw, F&& f) {
  const Shape sh = load_shape(tv, s);
  uint32_t g = probe_head(tv, s, sh, row);
  while (g != kNone) {
    const GroupRec rec = load_group(tv, g);
    if (rec.rem_len == 0 || remainder_matches(tv, rec, row, sh.len, nw)) f(g, rec);
    g = rec.next;
  }
}

// ------------------------------------------------------------------------------------------
// warp-level set union of the matched groups' sorted queue lists
// ------------------------------------------------------------------------------------------
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
  uint32_t kept = 0;
  for (uint32_t base = 0; base < n; base += 32) {
    const uint32_t i = base + lane;
    const uint32_t v = i < n ? a[i] : 0;
    const bool keep = i < n && (i == 0 || a[i - 1] != v);
    __syncwarp();
    const unsigned m = __ballot_sync(kFull, keep);
    if (keep) a[kept + __popc(m & ((1u << lane) - 1))] = v;
    kept += __popc(m);
    __syncwarp();
  }
  return kept;
}

// Lane j < k holds list j = qlist[my_off .. my_off+my_len).  The longest list is streamed from
// L2 in chunks of 32 while the others (together <= scap ids, staged and sorted in shared memory)
// are woven in; equal ids are emitted once.  Returns the size of the union; with WRITE the
// sorted union goes to out[pos ...), never past cap.
template <bool WRITE>
__device__ uint32_t warp_union(const uint32_t* __restrict__ qlist, uint32_t my_off, uint32_t my_len, int k,
                               uint32_t* sbuf, uint32_t* __restrict__ out, uint64_t pos, uint64_t cap, int lane) {
  // pick the longest list
  uint32_t len_k = lane < k ? my_len : 0;
  uint32_t mx = len_k;
#pragma unroll
  for (int d = 16; d; d >>= 1) mx = max(mx, __shfl_xor_sync(kFull, mx, d));
  const int big = __ffs(__ballot_sync(kFull, lane < k && len_k == mx)) - 1;
  const uint32_t big_off = __shfl_sync(kFull, my_off, big), nb = mx;
  // stage the rest
  uint32_t ns = 0;
  for (int j = 0; j < k; ++j) {
    const uint32_t o = __shfl_sync(kFull, my_off, j), l = __shfl_sync(kFull, my_len, j);
    if (j == big) continue;
    for (uint32_t i = lane; i < l; i += 32) sbuf[ns + i] = __ldg(qlist + o + i);
    ns += l;
  }
  __syncwarp();
  if (k > 2) {
    uint32_t np = 32;
    while (np < ns) np <<= 1;
    for (uint32_t i = ns + lane; i < np; i += 32) sbuf[i] = kNone;
    __syncwarp();
    warp_bitonic_sort(sbuf, np, lane);
    ns = warp_unique(sbuf, ns, lane);
  }
  // weave
  uint32_t t = 0, kept = 0;
  for (uint32_t c0 = 0; c0 < nb; c0 += 32) {
    const uint32_t i = c0 + lane;
    const bool valid = i < nb;
    const uint32_t b = valid ? __ldg(qlist + big_off + i) : kNone;
    const uint32_t bmax = __shfl_sync(kFull, b, min(31u, nb - c0 - 1));
    const uint32_t kept0 = kept;
    uint32_t shift = 0;
    while (t < ns) {
      const uint32_t s = sbuf[t];
      if (s > bmax) break;
      const unsigned lt = __ballot_sync(kFull, valid && b < s);
      const unsigned eq = __ballot_sync(kFull, valid && b == s);
      if (eq == 0) {
        const uint32_t p = __popc(lt);
        if (WRITE && lane == 0) {
          const uint64_t o = pos + c0 + p + kept;
          if (o < cap) out[o] = s;
        }
        if ((uint32_t)lane >= p) ++shift;
        ++kept;
      }
      ++t;
    }
    if (WRITE && valid) {
      const uint64_t o = pos + i + kept0 + shift;
      if (o < cap) out[o] = b;
    }
  }
  // whatever is left sorts after the whole long list
  if (WRITE) {
    for (uint32_t i = t + lane; i < ns; i += 32) {
      const uint64_t o = pos + nb + kept + (i - t);
      if (o < cap) out[o] = sbuf[i];
    }
  }
  return nb + kept + (ns - t);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// the fused route kernel
// ------------------------------------------------------------------------------------------
constexpr uint64_t kFlagAgg = 1ull << 62, kFlagIncl = 2ull << 62, kValMask = (1ull << 62) - 1;

struct __align__(16) RouteSmem {
  uint32_t match[kRouteSlots][kRouteThreads];   // matched group ids, slot-major
  uint32_t small[kRouteThreads / 32][kRouteSmall];
  uint64_t warp_sum[kRouteThreads / 32];
  uint64_t tile_base;
  uint32_t tile;
};

// Generic slow path for one message, by one warp: every matched list is appended to a piece of the
// global scratch arena, sorted there and de-duplicated.  Returns the count; *scr_off = where.
__device__ uint32_t warp_fallback(const RouteParams& p, const uint32_t* __restrict__ row, uint32_t nw, int lane,
                                  uint64_t* scr_off, uint32_t* any_flags) {
  const TableView& tv = p.tv;
  const uint32_t e0 = nw <= (uint32_t)kLenCap ? tv.exact_rng[2 * nw] : 0, e1 = nw <= (uint32_t)kLenCap ? tv.exact_rng[2 * nw + 1] : 0;
  const uint32_t n_exact = e1 - e0, n_shapes = n_exact + (tv.hashy_end - tv.hashy_begin);
  uint64_t base = 0;
  uint32_t np = 0, total = 0, flags = 0;
  for (int pass = 0; pass < 2; ++pass) {
    uint32_t wpos = 0;
    for (uint32_t s0 = 0; s0 < n_shapes; s0 += 32) {
      const uint32_t si = s0 + lane;
      uint32_t mine = 0;
      // a lane can meet several groups under one shape (chained remainders): walk them in rounds
      uint32_t g = kNone;
      uint32_t shape_len = 0;
      if (si < n_shapes) {
        const uint32_t s = si < n_exact ? e0 + si : tv.hashy_begin + (si - n_exact);
        const Shape sh = load_shape(tv, s);
        shape_len = sh.len;
        if (sh.len <= nw) g = probe_head(tv, s, sh, row);
      }
      while (__any_sync(kFull, g != kNone)) {
        GroupRec rec{};
        bool hit = false;
        if (g != kNone) {
          rec = load_group(tv, g);
          hit = rec.rem_len == 0 || remainder_matches(tv, rec, row, shape_len, nw);
        }
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
        g = (g != kNone) ? rec.next : kNone;
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

__global__ void __launch_bounds__(kRouteThreads) k_route(RouteParams p) {
  __shared__ RouteSmem sm;
  const TableView& tv = p.tv;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  volatile uint64_t* tstate = p.tile_state;

  for (;;) {
    __syncthreads();  // previous tile fully done with sm
    if (tid == 0) sm.tile = atomicAdd(p.ctrl + kCtrlTile, 1u);
    __syncthreads();
    const uint32_t tile = sm.tile;
    if (tile >= p.n_tiles) return;
    const uint32_t m = tile * kRouteThreads + tid;

    // ---- phase A: one thread per message: probe every compatible shape -------------------
    uint32_t k = 0, ub = 0, mxlen = 0, flags = 0, cnt = 0;
    uint32_t nw = 0;
    const uint32_t* row = p.wid + (size_t)m * tv.wstride;
    if (m < p.n) nw = p.nwords[m];
    if (nw > 0) {
      auto record = [&](uint32_t g, const GroupRec& rec) {
        if (k < p.max_slots) sm.match[k][tid] = g;
        ++k;
        ub += rec.list_len;
        mxlen = max(mxlen, rec.list_len);
        flags |= rec.flags;
      };
      if (tv.exchange_type == 1 /* fanout */) {
        if (tv.fanout_group != kNone) record(tv.fanout_group, load_group(tv, tv.fanout_group));
      } else {
        if (nw <= (uint32_t)kLenCap) {
          const uint32_t e1 = tv.exact_rng[2 * nw + 1];
          for (uint32_t s = tv.exact_rng[2 * nw]; s < e1; ++s) visit_shape(tv, s, row, nw, record);
        }
        for (uint32_t s = tv.hashy_begin; s < tv.hashy_end; ++s) {
          if (((__ldg(reinterpret_cast<const uint32_t*>(&tv.shapes[s]) + 3) >> 8) & 0xFFu) > nw) break;
          visit_shape(tv, s, row, nw, record);
        }
      }
    }
    // classify: 0 = count known, 1 = needs the de-duplicating merge to count, 2 = slow path
    int cls = 0;
    if (k <= 1) cnt = ub;
    else if (k > p.max_slots || ub - mxlen > p.max_small) cls = 2;
    else if (flags & kShared) cls = 1;
    else cnt = ub;

    uint64_t scr_off = 0;
    for (unsigned todo = __ballot_sync(kFull, cls == 2); todo; todo &= todo - 1) {
      const int src = __ffs(todo) - 1;
      const uint32_t mm = tile * kRouteThreads + warp * 32 + src;
      uint64_t so;
      uint32_t fl;
      const uint32_t c = warp_fallback(p, p.wid + (size_t)mm * tv.wstride, __shfl_sync(kFull, nw, src), lane, &so, &fl);
      if (lane == src) { cnt = c; scr_off = so; flags = fl; }
    }
    for (unsigned todo = __ballot_sync(kFull, cls == 1); todo; todo &= todo - 1) {
      const int src = __ffs(todo) - 1;
      const int kk = __shfl_sync(kFull, k, src);
      uint32_t o = 0, l = 0;
      if (lane < kk) {
        const uint32_t g = sm.match[lane][warp * 32 + src];
        o = __ldg(&tv.groups[g].list_off);
        l = __ldg(&tv.groups[g].list_len);
      }
      const uint32_t c = warp_union<false>(tv.qlist, o, l, kk, sm.small[warp], nullptr, 0, 0, lane);
      if (lane == src) cnt = c;
      __syncwarp();
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
    __syncthreads();
    uint64_t warp_base = 0, tile_total = 0;
#pragma unroll
    for (int w = 0; w < kRouteThreads / 32; ++w) {
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
          v = kFlagIncl; if (idx >= 0) v = tstate[idx];
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
    const uint64_t pos = sm.tile_base + warp_base + (incl - cnt);
    if (m < p.n) p.offsets[m] = (uint32_t)min(pos, (uint64_t)0xFFFFFFFFull);

    // ---- phase B: one warp per message writes the sorted union -----------------------------
    for (unsigned todo = __ballot_sync(kFull, cnt > 0); todo; todo &= todo - 1) {
      const int src = __ffs(todo) - 1;
      const uint32_t c = __shfl_sync(kFull, cnt, src);
      const int kk = __shfl_sync(kFull, k, src);
      const int cl = __shfl_sync(kFull, cls, src);
      const uint64_t o = __shfl_sync(kFull, pos, src);
      if (cl == 2) {
        const uint64_t so = __shfl_sync(kFull, scr_off, src);
        for (uint32_t i = lane; i < c; i += 32)
          if (o + i < p.ids_cap) p.ids[o + i] = p.scratch[so + i];
      } else if (kk == 1) {
        const uint32_t g = sm.match[0][warp * 32 + src];
        const uint32_t lo = __ldg(&tv.groups[g].list_off);
        for (uint32_t i = lane; i < c; i += 32)
          if (o + i < p.ids_cap) p.ids[o + i] = __ldg(tv.qlist + lo + i);
      } else {
        uint32_t lo = 0, ll = 0;
        if (lane < kk) {
          const uint32_t g = sm.match[lane][warp * 32 + src];
          lo = __ldg(&tv.groups[g].list_off);
          ll = __ldg(&tv.groups[g].list_len);
        }
        warp_union<true>(tv.qlist, lo, ll, kk, sm.small[warp], p.ids, o, p.ids_cap, lane);
        __syncwarp();
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
static unsigned long long g_launches = 0;
unsigned long long kernel_launches() { return g_launches; }

cudaError_t launch_tokenize(const uint8_t* blob, const uint32_t* off, uint32_t n, const TableView& tv, int split,
                            uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_tokenize<<<(n + 255) / 256, 256, 0, st>>>(blob, off, n, tv, split, wid, nwords, err);
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
  static int occ = 0;
  if (occ == 0) {
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_route, kRouteThreads, 0);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
  }
  const uint64_t want = (uint64_t)n_sms * occ;
  const uint32_t grid = (uint32_t)(p.n_tiles < want ? p.n_tiles : want);
  k_route<<<grid, kRouteThreads, 0, st>>>(p);
  ++g_launches;
  return cudaGetLastError();
}

}  // namespace chq
