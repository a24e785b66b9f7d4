// Device-side probing of the compiled table (included by kernels.cu only): word dictionary,
// tuple-space hash probe with exact verification, and the word-level NFA for '#' remainders.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "chq_common.h"

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

// What one key needs for all of its probes, computed once: the Zobrist term of each of its first
// eight words (two 128-bit loads of the word-id row) and which of them no binding mentions.
struct KeyCtx {
  uint32_t z[8];
  const uint32_t* row;
  uint32_t nw;
  uint32_t unknown;   // bit p: word p is kUnknownWord (can only meet a wildcard)
};
__device__ __forceinline__ void key_ctx_init(KeyCtx& k, const uint32_t* __restrict__ row, uint32_t nw, uint32_t wstride) {
  k.row = row;
  k.nw = nw;
  uint32_t w[8];
  const uint4 lo = ldg16(row);
  w[0] = lo.x; w[1] = lo.y; w[2] = lo.z; w[3] = lo.w;
  if (wstride >= 8 && nw > 4) {
    const uint4 hi = ldg16(row + 4);
    w[4] = hi.x; w[5] = hi.y; w[6] = hi.z; w[7] = hi.w;
  } else {
    w[4] = w[5] = w[6] = w[7] = 0;
  }
  k.unknown = 0;
#pragma unroll
  for (int p = 0; p < 8; ++p) {
    k.z[p] = zob(p, w[p]);
    if (w[p] == kUnknownWord || (uint32_t)p >= nw) k.unknown |= 1u << p;
  }
}

// Head group of the chain whose (shape, literal tuple) equals the key's words at the shape's
// literal positions, or kNone.  Exact: on a tag hit the stored tuple is compared word by word.
__device__ __forceinline__ uint32_t probe_head(const TableView& tv, uint32_t shape_idx, const Shape& sh, const KeyCtx& k) {
  if (sh.lit_mask & k.unknown & 0xFFu) return kNone;
  uint32_t h = (uint32_t)sh.seed;
#pragma unroll
  for (int p = 0; p < 8; ++p)
    if ((sh.lit_mask >> p) & 1u) h ^= k.z[p];
  for (uint32_t m = sh.lit_mask >> 8; m; m &= m - 1) {  // literal positions 8..15: rare, on the fly
    const uint32_t p = 8 + (__ffs(m) - 1);
    const uint32_t w = __ldg(k.row + p);
    if (w == kUnknownWord) return kNone;
    h ^= zob(p, w);
  }
  const uint32_t tag = tag32(h);
  uint32_t b = h & tv.bucket_mask;
  for (;;) {
    const uint4 s01 = ldg16(&tv.slots[(size_t)b * 4]);
    const uint4 s23 = ldg16(&tv.slots[(size_t)b * 4 + 2]);
    const bool any_hit = (s01.x == tag) | (s01.z == tag) | (s23.x == tag) | (s23.z == tag);
    if (any_hit) {
      const uint32_t tags[4] = {s01.x, s01.z, s23.x, s23.z};
      const uint32_t grps[4] = {s01.y, s01.w, s23.y, s23.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (tags[i] != tag) continue;
        const uint32_t g = grps[i];
        const Group* gp = &tv.groups[g];
        if (__ldg(&gp->shape) != shape_idx) continue;
        const uint32_t lit_off = __ldg(&gp->lit_off);
        bool same = true;
        uint32_t li = 0;
        for (uint32_t m = sh.lit_mask; m; m &= m - 1, ++li)
          if (__ldg(tv.lits + lit_off + li) != __ldg(k.row + (__ffs(m) - 1))) { same = false; break; }
        if (same) return g;
      }
    }
    if ((s01.x == 0) | (s01.z == 0) | (s23.x == 0) | (s23.z == 0)) return kNone;  // bucket not full: chain ends
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
__device__ __forceinline__ void visit_shape(const TableView& tv, uint32_t s, const KeyCtx& k, F&& f) {
  const Shape sh = load_shape(tv, s);
  uint32_t g = probe_head(tv, s, sh, k);
  while (g != kNone) {
    const GroupRec rec = load_group(tv, g);
    if (rec.rem_len == 0 || remainder_matches(tv, rec, k.row, sh.len, k.nw)) f(g, rec);
    g = rec.next;
  }
}

}  // namespace
}  // namespace chq
