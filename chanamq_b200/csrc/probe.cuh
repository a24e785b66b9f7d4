// Device-side probing of the compiled table (included by kernels.cu only): word dictionary,
// tuple-space hash probe with exact verification, and the word-level NFA for '#' remainders.
//
// Probing is latency work (random 16-byte reads of an L2-resident tag array), so it is organised
// for memory-level parallelism: a key hashes FOUR shapes at a time and issues their tag loads
// together; tag hits are only noted (slot, shape) and resolved afterwards — group record and
// literal tuple arrive in one 64-byte read — so no dependent load sits inside the probe loop.
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

// ---- bytes as aligned 32-bit words ---------------------------------------------------------------------
// Byte offsets are relative to a 4-byte aligned base `ab`; [rlo, rhi) is the range that may be read.  A word
// that straddles the edge of the range is assembled from byte loads, so nothing outside it is ever touched.
__device__ __forceinline__ uint32_t key_word32(const uint32_t* __restrict__ ab, uint32_t k, uint32_t rlo, uint32_t rhi) {
  const uint32_t lo = 4 * k;
  if (lo >= rhi) return 0;
  if (lo >= rlo && lo + 4 <= rhi) return __ldg(ab + k);
  uint32_t w = 0;
  const uint8_t* b = reinterpret_cast<const uint8_t*>(ab);
#pragma unroll
  for (uint32_t j = 0; j < 4; ++j)
    if (lo + j >= rlo && lo + j < rhi) w |= (uint32_t)__ldg(b + lo + j) << (8 * j);
  return w;
}
// eight bytes from byte offset q, little-endian (lo = bytes 0..3)
__device__ __forceinline__ void load8(const uint32_t* __restrict__ ab, uint32_t q, uint32_t rlo, uint32_t rhi, uint32_t& lo,
                                      uint32_t& hi) {
  const uint32_t k = q >> 2, sh = (q & 3u) * 8u;
  const uint32_t w0 = key_word32(ab, k, rlo, rhi), w1 = key_word32(ab, k + 1, rlo, rhi);
  const uint32_t w2 = sh ? key_word32(ab, k + 2, rlo, rhi) : 0u;
  lo = __funnelshift_r(w0, w1, sh);
  hi = __funnelshift_r(w1, w2, sh);
}
// keeps the first `take` (0..8) bytes of (lo, hi)
__device__ __forceinline__ void keep_bytes(uint32_t take, uint32_t& lo, uint32_t& hi) {
  if (take >= 8) return;
  if (take >= 4) hi &= (1u << (8 * (take - 4))) - 1u;
  else { hi = 0; lo &= (1u << (8 * take)) - 1u; }
}
// word_fin(chunks of word_chunk over bytes [q, q + len), len): the hash the host computes with bytes_hash();
// (f_lo, f_hi) = the first eight bytes, zero padded
__device__ __forceinline__ uint32_t hash_bytes(const uint32_t* __restrict__ ab, uint32_t q, uint32_t len, uint32_t rlo,
                                               uint32_t rhi, uint32_t seed, uint32_t& f_lo, uint32_t& f_hi) {
  uint32_t h = seed, done = 0;
  do {
    uint32_t lo, hi;
    load8(ab, q + done, rlo, rhi, lo, hi);
    keep_bytes(len - done, lo, hi);
    if (done == 0) { f_lo = lo; f_hi = hi; }
    h = word_chunk(h, lo, hi);
    done += 8;
  } while (done < len);
  return word_fin(h, len);
}

// Routing-key word -> id (topic / direct tables): one 16-byte load per probe step.  h = word_fin(...),
// (lo, hi) = the first 8 bytes, bytes = the whole word (read only to confirm a word longer than 8 bytes).
__device__ __forceinline__ uint32_t word_lookup(const TableView& tv, uint32_t h, uint32_t lo, uint32_t hi, uint32_t len,
                                                const uint8_t* bytes) {
  const uint32_t want = (h & 0xFFFFFF00u) | len;
  uint32_t i = h & tv.wdict_mask;
  for (;;) {
    const uint4 a = ldg16(&tv.wdict[i]);
    if (a.z == 0) return kUnknownWord;
    if (a.x == lo && a.y == hi && a.w == want) {
      if (len <= 8) return a.z;
      const uint8_t* d = tv.dict_bytes + __ldg(tv.wdict_off + i);
      bool same = true;
      for (uint32_t j = 8; j < len; ++j)
        if (__ldg(d + j) != __ldg(bytes + j)) { same = false; break; }
      if (same) return a.z;
    }
    i = (i + 1) & tv.wdict_mask;
  }
}

// ------------------------------------------------------------------------------------------
// table records
// ------------------------------------------------------------------------------------------
struct GroupRec {
  uint32_t list_off, list_len, next, shape, flags;
};
// The head 16 bytes of a group record (one lane-load).
__device__ __forceinline__ GroupRec load_group(const TableView& tv, uint32_t g) {
  const uint4 a = ldg16(&tv.groups[g]);
  GroupRec r;
  r.list_off = a.x; r.list_len = a.y; r.next = a.z;
  r.shape = a.w & 0xFFFFu; r.flags = a.w >> 24;
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
  uint32_t unknown;   // bit p: word p is kUnknownWord or past the key's end (can only meet a wildcard)
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

// Hash of the key's words at the shape's literal positions; false when one of them is a word no
// binding mentions (then nothing under this shape can match).
__device__ __forceinline__ bool shape_hash(const Shape& sh, const KeyCtx& k, uint32_t& h_out) {
  if (sh.lit_mask & k.unknown & 0xFFu) return false;
  uint32_t h = (uint32_t)sh.seed;
#pragma unroll
  for (int p = 0; p < 8; ++p)
    if ((sh.lit_mask >> p) & 1u) h ^= k.z[p];
  for (uint32_t m = sh.lit_mask >> 8; m; m &= m - 1) {  // literal positions 8..15: rare, on the fly
    const uint32_t p = 8 + (__ffs(m) - 1);
    const uint32_t w = __ldg(k.row + p);
    if (w == kUnknownWord) return false;
    h ^= zob(p, w);
  }
  h_out = h;
  return true;
}

// eight 16-bit tags against one: bit i of the result = slot i matches; full = no empty slot.
// Slots of a bucket are filled in order at build time and a tag is never 0, so the bucket is full exactly
// when its last slot is taken.  zero16() flags the zero halfwords of a word (bits 15 and 31), exactly.
__device__ __forceinline__ uint32_t zero16(uint32_t x) {
  const uint32_t y = (x & 0x7FFF7FFFu) + 0x7FFF7FFFu;
  return ~(y | x | 0x7FFF7FFFu);
}
__device__ __forceinline__ uint32_t match_tags(const uint4& t, uint32_t tag, bool& full) {
  const uint32_t tt = tag * 0x10001u;
  const uint32_t c0 = zero16(t.x ^ tt), c1 = zero16(t.y ^ tt), c2 = zero16(t.z ^ tt), c3 = zero16(t.w ^ tt);
  full = (t.w >> 16) != 0;
  if ((c0 | c1 | c2 | c3) == 0) return 0;
  return ((c0 >> 15) & 1u) | ((c0 >> 30) & 2u) | ((c1 >> 13) & 4u) | ((c1 >> 28) & 8u) | ((c2 >> 11) & 16u) |
         ((c2 >> 26) & 32u) | ((c3 >> 9) & 64u) | ((c3 >> 24) & 128u);
}

__device__ bool remainder_matches(const TableView& tv, uint32_t rem_off, uint32_t rem_len, const uint32_t* __restrict__ row,
                                  uint32_t from, uint32_t nw);

// Exact check of a candidate: the group must belong to the probing shape and its literal tuple
// must equal the key's words at the shape's literal positions.  Positions 0..7 are compared as
// zob values held in registers (zob(p, .) is a bijection, so this is word equality); the second
// half of zpos[] is only fetched when the shape has a literal there.
__device__ __forceinline__ bool verify_group(const TableView& tv, uint32_t g, const GroupRec& rec, uint32_t shape_idx,
                                             uint32_t lit_mask, const KeyCtx& k) {
  if (rec.shape != shape_idx) return false;
  const uint8_t* base = reinterpret_cast<const uint8_t*>(&tv.groups[g]);
  const uint4 lo = ldg16(base + 16);
  bool ok = true;
  if (lit_mask & 1u) ok = ok && lo.x == k.z[0];
  if (lit_mask & 2u) ok = ok && lo.y == k.z[1];
  if (lit_mask & 4u) ok = ok && lo.z == k.z[2];
  if (lit_mask & 8u) ok = ok && lo.w == k.z[3];
  if (lit_mask & 0xF0u) {
    const uint4 hi = ldg16(base + 32);
    if (lit_mask & 16u) ok = ok && hi.x == k.z[4];
    if (lit_mask & 32u) ok = ok && hi.y == k.z[5];
    if (lit_mask & 64u) ok = ok && hi.z == k.z[6];
    if (lit_mask & 128u) ok = ok && hi.w == k.z[7];
  }
  if (ok && (lit_mask >> 8)) {  // literals at positions 8..15: compare the word ids themselves
    const uint32_t lit_off = __ldg(&tv.groups[g].lit_off);
    uint32_t i = __popc(lit_mask & 0xFFu);
    for (uint32_t m = lit_mask >> 8; m && ok; m &= m - 1, ++i)
      ok = __ldg(tv.lits + lit_off + i) == __ldg(k.row + 8 + (__ffs(m) - 1));
  }
  return ok;
}
// the remainder of group g (if any) against the key's words from position `from`
__device__ __forceinline__ bool group_rem_ok(const TableView& tv, uint32_t g, const GroupRec& rec, const KeyCtx& k, uint32_t from) {
  if (!(rec.flags & kHasRem)) return true;
  return remainder_matches(tv, __ldg(&tv.groups[g].rem_off), __ldg(&tv.groups[g].rem_len), k.row, from, k.nw);
}

// Word-level NFA for a pattern remainder (literals, '*', '#') against key words [from, nw).
// State s = remainder items consumed so far; '#' may take zero or more words.
__device__ bool remainder_matches(const TableView& tv, uint32_t rem_off, uint32_t rem_len, const uint32_t* __restrict__ row,
                                  uint32_t from, uint32_t nw) {
  const uint32_t r = rem_len;
  const uint32_t* rem = tv.rems + rem_off;
  // reach bitset over states 0..r: a pattern is at most 255 bytes, so r <= kMaxWords - 1 items
  constexpr int kW = kMaxWords / 32 + 1;
  if (r > (uint32_t)kMaxWords) return false;   // (the host never builds such a remainder)
  uint32_t cur[kW], nxt[kW];
#pragma unroll
  for (int i = 0; i < kW; ++i) cur[i] = 0;
  cur[0] = 1u;
  auto close = [&](uint32_t* st) {  // epsilon: a '#' can be skipped
    for (uint32_t s = 0; s < r; ++s)
      if ((st[s >> 5] >> (s & 31)) & 1u)
        if (__ldg(rem + s) == kHash) st[(s + 1) >> 5] |= 1u << ((s + 1) & 31);
  };
  close(cur);
  for (uint32_t p = from; p < nw; ++p) {
    const uint32_t w = __ldg(row + p);
#pragma unroll
    for (int i = 0; i < kW; ++i) nxt[i] = 0;
    for (uint32_t s = 0; s < r; ++s) {
      if (!((cur[s >> 5] >> (s & 31)) & 1u)) continue;
      const uint32_t item = __ldg(rem + s);
      if (item == kHash) nxt[s >> 5] |= 1u << (s & 31);
      else if (item == kStar || (item == w && w != kUnknownWord)) nxt[(s + 1) >> 5] |= 1u << ((s + 1) & 31);
    }
    close(nxt);
#pragma unroll
    for (int i = 0; i < kW; ++i) cur[i] = nxt[i];
  }
  return (cur[r >> 5] >> (r & 31)) & 1u;
}

// Pass 1 of a key's lookup: probe every shape in [e0, e0+ne) and [h0, h0+nh), four at a time.
// on_cand(slot index, shape index, literal mask) is called for every tag hit; nothing is dereferenced here.
template <typename F>
__device__ __forceinline__ void probe_shapes(const TableView& tv, const KeyCtx& k, uint32_t e0, uint32_t ne, uint32_t h0,
                                             uint32_t nh, F&& on_cand) {
#ifndef CHQ_PROBE_BATCH
#define CHQ_PROBE_BATCH 1
#endif
  constexpr int kB = CHQ_PROBE_BATCH;
  const uint32_t total = ne + nh;
  for (uint32_t i = 0; i < total; i += kB) {
    uint32_t sidx[kB], hh[kB], lmask[kB];
    uint4 tg[kB];
    bool live[kB];
#pragma unroll
    for (int u = 0; u < kB; ++u) {
      const uint32_t ii = i + u;
      live[u] = ii < total;
      sidx[u] = ii < ne ? e0 + ii : h0 + (ii - ne);
      hh[u] = 0;
      lmask[u] = 0;
      if (live[u]) {
        const Shape sh = load_shape(tv, sidx[u]);
        lmask[u] = sh.lit_mask;
        live[u] = shape_hash(sh, k, hh[u]);
      }
      if (live[u]) tg[u] = ldg16(tv.tags + (size_t)(hh[u] & tv.bucket_mask) * kBucketSlots);
    }
#pragma unroll
    for (int u = 0; u < kB; ++u) {
      if (!live[u]) continue;
      const uint32_t tag = tag16(hh[u]);
      uint32_t b = hh[u] & tv.bucket_mask;
      uint4 t = tg[u];
      for (;;) {
        bool full;
        uint32_t m = match_tags(t, tag, full);
        for (; m; m &= m - 1) on_cand(b * kBucketSlots + (__ffs(m) - 1), sidx[u], lmask[u]);
        if (!full) break;
        b = (b + 1) & tv.bucket_mask;  // the bucket overflowed at build time: keep going (rare)
        t = ldg16(tv.tags + (size_t)b * kBucketSlots);
      }
    }
  }
}

// Pass 2: resolve one candidate.  f(record) is called for every group that really matches.
template <typename F>
__device__ __forceinline__ void resolve_candidate(const TableView& tv, const KeyCtx& k, uint32_t slot, uint32_t shape_idx,
                                                  uint32_t lit_mask, F&& f) {
  uint32_t g = __ldg(tv.slot_group + slot);
  GroupRec rec = load_group(tv, g);
  if (!verify_group(tv, g, rec, shape_idx, lit_mask, k)) return;
  for (;;) {  // a chain shares one literal tuple and differs in its remainders
    if (!(rec.flags & kHasRem)) f(rec);
    else if (group_rem_ok(tv, g, rec, k, (uint32_t)__ldg(&tv.shapes[shape_idx].len))) f(rec);
    g = rec.next;
    if (g == kNone) return;
    rec = load_group(tv, g);
  }
}

// Slow-path iterator: all groups matching this key under ONE shape, one per call (the warp-level
// slow path gives every lane its own shape and steps the lanes together).
struct LaneMatches {
  uint4 t;
  uint32_t shape, lit_mask, len, tag, b, m, g;
  bool live, full;
};
__device__ __forceinline__ void lane_matches_begin(LaneMatches& lm, const TableView& tv, const KeyCtx& k, bool active,
                                                   uint32_t shape_idx) {
  lm.live = active;
  lm.g = kNone;
  lm.m = 0;
  lm.full = false;
  if (!active) return;
  const Shape sh = load_shape(tv, shape_idx);
  uint32_t h = 0;
  lm.live = sh.len <= k.nw && shape_hash(sh, k, h);
  if (!lm.live) return;
  lm.shape = shape_idx;
  lm.lit_mask = sh.lit_mask;
  lm.len = sh.len;
  lm.tag = tag16(h);
  lm.b = h & tv.bucket_mask;
  lm.t = ldg16(tv.tags + (size_t)lm.b * kBucketSlots);
  lm.m = match_tags(lm.t, lm.tag, lm.full);
}
__device__ __forceinline__ bool lane_matches_next(LaneMatches& lm, const TableView& tv, const KeyCtx& k, GroupRec& rec) {
  while (lm.live) {
    if (lm.g != kNone) {  // further groups with the same literal tuple (other remainders)
      const uint32_t g = lm.g;
      rec = load_group(tv, g);
      lm.g = rec.next;
      if (group_rem_ok(tv, g, rec, k, lm.len)) return true;
    } else if (lm.m) {
      const uint32_t slot = lm.b * kBucketSlots + (__ffs(lm.m) - 1);
      lm.m &= lm.m - 1;
      const uint32_t g = __ldg(tv.slot_group + slot);
      rec = load_group(tv, g);
      if (verify_group(tv, g, rec, lm.shape, lm.lit_mask, k)) {
        lm.g = rec.next;
        if (group_rem_ok(tv, g, rec, k, lm.len)) return true;
      }
    } else if (lm.full) {
      lm.b = (lm.b + 1) & tv.bucket_mask;
      lm.t = ldg16(tv.tags + (size_t)lm.b * kBucketSlots);
      lm.m = match_tags(lm.t, lm.tag, lm.full);
    } else {
      lm.live = false;
    }
  }
  return false;
}

}  // namespace
}  // namespace chq
