// Headers-exchange matching on the device (K2 of SURVEY.md §2b; SPEC mode only — the reference has no
// headers matcher, ExchangeEntity.scala:210, and routes such exchanges by routing key, which is what
// COMPAT mode does through the topic path).
//
// k_headers_match: one thread per message walks the message's `headers` field table (raw AMQP bytes,
// ValueReader.scala:62-113 layout), interns each field into the atoms the bindings ask about
// ("field present", "field has this typed value"), and collects the queue lists that apply:
//   - the list of `any` bindings of every atom the message satisfies,
//   - every `all` group anchored at one of those atoms whose whole atom set is satisfied,
//   - the `all` bindings without arguments.
// The lists go to k_route (descriptor mode), which writes their sorted, de-duplicated union as CSR —
// the same union machinery as topic routing.
#include <cuda_runtime.h>
#include <stdint.h>

#include "chq_common.h"
#include "kernels.h"
#include "probe.cuh"

namespace chq {

namespace {

// (name id, type tag, value bytes [q, q + vl)) -> atom id, 0 if no binding asks for that value
__device__ uint32_t value_atom_lookup(const TableView& tv, uint32_t name_id, uint32_t tag, const uint32_t* __restrict__ ab,
                                      uint32_t q, uint32_t vl, uint32_t rlo, uint32_t rhi, const uint8_t* __restrict__ v) {
  const uint32_t key = (name_id << 8) | tag;
  uint32_t f_lo = 0, f_hi = 0;
  const uint32_t h = hash_bytes(ab, q, vl, rlo, rhi, kWordBasis ^ (key * 0x9E3779B1u), f_lo, f_hi);
  uint32_t i = h & tv.vdict_mask;
  for (;;) {
    const uint4 a = ldg16(&tv.vdict[i]);
    const uint4 b = ldg16(reinterpret_cast<const uint8_t*>(&tv.vdict[i]) + 16);
    if (b.y == 0) return 0;
    if (a.x == h && a.y == 0 && a.z == f_lo && a.w == f_hi && b.x == vl && b.w == key) {
      bool same = true;
      if (vl > 8) {
        const uint8_t* d = tv.vdict_bytes + b.z;
        for (uint32_t j = 8; j < vl; ++j)
          if (__ldg(d + j) != __ldg(v + j)) { same = false; break; }
      }
      if (same) return b.y;
    }
    i = (i + 1) & tv.vdict_mask;
  }
}

}  // namespace

__global__ void __launch_bounds__(128) k_headers_match(const uint8_t* __restrict__ blob, const uint32_t* __restrict__ off,
                                                       const uint32_t* __restrict__ hlen, uint32_t n, TableView tv,
                                                       uint32_t* __restrict__ desc_k, uint32_t* __restrict__ desc_start,
                                                       uint8_t* __restrict__ desc_flags, uint2* __restrict__ desc_lists,
                                                       uint32_t arena_cap, uint32_t* __restrict__ arena_used,
                                                       uint32_t* __restrict__ status_word, uint8_t* __restrict__ err) {
  const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n) return;  // only whole tail lanes leave: the warp-level allocation below uses the active mask
  const uint32_t b0 = off[m], b1 = hlen ? b0 + hlen[m] : off[m + 1];
  // names and values are read as aligned 32-bit words (probe.cuh): offsets relative to the aligned base
  const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(blob) & 3u);
  const uint32_t* ab = reinterpret_cast<const uint32_t*>(blob - mis);
  const uint32_t rlo = b0 + mis, rhi = b1 + mis;
  uint32_t atoms[2 * kMaxHeaderFields];
  uint32_t seen[kMaxHeaderFields];   // names with id >= 64 already met (the others are bits of seen_lo)
  uint64_t seen_lo = 0;
  int na = 0, ns = 0, ns_hi = 0;
  uint32_t status = 0;  // 3 = malformed table (the reference would have thrown while decoding), 4 = too many fields
  uint32_t p = b0;
  while (p < b1 && !status) {
    const uint32_t nl = __ldg(blob + p++);
    if (p + nl >= b1) { status = 3; break; }  // the name and at least a tag must fit
    uint32_t f_lo = 0, f_hi = 0;
    const uint32_t nh = hash_bytes(ab, p + mis, nl, rlo, rhi, kWordBasis, f_lo, f_hi);
    const uint32_t name_id = word_lookup(tv, nh, f_lo, f_hi, nl, blob + p);
    p += nl;
    const uint32_t tag = __ldg(blob + p++);
    uint32_t vl = 0;
    switch (tag) {
      case 'S': case 'x': case 'F': case 'A':
        if (p + 4 > b1) { status = 3; break; }
        vl = (__ldg(blob + p) << 24) | (__ldg(blob + p + 1) << 16) | (__ldg(blob + p + 2) << 8) | __ldg(blob + p + 3);
        p += 4;
        break;
      case 'I': case 'f': vl = 4; break;
      case 'D': vl = 5; break;
      case 'T': case 'd': case 'l': vl = 8; break;
      case 'b': case 't': vl = 1; break;
      case 's': vl = 2; break;
      case 'V': vl = 0; break;
      default: status = 3;
    }
    if (status) break;
    if ((uint64_t)p + vl > b1) { status = 3; break; }
    if (name_id != kUnknownWord) {
      bool dup = false;  // a name that occurs twice: the first occurrence counts (ValueReader.scala:71-72)
      if (name_id < 64) {
        dup = (seen_lo >> name_id) & 1ull;
        seen_lo |= 1ull << name_id;
      } else {
        for (int i = 0; i < ns_hi; ++i) dup = dup || seen[i] == name_id;
      }
      if (!dup) {
        if (ns == kMaxHeaderFields) { status = 4; break; }
        ++ns;
        if (name_id >= 64) seen[ns_hi++] = name_id;
        const uint32_t pa = __ldg(tv.presence_atom + name_id);
        if (pa) atoms[na++] = pa;
        const uint32_t va = value_atom_lookup(tv, name_id, tag, ab, p + mis, vl, rlo, rhi, blob + p);
        if (va) atoms[na++] = va;
      }
    }
    p += vl;
  }
  // the message's atoms in ascending order (insertion sort: a dozen entries) and their 64-bit signature
  uint32_t sig_lo = 0, sig_hi = 0;
  for (int i = 0; i < na; ++i) {
    const uint32_t x = atoms[i];
    if (x & 32u) sig_hi |= 1u << (x & 31u); else sig_lo |= 1u << (x & 31u);
    int j = i;
    for (; j > 0 && atoms[j - 1] > x; --j) atoms[j] = atoms[j - 1];
    atoms[j] = x;
  }
  // Every matching list becomes one (offset, length) descriptor.  The matches are enumerated twice:
  // once to count, so that the warp can take one contiguous piece of the descriptor arena with a
  // single atomic, once to write.
  uint32_t k = 0, flags = 0;
  uint2* lists = nullptr;
  constexpr uint32_t kKept = 24;  // descriptors found by the counting pass are kept, so most messages enumerate once
  uint2 kept[kKept];
  auto for_each_match = [&](bool write) {
    uint32_t j = 0;
    auto emit = [&](uint32_t o, uint32_t l, uint32_t fl) {
      if (l == 0) return;
      if (write) lists[j] = make_uint2(o, l);
      else if (j < kKept) kept[j] = make_uint2(o, l);
      ++j;
      flags |= fl;
    };
    emit(tv.always_off, tv.always_len, tv.always_flags);
    for (int i = 0; i < na; ++i) {
      const uint8_t* rec = reinterpret_cast<const uint8_t*>(&tv.atoms[atoms[i]]);
      const uint4 a = ldg16(rec);
      const uint32_t all_end = __ldg(reinterpret_cast<const uint32_t*>(rec) + 4);
      emit(a.x, a.y, a.z);
      for (uint32_t g = a.w; g < all_end; ++g) {  // `all` groups anchored here: is every atom of the group satisfied?
        const uint8_t* gr = reinterpret_cast<const uint8_t*>(&tv.all_groups[g]);
        const uint4 ga = ldg16(gr);
        const uint4 gb = ldg16(gr + 16);  // flags, signature
        bool ok = (gb.y & ~sig_lo) == 0 && (gb.z & ~sig_hi) == 0;
        // both atom lists ascend: one forward walk over the message's atoms finds them all or stops
        int q = 0;
        for (uint32_t t = 0; t < ga.y && ok; ++t) {
          const uint32_t need = __ldg(tv.all_atoms + ga.x + t);
          while (q < na && atoms[q] < need) ++q;
          ok = q < na && atoms[q] == need;
          ++q;
        }
        if (ok) emit(ga.z, ga.w, gb.x);
      }
    }
    return j;
  };
  if (!status) k = for_each_match(false);
  // warp-aggregated allocation (all 32 lanes arrive here: no early return above for m < n lanes of a full warp)
  __syncwarp();  // reconverge after the data-dependent loops (lanes that left at the top are exempt)
  const unsigned active = __activemask();
  const int lane = threadIdx.x & 31;
  uint32_t incl = k;
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t v = __shfl_up_sync(active, incl, d);
    if (lane >= d) incl += v;
  }
  const int last = 31 - __clz(active);
  const uint32_t warp_total = __shfl_sync(active, incl, last);
  uint32_t base = 0;
  if (lane == last && warp_total) base = atomicAdd(arena_used, warp_total);
  base = __shfl_sync(active, base, last);
  uint32_t start = base + incl - k;
  if (k && (uint64_t)start + k > arena_cap) {  // arena exhausted: the launch reports CHQ_E_SCRATCH
    atomicOr(status_word, kStatusScratch);
    status = 4;
    k = 0;
  }
  if (k) {
    lists = desc_lists + start;
    if (k <= kKept) {
      for (uint32_t j = 0; j < k; ++j) lists[j] = kept[j];
    } else {
      flags = 0;
      for_each_match(true);
    }
  }
  if (status) { k = 0; flags = 0; start = 0; }
  desc_k[m] = k;
  desc_start[m] = start;
  desc_flags[m] = (uint8_t)flags;
  err[m] = (uint8_t)status;
}

cudaError_t launch_headers_match(const uint8_t* blob, const uint32_t* off, const uint32_t* hlen, uint32_t n, const TableView& tv,
                                 uint32_t* desc_k, uint32_t* desc_start, uint8_t* desc_flags, uint2* desc_lists, uint32_t arena_cap,
                                 uint32_t* arena_used, uint32_t* status_word, uint8_t* err, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  k_headers_match<<<(n + 127) / 128, 128, 0, st>>>(blob, off, hlen, n, tv, desc_k, desc_start, desc_flags, desc_lists, arena_cap,
                                                    arena_used, status_word, err);
  note_launches(1);
  return cudaGetLastError();
}

}  // namespace chq
