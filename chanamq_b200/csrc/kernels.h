// Launch interface between the C-ABI layer and kernels.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "chq_common.h"

#ifndef CHQ_ROUTE_SMALL
#define CHQ_ROUTE_SMALL 128
#endif

#include <nvtx3/nvToolsExt.h>   // header-only: ranges cost nothing unless a profiler is attached

namespace chq {

// NVTX range for the scope (tracing hooks of the shim) — nsys / ncu show submit, kernels and copies by name
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

constexpr int kRouteThreads = 128;   // messages per tile; one thread per message while probing
constexpr int kRouteSlots = 8;       // matched groups remembered per message before the slow path
constexpr int kRouteSmall = CHQ_ROUTE_SMALL;     // ids (all lists but the longest) a warp merges in shared memory

// control words (device): zeroed before every launch
constexpr int kCtrlTile = 0 /* scan tile counter */, kCtrlStatus = 1, kCtrlScratchLo = 2 /* u64 bump pointer */,
              kCtrlChunk = 4 /* write-kernel work counter */, kCtrlDescBump = 5 /* headers descriptor arena */,
              kCtrlRestBig = 6 /* messages k_write leaves to k_write_rest: long results (done first) */,
              kCtrlRest = 7 /* ... and the others */, kCtrlRestNext = 8 /* k_write_rest's work counter */,
              kCtrlWords = 16;
constexpr int kRouteScanTile = 1024;  // counts per tile of the offsets scan
constexpr int kStateHead = kCtrlWords / 2;  // the control words take the first kStateHead u64 of the state array
inline uint32_t route_scan_tiles(uint32_t n) { return (n + kRouteScanTile - 1) / kRouteScanTile; }

constexpr uint32_t kStatusOverflow = 1, kStatusScratch = 2;

struct RouteParams {
  TableView tv;
  const uint32_t* wid;       // [n][tv.wstride] interned words
  const uint8_t* nwords;     // [n] word count after Java split (0 = error / nothing to route)
  uint32_t n, n_tiles;
  uint32_t* offsets;         // [n+1] out
  uint32_t* ids;             // out
  uint64_t ids_cap;
  uint8_t* durable;          // [n] out
  uint64_t* total;           // out: number of ids
  uint32_t* cnt;             // [n] scratch: ids per message (k_probe -> k_offsets, k_write)
  uint32_t* meta;            // [n] scratch: lists | class << 8 | flags
  uint2* mlist;              // [n][kRouteSlots] scratch: the message's lists (offset, length)
  uint32_t* rest;            // [n] scratch: the messages k_write leaves to k_write_rest (long results from the front,
                             //     the others from the back)
  uint64_t* tile_state;      // [n_tiles] scratch of the offsets scan, zeroed before launch
  uint32_t* ctrl;            // [kCtrlWords] scratch, zeroed before launch
  uint32_t* scratch;         // slow-path arena
  uint64_t scratch_cap;      // in ids
  uint32_t max_slots, max_small;
  // slice-merge mode (binding-slice sharding): instead of probing a table, message m's lists are slice r's
  // ids slice_ids[r * slice_cap + slice_offsets[r * (n+1) + m] ...] for r < n_slices; tv.qlist = slice_ids
  const uint32_t* slice_offsets;   // null = normal routing
  const uint8_t* slice_durable;    // [n_slices][n]
  uint32_t n_slices;
  uint64_t slice_cap;
  uint32_t slice_disjoint;   // no queue id occurs in two slices: the merge need not de-duplicate
  // descriptor mode (headers exchange): message m's lists are desc_lists[desc_start[m] + j], j < desc_k[m]
  const uint32_t* desc_k;          // null = not this mode
  const uint32_t* desc_start;      // message m's descriptors are desc_lists[desc_start[m] ...]
  const uint8_t* desc_flags;
  const uint2* desc_lists;
  uint32_t debug;            // experiment switches (env CHQ_DEBUG): 1 = skip phase B, 2 = skip probing
};

cudaError_t launch_tokenize(const uint8_t* blob, const uint32_t* off, uint32_t n, const TableView& tv, int split,
                            uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st);
cudaError_t launch_tokenize_keys(const uint8_t* bytes, const uint32_t* key_off, const uint32_t* key_len, uint32_t n,
                                 const TableView& tv, int split, uint32_t* wid, uint8_t* nwords, uint8_t* err, cudaStream_t st);
void note_launches(int n);
cudaError_t launch_headers_match(const uint8_t* blob, const uint32_t* off, const uint32_t* hlen, uint32_t n, const TableView& tv,
                                 uint32_t* desc_k, uint32_t* desc_start, uint8_t* desc_flags, uint2* desc_lists, uint32_t arena_cap,
                                 uint32_t* arena_used, uint32_t* status_word, uint8_t* err, cudaStream_t st);
// between: optional three events, recorded after k_probe, k_offsets and k_write (per-kernel timing)
cudaError_t launch_route(const RouteParams& p, int n_sms, cudaStream_t st, cudaEvent_t* between = nullptr);
cudaError_t launch_finish(const uint32_t* ctrl, int32_t* d_status, cudaStream_t st);
// pack.cu: the ids of a finished batch as 3-byte values (total_status = the launch's {ids, status} pair on the device)
cudaError_t launch_pack24(const uint32_t* ids, const uint64_t* total_status, uint64_t ids_cap, uint32_t* out, int n_sms,
                          cudaStream_t st);
unsigned long long kernel_launches();
cudaError_t read_clocks(unsigned long long* out32);

}  // namespace chq
