// The packed form of a batch's queue ids (no counterpart in the reference): 3 bytes per id instead of 4.
//
// End to end, chq_route is the device->host copy of the ids and nothing else (415 MB per 2^20-key batch
// on the headline table against 0.43 ms of kernels).  A table with fewer than 2^24 queues can ship a quarter less:
// k_pack24 rewrites the CSR's id array as little-endian 3-byte ids, back to back, and chq_route_wait copies that.
// The offsets still count ids, so message i's ids are bytes [3 offsets[i], 3 offsets[i+1]) of the packed array; the
// consumer reads them in place (b[3j] | b[3j+1] << 8 | b[3j+2] << 16) or expands them with chq_unpack_u24.
#include <cuda_runtime.h>
#include <stdint.h>

#include "chq_common.h"
#include "kernels.h"

namespace chq {

// Four ids (one 16-byte load) become three words.  `total_status` is the route launch's pair on the device:
// [0] = number of ids, [1] low word = status; a batch that failed (overflow: run again by chq_route_wait) is left alone.
__global__ void __launch_bounds__(256) k_pack24(const uint32_t* __restrict__ ids, const uint64_t* __restrict__ total_status,
                                                uint64_t ids_cap, uint32_t* __restrict__ out) {
  if ((uint32_t)total_status[1] != 0) return;
  const uint64_t total = total_status[0];
  if (total > ids_cap) return;
  const uint64_t groups = (total + 3) / 4, stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    uint4 v;
    if (4 * g + 4 <= total) {
      v = __ldcs(reinterpret_cast<const uint4*>(ids) + g);
    } else {  // the last, partial group: padded with zeros
      const uint64_t b = 4 * g;
      v.x = ids[b];
      v.y = b + 1 < total ? ids[b + 1] : 0u;
      v.z = b + 2 < total ? ids[b + 2] : 0u;
      v.w = 0u;
    }
    uint32_t* o = out + 3 * g;
    __stcs(o, (v.x & 0xFFFFFFu) | (v.y << 24));
    __stcs(o + 1, ((v.y >> 8) & 0xFFFFu) | (v.z << 16));
    __stcs(o + 2, ((v.z >> 16) & 0xFFu) | (v.w << 8));
  }
}

cudaError_t launch_pack24(const uint32_t* ids, const uint64_t* total_status, uint64_t ids_cap, uint32_t* out, int n_sm,
                          cudaStream_t st) {
  k_pack24<<<(n_sm > 0 ? n_sm : 148) * 8, 256, 0, st>>>(ids, total_status, ids_cap, out);
  note_launches(1);
  return cudaGetLastError();
}

}  // namespace chq
