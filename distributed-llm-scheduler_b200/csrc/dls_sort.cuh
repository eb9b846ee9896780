// dls_sort.cuh — CTA-level stable counting sort shared by the priority kernels.
#pragma once

#include "dls_common.cuh"

namespace dls {

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// dst = src stably sorted by key(element) ascending, key in [0, nkeys). s_hist has hist_cap words. Elements are
// 32-bit words that carry whatever the later passes sort by (the topology kernel packs task | depth << 16, so that a
// sort by depth looks nothing up).
// When nkeys * (warps per CTA) counters fit, every warp owns a contiguous chunk of src: it counts its keys, one
// block-wide scan over the [key][warp] table turns the counts into start offsets, and each warp then places its own
// chunk, 32 elements at a time, with __match_any giving each lane its rank among equal keys (stable). Otherwise the
// histogram is shared and warp 0 places everything.
// `ends_out` (optional, global memory): ends_out[k] = position after the last element with key k.
template <class KeyF>
__device__ inline void stable_counting_sort(const uint32_t* src, uint32_t* dst, int n, int nkeys, KeyF key, uint32_t* s_hist,
                                     int hist_cap, uint32_t* s_part, uint16_t* ends_out = nullptr) {
  const int nwarp = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool split = nkeys * nwarp + 1 <= hist_cap;
  const int stride = split ? nwarp : 1, col = split ? wid : 0;
  const int per = split ? (((n + nwarp - 1) / nwarp + 31) & ~31) : n;
  const int lo = split ? min(wid * per, n) : 0, hi = split ? min(lo + per, n) : n;
  for (int k = threadIdx.x; k <= nkeys * stride; k += blockDim.x) s_hist[k] = 0;
  __syncthreads();
  if (split) {
    for (int q = lo + lane; q < hi; q += 32) atomicAdd(&s_hist[key(src[q]) * stride + col], 1u);
  } else {
    for (int q = threadIdx.x; q < n; q += blockDim.x) atomicAdd(&s_hist[key(src[q])], 1u);
  }
  __syncthreads();
  block_exclusive_scan(s_hist, nkeys * stride, s_part);
  if (split || wid == 0) {
    for (int base = lo; base < hi; base += 32) {
      const int q = base + lane;
      const bool valid = q < hi;
      const unsigned vm = __ballot_sync(DLS_FULL_MASK, valid);
      if (valid) {
        const uint32_t i = src[q];
        const uint32_t k = key(i) * stride + col;
        const unsigned peers = __match_any_sync(vm, k);
        const uint32_t pos = s_hist[k] + __popc(peers & lanemask_lt());
        dst[pos] = i;
        __syncwarp(vm);
        if (lane == __ffs(peers) - 1) s_hist[k] += __popc(peers);
        __syncwarp(vm);
      }
    }
  }
  __syncthreads();
  // (after placement every counter has advanced to the end of its segment: the last column of a key is the key's end)
  if (ends_out)
    for (int k = threadIdx.x; k < nkeys; k += blockDim.x) ends_out[k] = (uint16_t)s_hist[k * stride + stride - 1];
}

// (Measured and rejected, round 2c: the same sort with one counter per (key, 32-element chunk) — no atomics, ranks kept
// in registers between the two passes. The scan over keys x chunks counters costs more than the atomics it replaces:
// 4.04 -> 4.30 ms for the topology kernel.)

}  // namespace dls
