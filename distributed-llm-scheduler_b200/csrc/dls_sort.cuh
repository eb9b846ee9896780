// dls_sort.cuh — CTA-level stable counting sort shared by the priority kernels.
#pragma once

#include "dls_common.cuh"

namespace dls {

__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// dst = src stably sorted by key(task) ascending, key in [0, nkeys). s_hist has hist_cap words.
// When nkeys * (warps per CTA) counters fit, every warp owns a contiguous chunk of src: it counts its keys, one
// block-wide scan over the [key][warp] table turns the counts into start offsets, and each warp then places its own
// chunk, 32 elements at a time, with __match_any giving each lane its rank among equal keys (stable). Otherwise the
// histogram is shared and warp 0 places everything.
template <class KeyF>
__device__ inline void stable_counting_sort(const uint16_t* src, uint16_t* dst, int n, int nkeys, KeyF key, uint32_t* s_hist,
                                     int hist_cap, uint32_t* s_part) {
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
        const int i = src[q];
        const uint32_t k = key(i) * stride + col;
        const unsigned peers = __match_any_sync(vm, k);
        const uint32_t pos = s_hist[k] + __popc(peers & lanemask_lt());
        dst[pos] = (uint16_t)i;
        __syncwarp(vm);
        if (lane == __ffs(peers) - 1) s_hist[k] += __popc(peers);
        __syncwarp(vm);
      }
    }
  }
  __syncthreads();
}

// The same sort without atomics, for up to kChunkSortMax * 32 * (warps per CTA) elements: one counter per (key,
// 32-element chunk of src). A warp takes every (warps per CTA)-th chunk; __match_any gives each lane the lanes of its
// chunk that hold the same key, the lowest of them stores their number (no atomic, no read-modify-write), and after
// one scan of the [key][chunk] table a lane's destination is its counter plus its rank among those lanes, both still
// in registers from the first pass. Measured round 2c on the topology kernel's three sorts: see DESIGN.md.
constexpr int kChunkSortMax = 8;   // chunks per thread kept in registers

__device__ __forceinline__ bool chunk_sort_fits(int n, int nkeys, int hist_cap) {
  const int nchunks = (n + 31) >> 5, nwarp = blockDim.x >> 5;
  return nchunks <= kChunkSortMax * nwarp && nkeys < 2048 && nkeys * (nchunks | 1) + 1 <= hist_cap;
}

template <class KeyF>
__device__ inline void chunk_counting_sort(const uint16_t* src, uint16_t* dst, int n, int nkeys, KeyF key, uint32_t* s_hist,
                                           uint32_t* s_part) {
  const int nwarp = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nchunks = (n + 31) >> 5;
  const int row = nchunks | 1;  // odd row length: the lanes of a warp (same chunk, different keys) fall into different banks
  for (int k = threadIdx.x; k <= nkeys * row; k += blockDim.x) s_hist[k] = 0;
  __syncthreads();
  uint32_t keep[kChunkSortMax];  // element (16 bits) | key (11 bits) << 16 | rank among equal keys of the chunk << 27
#pragma unroll
  for (int u = 0; u < kChunkSortMax; ++u) {
    const int c = wid + u * nwarp, q = c * 32 + lane;
    keep[u] = 0xffffffffu;
    if (c < nchunks) {  // (uniform per warp)
      const bool valid = q < n;
      const unsigned vm = __ballot_sync(DLS_FULL_MASK, valid);
      if (valid) {
        const uint32_t i = src[q];
        const uint32_t k = key((int)i);
        const unsigned peers = __match_any_sync(vm, k);
        const uint32_t rank = __popc(peers & lanemask_lt());
        if (rank == 0) s_hist[k * row + c] = __popc(peers);
        keep[u] = i | (k << 16) | (rank << 27);
      }
    }
  }
  __syncthreads();
  block_exclusive_scan(s_hist, nkeys * row, s_part);
#pragma unroll
  for (int u = 0; u < kChunkSortMax; ++u) {
    const int c = wid + u * nwarp;
    if (keep[u] != 0xffffffffu) {
      const uint32_t k = (keep[u] >> 16) & 0x7ffu;
      dst[s_hist[k * row + c] + (keep[u] >> 27)] = (uint16_t)(keep[u] & 0xffffu);
    }
  }
  __syncthreads();
}

}  // namespace dls
