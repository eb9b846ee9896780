// dls_tma.cuh — global -> shared staging through the bulk-copy engine (cp.async.bulk + mbarrier, SASS: UBLKCP / SYNCS).
//
// The prep kernels begin by bringing a DAG's arrays into shared memory. As LDG -> STS that is two trips through the
// L1 data pipe per element, in exactly the kernels that pipe bounds, and the whole load latency sits in front of the
// CTA's first useful instruction. A bulk copy is ONE instruction of one thread per array; the data never touches a
// register and its arrival is counted on an mbarrier the CTA waits on once.
//
// cp.async.bulk wants 16-byte aligned addresses and sizes; the batch's arrays are aligned to their element only (a
// row-pointer slice starts at task_off[t] + t). stage_array() therefore places an array in shared memory at the same
// offset modulo 16 as in global memory, bulk-copies the 16-byte aligned interior and lets threads copy the (at most
// 15) bytes on either side.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace dls {

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(arrivals) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");  // visible to the async proxy before any copy names it
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_addr(bar)),
      "r"(parity)
      : "memory");
}

// Room stage_array() needs for `bytes` of payload.
__host__ __device__ constexpr size_t staged_bytes(size_t bytes) { return ((bytes + 15) & ~(size_t)15) + 16; }

// Stage src[0..count) into the 16-byte aligned region `base` (staged_bytes(count * sizeof(T)) bytes). Every thread of
// the CTA calls it with the same arguments; thread 0 issues the bulk copy and adds its size to `tx` (the bytes the
// caller then announces with mbar_expect_tx — before the data can be relied on: mbar_wait, then __syncthreads for the
// edge elements). Returns where src[0] lands.
template <class T>
__device__ __forceinline__ T* stage_array(unsigned char* base, const T* src, int count, uint64_t* bar, uint32_t& tx) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(src);
  const uintptr_t end = a + (uintptr_t)count * sizeof(T);
  T* dst = reinterpret_cast<T*>(base + (a & 15));
  uintptr_t a0 = (a + 15) & ~(uintptr_t)15, a1 = end & ~(uintptr_t)15;
  if (a1 <= a0) { a0 = end; a1 = end; }  // too short for a whole chunk: threads copy everything
  if (threadIdx.x == 0 && a1 > a0) {
    bulk_copy_g2s(reinterpret_cast<unsigned char*>(dst) + (a0 - a), reinterpret_cast<const void*>(a0), (uint32_t)(a1 - a0), bar);
    tx += (uint32_t)(a1 - a0);
  }
  const int head = (int)((a0 - a) / sizeof(T)), tail0 = (int)((a1 - a) / sizeof(T));
  for (int i = threadIdx.x; i < head; i += blockDim.x) dst[i] = __ldg(src + i);
  for (int i = tail0 + threadIdx.x; i < count; i += blockDim.x) dst[i] = __ldg(src + i);
  return dst;
}

}  // namespace dls
