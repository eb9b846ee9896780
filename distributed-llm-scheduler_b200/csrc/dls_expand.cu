// dls_expand.cu — rebuilds on the device what would otherwise cross PCIe (no counterpart in the reference).
//
// Of the 32 KB a 1k-task instance occupies, 8 KB are the two u32 CSR row-pointer arrays and 2 KB the rank row. A row
// pointer is the prefix sum of the row lengths (one byte each for DAGs of up to 255 dependencies / params per
// task), and a homogeneous sweep (every instance names its tasks task_0..task_{n-1}) repeats one rank row. The
// end-to-end rate of a pipelined host batch is the host-to-device copy (bench.py: 46 GB/s of a 47.6 GB/s duplex
// link), so the host ships the bytes and these two kernels expand them next to where they are used.
#include "dls_common.cuh"

namespace dls {

namespace {

constexpr int kExpandWarps = 4;

// one warp per topology: ptr[to + t + i] = sum of deg[to .. to + i), i = 0..n
__global__ void __launch_bounds__(kExpandWarps * 32) rows_from_degrees_kernel(const uint8_t* deg, const int64_t* task_off,
                                                                              int T, uint32_t* ptr) {
  const int t = blockIdx.x * kExpandWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (t >= T) return;
  const int64_t to = task_off[t];
  const int n = (int)(task_off[t + 1] - to);
  const uint8_t* d = deg + to;
  uint32_t* out = ptr + to + t;
  uint32_t carry = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    const uint32_t x = i < n ? d[i] : 0u;
    uint32_t incl = x;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t v = __shfl_up_sync(DLS_FULL_MASK, incl, off);
      if (lane >= off) incl += v;
    }
    if (i < n) out[i] = carry + incl - x;
    carry += __shfl_sync(DLS_FULL_MASK, incl, 31);
  }
  if (lane == 0) out[n] = carry;
}

__global__ void __launch_bounds__(256) repeat_row_kernel(const uint16_t* row, int n, long long total, uint16_t* out) {
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < total; k += (long long)gridDim.x * blockDim.x)
    out[k] = row[k % n];
}

}  // namespace

cudaError_t launch_rows_from_degrees(const uint8_t* deg, const int64_t* task_off, int T, uint32_t* ptr, LaunchCtx& ctx) {
  rows_from_degrees_kernel<<<(T + kExpandWarps - 1) / kExpandWarps, kExpandWarps * 32, 0, ctx.st>>>(deg, task_off, T, ptr);
  ++ctx.launches;
  ctx.why = "rows_from_degrees_kernel launch";
  return cudaGetLastError();
}

cudaError_t launch_repeat_row(const uint16_t* row, int n, int T, uint16_t* out, LaunchCtx& ctx) {
  const long long total = (long long)n * T;
  const int grid = (int)min((total + 255) / 256, (long long)148 * 16);
  repeat_row_kernel<<<grid > 0 ? grid : 1, 256, 0, ctx.st>>>(row, n, total, out);
  ++ctx.launches;
  ctx.why = "repeat_row_kernel launch";
  return cudaGetLastError();
}

}  // namespace dls
