// dls_aggregate.cu — sweep-level aggregation on the device.
//
// analyze_results (ref simulation.py:452,518-524,552-563) groups the TestResults by (scheduler, memory regime, ...) and
// averages completion rate, makespan, node utilisation and load balance. For million-instance sweeps lifting every
// record to the host just to average it is the bottleneck; this kernel derives the per-instance metrics of
// run_single_test (ref simulation.py:339-347: completion_rate, avg_node_utilization = mean over used nodes of
// load/makespan, load_balance_score = 1/(1 + std/mean) of the used nodes' loads, ddof 0) straight from the device
// records and accumulates per-group sums. Sums are accumulated with fp64 atomics, so the summation order — unlike
// everything on the scheduling path — is not the reference's; group means agree to ~1e-12 relative.
#include <math.h>

#include "dls_common.cuh"

namespace dls {

namespace {

__global__ void __launch_bounds__(256) aggregate_kernel(dls_batch_t b, dls_result_t r, int nslots, const int32_t* group_of,
                                                        int n_groups, double* out) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)nslots * b.B) return;
  const int slot = (int)(idx / b.B), inst = (int)(idx % b.B);
  const int g = group_of ? group_of[inst] : 0;
  if (g < 0 || g >= n_groups) return;
  const dls_summary_t s = r.summary[(int64_t)slot * b.B + inst];
  int64_t k0;
  int K;
  if (b.node_off) { k0 = b.node_off[inst]; K = (int)(b.node_off[inst + 1] - k0); }
  else { k0 = (int64_t)inst * b.K_max; K = b.K_max; }
  const int t = b.T == 1 ? 0 : inst;
  const int n = (int)(b.task_off[t + 1] - b.task_off[t]);
  const double* fin = r.fin + (int64_t)slot * b.nk_inst + k0;
  const int32_t* cnt = r.cnt + (int64_t)slot * b.nk_inst + k0;
  double sum = 0.0, mx = 0.0;
  int used = 0;
  for (int k = 0; k < K; ++k)
    if (cnt[k] > 0) { sum += fin[k]; mx = fmax(mx, fin[k]); ++used; }
  double util = 0.0, balance = 0.0;
  if (used > 0) {
    const double mean = sum / used;
    if (s.makespan > 0) util = mean / s.makespan;  // mean over used nodes of (load / makespan)
    if (mx != 0.0 && mean > 0) {
      double var = 0.0;
      for (int k = 0; k < K; ++k)
        if (cnt[k] > 0) var += (fin[k] - mean) * (fin[k] - mean);
      balance = 1.0 / (1.0 + sqrt(var / used) / mean);
    }
  }
  const double rate = n > 0 ? (double)s.completed / (double)n * 100.0 : 0.0;
  double* o = out + ((int64_t)slot * n_groups + g) * DLS_AGG_FIELDS;
  atomicAdd(o + 0, 1.0);
  atomicAdd(o + 1, rate);
  atomicAdd(o + 2, s.makespan);
  atomicAdd(o + 3, util);
  atomicAdd(o + 4, balance);
  atomicAdd(o + 5, (double)s.hits);
  atomicAdd(o + 6, (double)s.misses);
  atomicAdd(o + 7, s.completed > 0 ? 1.0 : 0.0);
}

}  // namespace

cudaError_t launch_aggregate(const dls_batch_t& b, const dls_result_t& r, int nslots, const int32_t* group_of, int n_groups,
                             double* out, LaunchCtx& ctx) {
  ctx.why = "cudaMemsetAsync(aggregate)";
  cudaError_t e = cudaMemsetAsync(out, 0, (size_t)nslots * n_groups * DLS_AGG_FIELDS * sizeof(double), ctx.st);
  if (e != cudaSuccess) return e;
  const int64_t total = (int64_t)nslots * b.B;
  aggregate_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx.st>>>(b, r, nslots, group_of, n_groups, out);
  ++ctx.launches;
  ctx.why = "aggregate_kernel launch";
  return cudaGetLastError();
}

}  // namespace dls
