// dls_gen.cu — device-side DAG builder emitting the CSR batch of include/dls.h directly.
//
// Same distribution as the reference's DAGGenerator.generate_random_dag (ref simulation.py:91-114): task i depends on
// an ordered sample without replacement of randint(0, min(max_deps, i)) earlier tasks, has randint(1, 2) params
// of its own named 'param_{i}_{j}', memory ~ U(0.1, 0.5), time ~ U(0.05, 0.15); node sets follow
// calculate_total_memory_needed / create_nodes_with_memory_regime (ref simulation.py:161-214). The random stream is
// a counter-based hash of (seed, instance, task, draw), not CPython's Mersenne Twister, so the instances are
// distributed like the reference's but are not the same instances: parity is checked by running the oracle on
// these very arrays. Generating a 1k-task DAG costs the reference 35 ms of CPython; a million
// of them would be ten core-hours, which is why the sweep inputs are built where they are consumed.
#include "dls_common.cuh"

namespace dls {

namespace {

constexpr int kGenThreads = 256;

__device__ __forceinline__ uint64_t mix64(uint64_t x) {  // splitmix64 finaliser
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// uniform double in [0, 1) for (instance, task, draw): the draws of one task share the hash of (seed, instance, task)
__device__ __forceinline__ uint64_t task_key(uint64_t seed, uint32_t inst, uint32_t task) {
  return mix64(seed ^ ((uint64_t)inst << 32 | task));
}
__device__ __forceinline__ double u01(uint64_t key, uint32_t draw) {
  return (double)(mix64(key + draw) >> 11) * (1.0 / 9007199254740992.0);
}
__device__ __forceinline__ double u01(uint64_t seed, uint32_t inst, uint32_t task, uint32_t draw) {
  return u01(task_key(seed, inst, task), draw);
}

__device__ __forceinline__ int draw_ndeps(const dls_gen_t& g, uint64_t key, int i) {
  const int hi = min(g.max_deps, i);  // random.randint(0, min(max_deps, i))
  return min((int)(u01(key, 0) * (double)(hi + 1)), hi);
}
__device__ __forceinline__ int draw_nparams(uint64_t key) {
  return u01(key, 1) < 0.5 ? 1 : 2;  // random.randint(1, 2)
}

__global__ void __launch_bounds__(kGenThreads) gen_count_kernel(dls_gen_t g, int32_t* e_count, int32_t* r_count) {
  __shared__ int s_e[kGenThreads / 32], s_r[kGenThreads / 32];
  const uint32_t inst = blockIdx.x;
  int e = 0, r = 0;
  for (int i = threadIdx.x; i < g.n; i += blockDim.x) {
    const uint64_t key = task_key(g.seed, inst, i);
    e += draw_ndeps(g, key, i);
    r += draw_nparams(key);
  }
  for (int off = 16; off > 0; off >>= 1) {
    e += __shfl_xor_sync(DLS_FULL_MASK, e, off);
    r += __shfl_xor_sync(DLS_FULL_MASK, r, off);
  }
  if ((threadIdx.x & 31) == 0) { s_e[threadIdx.x >> 5] = e; s_r[threadIdx.x >> 5] = r; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int te = 0, tr = 0;
    for (int k = 0; k < kGenThreads / 32; ++k) { te += s_e[k]; tr += s_r[k]; }
    e_count[inst] = te;
    r_count[inst] = tr;
  }
}

// One CTA per instance. `out` holds DEVICE pointers; the offset tables are already final.
__global__ void __launch_bounds__(kGenThreads) gen_fill_kernel(dls_gen_t g, dls_batch_t out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  __shared__ double s_max[kGenThreads / 32];
  const uint32_t inst = blockIdx.x;
  const int n = g.n, tid = threadIdx.x;
  uint32_t* s_e = reinterpret_cast<uint32_t*>(smem_raw);  // [n+1] dependency row pointers
  uint32_t* s_r = s_e + (n + 1);                          // [n+1] param row pointers
  uint32_t* s_present = s_r + (n + 1);                    // [2n+1] present flags in param-name order -> ids
  const int64_t to = out.task_off[inst];
  uint32_t* pred_ptr = const_cast<uint32_t*>(out.pred_ptr) + to + inst;
  uint32_t* par_ptr = const_cast<uint32_t*>(out.par_ptr) + to + inst;
  uint16_t* pred_idx = const_cast<uint16_t*>(out.pred_idx) + out.pred_off[inst];
  uint16_t* par_idx = const_cast<uint16_t*>(out.par_idx) + out.par_off[inst];
  uint16_t* rank = const_cast<uint16_t*>(out.rank) + to;
  double* mem = const_cast<double*>(out.mem) + to;
  double* dur = const_cast<double*>(out.dur) + to;
  double* cap = const_cast<double*>(out.cap) + out.node_off[inst];
  double* spd = const_cast<double*>(out.spd) + out.node_off[inst];

  for (int x = tid; x <= 2 * n; x += blockDim.x) s_present[x] = 0;
  __syncthreads();
  double need_max = 0.0;
  for (int i = tid; i < n; i += blockDim.x) {
    const uint64_t key = task_key(g.seed, inst, i);
    const int nd = draw_ndeps(g, key, i), np = draw_nparams(key);
    s_e[i] = nd;
    s_r[i] = np;
    s_present[g.param_pos[2 * i]] = 1;
    if (np == 2) s_present[g.param_pos[2 * i + 1]] = 1;
    const double m = 0.1 + (0.5 - 0.1) * u01(key, 2);    // random.uniform(0.1, 0.5)
    const double c = 0.05 + (0.15 - 0.05) * u01(key, 3); // random.uniform(0.05, 0.15)
    mem[i] = m;
    dur[i] = c;
    rank[i] = g.rank[i];
    need_max = fmax(need_max, m + (double)np * 0.5);  // task_total = memory_required + len(params)*0.5
  }
  if (tid == 0) { s_e[n] = 0; s_r[n] = 0; }
  __syncthreads();
  block_exclusive_scan(s_e, n + 1, s_part);
  const uint32_t R = block_exclusive_scan(s_r, n + 1, s_part);
  block_exclusive_scan(s_present, 2 * n + 1, s_part);  // s_present[pos] = id of the param at name position pos
  for (int i = tid; i <= n; i += blockDim.x) { pred_ptr[i] = s_e[i]; par_ptr[i] = s_r[i]; }
  for (int i = tid; i < n; i += blockDim.x) {
    // ordered sample without replacement of nd values from range(i): draw r_j in [0, i-j), skip earlier picks
    const int nd = (int)(s_e[i + 1] - s_e[i]);
    const uint64_t key = task_key(g.seed, inst, i);
    int sorted[8];  // the picks so far in ascending order (registers: every index below is a compile-time constant)
#pragma unroll
    for (int a = 0; a < 8; ++a) sorted[a] = 0x7fffffff;
    for (int j = 0; j < nd; ++j) {
      const int span = i - j;
      int rj = min((int)(u01(key, 8 + j) * (double)span), span - 1);
#pragma unroll
      for (int a = 0; a < 8; ++a) rj += (a < j) && (rj >= sorted[a]);
      pred_idx[s_e[i] + j] = (uint16_t)rj;
      int carry = rj;  // insert: every slot keeps the smaller of (its value, what is carried), the larger moves on
#pragma unroll
      for (int a = 0; a < 8; ++a) {
        const int lo = min(sorted[a], carry), hi = max(sorted[a], carry);
        sorted[a] = lo;
        carry = hi;
      }
    }
    const int np = (int)(s_r[i + 1] - s_r[i]);
    const uint32_t id0 = s_present[g.param_pos[2 * i]];
    if (np == 1) {
      par_idx[s_r[i]] = (uint16_t)id0;
    } else {
      const uint32_t id1 = s_present[g.param_pos[2 * i + 1]];
      par_idx[s_r[i]] = (uint16_t)min(id0, id1);
      par_idx[s_r[i] + 1] = (uint16_t)max(id0, id1);
    }
  }
  // nodes: calculate_total_memory_needed + create_nodes_with_memory_regime
  for (int off = 16; off > 0; off >>= 1) need_max = fmax(need_max, __shfl_xor_sync(DLS_FULL_MASK, need_max, off));
  if ((tid & 31) == 0) s_max[tid >> 5] = need_max;
  __syncthreads();
  if (tid < g.K) {
    double mx = 0.0;
    for (int k = 0; k < kGenThreads / 32; ++k) mx = fmax(mx, s_max[k]);
    const double total = __dadd_rn(mx, __dmul_rn((double)R, 0.5));
    const double available = __dmul_rn(total, g.regime);
    double c, s;
    if (g.K == 2) {
      c = __dmul_rn(available, tid == 0 ? 0.6 : 0.4);
      s = tid == 0 ? 1.2 : 1.0;
    } else if (g.K == 4) {
      const double frac[4] = {0.35, 0.25, 0.25, 0.15}, sp[4] = {1.2, 1.0, 1.0, 0.8};
      c = __dmul_rn(available, frac[tid]);
      s = sp[tid];
    } else {
      c = __ddiv_rn(available, (double)g.K);
      s = 0.7 + (1.3 - 0.7) * u01(g.seed, inst, 0xFFFFFFFFu, tid);  // random.uniform(0.7, 1.3)
    }
    cap[tid] = c;
    spd[tid] = s;
  }
}

}  // namespace

cudaError_t launch_gen_count(const dls_gen_t& g, int32_t* e_count, int32_t* r_count, LaunchCtx& ctx) {
  gen_count_kernel<<<g.B, kGenThreads, 0, ctx.st>>>(g, e_count, r_count);
  ++ctx.launches;
  ctx.why = "gen_count_kernel launch";
  return cudaGetLastError();
}

cudaError_t launch_gen_fill(const dls_gen_t& g, const dls_batch_t& out, LaunchCtx& ctx) {
  const size_t smem = ((size_t)2 * (g.n + 1) + 2 * (size_t)g.n + 1) * sizeof(uint32_t);
  ctx.why = "cudaFuncSetAttribute(gen_fill_kernel)";
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(gen_fill_kernel), (size_t)smem);
  if (e != cudaSuccess) return e;
  gen_fill_kernel<<<g.B, kGenThreads, smem, ctx.st>>>(g, out);
  ++ctx.launches;
  ctx.why = "gen_fill_kernel launch";
  return cudaGetLastError();
}

}  // namespace dls
