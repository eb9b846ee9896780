// dls_common.cuh — shared device helpers for the sm_100a kernels behind include/dls.h.
//
// Numerics contract: every fp64 value that feeds a comparison is produced by the same IEEE operations, in
// the same order, as the reference's Python floats (SURVEY.md section 7 "fp64 replay"). The library is
// compiled with -fmad=false and the arithmetic below additionally uses the explicit round-to-nearest
// intrinsics, so no FMA contraction can change a rounding.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "dls.h"

#define DLS_FULL_MASK 0xffffffffu

namespace dls {

// Scratch produced by the priority pass (kernel P) and consumed by the scheduling kernels (kernel S).
struct Workspace {
  uint16_t* depth;      // [nt_topo]   longest path from a source = round index          ref schedulers.py:140-152
  uint16_t* outdeg;     // [nt_topo]   number of dependents (with multiplicity)           ref schedulers.py:468-469
  uint16_t* order_dg;   // [nt_topo]   (depth asc, rank asc)
  uint16_t* order_mru;  // [nt_topo]   (depth asc, outdeg desc, rank asc)
  uint16_t* order_cp;   // [nt_cp]     (depth asc, up desc, rank asc)
  uint16_t* lvl_end;    // [nt_topo]   lvl_end[d] = end of depth d in the orders (all three list depth 0, 1, 2, ...)
  double* up;           // [nt_cp]     upward rank                                        ref schedulers.py:301-321
  uint32_t* succ_ptr;   // [nt_topo+T] reverse adjacency (task_dependencies)              ref schedulers.py:39,47-48
  uint16_t* succ_idx;   // [e_total]
  int32_t* topo_status; // [T]         0 ok, DLS_ST_BAD_TOPOLOGY
  int cp_per_instance;  // order_cp/up indexed per instance (T==B or per-instance durations) or per topology
  int keep_up;          // the upward ranks are an output (dls_priorities); the scheduling call only needs the order they give
  // ---- streamed (thread-per-DAG) path; all null when that path is not used ----
  uint16_t* par_sid;    // [np_topo]   dense id of a param shared by >= 2 tasks, 0xFFFF if private to one task
  int32_t* nshared;     // [T]         shared params of the topology; kUnstreamable = leave the DAG to the general kernels
  unsigned char* stream[3];  // task blocks in processing order (see kBlkBytes): kOrderDG (DFS/Greedy), kOrderCP, kOrderMRU
  int8_t* node_pos;     // [A][nt_inst] verdict per order position (node, -1 failed, -2 pending)
  uint8_t* route;       // [A][B]      1 = instance left to the warp-per-DAG kernel
};

// What kernel S needs of a task for one step: its memory and duration (16 bytes, "hot") and one 32-bit word
// ("meta": task id, number of private params, depth parity, has-shared-params flag). A stream holds the tasks of one
// instance in one processing order, eight to a 160-byte block: 8 x 16 hot bytes, then 8 meta words. A scheduling
// thread reads a block with five 32-byte loads (one per two steps for the hot pairs, one per eight for the meta words):
// 0.625 requests and 20 bytes per step. Nothing in a block depends on the order it sits in: "a new round starts here"
// is the parity of the task's depth flipping (the orders list depths 0, 1, 2, ... without gaps), and dependencies are
// not in the stream at all — a task that fails or stays pending marks its dependents through the reverse CSR
// (dls_stream.cu). (Round 2b-2e: one 32-byte record per task with the first three dependencies inline, 12.6 GB
// written and read per 131k-instance sweep instead of 7.9; round 1: the same with dependencies translated to order
// positions per stream. Reading ONE array of records through the order arrays instead was measured in round 2a: each
// 32-byte read costs a 128-byte DRAM line and the scheduling kernels ran 2.4x slower.)
constexpr int kBlkTasks = 8;
constexpr int kBlkBytes = 160;
constexpr uint32_t kMetaParity = 1u << 24, kMetaShared = 1u << 25;   // meta = id | npriv << 16 | flags
constexpr int kOrderDG = 0, kOrderCP = 1, kOrderMRU = 2;  // which of kernel P's orders an algorithm slot walks
constexpr int kSharedCap = 32;   // shared params per DAG the streamed DFS/Greedy/Critical kernel keeps masks for
constexpr int kUnstreamable = 0x7fffffff;  // value of nshared[t] for DAGs a record cannot describe

// One DAG instance resolved against the batch descriptor.
struct Inst {
  int n, K, P, topo;
  uint32_t E, R;
  const uint32_t* pred_ptr;
  const uint16_t* pred_idx;
  const uint32_t* par_ptr;
  const uint16_t* par_idx;
  const uint16_t* rank;
  const double* mem;
  const double* dur;
  const double* cap;
  const double* spd;
  int64_t t_topo;  // offset of this topology in per-topology-task arrays
  int64_t t_inst;  // offset of this instance in per-instance-task arrays
  int64_t k_out;   // offset in per-instance-node outputs
  int64_t p_out;   // offset in per-instance-param outputs
};

__device__ __forceinline__ void bind_topology(const dls_batch_t& b, int t, Inst& v) {
  v.topo = t;
  int64_t to = b.task_off[t];
  v.t_topo = to;
  v.n = (int)(b.task_off[t + 1] - to);
  v.P = (int)(b.pid_off[t + 1] - b.pid_off[t]);
  v.pred_ptr = b.pred_ptr + to + t;
  v.pred_idx = b.pred_idx + b.pred_off[t];
  v.par_ptr = b.par_ptr + to + t;
  v.par_idx = b.par_idx + b.par_off[t];
  v.E = (uint32_t)(b.pred_off[t + 1] - b.pred_off[t]);
  v.R = (uint32_t)(b.par_off[t + 1] - b.par_off[t]);
  v.rank = b.rank + to;
}

__device__ __forceinline__ void bind_instance(const dls_batch_t& b, int inst, Inst& v) {
  int t = b.T == 1 ? 0 : inst;
  bind_topology(b, t, v);
  int64_t ti = b.T == 1 ? (int64_t)inst * v.n : v.t_topo;
  v.t_inst = ti;
  v.mem = b.mem + ((b.T == 1 && !(b.flags & DLS_F_MEM_PER_INSTANCE)) ? 0 : ti);
  v.dur = b.dur + ((b.T == 1 && !(b.flags & DLS_F_DUR_PER_INSTANCE)) ? 0 : ti);
  int64_t kin;
  if (b.node_off) {
    kin = b.node_off[inst];
    v.K = (int)(b.node_off[inst + 1] - kin);
    v.k_out = kin;
  } else {
    kin = 0;
    v.K = b.K_max;
    v.k_out = (int64_t)inst * b.K_max;
  }
  v.cap = b.cap + kin;
  v.spd = b.spd + kin;
  v.p_out = b.T == 1 ? (int64_t)inst * v.P : b.pid_off[t];
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// (key, lane) butterfly: winner = largest key among valid lanes, ties -> LOWEST lane. This is the reference's
// "for node in self.nodes.values(): if ... > best" scan with a strict '>' (ref schedulers.py:189-192,355-358,
// 500-502): the first node reaching the maximum keeps it. Returns -1 when no lane is valid.
__device__ __forceinline__ int warp_argmax_first(bool valid, double key) {
  unsigned any = __ballot_sync(DLS_FULL_MASK, valid);
  if (any == 0) return -1;
  int idx = valid ? lane_id() : 64;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    double k2 = __shfl_xor_sync(DLS_FULL_MASK, key, off);
    int i2 = __shfl_xor_sync(DLS_FULL_MASK, idx, off);
    bool v1 = idx < 64, v2 = i2 < 64;
    bool take = v2 && (!v1 || k2 > key || (k2 == key && i2 < idx));
    if (take) {
      key = k2;
      idx = i2;
    }
  }
  return idx;
}

// Smallest (score, id) among valid lanes: the head of `evictable.sort()` over (score, param-name) tuples
// (ref schedulers.py:417-420; param ids are lexicographic ranks of the names). Returns id or -1.
__device__ __forceinline__ int warp_argmin_pair(bool valid, double score, int id, double* win_score) {
  unsigned any = __ballot_sync(DLS_FULL_MASK, valid);
  if (any == 0) return -1;
  if (!valid) id = 0x7fffffff;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    double s2 = __shfl_xor_sync(DLS_FULL_MASK, score, off);
    int i2 = __shfl_xor_sync(DLS_FULL_MASK, id, off);
    bool v1 = id != 0x7fffffff, v2 = i2 != 0x7fffffff;
    bool take = v2 && (!v1 || s2 < score || (s2 == score && i2 < id));
    if (take) {
      score = s2;
      id = i2;
    }
  }
  if (win_score) *win_score = score;
  return id;
}

__device__ __forceinline__ double warp_bcast(double v, int src) { return __shfl_sync(DLS_FULL_MASK, v, src); }
__device__ __forceinline__ int warp_bcast(int v, int src) { return __shfl_sync(DLS_FULL_MASK, v, src); }

// calculate_memory_requirement: memory_needed = m; memory_needed += len(params_to_load) * 0.5  (ref schedulers.py:66-70)
__device__ __forceinline__ double memory_requirement(double m, int params_to_load) {
  return __dadd_rn(m, __dmul_rn((double)params_to_load, 0.5));
}

// Exclusive prefix sum of a[0..m) in shared memory by the whole CTA; returns the total. s_part: [blockDim/32 + 1].
__device__ inline uint32_t block_exclusive_scan(uint32_t* a, int m, uint32_t* s_part, bool odd_chunk = true) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  // An odd chunk length keeps the lanes of a warp (chunk words apart) on different banks: with chunk = 8 the two
  // passes below ran at 7 wavefronts per ideal one (ncu source view, round 1i). Measured per call site: 6 % off the
  // Critical-order kernel, 14 % off param_class — and 5 % ON the topology kernel, which therefore opts out.
  const int chunk = ((m + nthr - 1) / nthr) | (odd_chunk ? 1 : 0);
  const int lo = min(tid * chunk, m), hi = min(lo + chunk, m);
  uint32_t sum = 0;
  for (int i = lo; i < hi; ++i) sum += a[i];
  // scan of per-thread sums: warp scan + scan of warp totals
  uint32_t incl = sum;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    uint32_t t = __shfl_up_sync(DLS_FULL_MASK, incl, off);
    if ((tid & 31) >= off) incl += t;
  }
  const int wid = tid >> 5, nw = nthr >> 5;
  if ((tid & 31) == 31) s_part[wid] = incl;
  __syncthreads();
  if (tid < 32) {
    uint32_t v = tid < nw ? s_part[tid] : 0, inc = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      uint32_t t = __shfl_up_sync(DLS_FULL_MASK, inc, off);
      if (tid >= off) inc += t;
    }
    if (tid < nw) s_part[tid] = inc - v;  // exclusive warp offsets
    if (tid == 31) s_part[32] = inc;      // total
  }
  __syncthreads();
  uint32_t run = s_part[wid] + (incl - sum);
  for (int i = lo; i < hi; ++i) {
    uint32_t t = a[i];
    a[i] = run;
    run += t;
  }
  uint32_t total = s_part[32];
  __syncthreads();
  return total;
}

// for (i = threadIdx.x; i < n; i += blockDim.x) f(i), issued U iterations at a time with the bound as a predicate,
// so that the U independent global loads of a body are all in flight together. (A "#pragma unroll U" on the plain
// loop does not achieve that here: with ~8 iterations per thread all of them land in the compiler's sequential
// remainder loop — measured round 1f: 34 k cycles of exposed latency in the topology kernel's validation alone.)
template <int U, class F>
__device__ __forceinline__ void strided_for(int n, F f) {
  for (int base = threadIdx.x; base < n; base += U * blockDim.x) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * blockDim.x;
      if (i < n) f(i);
    }
  }
}

// f(e, p[e]) for every e < n, by the whole CTA, the 16-byte-aligned middle of the array eight elements per load (the
// batch's u16 arrays are aligned to their element only). One 128-bit load and eight calls instead of eight loads with
// their index arithmetic: the passes over param and rank indices are instruction-bound, not byte-bound.
template <class F>
__device__ __forceinline__ void for_each_u16(const uint16_t* p, int n, F f) {
  const int head = min(n, (int)(((16u - (uint32_t)((uintptr_t)p & 15u)) & 15u) >> 1));
  const int nvec = (n - head) >> 3;
  for (int e = threadIdx.x; e < head; e += blockDim.x) f(e, (int)__ldg(p + e));
  const uint4* pv = reinterpret_cast<const uint4*>(p + head);
  for (int q = threadIdx.x; q < nvec; q += blockDim.x) {
    const uint4 x = __ldg(pv + q);
    const int e0 = head + q * 8;
    f(e0, (int)(x.x & 0xffffu)); f(e0 + 1, (int)(x.x >> 16));
    f(e0 + 2, (int)(x.y & 0xffffu)); f(e0 + 3, (int)(x.y >> 16));
    f(e0 + 4, (int)(x.z & 0xffffu)); f(e0 + 5, (int)(x.z >> 16));
    f(e0 + 6, (int)(x.w & 0xffffu)); f(e0 + 7, (int)(x.w >> 16));
  }
  for (int e = head + nvec * 8 + threadIdx.x; e < n; e += blockDim.x) f(e, (int)__ldg(p + e));
}

// Optional event bracketing of every launch (dls_profile_enable / dls_profile_read).
struct Profiler {
  static constexpr int kMaxRec = 4096;
  bool on = false;
  int n = 0, created = 0;
  cudaEvent_t ev[kMaxRec][2];
  int kid[kMaxRec];
  // returns record index or -1
  int begin(int kernel_id, cudaStream_t st) {
    if (!on || n >= kMaxRec) return -1;
    if (n >= created) {
      if (cudaEventCreate(&ev[n][0]) != cudaSuccess || cudaEventCreate(&ev[n][1]) != cudaSuccess) return -1;
      created = n + 1;
    }
    kid[n] = kernel_id;
    cudaEventRecord(ev[n][0], st);
    return n++;
  }
  void end(int rec, cudaStream_t st) {
    if (rec >= 0) cudaEventRecord(ev[rec][1], st);
  }
};

struct LaunchCtx {
  cudaStream_t st;
  int64_t launches = 0;
  const char* why = "";
  Profiler* prof = nullptr;
};

// What replay_kernel (dls_replay.cu) reads: the results of the one-instance "template" run, per template slot.
struct ReplayArgs {
  int nslots;            // template slots (DFS / Greedy / MRU among the algorithms asked for)
  int n;                 // tasks of the shared topology
  int full_slot[4];      // template slot -> slot in the caller's result arrays
  const int8_t* t_node_of;     // [nslots][n]
  const uint16_t* t_seq_pos;   // [nslots][n]
  const dls_summary_t* t_summary;  // [nslots]
  const int32_t* t_cnt;        // [nslots][K]
  const double* t_avail;       // [nslots][K]
  const double* t_peak;        // [nslots][K] (with the peak output)
  const uint32_t* t_cached;    // [nslots][P] (with the cached output)
  uint16_t* seq_task;          // [nslots][n] scratch: task assigned q-th
  int8_t* seq_node;            // [nslots][n] scratch: its node
};

// cudaFuncAttributeMaxDynamicSharedMemorySize for `kernel`, set only when a call needs more than was granted to that
// kernel on the current device so far (the attribute is sticky; a scheduling call launches 13 kernels, and setting it
// on every launch was 13 runtime calls per call). dls_abi.cu.
cudaError_t ensure_dynamic_smem(const void* kernel, size_t bytes);

// Host-side launchers (one translation unit each).
cudaError_t launch_replay(const dls_batch_t& b, const dls_result_t& r, const ReplayArgs& a, LaunchCtx& ctx);
cudaError_t launch_priority(const dls_batch_t& b, const Workspace& w, int algo_mask, LaunchCtx& ctx);
cudaError_t launch_schedule(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                            LaunchCtx& ctx);
cudaError_t launch_rows_from_degrees(const uint8_t* deg, const int64_t* task_off, int T, uint32_t* ptr, LaunchCtx& ctx);
cudaError_t launch_repeat_row(const uint16_t* row, int n, int T, uint16_t* out, LaunchCtx& ctx);
cudaError_t launch_aggregate(const dls_batch_t& b, const dls_result_t& r, int nslots, const int32_t* group_of, int n_groups,
                             double* out, LaunchCtx& ctx);
cudaError_t launch_gen_count(const dls_gen_t& g, int32_t* e_count, int32_t* r_count, LaunchCtx& ctx);
cudaError_t launch_gen_fill(const dls_gen_t& g, const dls_batch_t& out, LaunchCtx& ctx);
cudaError_t launch_stream_build(const dls_batch_t& b, const Workspace& w, int algo_mask, LaunchCtx& ctx);
cudaError_t launch_stream_sched(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                                LaunchCtx& ctx);
bool stream_path_possible(const dls_batch_t& b, const dls_result_t* r, int smem_limit);
size_t priority_smem_param(int n_max, int P_max);
size_t priority_smem_topo(int n_max, int E_max);
size_t priority_smem_inst(int n_max, bool with_key);
size_t sched_smem_simple(int n_max, int P_max);
size_t sched_smem_mru(int n_max, int P_max, int K_max);

}  // namespace dls
