// dls_sched.cu — kernel S: the list-scheduling loops of the four algorithms, one warp per DAG instance.
//
// Lane k of the warp IS node k (K <= 32): it keeps Node.available_memory, compute_speed, the node's finish
// time, task count and peak in registers for the whole schedule. The per-DAG mutable state (which tasks are
// done, which params are cached where, and MRU's usage/recency/ready counters) lives in shared memory.
// Tasks are visited in the static order kernel P produced; a "round" (one iteration of the reference's
// while-loop, ref schedulers.py:167,252,335,451) is a maximal run of equal depth in that order.
//
// fp64 replay: need = m + l*0.5 ; need <= avail ; avail -= 0.5 (l times) ; avail -= m ;
// avail += m ; fin += c/s — each a separate IEEE operation, in this order, like the reference's Python floats
// (ref schedulers.py:66-70,76,89,95,126; simulation.py:261-264).
#include <math.h>

#include "dls_common.cuh"

namespace dls {

namespace {

constexpr int kWarpsPerCta = 4;  // instances per CTA of the warp-per-DAG kernels when their shared memory allows


struct OutPtrs {
  int8_t* node_of;
  uint16_t* seq_pos;
  dls_summary_t* summary;
  double *fin, *avail, *peak, *tfin;
  int32_t* cnt;
  uint32_t* cached;
};

__device__ __forceinline__ OutPtrs bind_out(const dls_batch_t& b, const dls_result_t& r, int slot, int inst, const Inst& v) {
  OutPtrs o;
  o.node_of = r.node_of + (int64_t)slot * b.nt_inst + v.t_inst;
  o.seq_pos = r.seq_pos + (int64_t)slot * b.nt_inst + v.t_inst;
  o.summary = r.summary + (int64_t)slot * b.B + inst;
  o.fin = r.fin + (int64_t)slot * b.nk_inst + v.k_out;
  o.cnt = r.cnt + (int64_t)slot * b.nk_inst + v.k_out;
  o.avail = r.avail + (int64_t)slot * b.nk_inst + v.k_out;
  o.peak = r.peak ? r.peak + (int64_t)slot * b.nk_inst + v.k_out : nullptr;
  o.cached = r.cached ? r.cached + (int64_t)slot * b.np_inst + v.p_out : nullptr;
  o.tfin = r.tfin ? r.tfin + (int64_t)slot * b.nk_inst + v.k_out : nullptr;
  return o;
}

__device__ void write_unscheduled(const OutPtrs& o, const Inst& v, int lane, int status) {
  for (int i = lane; i < v.n; i += 32) {
    o.node_of[i] = DLS_TASK_PENDING;
    o.seq_pos[i] = DLS_SEQ_NONE;
  }
  if (lane < v.K) {
    o.fin[lane] = 0.0;
    o.cnt[lane] = 0;
    o.avail[lane] = v.cap[lane];
    if (o.peak) o.peak[lane] = 0.0;
    if (o.tfin) o.tfin[lane] = 0.0;
  }
  if (o.cached)
    for (int p = lane; p < v.P; p += 32) o.cached[p] = 0;
  if (lane == 0) {
    dls_summary_t s;
    s.makespan = 0.0;
    s.peak = 0.0;
    s.completed = 0; s.failed = 0; s.pending = v.n; s.hits = 0; s.misses = 0; s.rounds = 0;
    s.status = status; s.reserved = 0;
    *o.summary = s;
  }
}

// Node state a lane carries. Lanes >= K hold inert values.
struct NodeRegs {
  double avail, spd, cap, fin, peak;
  double tfree;  // extension (dls_result_t.tfin): when the node is free on the dependency-aware timeline
  int cnt;
};

__device__ __forceinline__ void epilogue(const OutPtrs& o, const Inst& v, int lane, const NodeRegs& nd, int completed,
                                         int failed, int hits, int misses, int rounds, int status) {
  // makespan = max(node_finish_times.values()) over nodes that appear in the schedule (ref simulation.py:269-271)
  double mk = (lane < v.K && nd.cnt > 0) ? nd.fin : -INFINITY;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mk = fmax(mk, __shfl_xor_sync(DLS_FULL_MASK, mk, off));
  if (lane < v.K) {
    o.fin[lane] = nd.fin;
    o.cnt[lane] = nd.cnt;
    o.avail[lane] = nd.avail;
    if (o.peak) o.peak[lane] = nd.peak;
    if (o.tfin) o.tfin[lane] = nd.tfree;
  }
  double pk = (o.peak && lane < v.K) ? nd.peak : 0.0;  // dls_summary_t.peak: the largest node peak, when peaks are asked for
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) pk = fmax(pk, __shfl_xor_sync(DLS_FULL_MASK, pk, off));
  if (lane == 0) {
    dls_summary_t s;
    s.makespan = completed > 0 ? mk : 0.0;
    s.peak = pk;
    s.completed = completed; s.failed = failed; s.pending = v.n - completed - failed;
    s.hits = hits; s.misses = misses; s.rounds = rounds; s.status = status; s.reserved = 0;
    *o.summary = s;
  }
}

// assign_task_to_node + complete_task on the chosen lane (ref schedulers.py:86-126) and simulate_execution's
// running finish time (ref simulation.py:261-265).
__device__ __forceinline__ void apply_assignment(NodeRegs& nd, int params_to_load, double m, double dur) {
  for (int j = 0; j < params_to_load; ++j) nd.avail = __dsub_rn(nd.avail, 0.5);
  nd.avail = __dsub_rn(nd.avail, m);
  nd.peak = fmax(nd.peak, __dsub_rn(nd.cap, nd.avail));
  nd.avail = __dadd_rn(nd.avail, m);
  nd.fin = __dadd_rn(nd.fin, __ddiv_rn(dur, nd.spd));
  nd.cnt += 1;
}

// Extension, no reference counterpart (dls_result_t.tfin): task `i`, just assigned to node `best`, on a timeline that
// honours dependencies: it starts when the node is free and every dependency has finished. s_tf[j] = finish time of
// task j (dependencies are assigned, hence written, earlier).
__device__ __forceinline__ void timeline_step(NodeRegs& nd, double* s_tf, const Inst& v, int i, int best, int lane) {
  double dm = 0.0;
  for (uint32_t e = v.pred_ptr[i] + lane; e < v.pred_ptr[i + 1]; e += 32) dm = fmax(dm, s_tf[v.pred_idx[e]]);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) dm = fmax(dm, __shfl_xor_sync(DLS_FULL_MASK, dm, off));
  if (lane == best) nd.tfree = __dadd_rn(fmax(nd.tfree, dm), __ddiv_rn(v.dur[i], nd.spd));
  const double tf = __shfl_sync(DLS_FULL_MASK, nd.tfree, best);
  if (lane == 0) s_tf[i] = tf;
  __syncwarp();
}

// ===================================================================================================================
// DFS / Greedy / Critical
// ===================================================================================================================
template <int ALGO>
__global__ void __launch_bounds__(kWarpsPerCta * 32) sched_simple_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot,
                                                                        int warp_words, int tl_word) {
  extern __shared__ __align__(16) uint32_t smem_all[];
  // one WARP per instance; blockDim / 32 of them share a CTA (see launch_one), each with its own slice of shared memory
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + wic;
  if (inst >= b.B) return;
  uint32_t* smem = smem_all + (size_t)wic * warp_words;
  if (w.route && w.route[(int64_t)slot * b.B + inst] != 1) return;  // already done by the thread-per-DAG kernel
  Inst v;
  bind_instance(b, inst, v);
  const OutPtrs o = bind_out(b, r, slot, inst, v);
  int tstat = w.topo_status[v.topo];
  if (v.K < 0 || v.K > b.K_max) {  // node_off inconsistent with K_max, which is what the per-DAG state was sized for
    tstat = DLS_ST_BAD_TOPOLOGY;
    v.K = max(0, min(v.K, b.K_max));
  }
  if (tstat != DLS_ST_OK || v.n == 0) {
    write_unscheduled(o, v, lane, tstat);
    return;
  }
  const int n = v.n, K = v.K;
  const int nw = (n + 31) >> 5;
  uint32_t* s_done = smem;         // [nw]  bit i = task i completed
  uint32_t* s_cached = smem + nw;  // [P]   bit k = param cached on node k (Node.cached_params)
  for (int i = lane; i < nw; i += 32) s_done[i] = 0;
  for (int p = lane; p < v.P; p += 32) s_cached[p] = 0;
  for (int i = lane; i < n; i += 32) {
    o.node_of[i] = DLS_TASK_PENDING;
    o.seq_pos[i] = DLS_SEQ_NONE;
  }
  __syncwarp();

  const uint16_t* order = (ALGO == DLS_ALGO_CRITICAL) ? w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo)
                                                      : w.order_dg + v.t_topo;
  const uint16_t* depth = w.depth + v.t_topo;

  NodeRegs nd;
  nd.cap = lane < K ? v.cap[lane] : 0.0;
  nd.spd = lane < K ? v.spd[lane] : 1.0;
  nd.avail = nd.cap;
  nd.fin = 0.0;
  nd.peak = 0.0;
  nd.tfree = 0.0;
  nd.cnt = 0;
  double* s_tf = (o.tfin && tl_word >= 0) ? reinterpret_cast<double*>(smem + tl_word) : nullptr;
  int completed = 0, failed = 0, hits = 0, misses = 0, rounds = 0;
  int t = 0;
  while (completed + failed < n) {  // while self.pending_tasks
    ++rounds;
    if (t >= n) break;  // ready list empty: leftovers stay pending (ref schedulers.py:172-174)
    const int d = depth[order[t]];
    bool any_ready = false, any_assigned = false;
    for (; t < n; ++t) {
      const int i = order[t];
      if (depth[i] != d) break;
      // is_task_ready (ref schedulers.py:50-53); dependencies live in earlier rounds, so "at round start" == now
      const uint32_t pb = v.pred_ptr[i], pe = v.pred_ptr[i + 1];
      bool ok = true;
      for (uint32_t e = pb + lane; e < pe; e += 32) {
        const int p = v.pred_idx[e];
        ok = ok && ((s_done[p >> 5] >> (p & 31)) & 1u);
      }
      if (!__all_sync(DLS_FULL_MASK, ok)) continue;
      any_ready = true;
      // params_to_load on my node
      const uint32_t qb = v.par_ptr[i], qe = v.par_ptr[i + 1];
      int l = 0;
      for (uint32_t e = qb; e < qe; ++e) l += !((s_cached[v.par_idx[e]] >> lane) & 1u);
      const double m = v.mem[i];
      const double need = memory_requirement(m, l);
      const bool fits = lane < K && need <= nd.avail;  // can_fit_on_node (ref schedulers.py:74-76)
      int best;
      if (ALGO == DLS_ALGO_DFS) {
        // most available memory among fitting nodes; max_memory starts at -1 (ref schedulers.py:186-192)
        best = warp_argmax_first(fits && nd.avail > -1.0, nd.avail);
      } else if (ALGO == DLS_ALGO_GREEDY) {
        // fewest params to load, then most available memory (ref schedulers.py:267-281)
        const unsigned lmin = __reduce_min_sync(DLS_FULL_MASK, fits ? (unsigned)l : 0xffffffffu);
        best = warp_argmax_first(fits && (unsigned)l == lmin, nd.avail);
      } else {
        // fastest fitting node; best_speed starts at 0 (ref schedulers.py:352-358)
        best = warp_argmax_first(fits && nd.spd > 0.0, nd.spd);
      }
      if (best >= 0) {
        if (lane == best) apply_assignment(nd, l, m, v.dur[i]);
        if (s_tf) timeline_step(nd, s_tf, v, i, best, lane);
        const int lb = __shfl_sync(DLS_FULL_MASK, l, best);
        misses += lb;                   // replay of simulate_execution's seen-set: without evictions it equals
        hits += (int)(qe - qb) - lb;    // the cached set at assignment time (ref simulation.py:253-258)
        for (uint32_t e = qb + lane; e < qe; e += 32) s_cached[v.par_idx[e]] |= 1u << best;
        if (lane == 0) {
          s_done[i >> 5] |= 1u << (i & 31);
          o.node_of[i] = (int8_t)best;
          o.seq_pos[i] = (uint16_t)completed;
        }
        ++completed;
        any_assigned = true;
        __syncwarp();
      } else {
        if (lane == 0) o.node_of[i] = DLS_TASK_FAILED;  // fail_task (ref schedulers.py:128-131,200)
        ++failed;
      }
    }
    if (!any_ready) break;
    if (!any_assigned) {  // no progress: fail everything still pending (ref schedulers.py:202-206)
      __syncwarp();
      int c = 0;
      for (int i = lane; i < n; i += 32)
        if (o.node_of[i] == DLS_TASK_PENDING) {
          o.node_of[i] = DLS_TASK_FAILED;
          ++c;
        }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(DLS_FULL_MASK, c, off);
      failed += c;
      break;
    }
  }
  __syncwarp();
  if (o.cached)
    for (int p = lane; p < v.P; p += 32) o.cached[p] = s_cached[p];
  epilogue(o, v, lane, nd, completed, failed, hits, misses, rounds, DLS_ST_OK);
}

// ===================================================================================================================
// MRU
// ===================================================================================================================
struct MruState {
  uint8_t* state;     // [n]     0 pending, 1 done, 2 failed
  uint32_t* rem;      // [n]     dependency entries not yet completed (0 => task is ready right now)
  uint32_t* cset;     // [K][W]  Node.cached_params as bitsets
  uint32_t* seen;     // [K][W]  simulate_execution's per-node seen-set (ignores evictions, ref simulation.py:244-258)
  uint32_t* rowsave;  // [W]     row backup for the rollback path (ref schedulers.py:437-442)
  uint16_t* usage;    // [P]     param_usage_count
  uint16_t* last;     // [P]     param_last_used (0 = never; time steps start at 1)
  uint16_t* rcnt;     // [P]     number of pending tasks with all dependencies completed that need the param
  int W;
};

// calculate_eviction_score (ref schedulers.py:383-402). The reference rescans every pending task per param;
// rcnt[] holds the same count incrementally. The += 1000 is replayed one add at a time.
__device__ __forceinline__ double eviction_score(const MruState& ms, int p, int time_step) {
  double score = 0.0;
  score = __dadd_rn(score, (double)((int)ms.usage[p] * 10));
  const int lu = ms.last[p];
  if (lu != 0) score = __dadd_rn(score, __ddiv_rn(100.0, (double)(time_step - lu + 1)));
  const int c = ms.rcnt[p];
  for (int j = 0; j < c; ++j) score = __dadd_rn(score, 1000.0);
  return score;
}

// evict_params_for_task (ref schedulers.py:404-442) for node k, cooperatively by the warp. `avail_k` is
// node k's available memory (uniform across the warp) and is updated in place. Returns the reference's bool.
__device__ bool evict_params_for_task(const MruState& ms, int k, double need, double& avail_k, const uint16_t* task_par,
                                      int npar, int time_step, int lane) {
  const double shortage = __dsub_rn(need, avail_k);
  if (shortage <= 0.0) return true;
  uint32_t* row = ms.cset + (size_t)k * ms.W;
  for (int wd = lane; wd < ms.W; wd += 32) ms.rowsave[wd] = row[wd];
  __syncwarp();
  double freed = 0.0;
  int nev = 0;
  bool ok = false;
  // Local best of each lane over its words; recomputed by the owner after each eviction.
  double my_score = 0.0;
  int my_p = -1;
  bool dirty = true;
  while (true) {
    if (freed >= shortage) { ok = true; break; }
    if (dirty) {
      my_p = -1;
      for (int wd = lane; wd < ms.W; wd += 32) {
        uint32_t bits = row[wd];
        while (bits) {
          const int bpos = __ffs(bits) - 1;
          bits &= bits - 1;
          const int p = (wd << 5) + bpos;
          bool needed = false;  // "if param not in task.params_needed"
          for (int e = 0; e < npar; ++e) needed = needed || (task_par[e] == p);
          if (needed) continue;
          const double s = eviction_score(ms, p, time_step);
          if (my_p < 0 || s < my_score) {  // ids ascend inside a lane's scan, so ties keep the smaller id
            my_score = s;
            my_p = p;
          }
        }
      }
      dirty = false;
    }
    const int win = warp_argmin_pair(my_p >= 0, my_score, my_p, nullptr);
    if (win < 0) break;  // candidates exhausted
    if (win == my_p) {
      row[win >> 5] &= ~(1u << (win & 31));  // node.cached_params.remove(param)
      dirty = true;
    }
    avail_k = __dadd_rn(avail_k, 0.5);
    freed = __dadd_rn(freed, 0.5);
    ++nev;
    __syncwarp();
  }
  if (!ok) {  // rollback evictions
    for (int wd = lane; wd < ms.W; wd += 32) row[wd] = ms.rowsave[wd];
    for (int j = 0; j < nev; ++j) avail_k = __dsub_rn(avail_k, 0.5);
  }
  __syncwarp();
  return ok;
}

// A task left the pending set / became ready: maintain rcnt over its params (lanes over params, distinct ids).
__device__ __forceinline__ void rcnt_add(const MruState& ms, const Inst& v, int task, int delta, int lane) {
  for (uint32_t e = v.par_ptr[task] + lane; e < v.par_ptr[task + 1]; e += 32) {
    const int p = v.par_idx[e];
    ms.rcnt[p] = (uint16_t)((int)ms.rcnt[p] + delta);
  }
  __syncwarp();
}

__global__ void __launch_bounds__(kWarpsPerCta * 32) sched_mru_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot,
                                                                     int warp_words, int tl_word) {
  extern __shared__ __align__(16) uint32_t smem_all[];
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + wic;
  if (inst >= b.B) return;
  uint32_t* smem = smem_all + (size_t)wic * warp_words;
  if (w.route && w.route[(int64_t)slot * b.B + inst] != 1) return;  // already done by the thread-per-DAG kernel
  Inst v;
  bind_instance(b, inst, v);
  const OutPtrs o = bind_out(b, r, slot, inst, v);
  int tstat = w.topo_status[v.topo];
  if (v.K < 0 || v.K > b.K_max) {  // node_off inconsistent with K_max, which is what the per-DAG state was sized for
    tstat = DLS_ST_BAD_TOPOLOGY;
    v.K = max(0, min(v.K, b.K_max));
  }
  if (tstat != DLS_ST_OK || v.n == 0) {
    write_unscheduled(o, v, lane, tstat);
    return;
  }
  const int n = v.n, K = v.K, P = v.P;
  MruState ms;
  ms.W = ((P + 31) >> 5) | 1;  // odd stride: rows of different nodes fall into different banks
  {
    uint32_t* p32 = smem;
    ms.rem = p32; p32 += n;
    ms.cset = p32; p32 += (size_t)K * ms.W;
    ms.seen = p32; p32 += (size_t)K * ms.W;
    ms.rowsave = p32; p32 += ms.W;
    uint16_t* p16 = reinterpret_cast<uint16_t*>(p32);
    const int P_al = (P + 1) & ~1;
    ms.usage = p16; p16 += P_al;
    ms.last = p16; p16 += P_al;
    ms.rcnt = p16; p16 += P_al;
    ms.state = reinterpret_cast<uint8_t*>(p16);
  }
  for (int i = lane; i < n; i += 32) {
    ms.state[i] = 0;
    ms.rem[i] = v.pred_ptr[i + 1] - v.pred_ptr[i];
    o.node_of[i] = DLS_TASK_PENDING;
    o.seq_pos[i] = DLS_SEQ_NONE;
  }
  for (int x = lane; x < 2 * K * ms.W; x += 32) ms.cset[x] = 0;  // cset and seen are contiguous
  for (int p = lane; p < P; p += 32) {
    ms.usage[p] = 0;
    ms.last[p] = 0;
    ms.rcnt[p] = 0;
  }
  __syncwarp();
  for (int i = 0; i < n; ++i)  // tasks without dependencies are ready from the start
    if (ms.rem[i] == 0) rcnt_add(ms, v, i, +1, lane);

  const uint16_t* order = w.order_mru + v.t_topo;
  const uint16_t* depth = w.depth + v.t_topo;
  const uint32_t* succ_ptr = w.succ_ptr + v.t_topo + v.topo;
  const uint16_t* succ_idx = w.succ_idx + b.pred_off[v.topo];

  NodeRegs nd;
  nd.cap = lane < K ? v.cap[lane] : 0.0;
  nd.spd = lane < K ? v.spd[lane] : 1.0;
  nd.avail = nd.cap;
  nd.fin = 0.0;
  nd.peak = 0.0;
  nd.tfree = 0.0;
  nd.cnt = 0;
  double* s_tf = (o.tfin && tl_word >= 0) ? reinterpret_cast<double*>(smem + tl_word) : nullptr;
  int completed = 0, failed = 0, hits = 0, misses = 0, rounds = 0, status = DLS_ST_OK, carried = 0;
  int t = 0;
  while (completed + failed + carried < n) {
    ++rounds;  // self.time_step += 1 (ref schedulers.py:453); time_step == rounds
    if (t >= n) break;
    const int d = depth[order[t]];
    bool any_ready = false, any_assigned = false;
    for (; t < n; ++t) {
      const int i = order[t];
      if (depth[i] != d) break;
      if (ms.rem[i] != 0) continue;  // some dependency never completed
      any_ready = true;
      const uint32_t qb = v.par_ptr[i], qe = v.par_ptr[i + 1];
      const int npar = (int)(qe - qb);
      const uint16_t* tpar = v.par_idx + qb;
      const uint32_t* myrow = ms.cset + (size_t)(lane < K ? lane : 0) * ms.W;
      int ncached = 0;  // len(task.params_needed & node.cached_params)
      for (int e = 0; e < npar; ++e) {
        const int p = tpar[e];
        ncached += (myrow[p >> 5] >> (p & 31)) & 1u;
      }
      const int l = npar - ncached;
      const double m = v.mem[i];
      const double need = memory_requirement(m, l);
      const bool fits = lane < K && need <= nd.avail;
      // nodes that do not fit try eviction, which mutates them even if another node wins (ref schedulers.py:490-495)
      unsigned evict_ok = 0;
      unsigned todo = __ballot_sync(DLS_FULL_MASK, lane < K && !fits);
      while (todo) {
        const int k = __ffs(todo) - 1;
        todo &= todo - 1;
        double avail_k = __shfl_sync(DLS_FULL_MASK, nd.avail, k);
        const double need_k = __shfl_sync(DLS_FULL_MASK, need, k);
        if (evict_params_for_task(ms, k, need_k, avail_k, tpar, npar, rounds, lane)) evict_ok |= 1u << k;
        if (lane == k) nd.avail = avail_k;
      }
      double score = 0.0;
      score = __dadd_rn(score, (double)(ncached * 20));
      bool cand = lane < K;
      if (fits) score = __dadd_rn(score, nd.avail);
      else if ((evict_ok >> lane) & 1u) score = __dadd_rn(score, 5.0);
      else cand = false;
      score = __dsub_rn(score, __dmul_rn((double)nd.cnt, 0.5));
      const int best = warp_argmax_first(cand, score);
      if (best >= 0) {
        double avail_b = __shfl_sync(DLS_FULL_MASK, nd.avail, best);
        const double need_b = __shfl_sync(DLS_FULL_MASK, need, best);
        const int lb = __shfl_sync(DLS_FULL_MASK, l, best);
        if (!(need_b <= avail_b)) {  // "if not self.can_fit_on_node(task, best_node): evict" (ref schedulers.py:506-507)
          evict_params_for_task(ms, best, need_b, avail_b, tpar, npar, rounds, lane);
          if (lane == best) nd.avail = avail_b;
        }
        if (need_b > avail_b) {
          // assign_task_to_node returns False (ref schedulers.py:82-83): the reference leaves the task pending and
          // retries it next round. Not replayed here; the instance is flagged.
          status |= DLS_ST_CARRY;
          ++carried;
          continue;
        }
        uint32_t* brow = ms.cset + (size_t)best * ms.W;
        uint32_t* srow = ms.seen + (size_t)best * ms.W;
        int h = 0;
        if (lane == 0) {  // a task has few params and two of them may share a bitset word: one lane, no races
          for (int e = 0; e < npar; ++e) {
            const int p = tpar[e];
            const uint32_t bit = 1u << (p & 31);
            brow[p >> 5] |= bit;
            if (srow[p >> 5] & bit) ++h;
            else srow[p >> 5] |= bit;
            ms.usage[p] = (uint16_t)(ms.usage[p] + 1);  // ref schedulers.py:514-516
            ms.last[p] = (uint16_t)rounds;
          }
        }
        h = __shfl_sync(DLS_FULL_MASK, h, 0);
        hits += h;
        misses += npar - h;
        if (lane == best) apply_assignment(nd, lb, m, v.dur[i]);
        if (s_tf) timeline_step(nd, s_tf, v, i, best, lane);
        if (lane == 0) {
          ms.state[i] = 1;
          o.node_of[i] = (int8_t)best;
          o.seq_pos[i] = (uint16_t)completed;
        }
        ++completed;
        any_assigned = true;
        __syncwarp();
        rcnt_add(ms, v, i, -1, lane);  // i left the pending set
        // dependents: one dependency fewer; those reaching zero are ready from now on (LIVE readiness used by the
        // eviction score, ref schedulers.py:396-397)
        const uint32_t sb = succ_ptr[i], se = succ_ptr[i + 1];
        for (uint32_t base = sb; base < se; base += 32) {
          const uint32_t e = base + lane;
          int s = -1;
          bool now_ready = false;
          if (e < se) {
            s = succ_idx[e];
            now_ready = atomicSub(&ms.rem[s], 1u) == 1u;
          }
          unsigned rdy = __ballot_sync(DLS_FULL_MASK, now_ready);
          while (rdy) {
            const int src = __ffs(rdy) - 1;
            rdy &= rdy - 1;
            rcnt_add(ms, v, __shfl_sync(DLS_FULL_MASK, s, src), +1, lane);
          }
        }
        __syncwarp();
      } else {
        if (lane == 0) {
          ms.state[i] = 2;
          o.node_of[i] = DLS_TASK_FAILED;
        }
        ++failed;
        __syncwarp();
        rcnt_add(ms, v, i, -1, lane);
      }
    }
    if (!any_ready) break;
    if (!any_assigned) {
      __syncwarp();
      int c = 0;
      for (int i = lane; i < n; i += 32)
        if (o.node_of[i] == DLS_TASK_PENDING) {
          o.node_of[i] = DLS_TASK_FAILED;
          ++c;
        }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(DLS_FULL_MASK, c, off);
      failed += c;
      break;
    }
  }
  __syncwarp();
  if (o.cached)
    for (int p = lane; p < P; p += 32) {
      uint32_t msk = 0;
      for (int k = 0; k < K; ++k) msk |= ((ms.cset[(size_t)k * ms.W + (p >> 5)] >> (p & 31)) & 1u) << k;
      o.cached[p] = msk;
    }
  epilogue(o, v, lane, nd, completed, failed, hits, misses, rounds, status);
}

}  // namespace

size_t sched_smem_simple(int n_max, int P_max) {
  return ((size_t)((n_max + 31) >> 5) + (size_t)P_max) * sizeof(uint32_t);
}

size_t sched_smem_mru(int n_max, int P_max, int K_max) {
  size_t W = (size_t)(((P_max + 31) >> 5) | 1);
  size_t P_al = ((size_t)P_max + 1) & ~(size_t)1;
  size_t bytes = ((size_t)n_max + 2 * (size_t)K_max * W + W) * sizeof(uint32_t) + 3 * P_al * sizeof(uint16_t) + (size_t)n_max;
  return (bytes + 15) & ~(size_t)15;
}

template <class Kern>
static cudaError_t launch_one(Kern kern, int kernel_id, size_t smem, const dls_batch_t& b, const dls_result_t& r,
                              const Workspace& w, int slot, LaunchCtx& ctx) {
  // Four warps (instances) to a CTA while four slices of shared memory still leave room for several CTAs per SM,
  // else one. Nothing is shared between the warps; it only makes the grid four times smaller, which is what an
  // (almost) empty pick-up launch after the thread-per-DAG kernels costs: 0.07 ms per 131k one-warp CTAs.
  const size_t base = (smem + 15) & ~(size_t)15;  // the timeline extension's per-task finish times follow the rest
  const size_t slice = base + (r.tfin ? (((size_t)b.n_max * sizeof(double) + 15) & ~(size_t)15) : 0);
  const int wpc = slice * kWarpsPerCta <= 100 * 1024 ? kWarpsPerCta : 1;
  ctx.why = "cudaFuncSetAttribute(sched kernel)";
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), (size_t)(slice * wpc));
  if (e != cudaSuccess) return e;
  int rec = ctx.prof ? ctx.prof->begin(kernel_id, ctx.st) : -1;
  kern<<<(b.B + wpc - 1) / wpc, 32 * wpc, slice * wpc, ctx.st>>>(b, r, w, slot, (int)(slice / 4), r.tfin ? (int)(base / 4) : -1);
  if (ctx.prof) ctx.prof->end(rec, ctx.st);
  ++ctx.launches;
  ctx.why = "sched kernel launch";
  return cudaGetLastError();
}

cudaError_t launch_schedule(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                            LaunchCtx& ctx) {
  int slot = 0;
  cudaError_t e = cudaSuccess;
  const size_t sm_s = sched_smem_simple(b.n_max, b.P_max);
  if (algo_mask & (1 << DLS_ALGO_DFS)) {
    e = launch_one(sched_simple_kernel<DLS_ALGO_DFS>, DLS_KERNEL_SCHED_DFS, sm_s, b, r, w, slot++, ctx);
    if (e != cudaSuccess) return e;
  }
  if (algo_mask & (1 << DLS_ALGO_GREEDY)) {
    e = launch_one(sched_simple_kernel<DLS_ALGO_GREEDY>, DLS_KERNEL_SCHED_GREEDY, sm_s, b, r, w, slot++, ctx);
    if (e != cudaSuccess) return e;
  }
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) {
    e = launch_one(sched_simple_kernel<DLS_ALGO_CRITICAL>, DLS_KERNEL_SCHED_CRITICAL, sm_s, b, r, w, slot++, ctx);
    if (e != cudaSuccess) return e;
  }
  if (algo_mask & (1 << DLS_ALGO_MRU)) {
    e = launch_one(sched_mru_kernel, DLS_KERNEL_SCHED_MRU, sched_smem_mru(b.n_max, b.P_max, b.K_max), b, r, w, slot++, ctx);
    if (e != cudaSuccess) return e;
  }
  return e;
}

}  // namespace dls
