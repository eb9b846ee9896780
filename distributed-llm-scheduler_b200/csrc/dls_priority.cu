// dls_priority.cu — kernel P: priorities and the static processing orders.
//
// What the reference recomputes inside every schedule() call, once per DAG:
//   depth[i]  = 0 if no deps else 1 + max(depth[dep])                      ref schedulers.py:140-152
//   up[i]     = dur[i] (+ max(up[dependents]) if any)                       ref schedulers.py:301-321
//   outdeg[i] = len(task_dependencies[i])                                   ref schedulers.py:468-469
// and what its round structure implies: round r handles exactly the ready tasks of
// depth r, in the order of the pending set (rank) after a stable sort by the algorithm's key, so the whole
// processing order is static:  order = sort by (depth asc, key desc, rank asc).
//
// priority_topology_kernel: one CTA per topology. Reverse CSR by counting sort; depth by one in-order pass over
// CTA-wide chunks when the tasks are listed topologically (else level-synchronous relaxation); stable counting sorts
// (warp __match_any placement) for the DFS/Greedy and MRU orders.
// param_class_kernel: one CTA per topology: which params are shared between tasks.
// priority_critical_kernel: one CTA per instance. Upward rank level by level from the sinks, then inside each level
// value buckets and pair counting inside a bucket.
#include "dls_common.cuh"
#include "dls_sort.cuh"
#include "dls_tma.cuh"

namespace dls {

namespace {

// 11 CTAs per SM is what the topology kernel's shared memory allows at 1k tasks; without the bound ptxas takes 56-64
// registers (8-9 CTAs): 4.65 ms instead of 4.01 ms per 131k instances (round 2c)
#ifndef DLS_TOPO_MINB
#define DLS_TOPO_MINB 11
#endif
#ifndef DLS_CRIT_MINB
#define DLS_CRIT_MINB 1
#endif
constexpr int kThreadsP = 128;   // measured r1d: 64 and 256 threads per CTA are both slower for the topology kernel

__global__ void __launch_bounds__(kThreadsP, DLS_TOPO_MINB) priority_topology_kernel(dls_batch_t b, Workspace w, int want_mru, int stage,
                                                                      int check_params) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  __shared__ int s_maxd, s_maxdeg;
  const int t = blockIdx.x, tid = threadIdx.x;
  Inst v;
  bind_topology(b, t, v);
  const int n = v.n;
  const int n_al = (n + 3) & ~3;
  auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
  unsigned char* r_cnt = smem_raw;                                      // every region 16-byte aligned
  unsigned char* r_hist = r_cnt + up16((size_t)(n + 2) * 4);
  unsigned char* r_ord = r_hist + staged_bytes((size_t)(n + 2) * 4);
  unsigned char* r_tmp = r_ord + up16((size_t)n_al * 4);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(r_cnt);                 // [n+2]
  uint32_t* s_hist = reinterpret_cast<uint32_t*>(r_hist);               // [n+2]
  uint32_t* s_ord = reinterpret_cast<uint32_t*>(r_ord);                 // [n_al] order elements: task | depth << 16
  uint32_t* s_tmp = reinterpret_cast<uint32_t*>(r_tmp);                 // [n_al] the sorts' second buffer ...
  uint16_t* s_depth = reinterpret_cast<uint16_t*>(r_tmp);               // [n_al] ... after the depths moved into the elements
  // Until the orders are built (which is when s_hist / s_ord / s_tmp are first needed) the same bytes hold a copy of
  // the dependency CSR: row pointers over s_hist, indices over s_ord (if they fit: E <= 2 n). The depth pass
  // walks the CSR with dependent loads, one 32-task chunk after the other; from global memory that chain cost 48 %
  // of this kernel (measured with clock64 per phase, round 1e). The copy is made by the bulk-copy engine
  // (dls_tma.cuh) while the threads check the param row pointers and the ranks, which stay in global memory.
  stage = stage && staged_bytes((size_t)v.E * 2) <= (size_t)n_al * 4;
  __shared__ __align__(8) uint64_t s_bar;
  const uint32_t* pptr = v.pred_ptr;
  const uint16_t* pidx = v.pred_idx;

  uint16_t* g_depth = w.depth + v.t_topo;
  uint16_t* g_outdeg = w.outdeg + v.t_topo;
  uint32_t* g_succ_ptr = w.succ_ptr + v.t_topo + t;
  uint16_t* g_succ_idx = w.succ_idx + b.pred_off[t];

  // With the streamed path on, param_class_kernel has run first and has checked the param CSR against the
  // descriptor; a topology it rejected is not looked at again.
  if (w.nshared != nullptr && w.topo_status[t] != DLS_ST_OK) return;
  if (tid == 0) { s_maxd = 0; s_maxdeg = 0; }
  __syncthreads();
  // ---- validate the descriptor slice (never trust indices that address shared memory) ----
  {
    int bad = 0;
    if (n > b.n_max || v.P > b.P_max || v.E > (uint32_t)b.E_max || v.R > (uint32_t)b.R_max) {
      if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;  // descriptor maxima are what shared memory was sized for
      return;
    }
    if (n > 0 && (v.pred_ptr[0] != 0 || v.par_ptr[0] != 0 || v.pred_ptr[n] != v.E || v.par_ptr[n] != v.R)) bad = 1;
    if (stage) {
      if (tid == 0) mbar_init(&s_bar, 1);
      __syncthreads();
      uint32_t tx = 0;
      pptr = stage_array(r_hist, v.pred_ptr, n + 1, &s_bar, tx);
      pidx = stage_array(r_ord, v.pred_idx, (int)v.E, &s_bar, tx);
      if (tid == 0) mbar_expect_tx(&s_bar, tx);
    }
    // every load below is independent of the others: a thread keeps 8 iterations' worth in flight at once
    strided_for<8>(n, [&](int i) {
      const uint32_t q0 = __ldg(v.par_ptr + i), q1 = __ldg(v.par_ptr + i + 1);
      if (q0 > q1) bad = 1;
    });
    for_each_u16(v.rank, n, [&](int, int r) { bad |= r >= n; });
    if (stage) {
      if (tid < 32) mbar_wait(&s_bar, 0);  // one warp polls; the others sleep at the barrier (polling took 6 % of the issue slots)
      __syncthreads();  // (also: the few edge elements were copied by threads)
    }
    for (int i = tid; i < n; i += blockDim.x) {
      const uint32_t p0 = pptr[i], p1 = pptr[i + 1];
      if (p0 > p1 || p1 > v.E) bad = 1;
    }
    for (int e = tid; e < (int)v.E; e += blockDim.x)
      if (pidx[e] >= n) bad = 1;
    if (check_params) {  // (with the streamed path on, param_class_kernel checks them as it reads them anyway)
      for_each_u16(v.par_idx, (int)v.R, [&](int, int p) { bad |= p >= v.P; });
    }
    if (__syncthreads_or(bad)) {
      if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;
      return;
    }
  }
  // ---- two things at once: warp 0 computes the depths, the other warps build out-degrees and the reverse CSR ----
  // (with one warp per CTA, small DAGs, the same warp does one after the other)
  __shared__ int s_unsorted;
  const int nthr = blockDim.x;
  const bool split = nthr >= 64;                 // warp 0 | warps 1..: each side syncs by itself
  const int gtid = split ? tid - 32 : tid;       // thread index inside the CSR group (negative: warp 0 of a split CTA)
  const int gthr = split ? nthr - 32 : nthr;
  auto group_sync = [&]() {  // the CSR group's own barrier
    if (split) asm volatile("bar.sync 1, %0;" ::"r"(gthr) : "memory");
    else __syncthreads();
  };
  for (int i = tid; i < n + 2; i += nthr) s_cnt[i] = 0;
  for (int i = tid; i < n; i += nthr) s_depth[i] = 0;
  if (tid == 0) s_unsorted = 0;
  __syncthreads();
  if (gtid >= 0) {
    // out-degree and reverse CSR (task_dependencies, ref schedulers.py:47-48)
    for (uint32_t e = gtid; e < v.E; e += gthr) atomicAdd(&s_cnt[pidx[e]], 1u);
    group_sync();
    {
      int mdg = 0;
      for (int i = gtid; i < n; i += gthr) {
        const uint32_t c = min(s_cnt[i], 65535u);
        if (w.keep_up) g_outdeg[i] = (uint16_t)c;  // (an output of dls_priorities only; the MRU order below reads shared memory)
        mdg = max(mdg, (int)c);
      }
      atomicMax(&s_maxdeg, mdg);
    }
    group_sync();
    {  // exclusive scan of s_cnt[0..n] by the group: s_cnt[i] = start of i's dependents, s_cnt[n] = E
      const int m = n + 1, chunk = (m + gthr - 1) / gthr;
      const int lo = min(gtid * chunk, m), hi = min(lo + chunk, m);
      uint32_t sum = 0;
      for (int i = lo; i < hi; ++i) sum += s_cnt[i];
      uint32_t incl = sum;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const uint32_t x = __shfl_up_sync(DLS_FULL_MASK, incl, off);
        if ((gtid & 31) >= off) incl += x;
      }
      const int gw = gtid >> 5, ngw = gthr >> 5;
      if ((gtid & 31) == 31) s_part[gw] = incl;
      group_sync();
      uint32_t before = 0;
      for (int k = 0; k < gw && k < ngw; ++k) before += s_part[k];
      uint32_t run = before + (incl - sum);
      for (int i = lo; i < hi; ++i) {
        const uint32_t x = s_cnt[i];
        s_cnt[i] = run;
        run += x;
      }
    }
    group_sync();
    for (int i = gtid; i <= n; i += gthr) g_succ_ptr[i] = s_cnt[i];
    group_sync();
    for (int i = gtid; i < n; i += gthr)
      for (uint32_t e = pptr[i]; e < pptr[i + 1]; ++e) {
        const uint32_t p = pidx[e];
        uint32_t pos = atomicAdd(&s_cnt[p], 1u);  // order inside a list is irrelevant to max() and counts
        g_succ_idx[pos] = (uint16_t)i;
      }
  }
  if (!split) __syncthreads();
  // Tasks are usually listed in a topological order already (every dependency has a smaller index: all of the
  // reference's generators emit them that way). Then one in-order pass by ONE warp, 32 tasks at a time, is exact:
  // dependencies below the chunk are final and are folded in once; a chunk is repeated only while one of its tasks
  // depends on another task of the same chunk whose depth moved (~2 rounds per chunk; CTA-wide chunks of 128 took ~4
  // rounds, two CTA barriers each, and four times the lanes: 5.3 k -> 2 k warp-instructions per 1k-task DAG). A
  // dependency that does not point backwards ends the pass: the relaxation below takes over.
  if (tid < 32) {
    const int lane = tid;
    bool out_of_order = false;
    for (int base = 0; base < n && !out_of_order; base += 32) {
      const int i = base + lane;
      const bool valid = i < n;
      const uint32_t eb = valid ? pptr[i] : 0, ee = valid ? pptr[i + 1] : 0, np = ee - eb;
      const uint32_t p0 = np > 0 ? pidx[eb] : 0, p1 = np > 1 ? pidx[eb + 1] : 0, p2 = np > 2 ? pidx[eb + 2] : 0;
      bool fwd = (np > 0 && (int)p0 >= i) || (np > 1 && (int)p1 >= i) || (np > 2 && (int)p2 >= i);
      const bool in0 = np > 0 && (int)p0 >= base, in1 = np > 1 && (int)p1 >= base, in2 = np > 2 && (int)p2 >= base;
      int d_out = 0;  // what the dependencies below the chunk ask for
      if (np > 0 && !in0) d_out = (int)s_depth[p0] + 1;
      if (np > 1 && !in1) d_out = max(d_out, (int)s_depth[p1] + 1);
      if (np > 2 && !in2) d_out = max(d_out, (int)s_depth[p2] + 1);
      bool in_more = false;
      for (uint32_t e = eb + 3; e < ee; ++e) {
        const int p = pidx[e];
        fwd |= p >= i;
        if (p < base) d_out = max(d_out, (int)s_depth[p] + 1);
        else in_more = true;
      }
      if (__any_sync(DLS_FULL_MASK, fwd)) { out_of_order = true; break; }
      int cur = min(d_out, 65535);
      if (valid) s_depth[i] = (uint16_t)cur;
      if (__any_sync(DLS_FULL_MASK, in0 || in1 || in2 || in_more)) {
        bool moved;
        do {
          __syncwarp();
          int d = d_out;
          if (in0) d = max(d, (int)s_depth[p0] + 1);
          if (in1) d = max(d, (int)s_depth[p1] + 1);
          if (in2) d = max(d, (int)s_depth[p2] + 1);
          if (in_more)
            for (uint32_t e = eb + 3; e < ee; ++e) {
              const int p = pidx[e];
              if (p >= base) d = max(d, (int)s_depth[p] + 1);
            }
          d = min(d, 65535);
          moved = valid && d != cur;
          cur = d;
          __syncwarp();  // everybody has read this round's depths
          if (moved) s_depth[i] = (uint16_t)d;
        } while (__any_sync(DLS_FULL_MASK, moved));
      }
    }
    if (lane == 0 && out_of_order) s_unsorted = 1;
  }
  __syncthreads();
  // ---- depth of DAGs listed in any other order: relax until a fixed point (<= levels+1 sweeps; more than n+1 sweeps
  // means a cycle). The dependency CSR is staged in shared memory once: every sweep walks it again.
  int sweeps = 0, changed = 0;
  if (s_unsorted) {
    for (int i = tid; i < n; i += nthr) s_depth[i] = 0;
    __syncthreads();
    do {
      changed = 0;
      for (int i = tid; i < n; i += blockDim.x) {
        int d = 0;
        for (uint32_t e = pptr[i]; e < pptr[i + 1]; ++e) d = max(d, (int)s_depth[pidx[e]] + 1);
        d = min(d, 65535);
        if (d != s_depth[i]) {
          s_depth[i] = (uint16_t)d;
          changed = 1;
        }
      }
      changed = __syncthreads_or(changed);
      ++sweeps;
    } while (changed && sweeps <= n + 1);
  }
  if (changed) {
    if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;
    return;
  }
  if (tid == 0) w.topo_status[t] = DLS_ST_OK;
  {
    int md = 0;
    for (int i = tid; i < n; i += blockDim.x) {
      g_depth[i] = s_depth[i];
      md = max(md, (int)s_depth[i]);
    }
    atomicMax(&s_maxd, md);
  }
  // ---- pi order: s_ord[r] = task at position r of list(pending_tasks) ----
  __syncthreads();  // the staged CSR (aliased with s_hist / s_ord / s_tmp) is dead from here on
  // Each element carries its task's depth along: the two sorts by depth then read their key out of the element instead
  // of gathering it (those gathers were 9 % of this kernel's shared-memory wavefronts at 3x the conflict-free count).
  for_each_u16(v.rank, n, [&](int i, int r) { s_ord[r] = (uint32_t)i | ((uint32_t)s_depth[i] << 16); });
  __syncthreads();  // (from here on s_depth's bytes are s_tmp)
  {  // rank[] must be a permutation: a repeated value would leave a position of every order unwritten
    // (a bitmap filled with atomicOr while the ranks are first read was measured slower than this second look)
    int dup = 0;
    for_each_u16(v.rank, n, [&](int i, int r) { dup |= (s_ord[r] & 0xffffu) != (uint32_t)i; });
    if (__syncthreads_or(dup)) {
      if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;
      return;
    }
  }
  const int maxd = s_maxd, maxdeg = s_maxdeg;
  // DFS / Greedy: (depth asc, rank asc)  — ref schedulers.py:171-177 (the depth sort is a no-op inside a round)
  auto by_depth = [](uint32_t e) { return e >> 16; };
  stable_counting_sort(s_ord, s_tmp, n, maxd + 1, by_depth, s_hist, n + 2, s_part, w.lvl_end + v.t_topo);
  for (int q = tid; q < n; q += blockDim.x) w.order_dg[v.t_topo + q] = (uint16_t)s_tmp[q];
  __syncthreads();
  if (want_mru) {
    // MRU: (depth asc, urgency desc, ready-index asc) — ref schedulers.py:462-475. LSD: urgency first, then depth.
    // out-degree of task i: after the fill above s_cnt[i] is the END of i's dependents in the reverse CSR
    auto deg_of = [&](uint32_t e) {
      const uint32_t i = e & 0xffffu;
      return min(s_cnt[i] - (i ? s_cnt[i - 1] : 0u), 65535u);
    };
    // (one pass on the combined key (depth, degree) needs ~200 x 4 counters and was measured slower than these two)
    auto by_degree = [&](uint32_t e) { return (uint32_t)maxdeg - deg_of(e); };
    if (maxdeg + 2 <= n + 2) {
      stable_counting_sort(s_ord, s_tmp, n, maxdeg + 1, by_degree, s_hist, n + 2, s_part);
    } else {
      // duplicate dependency entries can push an out-degree beyond n: more keys than the histogram holds.
      // Rare, so place by counting: position = #(larger degree) + #(equal degree earlier in pi order).
      for (int q = tid; q < n; q += blockDim.x) {
        const uint32_t dq = deg_of(s_ord[q]);
        int ahead = 0;
        for (int q2 = 0; q2 < n; ++q2) {
          const uint32_t d2 = deg_of(s_ord[q2]);
          ahead += (d2 > dq) || (d2 == dq && q2 < q);
        }
        s_tmp[ahead] = s_ord[q];
      }
      __syncthreads();
    }
    stable_counting_sort(s_tmp, s_ord, n, maxd + 1, by_depth, s_hist, n + 2, s_part);
    for (int q = tid; q < n; q += blockDim.x) w.order_mru[v.t_topo + q] = (uint16_t)s_ord[q];
  }
}

// Classify params of one topology (one CTA each); the first kernel of a scheduling call on the streamed path:
//   topo_status[t] = OK, or BAD_TOPOLOGY when the param CSR is inconsistent with the descriptor;
//   nshared[t]    = number of params referenced by >= 2 tasks ("shared");
//   par_sid[p]    = dense id among the shared params, 0xFFFF for params private to one task (written only if any).
// Tasks whose params are all private cost the same on every node (nothing of theirs is ever cached when they are
// looked at), which is what the thread-per-DAG kernels build their short cuts on.
// Most DAGs of a sweep have no shared param at all. That is settled without counting: every reference writes its
// own index into its param's slot, and a second look finds a foreign index exactly where a param has two references.
__global__ void __launch_bounds__(kThreadsP) param_class_kernel(dls_batch_t b, Workspace w) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  const int t = blockIdx.x, tid = threadIdx.x;
  Inst v;
  bind_topology(b, t, v);
  const int P = v.P;
  if (tid == 0) w.nshared[t] = 0;
  if (v.n > b.n_max || P > b.P_max || P < 0 || v.n < 0 || v.E > (uint32_t)b.E_max || v.R > (uint32_t)b.R_max) {
    if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;  // descriptor maxima are what shared memory was sized for
    return;
  }
  uint32_t* s_rc = reinterpret_cast<uint32_t*>(smem_raw);             // [P+2]
  uint16_t* g_sid = w.par_sid + b.pid_off[t];
  int bad = 0;
  for_each_u16(v.par_idx, (int)v.R, [&](int e, int p) {
    if (p < P) s_rc[p] = (uint32_t)e;
    else bad = 1;  // never trust an index that addresses shared memory
  });
  if (__syncthreads_or(bad)) {
    if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;
    return;
  }
  if (tid == 0) w.topo_status[t] = DLS_ST_OK;
  int twice = 0;
  for_each_u16(v.par_idx, (int)v.R, [&](int e, int p) { twice |= s_rc[p] != (uint32_t)e; });
  if (!__syncthreads_or(twice)) return;  // nshared[t] = 0 stands; par_sid is not read for such a DAG
  for (int p = tid; p < P + 2; p += blockDim.x) s_rc[p] = 0;
  __syncthreads();
  strided_for<8>((int)v.R, [&](int e) { atomicAdd(&s_rc[__ldg(v.par_idx + e)], 1u); });
  __syncthreads();
  for (int p = tid; p < P; p += blockDim.x) s_rc[p] = s_rc[p] >= 2 ? 1u : 0u;
  __syncthreads();
  // keep the flags: after the scan, flag[p] = (scan[p+1] != scan[p])
  const uint32_t nsh = block_exclusive_scan(s_rc, P + 1, s_part);
  for (int p = tid; p < P; p += blockDim.x) g_sid[p] = (s_rc[p + 1] != s_rc[p]) ? (uint16_t)s_rc[p] : (uint16_t)0xFFFF;
  if (tid == 0) w.nshared[t] = (int)nsh;
}

__global__ void __launch_bounds__(kThreadsP, DLS_CRIT_MINB) priority_critical_kernel(dls_batch_t b, Workspace w, int use_key, int have_mru) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_maxd;
  __shared__ uint32_t s_part[33];
  __shared__ double s_red[2 * (kThreadsP / 32)];
  const int inst = blockIdx.x, tid = threadIdx.x;
  Inst v;
  bind_instance(b, inst, v);
  if (w.topo_status[v.topo] != DLS_ST_OK) return;
  const int n = v.n;
  if (n == 0) return;
  const int n_al = (n + 3) & ~3;
  // shared memory, every region 16-byte aligned; the staged arrays sit at their global offset modulo 16 (dls_tma.cuh)
  const size_t nkey = use_key ? ((n + 1) & ~1) : 0;
  unsigned char* r_up = smem_raw;                                        // durations, then upward ranks [n]
  unsigned char* r_key = r_up + staged_bytes((size_t)n * 8);             // (use_key) keys in bucket order [n]; before: reverse CSR
  unsigned char* r_lvl = r_key + nkey * 8;                               // [n+2] first position of each level
  unsigned char* r_ord = r_lvl + (((size_t)(n + 2) * 4 + 15) & ~(size_t)15);  // order_dg [n]
  unsigned char* r_depth = r_ord + staged_bytes((size_t)n * 2);          // [n]
  unsigned char* r_pq = r_depth + staged_bytes((size_t)n * 2);           // (use_key) pi position of each bucketed key [n_al];
                                                                         // before the sort: the MRU order (see the level loop)
  double* s_key = reinterpret_cast<double*>(r_key);
  uint32_t* s_lvl = reinterpret_cast<uint32_t*>(r_lvl);
  uint16_t* s_pq = reinterpret_cast<uint16_t*>(r_pq);
  const uint16_t* g_depth = w.depth + v.t_topo;
  const uint32_t* succ_ptr = w.succ_ptr + v.t_topo + v.topo;
  const uint16_t* succ_idx = w.succ_idx + b.pred_off[v.topo];
  const int64_t cp_off = w.cp_per_instance ? v.t_inst : v.t_topo;

  // Everything the level loop touches comes in by bulk copies (one instruction of one thread per array, arrival counted
  // on an mbarrier): durations land in s_up (a task's slot holds its own duration until its level is processed), and
  // the reverse adjacency is copied over the key array, which is not needed before the sort (if it fits). From global
  // memory the dependent loads of the level loop were 53 % of this kernel (clock64 per phase, round 1e); as LDG -> STS
  // the staging was 8 % of its instructions and a fifth of its stall samples (round 2b).
  const uint32_t E = (uint32_t)(b.pred_off[v.topo + 1] - b.pred_off[v.topo]);
  const bool staged = use_key && staged_bytes((size_t)4 * (n + 1)) + staged_bytes((size_t)2 * E) <= nkey * 8;
  __shared__ __align__(8) uint64_t s_bar;
  if (tid == 0) mbar_init(&s_bar, 1);
  __syncthreads();
  uint32_t tx = 0;
  uint16_t* s_ord = stage_array(r_ord, w.order_dg + v.t_topo, n, &s_bar, tx);
  // The level loop may visit a level's tasks in any order. In the MRU order (same levels, out-degree descending) the 32
  // tasks of a warp have about the same number of dependents, which is what its inner loop runs over.
  const uint16_t* s_walk = s_ord;
  if (use_key && have_mru) s_walk = stage_array(r_pq, w.order_mru + v.t_topo, n, &s_bar, tx);
  uint16_t* s_depth = stage_array(r_depth, g_depth, n, &s_bar, tx);
  double* s_up = stage_array(r_up, v.dur, n, &s_bar, tx);
  uint32_t* s_sptr = nullptr;
  uint16_t* s_sidx = nullptr;
  if (staged) {
    s_sptr = stage_array(r_key, succ_ptr, n + 1, &s_bar, tx);
    s_sidx = stage_array(r_key + staged_bytes((size_t)4 * (n + 1)), succ_idx, (int)E, &s_bar, tx);
  }
  if (tid == 0) mbar_expect_tx(&s_bar, tx);  // (this thread's one arrival, with the byte count the copies will report)
  if (tid < 32) mbar_wait(&s_bar, 0);        // one warp polls; the others sleep at the barrier below
  const uint32_t* sptr = staged ? s_sptr : succ_ptr;
  const uint16_t* sidx = staged ? s_sidx : succ_idx;
  __syncthreads();
  // where each level starts in the orders: the topology kernel's depth sort left the ends (a few dozen values)
  if (tid == 0) s_maxd = s_depth[s_ord[n - 1]];
  __syncthreads();
  const int maxd = s_maxd;
  {
    const uint16_t* lvl_end = w.lvl_end + v.t_topo;
    for (int d = tid; d <= maxd; d += blockDim.x) s_lvl[d + 1] = __ldg(lvl_end + d);
    if (tid == 0) s_lvl[0] = 0;
  }
  __syncthreads();
  // upward rank, sinks first: every dependent lies on a deeper level, so it is final when read
  for (int d = maxd; d >= 0; --d) {
    for (int q = s_lvl[d] + tid; q < (int)s_lvl[d + 1]; q += blockDim.x) {
      const int i = s_walk[q];
      const uint32_t sb = sptr[i], se = sptr[i + 1];
      double path = s_up[i];  // == dur[i]
      if (se > sb) {
        double mx = s_up[sidx[sb]];
        for (uint32_t e = sb + 1; e < se; ++e) {
          double u = s_up[sidx[e]];
          if (u > mx) mx = u;
        }
        path = __dadd_rn(path, mx);  // task.compute_time + max(...)  ref schedulers.py:314
      }
      s_up[i] = path;
    }
    __syncthreads();
  }
  // (the last level's barrier separates the staged adjacency from the keys that now overwrite it)
  // Within-level stable sort by up desc: ready_tasks.sort(key=critical_path, reverse=True), ref schedulers.py:343.
  // Position of q inside its level = #{q2: key[q2] > key[q]} + #{q2 < q: key[q2] == key[q]}.
  if (use_key) {
    // Counting all pairs of a level costs sum(level^2) compares (a random 1k-task DAG has levels of 250 tasks) and was
    // half of this kernel. So the levels are first cut into NB value buckets by a monotone function of the key
    // (an unstable counting sort by (level, bucket): pi positions travel with the keys), and pairs are counted
    // inside a bucket only. Monotone: a <= b implies bucket(a) <= bucket(b), so buckets never reorder keys.
    // (NaN keys: the reference's own order is undefined for them. Here they land in some valid bucket and rank as
    // -inf among themselves in pi order, so that the result is a permutation whatever the input.)
    double lo = INFINITY, hi = -INFINITY;
    for (int i = tid; i < n; i += blockDim.x) {
      const double u = s_up[i];
      if (w.keep_up) w.up[cp_off + i] = u;
      lo = fmin(lo, u);
      hi = fmax(hi, u);
    }
    for (int off = 16; off > 0; off >>= 1) {
      lo = fmin(lo, __shfl_xor_sync(DLS_FULL_MASK, lo, off));
      hi = fmax(hi, __shfl_xor_sync(DLS_FULL_MASK, hi, off));
    }
    if ((tid & 31) == 0) { s_red[2 * (tid >> 5)] = lo; s_red[2 * (tid >> 5) + 1] = hi; }
    uint32_t* s_hist = s_lvl;  // the level starts are not needed any more: [nlev * NB + 1] <= [n + 2]
    __syncthreads();
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { lo = fmin(lo, s_red[2 * k]); hi = fmax(hi, s_red[2 * k + 1]); }
    const int nlev = maxd + 1;
    const int NB = max(1, min(64, (n + 1) / nlev));
    // Buckets are linear in the bit pattern of the key (sign-corrected so that it orders like the value), i.e. roughly
    // logarithmic in the value: 40 % of the tasks of a random DAG are sinks whose key is just their duration, and a
    // scale linear in the value put all of them into the lowest three buckets of every level.
    auto mono = [](double u) -> uint32_t {
      const long long bits = __double_as_longlong(u + 0.0);  // -0.0 -> +0.0: equal keys must share a bucket
      const unsigned long long m = bits < 0 ? ~(unsigned long long)bits : ((unsigned long long)bits | 0x8000000000000000ull);
      return (uint32_t)(m >> 32);
    };
    const uint32_t mlo = mono(lo), mhi = mono(hi);
    const float scale = (float)NB / ((float)(mhi - mlo) + 1.0f);
    auto group_of = [&](int d, double u) {
      const int bk = max(0, min(NB - 1, (int)((float)(mono(u) - mlo) * scale)));
      return d * NB + (NB - 1 - bk);  // deeper levels later, larger keys first
    };
    const int ngroups = nlev * NB;
    __syncthreads();  // every warp has read the level starts (maxd) before the histogram reuses them
    for (int k = tid; k <= ngroups; k += blockDim.x) s_hist[k] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {  // (in task order: the count does not care, and the reads do not scatter)
      const int g = group_of(s_depth[i], s_up[i]);
      s_depth[i] = (uint16_t)g;  // the depth is not needed again: from here on this array holds the task's group
      atomicAdd(&s_hist[g], 1u);
    }
    __syncthreads();
    block_exclusive_scan(s_hist, ngroups + 1, s_part);
    for (int q = tid; q < n; q += blockDim.x) {
      const int i = s_ord[q];
      const double u = s_up[i];
      const uint32_t pos = atomicAdd(&s_hist[s_depth[i]], 1u);  // afterwards s_hist[g] = end of group g
      s_key[pos] = u == u ? u : -INFINITY;  // NaN keys: any total order will do, but it must be one (see below)
      s_pq[pos] = (uint16_t)q;
    }
    __syncthreads();
    for (int p = tid; p < n; p += blockDim.x) {
      const double u = s_key[p];
      const int q = s_pq[p], i = s_ord[q];
      const int g = s_depth[i];
      const int gs = g ? (int)s_hist[g - 1] : 0, ge = (int)s_hist[g];
      int ahead = 0, same = 0;
      for (int p2 = gs; p2 < ge; ++p2) {
        const double u2 = s_key[p2];
        ahead += u2 > u;
        same += u2 == u;
      }
      if (same > 1)  // equal keys (the task itself is one of them) keep their pi order; rare outside the LLM DAGs
        for (int p2 = gs; p2 < ge; ++p2) ahead += (s_key[p2] == u) && (s_pq[p2] < q);
      w.order_cp[cp_off + gs + ahead] = (uint16_t)i;
    }
    return;
  }
  if (w.keep_up)
    for (int i = tid; i < n; i += blockDim.x) w.up[cp_off + i] = s_up[i];
  __syncthreads();
  for (int q = tid; q < n; q += blockDim.x) {
    const int i = s_ord[q];
    const int d = s_depth[i];
    const int ls = s_lvl[d], le = s_lvl[d + 1];
    auto key_of = [&](int j) { const double x = s_up[j]; return x == x ? x : -INFINITY; };  // NaN: as above
    const double u = key_of(i);
    int ahead = 0;
    {  // very large DAGs: no room for a key array
      for (int q2 = ls; q2 < q; ++q2) ahead += key_of(s_ord[q2]) >= u;
      for (int q2 = q + 1; q2 < le; ++q2) ahead += key_of(s_ord[q2]) > u;
    }
    w.order_cp[cp_off + ls + ahead] = (uint16_t)i;
  }
}

}  // namespace

size_t priority_smem_topo(int n_max, int /*E_max*/) {  // the staged dependency CSR aliases the sort scratch
  size_t n = (size_t)n_max, n_al = (n + 3) & ~(size_t)3;
  auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
  return up16((n + 2) * 4) + staged_bytes((n + 2) * 4) + 2 * up16(n_al * 4);
}

size_t priority_smem_param(int /*n_max*/, int P_max) {
  return ((size_t)P_max + 2) * sizeof(uint32_t);
}

size_t priority_smem_inst(int n_max, bool with_key) {  // (the carve-up at the top of priority_critical_kernel)
  size_t n = (size_t)n_max, n_al = (n + 3) & ~(size_t)3;
  const size_t nkey = with_key ? ((n + 1) & ~(size_t)1) : 0;
  return staged_bytes(n * 8) + nkey * 8 + (((n + 2) * 4 + 15) & ~(size_t)15) + 2 * staged_bytes(n * 2) +
         (with_key ? staged_bytes(n_al * sizeof(uint16_t)) : 0);
}

// Small DAGs (GPT-2: 99 tasks) leave most of a 128-thread CTA idle at every barrier: one or two warps then.
static int threads_for(int n_max) { return n_max <= 128 ? 32 : n_max <= 384 ? 64 : kThreadsP; }

cudaError_t launch_priority(const dls_batch_t& b, const Workspace& w, int algo_mask, LaunchCtx& ctx) {
  cudaError_t e;
  int rec;
  const int threads = threads_for(b.n_max);
  if (w.par_sid) {
    size_t smp = priority_smem_param(b.n_max, b.P_max);
    ctx.why = "cudaFuncSetAttribute(param_class_kernel)";
    e = ensure_dynamic_smem(reinterpret_cast<const void*>(param_class_kernel), (size_t)smp);
    if (e != cudaSuccess) return e;
    rec = ctx.prof ? ctx.prof->begin(DLS_KERNEL_PARAM_CLASS, ctx.st) : -1;
    param_class_kernel<<<b.T, threads_for(b.P_max), smp, ctx.st>>>(b, w);
    if (ctx.prof) ctx.prof->end(rec, ctx.st);
    ++ctx.launches;
    ctx.why = "param_class_kernel launch";
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  size_t sm1 = priority_smem_topo(b.n_max, b.E_max);
  const int stage = 1;
  ctx.why = "cudaFuncSetAttribute(priority_topology_kernel)";
  e = ensure_dynamic_smem(reinterpret_cast<const void*>(priority_topology_kernel), (size_t)sm1);
  if (e != cudaSuccess) return e;
  rec = ctx.prof ? ctx.prof->begin(DLS_KERNEL_PRIO_TOPOLOGY, ctx.st) : -1;
  priority_topology_kernel<<<b.T, threads, sm1, ctx.st>>>(b, w, (algo_mask >> DLS_ALGO_MRU) & 1, stage,
                                                            w.par_sid ? 0 : 1);
  if (ctx.prof) ctx.prof->end(rec, ctx.st);
  ++ctx.launches;
  ctx.why = "priority_topology_kernel launch";
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) {
    size_t sm2 = priority_smem_inst(b.n_max, true);
    const int use_key = sm2 <= 96 * 1024;
    if (!use_key) sm2 = priority_smem_inst(b.n_max, false);
    ctx.why = "cudaFuncSetAttribute(priority_critical_kernel)";
    e = ensure_dynamic_smem(reinterpret_cast<const void*>(priority_critical_kernel), (size_t)sm2);
    if (e != cudaSuccess) return e;
    int grid = w.cp_per_instance ? b.B : 1;
    rec = ctx.prof ? ctx.prof->begin(DLS_KERNEL_PRIO_CRITICAL, ctx.st) : -1;
    priority_critical_kernel<<<grid, threads, sm2, ctx.st>>>(b, w, use_key, (algo_mask >> DLS_ALGO_MRU) & 1);
    if (ctx.prof) ctx.prof->end(rec, ctx.st);
    ++ctx.launches;
    ctx.why = "priority_critical_kernel launch";
    e = cudaGetLastError();
  }
  return e;
}

}  // namespace dls
