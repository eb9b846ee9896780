// dls_prep.cu — fused preparation kernel: everything between the input batch and the record streams, for one
// DAG instance per CTA, entirely out of shared memory.
//
// The separate kernels of dls_priority.cu / stream_build (round 1b/1c profiles) spend their time waiting: every
// one of their ~30 barrier-separated phases starts with dependent global loads, so a 1k-task DAG costs tens of
// microseconds of latency for a few thousand instructions of work. Here the instance's arrays are staged ONCE with
// coalesced loads (32 KB for 1k tasks) and every later phase — validation, depth, out-degree, param classification,
// the three static orders, upward rank, and the three 32-byte record streams — reads shared memory only; global
// memory is touched again only to write results. Used when topologies are per instance (T == B) and the staged
// arrays fit (n up to ~2.5k); other batches take the separate kernels, which compute the same things.
//
// What is computed, and the reference lines it replaces, is documented in dls_priority.cu and dls_stream.cu.
#include <math.h>

#include "dls_common.cuh"
#include "dls_sort.cuh"

namespace dls {

namespace {

constexpr int kPrepThreads = 256;

// order-preserving map double -> uint64 (so atomicMax on the integer is max on the double); 0 is below every double
__device__ __forceinline__ unsigned long long f64_key(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_f64(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}

struct PrepSmem {
  // staged inputs
  uint32_t *pptr, *qptr;
  uint16_t *pidx, *qidx, *rank;
  double *mem, *dur;
  // working arrays
  double* up;                 // [n]
  unsigned long long* mx;     // [n]  max over dependents (ordered key); later reused as level-ordered sort keys
  uint32_t *cnt, *hist, *lvl; // [n+2], [max(n,P)+2], [n+2]
  uint16_t *depth, *ord, *tmp, *odg, *omru, *ocp, *pos, *sid, *erank;
};

__host__ __device__ inline size_t prep_align(size_t x) { return (x + 15) & ~(size_t)15; }

__host__ __device__ inline size_t prep_layout(int n, int E, int R, int P, size_t* offs /*[24]*/) {
  size_t o = 0;
  int k = 0;
  auto take = [&](size_t bytes) { offs[k++] = o; o = prep_align(o + bytes); };
  const size_t np = (size_t)(n > P ? n : P);
  take(8 * (size_t)n);            // 0 mem
  take(8 * (size_t)n);            // 1 dur
  take(8 * (size_t)n);            // 2 up
  take(8 * (size_t)n);            // 3 mx / key
  take(4 * ((size_t)n + 1));      // 4 pptr
  take(4 * ((size_t)n + 1));      // 5 qptr
  take(4 * ((size_t)n + 2));      // 6 cnt
  take(4 * (np + 2));             // 7 hist
  take(4 * ((size_t)n + 2));      // 8 lvl (also param refcounts: max(n,P)+2 needed -> sized below)
  take(4 * (np + 2));             // 9 rc
  take(2 * (size_t)E + 2);        // 10 pidx
  take(2 * (size_t)R + 2);        // 11 qidx
  take(2 * (size_t)n);            // 12 rank
  take(2 * (size_t)n);            // 13 depth
  take(2 * np);                   // 14 ord
  take(2 * np);                   // 15 tmp
  take(2 * (size_t)n);            // 16 odg
  take(2 * (size_t)n);            // 17 omru
  take(2 * (size_t)n);            // 18 ocp
  take(2 * (size_t)n);            // 19 pos
  take(2 * (size_t)P + 2);        // 20 sid
  take(2 * (size_t)P + 2);        // 21 erank
  take(2 * (size_t)P + 2);        // 22 ownd
  return o;
}

__global__ void __launch_bounds__(kPrepThreads) prep_fused_kernel(dls_batch_t b, Workspace w, int algo_mask) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  __shared__ int s_maxd, s_maxdeg;
  const int inst = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
  Inst v;
  bind_instance(b, inst, v);  // T == B: topology == instance
  const int t = v.topo;
  const int n = v.n, P = v.P;
  const uint32_t E = v.E, R = v.R;
  if (n > b.n_max || P > b.P_max || E > (uint32_t)b.E_max || R > (uint32_t)b.R_max) {
    if (tid == 0) w.topo_status[t] = DLS_ST_BAD_TOPOLOGY;  // descriptor maxima are what shared memory was sized for
    return;
  }
  size_t off[24];
  prep_layout(b.n_max, b.E_max, b.R_max, b.P_max, off);
  double* s_mem = reinterpret_cast<double*>(smem_raw + off[0]);
  double* s_dur = reinterpret_cast<double*>(smem_raw + off[1]);
  double* s_up = reinterpret_cast<double*>(smem_raw + off[2]);
  unsigned long long* s_mx = reinterpret_cast<unsigned long long*>(smem_raw + off[3]);
  double* s_key = reinterpret_cast<double*>(smem_raw + off[3]);
  uint32_t* s_pptr = reinterpret_cast<uint32_t*>(smem_raw + off[4]);
  uint32_t* s_qptr = reinterpret_cast<uint32_t*>(smem_raw + off[5]);
  uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + off[6]);
  uint32_t* s_hist = reinterpret_cast<uint32_t*>(smem_raw + off[7]);
  uint32_t* s_lvl = reinterpret_cast<uint32_t*>(smem_raw + off[8]);
  uint32_t* s_rc = reinterpret_cast<uint32_t*>(smem_raw + off[9]);
  uint16_t* s_pidx = reinterpret_cast<uint16_t*>(smem_raw + off[10]);
  uint16_t* s_qidx = reinterpret_cast<uint16_t*>(smem_raw + off[11]);
  uint16_t* s_rank = reinterpret_cast<uint16_t*>(smem_raw + off[12]);
  uint16_t* s_depth = reinterpret_cast<uint16_t*>(smem_raw + off[13]);
  uint16_t* s_ord = reinterpret_cast<uint16_t*>(smem_raw + off[14]);
  uint16_t* s_tmp = reinterpret_cast<uint16_t*>(smem_raw + off[15]);
  uint16_t* s_odg = reinterpret_cast<uint16_t*>(smem_raw + off[16]);
  uint16_t* s_omru = reinterpret_cast<uint16_t*>(smem_raw + off[17]);
  uint16_t* s_ocp = reinterpret_cast<uint16_t*>(smem_raw + off[18]);
  uint16_t* s_pos = reinterpret_cast<uint16_t*>(smem_raw + off[19]);
  uint16_t* s_sid = reinterpret_cast<uint16_t*>(smem_raw + off[20]);
  uint16_t* s_erank = reinterpret_cast<uint16_t*>(smem_raw + off[21]);
  uint16_t* s_ownd = reinterpret_cast<uint16_t*>(smem_raw + off[22]);
  const int hist_cap = (n > P ? n : P) + 2;
  const bool want_cp = (algo_mask >> DLS_ALGO_CRITICAL) & 1, want_mru = (algo_mask >> DLS_ALGO_MRU) & 1;
  const bool want_dg = (algo_mask & 3) != 0;

  // ---- stage: one coalesced pass over the instance ----
  for (int i = tid; i <= n; i += blockDim.x) { s_pptr[i] = v.pred_ptr[i]; s_qptr[i] = v.par_ptr[i]; }
  for (uint32_t e = tid; e < E; e += blockDim.x) s_pidx[e] = v.pred_idx[e];
  for (uint32_t e = tid; e < R; e += blockDim.x) s_qidx[e] = v.par_idx[e];
  for (int i = tid; i < n; i += blockDim.x) {
    s_rank[i] = v.rank[i];
    s_mem[i] = v.mem[i];
    s_dur[i] = v.dur[i];
    s_depth[i] = 0;
    s_mx[i] = 0;
    s_up[i] = __longlong_as_double(-1ll);  // NaN pattern: "not computed yet"
  }
  for (int i = tid; i < n + 2; i += blockDim.x) s_cnt[i] = 0;
  for (int p = tid; p < P + 2; p += blockDim.x) s_rc[p] = 0;
  if (tid == 0) { s_maxd = 0; s_maxdeg = 0; }
  __syncthreads();
  // ---- validate (never trust indices that address shared memory) ----
  {
    int bad = 0;
    if (n == 0) bad = 1;
    if (n > 0 && (s_pptr[0] != 0 || s_qptr[0] != 0 || s_pptr[n] != E || s_qptr[n] != R)) bad = 1;
    for (int i = tid; i < n; i += blockDim.x) {
      if (s_pptr[i] > s_pptr[i + 1] || s_qptr[i] > s_qptr[i + 1] || s_pptr[i + 1] > E || s_qptr[i + 1] > R) bad = 1;
      if (s_rank[i] >= n) bad = 1;
    }
    for (uint32_t e = tid; e < E; e += blockDim.x) if (s_pidx[e] >= n) bad = 1;
    for (uint32_t e = tid; e < R; e += blockDim.x) if (s_qidx[e] >= P) bad = 1;
    if (__syncthreads_or(bad)) {
      // n == 0 is not an error: the general kernels write the empty result; anything else is a bad descriptor
      if (tid == 0) w.topo_status[t] = n == 0 ? DLS_ST_OK : DLS_ST_BAD_TOPOLOGY;
      if (tid == 0 && w.nshared) w.nshared[t] = 0x7fffffff;  // keeps the streamed kernels away from this instance
      return;
    }
  }
  // ---- out-degree, param reference counts, "already topologically listed?" ----
  int unsorted = 0;
  for (int i = tid; i < n; i += blockDim.x)
    for (uint32_t e = s_pptr[i]; e < s_pptr[i + 1]; ++e) {
      const int p = s_pidx[e];
      unsorted |= p >= i;
      atomicAdd(&s_cnt[p], 1u);
    }
  for (uint32_t e = tid; e < R; e += blockDim.x) atomicAdd(&s_rc[s_qidx[e]], 1u);
  unsorted = __syncthreads_or(unsorted);
  // ---- depth (ref schedulers.py:140-152) ----
  int changed = 0;
  if (!unsorted) {
    if (tid < 32) {
      for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const uint32_t eb = i < n ? s_pptr[i] : 0, ee = i < n ? s_pptr[i + 1] : 0;
        bool again;
        do {
          int d = 0;
          bool inside = false;
          for (uint32_t e = eb; e < ee; ++e) {
            const int p = s_pidx[e];
            d = max(d, (int)s_depth[p] + 1);
            inside |= p >= base;
          }
          const bool moved = i < n && d != s_depth[i];
          if (moved) s_depth[i] = (uint16_t)d;
          __syncwarp();
          again = __any_sync(DLS_FULL_MASK, moved) && __any_sync(DLS_FULL_MASK, inside);
        } while (again);
      }
    }
    __syncthreads();
  } else {
    int sweeps = 0;
    do {
      changed = 0;
      for (int i = tid; i < n; i += blockDim.x) {
        int d = 0;
        for (uint32_t e = s_pptr[i]; e < s_pptr[i + 1]; ++e) d = max(d, (int)s_depth[s_pidx[e]] + 1);
        d = min(d, 65535);
        if (d != s_depth[i]) { s_depth[i] = (uint16_t)d; changed = 1; }
      }
      changed = __syncthreads_or(changed);
      ++sweeps;
    } while (changed && sweeps <= n + 1);
    if (changed) {  // dependency cycle
      if (tid == 0) { w.topo_status[t] = DLS_ST_BAD_TOPOLOGY; if (w.nshared) w.nshared[t] = 0x7fffffff; }
      return;
    }
  }
  if (tid == 0) w.topo_status[t] = DLS_ST_OK;
  {
    int md = 0, mdg = 0;
    for (int i = tid; i < n; i += blockDim.x) {
      const int d = s_depth[i], dg = (int)min(s_cnt[i], 65535u);
      w.depth[v.t_topo + i] = (uint16_t)d;
      w.outdeg[v.t_topo + i] = (uint16_t)dg;
      md = max(md, d);
      mdg = max(mdg, dg);
    }
    for (int o2 = 16; o2 > 0; o2 >>= 1) {
      md = max(md, __shfl_xor_sync(DLS_FULL_MASK, md, o2));
      mdg = max(mdg, __shfl_xor_sync(DLS_FULL_MASK, mdg, o2));
    }
    if (lane == 0) { atomicMax(&s_maxd, md); atomicMax(&s_maxdeg, mdg); }
  }
  // ---- params: shared ids (ref count >= 2) ----
  for (int p = tid; p < P; p += blockDim.x) s_rc[p] = s_rc[p] >= 2 ? 1u : 0u;
  if (tid == 0) s_rc[P] = 0;
  __syncthreads();
  const int maxd = s_maxd, maxdeg = s_maxdeg;
  const uint32_t nsh = block_exclusive_scan(s_rc, P + 1, s_part);
  for (int p = tid; p < P; p += blockDim.x) {
    const uint16_t sd = (s_rc[p + 1] != s_rc[p]) ? (uint16_t)s_rc[p] : (uint16_t)0xFFFF;
    s_sid[p] = sd;
    w.par_sid[b.pid_off[t] + p] = sd;
  }
  if (tid == 0) w.nshared[t] = (int)nsh;
  // ---- MRU eviction order of private params (see param_class_kernel) ----
  if (nsh == 0 && want_mru) {
    for (int i = tid; i < n; i += blockDim.x)
      for (uint32_t e = s_qptr[i]; e < s_qptr[i + 1]; ++e) s_ownd[s_qidx[e]] = s_depth[i];
    for (int p = tid; p < P; p += blockDim.x) s_ord[p] = (uint16_t)p;
    __syncthreads();
    stable_counting_sort(s_ord, s_tmp, P, maxd + 1, [&](int p) { return (uint32_t)s_ownd[p]; }, s_hist, hist_cap, s_part);
    for (int q = tid; q < P; q += blockDim.x) s_erank[s_tmp[q]] = (uint16_t)q;
  }
  __syncthreads();
  // ---- static orders ----
  for (int i = tid; i < n; i += blockDim.x) s_ord[s_rank[i]] = (uint16_t)i;  // pi order
  __syncthreads();
  stable_counting_sort(s_ord, s_odg, n, maxd + 1, [&](int i) { return (uint32_t)s_depth[i]; }, s_hist, hist_cap, s_part);
  for (int q = tid; q < n; q += blockDim.x) w.order_dg[v.t_topo + q] = s_odg[q];
  if (want_mru) {
    if (maxdeg + 2 <= hist_cap) {
      stable_counting_sort(s_ord, s_tmp, n, maxdeg + 1, [&](int i) { return (uint32_t)maxdeg - min(s_cnt[i], 65535u); }, s_hist,
                           hist_cap, s_part);
    } else {  // out-degrees beyond the histogram (duplicate dependency entries): place by counting
      for (int q = tid; q < n; q += blockDim.x) {
        const uint32_t dq = min(s_cnt[s_ord[q]], 65535u);
        int ahead = 0;
        for (int q2 = 0; q2 < n; ++q2) {
          const uint32_t d2 = min(s_cnt[s_ord[q2]], 65535u);
          ahead += (d2 > dq) || (d2 == dq && q2 < q);
        }
        s_tmp[ahead] = s_ord[q];
      }
      __syncthreads();
    }
    stable_counting_sort(s_tmp, s_omru, n, maxd + 1, [&](int i) { return (uint32_t)s_depth[i]; }, s_hist, hist_cap, s_part);
    for (int q = tid; q < n; q += blockDim.x) w.order_mru[v.t_topo + q] = s_omru[q];
  }
  // ---- reverse adjacency for the general MRU kernel: only instances it will actually take (shared params) ----
  if (nsh != 0) {
    __syncthreads();
    for (int i = tid; i <= n; i += blockDim.x) s_hist[i] = i < n ? s_cnt[i] : 0;
    __syncthreads();
    block_exclusive_scan(s_hist, n + 1, s_part);
    uint32_t* g_succ_ptr = w.succ_ptr + v.t_topo + t;
    uint16_t* g_succ_idx = w.succ_idx + b.pred_off[t];
    for (int i = tid; i <= n; i += blockDim.x) g_succ_ptr[i] = s_hist[i];
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x)
      for (uint32_t e = s_pptr[i]; e < s_pptr[i + 1]; ++e) g_succ_idx[atomicAdd(&s_hist[s_pidx[e]], 1u)] = (uint16_t)i;
  }
  // ---- Critical: upward rank (ref schedulers.py:301-321) and the (depth, up desc, rank) order ----
  if (want_cp) {
    __syncthreads();
    if (!unsorted) {
      // Dependents have larger indices: walk the tasks from the back, 32 at a time; a chunk is repeated only while
      // a task of the chunk still receives a larger value from a dependent inside the same chunk.
      if (tid < 32) {
        for (int base = (n - 1) & ~31; base >= 0; base -= 32) {
          const int i = base + lane;
          const uint32_t eb = i < n ? s_pptr[i] : 0, ee = i < n ? s_pptr[i + 1] : 0;
          bool inside = false;
          for (uint32_t e = eb; e < ee; ++e) inside |= s_pidx[e] >= base;
          const bool any_inside = __any_sync(DLS_FULL_MASK, inside);
          bool again;
          do {
            bool moved = false;
            if (i < n) {
              const unsigned long long mk = s_mx[i];
              const double u = mk == 0 ? s_dur[i] : __dadd_rn(s_dur[i], key_f64(mk));  // compute_time + max(...)
              moved = __double_as_longlong(u) != __double_as_longlong(s_up[i]);
              if (moved) {
                s_up[i] = u;
                const unsigned long long uk = f64_key(u);
                for (uint32_t e = eb; e < ee; ++e) atomicMax(&s_mx[s_pidx[e]], uk);
              }
            }
            __syncwarp();
            again = any_inside && __any_sync(DLS_FULL_MASK, moved);
          } while (again);
        }
      }
      __syncthreads();
    } else {
      int sweeps = 0;
      do {
        changed = 0;
        for (int i = tid; i < n; i += blockDim.x) {
          const unsigned long long mk = s_mx[i];
          const double u = mk == 0 ? s_dur[i] : __dadd_rn(s_dur[i], key_f64(mk));
          if (__double_as_longlong(u) != __double_as_longlong(s_up[i])) {
            s_up[i] = u;
            changed = 1;
            const unsigned long long uk = f64_key(u);
            for (uint32_t e = s_pptr[i]; e < s_pptr[i + 1]; ++e) atomicMax(&s_mx[s_pidx[e]], uk);
          }
        }
        changed = __syncthreads_or(changed);
        ++sweeps;
      } while (changed && sweeps <= n + 1);
    }
    const int64_t cp_off = v.t_inst;
    // level starts in DFS/Greedy order, keys in that order, then stable count inside each level
    for (int q = tid; q < n; q += blockDim.x) {
      const int d = s_depth[s_odg[q]];
      if (q == 0 || s_depth[s_odg[q - 1]] != d) s_lvl[d] = q;
      w.up[cp_off + q] = s_up[q];
    }
    if (tid == 0) s_lvl[maxd + 1] = n;
    __syncthreads();  // s_mx is dead from here on: its storage becomes the key array
    for (int q = tid; q < n; q += blockDim.x) s_key[q] = s_up[s_odg[q]];
    __syncthreads();
    for (int q = tid; q < n; q += blockDim.x) {
      const int i = s_odg[q];
      const int d = s_depth[i];
      const int ls = s_lvl[d], le = s_lvl[d + 1];
      const double u = s_key[q];
      int ahead = 0, q2 = ls;
      if ((q2 & 1) && q2 < q) { ahead += s_key[q2] >= u; ++q2; }
#pragma unroll 4
      for (; q2 + 1 < q; q2 += 2) {
        const double2 kk = *reinterpret_cast<const double2*>(s_key + q2);
        ahead += (kk.x >= u) + (kk.y >= u);
      }
      for (; q2 < q; ++q2) ahead += s_key[q2] >= u;
      q2 = q + 1;
      if ((q2 & 1) && q2 < le) { ahead += s_key[q2] > u; ++q2; }
#pragma unroll 4
      for (; q2 + 1 < le; q2 += 2) {
        const double2 kk = *reinterpret_cast<const double2*>(s_key + q2);
        ahead += (kk.x > u) + (kk.y > u);
      }
      for (; q2 < le; ++q2) ahead += s_key[q2] > u;
      s_ocp[ls + ahead] = (uint16_t)i;
    }
    __syncthreads();
    for (int q = tid; q < n; q += blockDim.x) w.order_cp[cp_off + q] = s_ocp[q];
  }
  // ---- record streams (see dls_stream.cu) ----
  if (!w.stream[0] && !w.stream[1] && !w.stream[2]) return;
  bool any_ext = false;
  for (int i = tid; i < n; i += blockDim.x) any_ext |= (s_pptr[i + 1] - s_pptr[i] > 3) || (s_qptr[i + 1] - s_qptr[i] > 2);
  const bool ext_needed = __syncthreads_or(any_ext);
  for (int kind = 0; kind < 3; ++kind) {
    if (!w.stream[kind]) continue;
    if (kind == kStreamDG && !want_dg) continue;
    if (kind == kStreamCP && !want_cp) continue;
    if (kind == kStreamMRU && !want_mru) continue;
    if (kind == kStreamMRU ? nsh != 0 : nsh > (uint32_t)kSharedCap) continue;  // routed to the general kernels
    const uint16_t* order = kind == kStreamDG ? s_odg : kind == kStreamMRU ? s_omru : s_ocp;
    Rec* out = w.stream[kind] + v.t_inst;
    uint16_t* ext = w.ext[kind] + (b.pred_off[inst] + b.par_off[inst]);
    __syncthreads();
    for (int q = tid; q < n; q += blockDim.x) s_pos[order[q]] = (uint16_t)q;
    if (ext_needed) {
      for (int q = tid; q < n; q += blockDim.x) {
        const int i = order[q];
        int extra = max(0, (int)(s_pptr[i + 1] - s_pptr[i]) - 3);
        if (kind == kStreamMRU) {
          extra += max(0, (int)(s_qptr[i + 1] - s_qptr[i]) - 2);
        } else if (nsh > 0) {
          int c2 = 0;
          for (uint32_t e = s_qptr[i]; e < s_qptr[i + 1]; ++e) c2 += s_sid[s_qidx[e]] != 0xFFFF;
          extra += max(0, c2 - 2);
        }
        s_hist[q] = (uint32_t)extra;
      }
      if (tid == 0) s_hist[n] = 0;
      __syncthreads();
      block_exclusive_scan(s_hist, n + 1, s_part);
    } else {
      __syncthreads();
    }
    for (int q = tid; q < n; q += blockDim.x) {
      const int i = order[q];
      Rec rc;
      rc.m = s_mem[i];
      rc.c = s_dur[i];
      const uint32_t pb = s_pptr[i], pe = s_pptr[i + 1], qb = s_qptr[i], qe = s_qptr[i + 1];
      uint32_t xo = ext_needed ? s_hist[q] : 0;
      rc.npred = (uint16_t)min(pe - pb, 65535u);
      rc.pred[0] = rc.pred[1] = rc.pred[2] = 0;
      for (uint32_t e = pb; e < pe; ++e) {
        const uint16_t pp = s_pos[s_pidx[e]];
        if (e - pb < 3) rc.pred[e - pb] = pp;
        else ext[xo++] = pp;
      }
      rc.a0 = rc.a1 = 0;
      if (kind == kStreamMRU) {
        rc.npriv = (uint16_t)(qe - qb);
        rc.nsh = 0;
        for (uint32_t e = qb; e < qe; ++e) {
          const uint16_t sl = s_erank[s_qidx[e]];
          if (e - qb == 0) rc.a0 = sl;
          else if (e - qb == 1) rc.a1 = sl;
          else ext[xo++] = sl;
        }
      } else if (nsh == 0) {
        rc.npriv = (uint16_t)(qe - qb);
        rc.nsh = 0;
      } else {
        int c2 = 0, npriv = 0;
        for (uint32_t e = qb; e < qe; ++e) {
          const uint16_t sd = s_sid[s_qidx[e]];
          if (sd == 0xFFFF) { ++npriv; continue; }
          if (c2 == 0) rc.a0 = sd;
          else if (c2 == 1) rc.a1 = sd;
          else ext[xo++] = sd;
          ++c2;
        }
        rc.npriv = (uint16_t)npriv;
        rc.nsh = (uint8_t)min(c2, 255);
      }
      rc.flags = (q == 0 || s_depth[order[q - 1]] != s_depth[i]) ? kRecFirst : 0;
      out[q] = rc;
    }
  }
}

}  // namespace

size_t prep_fused_smem(const dls_batch_t& b) {
  size_t off[24];
  return prep_layout(b.n_max, b.E_max, b.R_max, b.P_max, off);
}

bool prep_fused_possible(const dls_batch_t& b, int smem_limit) {
  return b.T == b.B && b.n_max >= 1 && prep_fused_smem(b) <= (size_t)smem_limit / 2;  // at least two CTAs per SM
}

cudaError_t launch_prep_fused(const dls_batch_t& b, const Workspace& w, int algo_mask, LaunchCtx& ctx) {
  const size_t smem = prep_fused_smem(b);
  ctx.why = "cudaFuncSetAttribute(prep_fused_kernel)";
  cudaError_t e = cudaFuncSetAttribute(prep_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int rec = ctx.prof ? ctx.prof->begin(DLS_KERNEL_PREP_FUSED, ctx.st) : -1;
  prep_fused_kernel<<<b.B, kPrepThreads, smem, ctx.st>>>(b, w, algo_mask);
  if (ctx.prof) ctx.prof->end(rec, ctx.st);
  ++ctx.launches;
  ctx.why = "prep_fused_kernel launch";
  return cudaGetLastError();
}

}  // namespace dls
