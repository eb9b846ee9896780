// dls_stream.cu — the throughput path: one THREAD per DAG instance over an order-permuted record stream.
//
// Why: the list-scheduling loop is a serial chain of n dependent steps per DAG, but a sweep has ~10^5..10^6
// independent DAGs. With one warp per DAG every instruction of the chain serves one DAG (≈340 warp-instructions per
// task-step, profiles/r1a_*); with one thread per DAG the same instruction serves 32. To make that work the
// per-thread memory accesses must not scatter: stream_build_kernel rewrites each DAG, once per processing order, as
// a contiguous array of 32-byte records (one DRAM sector per step, dependencies already translated to order
// positions), so every step of every thread reads exactly one fully used sector, and per-thread mutable state lives
// in shared memory laid out [word][lane] (each lane owns a bank: no conflicts whatever the index).
//
// Verdicts are written in order-position space (sequential bytes per thread); unpermute_kernel maps them back to
// task ids and derives seq_pos by a prefix count. Instances the fast kernels cannot take (more than 8 nodes, more
// than 32 shared params, MRU with any shared param, bad topology) are flagged in `route` and picked up by the
// warp-per-DAG kernels of dls_sched.cu, which remain the general implementation.
//
// All fp64 arithmetic is the same op-by-op replay as in dls_sched.cu (see the header comment there).
#include <math.h>

#include "dls_common.cuh"

namespace dls {

namespace {

// route byte of an (algorithm slot, instance): 0 done here, 1 left to the warp-per-DAG kernel, kRouteFailRest = done and
// every still-pending task counts as failed, kRouteTwin = done by the DFS launch, whose verdict bytes are the ones to
// read, kRouteOpen = the DFS launch could not vouch for Greedy, which therefore runs this instance itself.
constexpr int kRouteGeneral = 1, kRouteFailRest = 2, kRouteTwin = 4, kRouteOpen = 0xFF;
constexpr int kTPB = 64;        // threads per CTA of the thread-per-DAG DFS/Greedy/Critical kernels
constexpr int kTPBM = 64;       // ... of the MRU kernel (measured r1e: one-warp CTAs, 7 instead of 6 warps/SM, are slower)
constexpr int kBuildThreads = 256;   // measured r1c: 128 -> 7.6 ms, 256 -> 5.7 ms per 131k instances

__device__ __forceinline__ int64_t ext_offset(const dls_batch_t& b, int inst, const Inst& v) {
  return b.T == 1 ? (int64_t)inst * (int64_t)(v.E + v.R) : b.pred_off[inst] + b.par_off[inst];
}

// -------------------------------------------------------------------------------------------------------------------
// stream_build_kernel: grid (B, 3). One CTA rewrites one instance for one processing order.
// -------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBuildThreads) stream_build_kernel(dls_batch_t b, Workspace w, int kind_mask) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  const int inst = blockIdx.x, kind = blockIdx.y, tid = threadIdx.x;
  if (!((kind_mask >> kind) & 1)) return;
  Inst v;
  bind_instance(b, inst, v);
  if (w.topo_status[v.topo] != DLS_ST_OK) return;
  const int n = v.n;
  if (n == 0) return;
  const int nsh_topo = w.nshared[v.topo];
  if (kind == kStreamMRU ? nsh_topo != 0 : nsh_topo > kSharedCap) return;  // routed to the general kernels
  uint32_t* s_extoff = reinterpret_cast<uint32_t*>(smem_raw);           // [n+1]
  uint16_t* s_pos = reinterpret_cast<uint16_t*>(s_extoff + (n + 1));    // [n] task -> order position
  const uint16_t* order = kind == kStreamDG ? w.order_dg + v.t_topo
                        : kind == kStreamMRU ? w.order_mru + v.t_topo
                                             : w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo);
  const uint16_t* depth = w.depth + v.t_topo;
  const uint16_t* sid = w.par_sid + b.pid_off[v.topo];
  Rec* out = w.stream[kind] + v.t_inst;
  uint16_t* ext = w.ext[kind] + ext_offset(b, inst, v);

  for (int t = tid; t < n; t += blockDim.x) {
    const int i = order[t];
    s_pos[i] = (uint16_t)t;
    const int npred = (int)(v.pred_ptr[i + 1] - v.pred_ptr[i]);
    int extra = max(0, npred - 3);
    const uint32_t qb = v.par_ptr[i], qe = v.par_ptr[i + 1];
    if (kind != kStreamMRU && nsh_topo > 0) {  // (no shared param in the DAG: nothing to look up, par_sid not even written)
      int nsh = 0;
      for (uint32_t e = qb; e < qe; ++e) nsh += sid[v.par_idx[e]] != 0xFFFF;
      extra += max(0, nsh - 2);
    }
    s_extoff[t] = (uint32_t)extra;
  }
  if (tid == 0) s_extoff[n] = 0;
  __syncthreads();
  block_exclusive_scan(s_extoff, n + 1, s_part);
  for (int t = tid; t < n; t += blockDim.x) {
    const int i = order[t];
    Rec rc;
    rc.m = v.mem[i];
    rc.c = v.dur[i];
    const uint32_t pb = v.pred_ptr[i], pe = v.pred_ptr[i + 1];
    uint32_t xo = s_extoff[t];
    rc.pred[0] = rc.pred[1] = rc.pred[2] = 0;
    rc.npred = (uint16_t)min(pe - pb, 65535u);
    for (uint32_t e = pb; e < pe; ++e) {
      const uint16_t pp = s_pos[v.pred_idx[e]];
      if (e - pb < 3) rc.pred[e - pb] = pp;
      else ext[xo++] = pp;
    }
    const uint32_t qb = v.par_ptr[i], qe = v.par_ptr[i + 1];
    rc.a0 = rc.a1 = 0;
    if (kind == kStreamMRU || nsh_topo == 0) {
      rc.npriv = (uint16_t)(qe - qb);  // all private: the count is all the kernels need
      rc.nsh = 0;
    } else {
      int nsh = 0, npriv = 0;
      for (uint32_t e = qb; e < qe; ++e) {
        const uint16_t s = sid[v.par_idx[e]];
        if (s == 0xFFFF) { ++npriv; continue; }
        if (nsh == 0) rc.a0 = s;
        else if (nsh == 1) rc.a1 = s;
        else ext[xo++] = s;
        ++nsh;
      }
      rc.npriv = (uint16_t)npriv;
      rc.nsh = (uint8_t)min(nsh, 255);
    }
    rc.flags = (t == 0 || depth[order[t - 1]] != depth[i]) ? kRecFirst : 0;
    out[t] = rc;
  }
}

// -------------------------------------------------------------------------------------------------------------------
// stream_build_staged_kernel: grid B. One CTA per instance builds all requested streams.
// Phase 1 reads the instance's arrays ONCE, in task order (fully coalesced), into task-indexed 32-byte records in
// shared memory. Phase 2, per processing order, permutes them through shared memory and writes the stream
// coalesced. Used when the staged records fit shared memory (n up to ~5.5k); otherwise stream_build_kernel.
// -------------------------------------------------------------------------------------------------------------------
// A task as staged in shared memory, in two 16-byte halves kept in two arrays: with 32-byte records every first half
// starts on one of only four bank groups, and the gather by processing order ran at 3.5 wavefronts per ideal one
// (ncu source view, round 1i); 16-byte strides spread the same gather over all eight.
struct __align__(16) TaskLo {
  double m, c;
};
struct __align__(16) TaskHi {  // dependencies still as task ids
  uint16_t pred[3];
  uint16_t npred;
  uint16_t npar;
  uint16_t depth;
  uint32_t pad;
};
struct TaskRec {
  TaskLo lo;
  TaskHi hi;
};
static_assert(sizeof(TaskLo) == 16 && sizeof(TaskHi) == 16, "two 16-byte halves");

// The record of task `i` (staged as `tr`) for one stream, all but the round flag. `posk`: task -> position in that
// order; ext lists (dependencies beyond three, shared params beyond two) start at ext[xo].
__device__ __forceinline__ Rec make_rec(int kind, const TaskRec& tr, int i, const Inst& v, int nsh_topo, const uint16_t* sid,
                                        const uint16_t* posk, uint16_t* ext, uint32_t xo) {
  Rec rc;
  rc.m = tr.lo.m;
  rc.c = tr.lo.c;
  rc.npred = tr.hi.npred;
  rc.pred[0] = posk[tr.hi.pred[0]];  // unused slots hold task 0: a valid index, masked by the count
  rc.pred[1] = posk[tr.hi.pred[1]];
  rc.pred[2] = posk[tr.hi.pred[2]];
  if (tr.hi.npred > 3) {
    const uint32_t pb = v.pred_ptr[i], pe = v.pred_ptr[i + 1];
    for (uint32_t e = pb + 3; e < pe; ++e) ext[xo++] = posk[v.pred_idx[e]];
  }
  rc.a0 = rc.a1 = 0;
  if (kind == kStreamMRU || nsh_topo == 0) {
    rc.npriv = tr.hi.npar;  // every param private: the count is all the kernels need
    rc.nsh = 0;
  } else {
    int nsh = 0, npriv = 0;
    const uint32_t qb = v.par_ptr[i], qe = v.par_ptr[i + 1];
    for (uint32_t e = qb; e < qe; ++e) {
      const uint16_t sd = sid[v.par_idx[e]];
      if (sd == 0xFFFF) { ++npriv; continue; }
      if (nsh == 0) rc.a0 = sd;
      else if (nsh == 1) rc.a1 = sd;
      else ext[xo++] = sd;
      ++nsh;
    }
    rc.npriv = (uint16_t)npriv;
    rc.nsh = (uint8_t)min(nsh, 255);
  }
  rc.flags = 0;
  return rc;
}

__global__ void __launch_bounds__(kBuildThreads) stream_build_staged_kernel(dls_batch_t b, Workspace w, int kind_mask) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint32_t s_part[33];
  const int inst = blockIdx.x, tid = threadIdx.x;
  Inst v;
  bind_instance(b, inst, v);
  if (w.topo_status[v.topo] != DLS_ST_OK) return;
  const int n = v.n;
  if (n == 0) return;
  const int nsh_topo = w.nshared[v.topo];
  const int n_al = (n + 1) & ~1;
  TaskLo* s_lo = reinterpret_cast<TaskLo*>(smem_raw);                        // [n]
  TaskHi* s_hi = reinterpret_cast<TaskHi*>(s_lo + n);                        // [n]
  uint32_t* s_extoff = reinterpret_cast<uint32_t*>(s_hi + n);                // [n+1]
  uint16_t* s_pos = reinterpret_cast<uint16_t*>(s_extoff + (n + 1));         // [3][n_al] task -> position, per order
  uint16_t* s_depth = s_pos + 3 * n_al;                                      // [n_al] compact copy for the round starts
  const uint16_t* depth = w.depth + v.t_topo;
  const uint16_t* sid = w.par_sid + b.pid_off[v.topo];
  const uint16_t* g_order[3] = {w.order_dg + v.t_topo, w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo),
                                w.order_mru + v.t_topo};
  static_assert(kStreamDG == 0 && kStreamCP == 1 && kStreamMRU == 2, "order table follows the stream kinds");
  bool use[3];
#pragma unroll
  for (int kind = 0; kind < 3; ++kind)  // DAGs the thread-per-DAG kernels leave to the general ones get no stream
    use[kind] = ((kind_mask >> kind) & 1) && (kind == kStreamMRU ? nsh_topo == 0 : nsh_topo <= kSharedCap);
  // ---- phase 1: task-order gather (a thread keeps several tasks' loads in flight), and the orders inverted ----
  int any_ext = 0;
  strided_for<4>(n, [&](int i) {
#pragma unroll
    for (int kind = 0; kind < 3; ++kind)
      if (use[kind]) s_pos[kind * n_al + __ldg(g_order[kind] + i)] = (uint16_t)i;
    TaskLo lo;
    TaskHi tr;
    lo.m = __ldg(v.mem + i);
    lo.c = __ldg(v.dur + i);
    const uint32_t pb = __ldg(v.pred_ptr + i), pe = __ldg(v.pred_ptr + i + 1);
    const uint32_t qb = __ldg(v.par_ptr + i), qe = __ldg(v.par_ptr + i + 1);
    tr.npred = (uint16_t)min(pe - pb, 65535u);
    tr.npar = (uint16_t)min(qe - qb, 65535u);
    tr.pred[0] = pe - pb > 0 ? __ldg(v.pred_idx + pb) : 0;
    tr.pred[1] = pe - pb > 1 ? __ldg(v.pred_idx + pb + 1) : 0;
    tr.pred[2] = pe - pb > 2 ? __ldg(v.pred_idx + pb + 2) : 0;
    tr.pad = 0;
    tr.depth = __ldg(depth + i);
    s_lo[i] = lo;
    s_hi[i] = tr;
    s_depth[i] = tr.depth;
    any_ext |= (pe - pb > 3) || (nsh_topo > 0 && qe - qb > 2);
  });
  const bool ext_needed = __syncthreads_or(any_ext) != 0;
  if (!ext_needed) {
    // Where each round starts is the same in all three orders (each sorts by depth first). Without ext lists the
    // offset array is free to hold those starts: "is this the first task of a round" is one lookup.
    uint16_t* s_lvl = reinterpret_cast<uint16_t*>(s_extoff);
    const uint16_t* odg = w.order_dg + v.t_topo;
    strided_for<4>(n, [&](int t) {
      const int d = s_depth[__ldg(odg + t)];
      const int dprev = t > 0 ? (int)s_depth[__ldg(odg + t - 1)] : -1;
      if (d != dprev) s_lvl[d] = (uint16_t)t;
    });
    __syncthreads();
    // ---- phase 2: position t of all the orders in one iteration: three independent gathers per thread ----
    strided_for<2>(n, [&](int t) {
#pragma unroll
      for (int kind = 0; kind < 3; ++kind) {
        if (!use[kind]) continue;
        const int i = __ldg(g_order[kind] + t);
        TaskRec tr;
        tr.lo = s_lo[i];
        tr.hi = s_hi[i];
        Rec rc = make_rec(kind, tr, i, v, nsh_topo, sid, s_pos + kind * n_al, w.ext[kind], 0);
        rc.flags = t == (int)s_lvl[tr.hi.depth] ? kRecFirst : 0;
        (w.stream[kind] + v.t_inst)[t] = rc;
      }
    });
    return;
  }
  // Some task carries an ext list: its offset is a prefix sum over the positions of the order, one order at a time.
  for (int kind = 0; kind < 3; ++kind) {
    if (!use[kind]) continue;
    const uint16_t* order = g_order[kind];
    const uint16_t* posk = s_pos + kind * n_al;
    Rec* out = w.stream[kind] + v.t_inst;
    uint16_t* ext = w.ext[kind] + ext_offset(b, inst, v);
    __syncthreads();  // the previous order is done with s_extoff
    for (int t = tid; t < n; t += blockDim.x) {
      const int i = order[t];
      int extra = max(0, (int)s_hi[i].npred - 3);
      if (kind != kStreamMRU && nsh_topo > 0) {
        int nsh = 0;
        for (uint32_t e = v.par_ptr[i]; e < v.par_ptr[i + 1]; ++e) nsh += sid[v.par_idx[e]] != 0xFFFF;
        extra += max(0, nsh - 2);
      }
      s_extoff[t] = (uint32_t)extra;
    }
    if (tid == 0) s_extoff[n] = 0;
    __syncthreads();
    block_exclusive_scan(s_extoff, n + 1, s_part);
    for (int t = tid; t < n; t += blockDim.x) {
      const int i = order[t];
      TaskRec tr;
      tr.lo = s_lo[i];
      tr.hi = s_hi[i];
      Rec rc = make_rec(kind, tr, i, v, nsh_topo, sid, posk, ext, s_extoff[t]);
      rc.flags = (t == 0 || s_depth[order[t - 1]] != tr.hi.depth) ? kRecFirst : 0;
      out[t] = rc;
    }
  }
}

// -------------------------------------------------------------------------------------------------------------------
// Shared pieces of the thread-per-DAG kernels
// -------------------------------------------------------------------------------------------------------------------
struct RecRegs {
  uint4 lo, hi;
};
__device__ __forceinline__ RecRegs load_rec(const Rec* p) {
  RecRegs r;
  const uint4* q = reinterpret_cast<const uint4*>(p);
  // L2::128B: a miss in L2 brings in the whole 128-byte line, i.e. this DAG's next three records as well
  // and one 256-bit load instead of two 128-bit ones halves the requests of the 32 lanes, each on its own line
  asm volatile("ld.global.nc.L2::128B.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w)
               : "l"(q));
  return r;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// A thread's verdict bytes (node or DLS_TASK_*) by order position. They leave in aligned words: 32 lanes storing one
// byte each to 32 different lines cost the L1-to-L2 request path — what bounds these kernels — as much as a record
// load. The (up to three) positions not yet stored are answered from the register.
struct VerdictOut {
  int8_t* base;
  uint32_t mis, acc;
  int flushed;  // positions below are in global memory
  __device__ __forceinline__ void init(int8_t* p) {
    base = p;
    mis = (uint32_t)(uintptr_t)p & 3u;
    acc = 0;
    flushed = 0;
  }
  __device__ __forceinline__ void put(int t, int verdict, bool last) {
    const uint32_t q = ((uint32_t)t + mis) & 3u;
    acc |= (uint32_t)(uint8_t)verdict << (8 * q);
    if (q == 3 || last) {
      if (q == 3 && t - flushed == 3) {
        *reinterpret_cast<uint32_t*>(base + t - 3) = acc;
      } else {
        for (int p = flushed; p <= t; ++p) base[p] = (int8_t)(acc >> (8 * (((uint32_t)p + mis) & 3u)));
      }
      flushed = t + 1;
      acc = 0;
    }
  }
  __device__ __forceinline__ int get(uint32_t p) const {  // p below the position being decided
    if ((int)p >= flushed) return (int8_t)(acc >> (8 * ((p + mis) & 3u)));
    return __ldcg(base + p);
  }
};

// Per-warp shared-memory slab: word `wd` of the calling lane.
struct Slab {
  uint32_t* u;  // [words][32]
  double* d;    // [dwords][32]
  int lane;
  __device__ __forceinline__ uint32_t& w(int wd) const { return u[wd * 32 + lane]; }
  __device__ __forceinline__ double& f(int k) const { return d[k * 32 + lane]; }
};

template <int KMAX>
__device__ __forceinline__ double pick(const double (&a)[KMAX], int k) {
  double x = a[0];
#pragma unroll
  for (int j = 1; j < KMAX; ++j) x = (k == j) ? a[j] : x;
  return x;
}

struct ThreadOut {
  int completed, failed, hits, misses, rounds, status;
  bool fail_rest;
};

template <int KMAX>
__device__ __forceinline__ void finish_thread(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int slot,
                                              int inst, const Inst& v, const double (&avail)[KMAX], const Slab& sl,
                                              int fin_base, int cnt_base, const ThreadOut& to, int route_or = 0) {
  const int64_t ko = (int64_t)slot * b.nk_inst + v.k_out;
  double mk = 0.0;
  bool any = false;
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    if (k < v.K) {
      const double f = sl.f(fin_base + k);
      const int c = (int)sl.w(cnt_base + k);
      r.fin[ko + k] = f;
      r.cnt[ko + k] = c;
      r.avail[ko + k] = avail[k];
      if (c > 0 && (!any || f > mk)) { mk = f; any = true; }
    }
  }
  dls_summary_t s;
  s.makespan = any ? mk : 0.0;
  s.completed = to.completed;
  s.failed = to.fail_rest ? v.n - to.completed : to.failed;
  s.pending = v.n - s.completed - s.failed;
  s.hits = to.hits;
  s.misses = to.misses;
  s.rounds = to.rounds;
  s.status = to.status;
  s.reserved = 0;
  r.summary[(int64_t)slot * b.B + inst] = s;
  w.route[(int64_t)slot * b.B + inst] = (uint8_t)((to.fail_rest ? kRouteFailRest : 0) | route_or);
}

// -------------------------------------------------------------------------------------------------------------------
// DFS / Greedy / Critical, one thread per DAG
// -------------------------------------------------------------------------------------------------------------------
template <int ALGO, int KMAX>
__global__ void __launch_bounds__(kTPB) sched_stream_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot,
                                                       int twin_mode, int twin_slot) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  // slab layout per warp: doubles fin[KMAX], spd[KMAX]; words smask[kSharedCap/4], cnt[KMAX]
  constexpr int kD = 2 * KMAX;
  const int kW = kSharedCap / 4 + KMAX;
  const size_t warp_bytes = (size_t)kD * 32 * 8 + (size_t)kW * 32 * 4;
  Slab sl;
  sl.d = reinterpret_cast<double*>(smem_raw + wib * warp_bytes);
  sl.u = reinterpret_cast<uint32_t*>(smem_raw + wib * warp_bytes + (size_t)kD * 32 * 8);
  sl.lane = lane;
  const int FIN = 0, SPD = KMAX, SMASK = 0, CNT = kSharedCap / 4;

  const int inst = blockIdx.x * kTPB + threadIdx.x;
  const bool have = inst < b.B;
  Inst v;
  v.n = 0; v.K = 0; v.topo = 0; v.t_inst = 0;
  // DFS and Greedy walk the same order, and on a DAG without shared params they are the same function: every param
  // must be loaded whichever node is taken, so Greedy's (fewest loads, then most memory) is DFS's (most memory), see
  // the node choice below. The DFS launch (twin_mode 1) then stores its result in Greedy's slot as well, and the Greedy
  // launch (twin_mode 2) leaves those instances alone. The one case where the two differ even so (no node has more
  // than -1 available: DFS starts from max_memory = -1, Greedy from the first fitting node) withdraws the offer.
  bool elig = false, twin_ok = false;
  if (have) {
    bind_instance(b, inst, v);
    bool taken = false;
    if (twin_mode == 2) {
      const int rt = w.route[(int64_t)slot * b.B + inst];
      taken = rt != kRouteOpen && (rt & kRouteTwin);
    }
    elig = !taken && w.topo_status[v.topo] == DLS_ST_OK && v.n > 0 && v.K <= KMAX && w.nshared[v.topo] <= kSharedCap;
    if (!elig && !taken) w.route[(int64_t)slot * b.B + inst] = kRouteGeneral;
    twin_ok = ALGO == DLS_ALGO_DFS && twin_mode == 1 && elig && w.nshared[v.topo] == 0;
  }
  const int n = elig ? v.n : 0, K = v.K;
  const int n_warp = __reduce_max_sync(DLS_FULL_MASK, n);
  if (n_warp == 0) {
    if (have && twin_mode == 1) w.route[(int64_t)twin_slot * b.B + inst] = kRouteOpen;
    return;
  }
  constexpr int kind = ALGO == DLS_ALGO_CRITICAL ? kStreamCP : kStreamDG;
  const Rec* rec = w.stream[kind] + v.t_inst;
  const uint16_t* ext = w.ext[kind] + (elig ? ext_offset(b, inst, v) : 0);
  VerdictOut vout;
  vout.init(w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst);

  // Node state. avail[k] of a node that does not exist is -inf: it never fits and never wins a max.
  double avail[KMAX];
  // Critical: nodes ranked by (speed desc, index asc) once; key = rank*KMAX + k, nodes with speed <= 0 never chosen
  // ("node.compute_speed > best_speed" with best_speed = 0, ref schedulers.py:353-358).
  int ckey[ALGO == DLS_ALGO_CRITICAL ? KMAX : 1];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const bool on = elig && k < K;
    avail[k] = on ? v.cap[k] : -INFINITY;
    const double sp = on ? v.spd[k] : 1.0;
    sl.f(FIN + k) = 0.0;
    sl.f(SPD + k) = sp;
    sl.w(CNT + k) = 0;
  }
  if constexpr (ALGO == DLS_ALGO_CRITICAL) {
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const double sk = sl.f(SPD + k);
      int rank = 0;
#pragma unroll
      for (int j = 0; j < KMAX; ++j) {
        const double sj = sl.f(SPD + j);
        rank += (j < K) && (sj > sk || (sj == sk && j < k));
      }
      ckey[k] = (k < K && sk > 0.0) ? rank * KMAX + k : 0x7fffffff;
    }
  }
  for (int x = 0; x < kSharedCap / 4; ++x) sl.w(SMASK + x) = 0;

  ThreadOut to;
  to.completed = to.failed = to.hits = to.misses = to.rounds = 0;
  to.status = DLS_ST_OK;
  to.fail_rest = false;
  bool running = elig, any_ready = false, any_assigned = false;
  uint32_t first_nd = 0xffffffffu;  // smallest order position whose task is not completed (none so far)
  uint32_t xpos = 0;
  RecRegs cur;
  cur.lo = cur.hi = make_uint4(0, 0, 0, 0);
  if (n > 0) cur = load_rec(rec);
  for (int t = 0; t < n_warp; ++t) {
    RecRegs nxt = cur;
    if (t + 1 < n) nxt = load_rec(rec + t + 1);
    if ((t & 3) == 0 && t + 8 < n) prefetch_l2(rec + t + 8);  // once per 128-byte line; (measured: of no use to MRU)
    if (t < n) {
      const double m = __hiloint2double((int)cur.lo.y, (int)cur.lo.x);
      const double c = __hiloint2double((int)cur.lo.w, (int)cur.lo.z);
      const int npred = cur.hi.y >> 16, npriv = cur.hi.w & 0xffff, nsh = (cur.hi.w >> 16) & 0xff;
      const int flags = cur.hi.w >> 24;
      if (flags & kRecFirst) {
        if (t > 0 && running) {
          if (!any_ready) running = false;                       // ready list empty (ref schedulers.py:172-174)
          else if (!any_assigned) { running = false; to.fail_rest = true; }  // no progress (ref :202-206)
        }
        if (running) { ++to.rounds; any_ready = any_assigned = false; }
      }
      int verdict = DLS_TASK_PENDING;
      const uint32_t x_pred = xpos, x_sh = xpos + (uint32_t)max(0, npred - 3);
      xpos = x_sh + (uint32_t)max(0, nsh - 2);
      if (running) {
        // is_task_ready (ref :50-53) without a done-bitmap: every position below `first_nd` is done, so a task
        // whose dependencies all sit below it is ready with no memory access at all; only after the first failed or
        // blocked task do the (few) later dependencies get looked up in this thread's own verdict bytes.
        bool ready = true;
        if (npred > 0) {
          const uint32_t p0 = cur.hi.x & 0xffff, p1 = cur.hi.x >> 16, p2 = cur.hi.y & 0xffff;
          uint32_t pm = p0;
          if (npred > 1) pm = max(pm, p1);
          if (npred > 2) pm = max(pm, p2);
          if (pm >= first_nd || npred > 3) {
            auto done_at = [&](uint32_t p) -> bool { return p < first_nd || vout.get(p) >= 0; };
            ready = done_at(p0) && (npred < 2 || done_at(p1)) && (npred < 3 || done_at(p2));
            for (int j = 3; j < npred; ++j) ready = ready && done_at(ext[x_pred + j - 3]);
          }
        }
        if (ready) {
          any_ready = true;
          int best = -1, lbest = npriv;
          double a = 0.0;  // available memory of the chosen node
          if (nsh == 0) {
            // Every param of the task is private: params_to_load and hence the requirement are the same on every
            // node, so "fits" is avail >= need and the best fitting node is the first maximum, if that one fits.
            const double need = memory_requirement(m, npriv);
            if constexpr (ALGO == DLS_ALGO_CRITICAL) {
              int bk = 0x7fffffff;
#pragma unroll
              for (int k = 0; k < KMAX; ++k) bk = min(bk, need <= avail[k] ? ckey[k] : 0x7fffffff);
              if (bk != 0x7fffffff) { best = bk & (KMAX - 1); a = pick<KMAX>(avail, best); }
            } else {
              // DFS: strict '>' from max_memory = -1 (ref :186-192). Greedy: first fitting node, then strictly more
              // available memory among equal params_to_load (ref :267-281) == first maximum among fitting nodes.
              double bv = ALGO == DLS_ALGO_DFS ? -1.0 : -INFINITY;
#pragma unroll
              for (int k = 0; k < KMAX; ++k)
                if (avail[k] > bv) { bv = avail[k]; best = k; }
              if constexpr (ALGO == DLS_ALGO_DFS) { if (best < 0) twin_ok = false; }
              if (!(need <= bv)) best = -1;
              a = bv;
            }
          } else {
            // shared params not yet cached, per node, packed one byte per node
            unsigned long long miss = 0;
            for (int j = 0; j < nsh; ++j) {
              const int sd = j == 0 ? (int)(cur.hi.z & 0xffff) : j == 1 ? (int)(cur.hi.z >> 16) : (int)ext[x_sh + j - 2];
              const uint32_t mask = (sl.w(SMASK + (sd >> 2)) >> ((sd & 3) * 8)) & 0xffu;
              const unsigned long long spread = ((unsigned long long)(mask ^ 0xffu) * 0x0101010101010101ull) & 0x8040201008040201ull;
              miss += ((spread + 0x7f7f7f7f7f7f7f7full) >> 7) & 0x0101010101010101ull;
            }
            double bv = ALGO == DLS_ALGO_DFS ? -1.0 : 0.0;  // max_memory = -1 (ref :187) / best_available_memory = 0 (ref :269)
            int bl = 0x7fffffff, bk = 0x7fffffff;             // min_params_to_load = inf (ref :268)
#pragma unroll
            for (int k = 0; k < KMAX; ++k) {
              const int lk = npriv + (int)((miss >> (8 * k)) & 0xffull);
              const bool fits = memory_requirement(m, lk) <= avail[k];  // can_fit_on_node (ref :74-76)
              bool take;
              if constexpr (ALGO == DLS_ALGO_DFS) take = fits && avail[k] > bv;                                     // ref :190
              else if constexpr (ALGO == DLS_ALGO_GREEDY) take = fits && (lk < bl || (lk == bl && avail[k] > bv));  // ref :276-278
              else take = fits && ckey[k] < bk;                                                                     // ref :356
              if (take) {
                best = k;
                lbest = lk;
                bl = lk;
                bv = avail[k];
                if constexpr (ALGO == DLS_ALGO_CRITICAL) bk = ckey[k];
              }
            }
            a = bv;
            if constexpr (ALGO == DLS_ALGO_CRITICAL) { if (best >= 0) a = pick<KMAX>(avail, best); }
          }
          if (best >= 0) {
            if (lbest > 0) a = __dsub_rn(a, 0.5);  // node.available_memory -= 0.5 per loaded param (ref :89)
            if (lbest > 1) a = __dsub_rn(a, 0.5);
            for (int j = 2; j < lbest; ++j) a = __dsub_rn(a, 0.5);
            a = __dsub_rn(a, m);                   // ref :95
            a = __dadd_rn(a, m);                   // complete_task frees it again (ref :126)
#pragma unroll
            for (int k = 0; k < KMAX; ++k) avail[k] = (k == best) ? a : avail[k];
            sl.f(FIN + best) = __dadd_rn(sl.f(FIN + best), __ddiv_rn(c, sl.f(SPD + best)));  // ref simulation.py:261-264
            sl.w(CNT + best) += 1;
            to.misses += lbest;
            to.hits += npriv + nsh - lbest;
            for (int j = 0; j < nsh; ++j) {
              const int sd = j == 0 ? (int)(cur.hi.z & 0xffff) : j == 1 ? (int)(cur.hi.z >> 16) : (int)ext[x_sh + j - 2];
              sl.w(SMASK + (sd >> 2)) |= (1u << best) << ((sd & 3) * 8);
            }
            verdict = best;
            ++to.completed;
            any_assigned = true;
          } else {
            verdict = DLS_TASK_FAILED;  // fail_task (ref :128-131)
            ++to.failed;
          }
        }
      }
      if (verdict < 0) first_nd = min(first_nd, (uint32_t)t);
      vout.put(t, verdict, t == n - 1);
    }
    cur = nxt;
  }
  if (elig) {
    if (running) {
      if (any_ready && !any_assigned) to.fail_rest = true;
      else if (any_ready && to.completed + to.failed < n) ++to.rounds;  // one more iteration finds nothing ready
    }
    finish_thread<KMAX>(b, r, w, slot, inst, v, avail, sl, FIN, CNT, to);
    if (twin_ok) finish_thread<KMAX>(b, r, w, twin_slot, inst, v, avail, sl, FIN, CNT, to, kRouteTwin);
  }
  if (have && twin_mode == 1 && !twin_ok) w.route[(int64_t)twin_slot * b.B + inst] = kRouteOpen;
}

// -------------------------------------------------------------------------------------------------------------------
// MRU for DAGs whose params are all private to one task, one thread per DAG.
//
// A private param is loaded when its only user is assigned and is never asked for again, so no task ever finds it
// cached: what MRU's cache does for such a DAG is hold memory until evict_params_for_task frees it, 0.5 per param.
// Every cached param of the node is a candidate (none is in the task's params_needed), so WHICH ones the
// (score, name) sort of ref schedulers.py:420 puts first changes node.cached_params but none of the outputs this
// path produces (verdicts, counters, finish times, available memory): those depend on HOW MANY are cached per node.
// The kernel therefore keeps one count per node and nothing per param. (The `cached` extension output needs the
// identities; asking for it selects the warp-per-DAG kernel, which keeps them.) Rounds 1b-1f kept 4 bits per param in
// shared memory to replay the eviction order: 24 KB per warp, 6 warps per SM, 7 ms per 131k instances.
// -------------------------------------------------------------------------------------------------------------------

// evict_params_for_task (ref schedulers.py:404-442) on one node, as arithmetic. Each of the node's `ncached` params
// frees exactly 0.5 and the reference walks them until `freed >= shortage`: it needs j = ceil(shortage / 0.5) of
// them and succeeds iff j <= ncached. On success the node's available memory goes up by 0.5, j times (ref :429); on
// failure all ncached are evicted and rolled back, which leaves only the +0.5 / -0.5 rounding trace (ref :438-441).
__device__ __forceinline__ bool plan_evict(double need, double& avail_k, uint32_t& ncached) {
  const double shortage = __dsub_rn(need, avail_k);
  if (shortage <= 0.0) return true;
  const double want = ceil(__dmul_rn(shortage, 2.0));  // exact: power-of-two scaling, then ceil
  if (want <= (double)ncached) {
    const uint32_t j = (uint32_t)want;
    for (uint32_t q = 0; q < j; ++q) avail_k = __dadd_rn(avail_k, 0.5);
    ncached -= j;
    return true;
  }
  for (uint32_t q = 0; q < ncached; ++q) avail_k = __dadd_rn(avail_k, 0.5);
  for (uint32_t q = 0; q < ncached; ++q) avail_k = __dsub_rn(avail_k, 0.5);
  return false;
}

template <int KMAX>
__global__ void __launch_bounds__(kTPBM) sched_stream_mru_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  // slab per warp: doubles fin[KMAX], spd[KMAX], hcnt[KMAX]; words cnt[KMAX], ncached[KMAX]
  constexpr int kD = 3 * KMAX;
  constexpr int kW = 2 * KMAX;
  const size_t warp_bytes = (size_t)kD * 32 * 8 + (size_t)kW * 32 * 4;
  Slab sl;
  sl.d = reinterpret_cast<double*>(smem_raw + wib * warp_bytes);
  sl.u = reinterpret_cast<uint32_t*>(smem_raw + wib * warp_bytes + (size_t)kD * 32 * 8);
  sl.lane = lane;
  // HCNT[k] = len(node.completed_tasks) * 0.5 kept as a double: every += 0.5 is exact (ref schedulers.py:498)
  const int FIN = 0, SPD = KMAX, HCNT = 2 * KMAX, CNT = 0, NC = KMAX;

  const int inst = blockIdx.x * kTPBM + threadIdx.x;
  const bool have = inst < b.B;
  Inst v;
  v.n = 0; v.K = 0; v.topo = 0; v.P = 0; v.t_inst = 0;
  bool elig = false;
  if (have) {
    bind_instance(b, inst, v);
    elig = w.topo_status[v.topo] == DLS_ST_OK && v.n > 0 && v.K <= KMAX && w.nshared[v.topo] == 0;
    if (!elig) w.route[(int64_t)slot * b.B + inst] = kRouteGeneral;
  }
  const int n = elig ? v.n : 0, K = v.K;
  const int n_warp = __reduce_max_sync(DLS_FULL_MASK, n);
  if (n_warp == 0) return;
  const Rec* rec = w.stream[kStreamMRU] + v.t_inst;
  const uint16_t* ext = w.ext[kStreamMRU] + (elig ? ext_offset(b, inst, v) : 0);
  VerdictOut vout;
  vout.init(w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst);

  double avail[KMAX];  // -inf for nodes that do not exist: they never fit
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const bool on = elig && k < K;
    avail[k] = on ? v.cap[k] : -INFINITY;
    sl.f(FIN + k) = 0.0;
    sl.f(SPD + k) = on ? v.spd[k] : 1.0;
    sl.f(HCNT + k) = 0.0;
    sl.w(CNT + k) = 0;
    sl.w(NC + k) = 0;  // len(node.cached_params)
  }

  ThreadOut to;
  to.completed = to.failed = to.hits = to.misses = to.rounds = 0;
  to.status = DLS_ST_OK;
  to.fail_rest = false;
  int carried = 0;
  bool running = elig, any_ready = false, any_assigned = false;
  uint32_t first_nd = 0xffffffffu, xpos = 0;
  RecRegs cur;
  cur.lo = cur.hi = make_uint4(0, 0, 0, 0);
  if (n > 0) cur = load_rec(rec);
  for (int t = 0; t < n_warp; ++t) {
    RecRegs nxt = cur;
    if (t + 1 < n) nxt = load_rec(rec + t + 1);
    if (t < n) {
      const double m = __hiloint2double((int)cur.lo.y, (int)cur.lo.x);
      const double c = __hiloint2double((int)cur.lo.w, (int)cur.lo.z);
      const int npred = cur.hi.y >> 16, npar = cur.hi.w & 0xffff;
      const int flags = cur.hi.w >> 24;
      if (flags & kRecFirst) {
        if (t > 0 && running) {
          if (!any_ready) running = false;
          else if (!any_assigned) { running = false; to.fail_rest = true; }
        }
        if (running) { ++to.rounds; any_ready = any_assigned = false; }  // self.time_step += 1 (ref :453)
      }
      int verdict = DLS_TASK_PENDING;
      const uint32_t x_pred = xpos;
      xpos += (uint32_t)max(0, npred - 3);
      if (running) {
        // is_task_ready (ref :50-53) without a done-bitmap: every position below `first_nd` is done, so a task
        // whose dependencies all sit below it is ready with no memory access at all; only after the first failed or
        // blocked task do the (few) later dependencies get looked up in this thread's own verdict bytes.
        bool ready = true;
        if (npred > 0) {
          const uint32_t p0 = cur.hi.x & 0xffff, p1 = cur.hi.x >> 16, p2 = cur.hi.y & 0xffff;
          uint32_t pm = p0;
          if (npred > 1) pm = max(pm, p1);
          if (npred > 2) pm = max(pm, p2);
          if (pm >= first_nd || npred > 3) {
            auto done_at = [&](uint32_t p) -> bool { return p < first_nd || vout.get(p) >= 0; };
            ready = done_at(p0) && (npred < 2 || done_at(p1)) && (npred < 3 || done_at(p2));
            for (int j = 3; j < npred; ++j) ready = ready && done_at(ext[x_pred + j - 3]);
          }
        }
        if (ready) {
          any_ready = true;
          // none of the task's (private) params is cached anywhere: params_to_load = npar on every node
          const double need = memory_requirement(m, npar);
          uint32_t fits = 0;  // can_fit_on_node per node; a node that does not exist holds -inf and never fits
#pragma unroll
          for (int k = 0; k < KMAX; ++k) fits |= need <= avail[k] ? 1u << k : 0u;
          int best = -1;
          double best_score = -INFINITY, a = 0.0;  // a: available memory of the chosen node
          if (fits == (1u << K) - 1u) {
            // no memory pressure: score = 0.0 + cached*20 + available_memory - len(completed_tasks)*0.5 with cached == 0.
            // (0.0 + x is x, but for x = -0.0, and then only the sign of a zero score differs: no comparison sees it)
            double hc[KMAX];
#pragma unroll
            for (int k = 0; k < KMAX; ++k) hc[k] = sl.f(HCNT + k);
#pragma unroll
            for (int k = 0; k < KMAX; ++k) {
              const double score = __dsub_rn(avail[k], hc[k]);  // ref :483-498 (avail = -inf off the end)
              if (score > best_score) { best_score = score; best = k; a = avail[k]; }
            }
          } else {
#pragma unroll
            for (int k = 0; k < KMAX; ++k) {
              if (k < K) {
                double score = 0.0;
                bool cand = true;
                if ((fits >> k) & 1u) {
                  score = __dadd_rn(score, avail[k]);  // ref :490-491
                } else {
                  // evict_params_for_task mutates the node even if another node wins (ref :492)
                  uint32_t nc = sl.w(NC + k);
                  const bool ok = plan_evict(need, avail[k], nc);
                  sl.w(NC + k) = nc;
                  if (ok) score = __dadd_rn(score, 5.0);
                  else cand = false;
                }
                score = __dsub_rn(score, sl.f(HCNT + k));  // ref :498
                if (cand && score > best_score) { best_score = score; best = k; }
              }
            }
          }
          if (best >= 0 && fits != (1u << K) - 1u) {
            a = pick<KMAX>(avail, best);
            if (!(need <= a)) {  // ref :506-507 (only after a rounding mishap in the first eviction)
              uint32_t nc = sl.w(NC + best);
              plan_evict(need, a, nc);
              sl.w(NC + best) = nc;
            }
          }
          if (best >= 0) {
            if (need > a) {
              // assign_task_to_node refuses (ref :82-83): the reference retries next round; flagged, not replayed
              to.status |= DLS_ST_CARRY;
              ++carried;
#pragma unroll
              for (int k = 0; k < KMAX; ++k) avail[k] = (k == best) ? a : avail[k];
            } else {
              if (npar > 0) a = __dsub_rn(a, 0.5);
              if (npar > 1) a = __dsub_rn(a, 0.5);
              for (int j = 2; j < npar; ++j) a = __dsub_rn(a, 0.5);
              a = __dsub_rn(a, m);
              a = __dadd_rn(a, m);
#pragma unroll
              for (int k = 0; k < KMAX; ++k) avail[k] = (k == best) ? a : avail[k];
              sl.f(FIN + best) = __dadd_rn(sl.f(FIN + best), __ddiv_rn(c, sl.f(SPD + best)));
              sl.f(HCNT + best) = __dadd_rn(sl.f(HCNT + best), 0.5);
              sl.w(CNT + best) += 1;
              to.misses += npar;
              sl.w(NC + best) += (uint32_t)npar;  // node.cached_params.add(param) for each loaded param (ref :90)
              verdict = best;
              ++to.completed;
              any_assigned = true;
            }
          } else {
            verdict = DLS_TASK_FAILED;
            ++to.failed;
          }
        }
      }
      if (verdict < 0) first_nd = min(first_nd, (uint32_t)t);
      vout.put(t, verdict, t == n - 1);
    }
    cur = nxt;
  }
  if (elig) {
    if (running) {
      if (any_ready && !any_assigned) to.fail_rest = true;
      else if (any_ready && to.completed + to.failed + carried < n) ++to.rounds;
    }
    finish_thread<KMAX>(b, r, w, slot, inst, v, avail, sl, FIN, CNT, to);
  }
}

// -------------------------------------------------------------------------------------------------------------------
// unpermute_kernel: one warp per instance. node_of[order[t]] = verdict[t]; seq_pos = prefix count of assignments.
// -------------------------------------------------------------------------------------------------------------------
struct SlotKinds {
  int kind[4];
  int follower[4];  // slot whose kRouteTwin instances take this slot's verdicts (-1: none)
};

// grid (ceil(B/4), A): one warp per (instance, algorithm slot). With `staged`, verdicts are scattered into shared
// memory first and written out coalesced; without (per-instance arrays too large), they are scattered directly.
__global__ void __launch_bounds__(128) unpermute_kernel(dls_batch_t b, dls_result_t r, Workspace w, SlotKinds sk, int n_al,
                                                        int staged) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int inst = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31, slot = blockIdx.y;
  if (inst >= b.B) return;
  const int route = w.route[(int64_t)slot * b.B + inst];
  if (route == kRouteGeneral) return;  // the warp-per-DAG kernel writes this instance
  if (route & kRouteTwin) return;      // written together with the slot it shares its verdicts with, below
  const int fol = sk.follower[slot];
  const bool dup = fol >= 0 && (w.route[(int64_t)fol * b.B + inst] & kRouteTwin);
  const int kind = sk.kind[slot];
  Inst v;
  bind_instance(b, inst, v);
  const uint16_t* order = kind == kStreamDG ? w.order_dg + v.t_topo
                        : kind == kStreamMRU ? w.order_mru + v.t_topo
                                             : w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo);
  const int8_t* npos = w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst;
  int8_t* node_of = r.node_of + (int64_t)slot * b.nt_inst + v.t_inst;
  uint16_t* seq_pos = r.seq_pos + (int64_t)slot * b.nt_inst + v.t_inst;
  int8_t* node_of2 = r.node_of + (int64_t)(dup ? fol : slot) * b.nt_inst + v.t_inst;
  uint16_t* seq_pos2 = r.seq_pos + (int64_t)(dup ? fol : slot) * b.nt_inst + v.t_inst;
  int8_t* s_node = reinterpret_cast<int8_t*>(smem_raw) + (size_t)(threadIdx.x >> 5) * 3 * n_al;
  uint16_t* s_seq = reinterpret_cast<uint16_t*>(s_node + n_al);
  int base = 0;
  constexpr int kU = 4;  // chunks of 32 positions whose (independent) loads are issued together
  for (int t0 = 0; t0 < v.n; t0 += 32 * kU) {
    int verdicts[kU], tasks[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int t = t0 + 32 * u + lane;
      verdicts[u] = t < v.n ? (int)npos[t] : DLS_TASK_PENDING;
      tasks[u] = t < v.n ? (int)order[t] : 0;
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int t = t0 + 32 * u + lane;
      int verdict = verdicts[u];
      const unsigned asg = __ballot_sync(DLS_FULL_MASK, verdict >= 0);
      if (t < v.n) {
        const int i = tasks[u];
        if ((route & kRouteFailRest) && verdict == DLS_TASK_PENDING) verdict = DLS_TASK_FAILED;  // "fail all pending" (ref :202-206)
        const uint16_t sq = verdict >= 0 ? (uint16_t)(base + __popc(asg & ((1u << lane) - 1u))) : (uint16_t)DLS_SEQ_NONE;
        if (staged) { s_node[i] = (int8_t)verdict; s_seq[i] = sq; }
        else {
          node_of[i] = (int8_t)verdict; seq_pos[i] = sq;
          if (dup) { node_of2[i] = (int8_t)verdict; seq_pos2[i] = sq; }
        }
      }
      base += __popc(asg);
    }
  }
  if (staged) {
    __syncwarp();
    for (int i = lane; i < v.n; i += 32) {
      const int8_t nd = s_node[i];
      const uint16_t sq = s_seq[i];
      node_of[i] = nd;
      seq_pos[i] = sq;
      if (dup) { node_of2[i] = nd; seq_pos2[i] = sq; }
    }
  }
}

template <int KMAX>
size_t smem_simple() {
  return (size_t)(kTPB / 32) * ((size_t)2 * KMAX * 32 * 8 + (size_t)(kSharedCap / 4 + KMAX) * 32 * 4);
}
template <int KMAX>
size_t smem_mru() {
  return (size_t)(kTPBM / 32) * ((size_t)3 * KMAX * 32 * 8 + (size_t)(2 * KMAX) * 32 * 4);
}

}  // namespace

// The streamed path needs K <= 8 nodes and per-thread state small enough for a useful number of resident warps.
bool stream_path_possible(const dls_batch_t& b, const dls_result_t* r, int smem_limit) {
  if (b.K_max > 8) return false;
  if (r && (r->peak || r->cached || r->tfin)) return false;  // extension outputs are produced by the general kernels only
  return smem_simple<8>() <= (size_t)smem_limit / 2 && b.n_max >= 1;
}

template <class Kern, class... Args>
static cudaError_t launch_fast(Kern kern, int kernel_id, dim3 grid, int threads, size_t smem, LaunchCtx& ctx, Args... args) {
  ctx.why = "cudaFuncSetAttribute(stream kernel)";
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int rec = ctx.prof ? ctx.prof->begin(kernel_id, ctx.st) : -1;
  kern<<<grid, threads, smem, ctx.st>>>(args...);
  if (ctx.prof) ctx.prof->end(rec, ctx.st);
  ++ctx.launches;
  ctx.why = "stream kernel launch";
  return cudaGetLastError();
}

cudaError_t launch_stream_build(const dls_batch_t& b, const Workspace& w, int algo_mask, LaunchCtx& ctx) {
  int kinds = 0;
  if (algo_mask & 3) kinds |= 1 << kStreamDG;
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) kinds |= 1 << kStreamCP;
  if (algo_mask & (1 << DLS_ALGO_MRU)) kinds |= 1 << kStreamMRU;
  const size_t staged = (size_t)b.n_max * 32 + ((size_t)b.n_max + 1) * sizeof(uint32_t) +
                        4 * (((size_t)b.n_max + 1) & ~(size_t)1) * sizeof(uint16_t) + 16;
  if (staged <= 180 * 1024)
    return launch_fast(stream_build_staged_kernel, DLS_KERNEL_STREAM_BUILD, dim3(b.B), kBuildThreads, staged, ctx, b, w, kinds);
  const size_t smem = ((size_t)b.n_max + 1) * sizeof(uint32_t) + (((size_t)b.n_max + 1) & ~(size_t)1) * sizeof(uint16_t);
  return launch_fast(stream_build_kernel, DLS_KERNEL_STREAM_BUILD, dim3(b.B, 3), kBuildThreads, smem, ctx, b, w, kinds);
}

template <int KMAX>
static cudaError_t launch_stream_sched_k(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                                         LaunchCtx& ctx) {
  const dim3 grid((b.B + kTPB - 1) / kTPB);
  int slot = 0;
  SlotKinds sk = {{0, 0, 0, 0}, {-1, -1, -1, -1}};
  cudaError_t e = cudaSuccess;
  const bool twins = (algo_mask & 3) == 3;  // DFS and Greedy both asked for: slots 0 and 1
  if (algo_mask & (1 << DLS_ALGO_DFS)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_DFS, KMAX>, DLS_KERNEL_STREAM_DFS, grid, kTPB, smem_simple<KMAX>(),
                    ctx, b, r, w, slot, twins ? 1 : 0, slot + 1);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kStreamDG;
  }
  if (algo_mask & (1 << DLS_ALGO_GREEDY)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_GREEDY, KMAX>, DLS_KERNEL_STREAM_GREEDY, grid, kTPB,
                    smem_simple<KMAX>(), ctx, b, r, w, slot, twins ? 2 : 0, slot - 1);
    if (e != cudaSuccess) return e;
    if (twins) sk.follower[slot - 1] = slot;
    sk.kind[slot++] = kStreamDG;
  }
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_CRITICAL, KMAX>, DLS_KERNEL_STREAM_CRITICAL, grid, kTPB,
                    smem_simple<KMAX>(), ctx, b, r, w, slot, 0, 0);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kStreamCP;
  }
  if (algo_mask & (1 << DLS_ALGO_MRU)) {
    e = launch_fast(sched_stream_mru_kernel<KMAX>, DLS_KERNEL_STREAM_MRU, dim3((b.B + kTPBM - 1) / kTPBM), kTPBM,
                    smem_mru<KMAX>(), ctx, b, r, w, slot);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kStreamMRU;
  }
  // verdicts of all slots back to task ids in one launch
  const int n_al = (b.n_max + 3) & ~3;
  const size_t usm = (size_t)4 * 3 * n_al;
  const int staged = usm <= 160 * 1024;
  return launch_fast(unpermute_kernel, DLS_KERNEL_UNPERMUTE, dim3((b.B + 3) / 4, slot), 128, staged ? usm : 0, ctx, b, r, w, sk,
                     n_al, staged);
}

cudaError_t launch_stream_sched(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                                LaunchCtx& ctx) {
  if (b.K_max <= 2) return launch_stream_sched_k<2>(b, r, w, algo_mask, ctx);
  if (b.K_max <= 4) return launch_stream_sched_k<4>(b, r, w, algo_mask, ctx);
  return launch_stream_sched_k<8>(b, r, w, algo_mask, ctx);
}

}  // namespace dls
