// dls_stream.cu — the throughput path: one THREAD per DAG instance over order-permuted task streams.
//
// Why: the list-scheduling loop is a serial chain of n dependent steps per DAG, but a sweep has ~10^5..10^6
// independent DAGs. With one warp per DAG every instruction of the chain serves one DAG (≈340 warp-instructions per
// task-step, profiles/r1a_*); with one thread per DAG the same instruction serves 32. To make that work the
// per-thread memory accesses must not scatter: stream_build_kernel writes each DAG, once per processing order, as
// contiguous 160-byte blocks of eight tasks (dls_common.cuh: 8 x {memory, duration}, then 8 meta words), so every
// thread reads fully used 32-byte sectors and nothing else, and per-thread mutable state lives in shared memory laid
// out [word][lane] (each lane owns a bank: no conflicts whatever the index), including one mark bit per task.
//
// A task's stream bits do not depend on the order they are placed in (a new round is the depth parity flipping;
// dependencies are not in the stream: a task that ends without a node marks its dependents through the reverse CSR,
// see BlockedSet / MarkPipe): the build kernel makes them once, in task order with coalesced reads, and then only
// permutes. (Round 1 translated dependencies to order positions per stream: an inverted order per stream, three
// look-ups per record and a 4.3 ms kernel. Round 2a tried no streams at all — one record array read through the
// order arrays — and found every 32-byte read costing a 128-byte DRAM line: the scheduling kernels ran 2.4x slower.
// Rounds 2b-2e: 32-byte records with the first three dependencies inline, build kernel 3.26 ms against 2.18.)
//
// Verdicts are written in order-position space (sequential bytes per thread, stored a word at a time);
// unpermute_kernel maps them back to task ids and derives seq_pos by a prefix count. Instances the fast kernels
// cannot take (more than 8 nodes, more than 32 shared params, MRU with any shared param, a task with 256+
// private params, bad topology) are flagged in `route` and picked up by the warp-per-DAG kernels of dls_sched.cu, which
// remain the general implementation.
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
#ifndef DLS_S_MINB
#define DLS_S_MINB 8
#endif
#ifdef DLS_S_MAXREG
#define DLS_S_BOUNDS(tpb) __maxnreg__(DLS_S_MAXREG)
#else
#define DLS_S_BOUNDS(tpb) __launch_bounds__(tpb, DLS_S_MINB)
#endif
#ifndef DLS_BUILD_U
#define DLS_BUILD_U 4  // positions per thread whose order loads are in flight together (2: 2.17 ms, 4: 2.12 ms)
#endif
constexpr int kBuildThreads = 256;   // measured r1c: 128 -> 7.6 ms, 256 -> 5.7 ms per 131k instances
constexpr int kBitmapWordsMax = 64;  // mark bits in shared memory up to 2048 tasks per DAG; beyond, mark bytes in HBM

// -------------------------------------------------------------------------------------------------------------------
// stream_build_kernel: grid B. One CTA per instance makes the instance's records and writes them in every requested
// processing order.
// -------------------------------------------------------------------------------------------------------------------
struct RecRegs {
  uint4 lo, hi;
};

// First block of an instance's stream: blocks hold eight tasks, instances have any number of them, so an instance
// starts at (its first task >> 3) + (its index) — never inside the previous instance's last block.
__device__ __forceinline__ size_t stream_block(int inst, const Inst& v) { return (size_t)(v.t_inst >> 3) + (size_t)inst; }

struct TaskBits {
  uint4 hot;      // m, c
  uint32_t meta;  // id | npriv << 16 | parity | has-shared
};

// What the stream holds of task i, from the instance's arrays. Sets `unstreamable` when a count does not fit its field.
__device__ __forceinline__ TaskBits make_task(const Inst& v, int i, const uint16_t* depth, int nsh_topo, const uint16_t* sid,
                                              int& unstreamable) {
  const double m = __ldg(v.mem + i), c = __ldg(v.dur + i);
  const uint32_t qb = __ldg(v.par_ptr + i), qe = __ldg(v.par_ptr + i + 1);
  const uint32_t d = __ldg(depth + i);
  uint32_t nsh = 0, npriv = qe - qb;
  if (nsh_topo > 0) {  // (no shared param in the DAG: nothing to look up, par_sid not even written)
    npriv = 0;
    for (uint32_t e = qb; e < qe; ++e) {
      if (sid[__ldg(v.par_idx + e)] == 0xFFFF) ++npriv;
      else ++nsh;
    }
  }
  if (npriv > 255u) { unstreamable = 1; npriv = 255u; }
  TaskBits tb;
  tb.hot.x = (uint32_t)__double2loint(m); tb.hot.y = (uint32_t)__double2hiint(m);
  tb.hot.z = (uint32_t)__double2loint(c); tb.hot.w = (uint32_t)__double2hiint(c);
  tb.meta = (uint32_t)i | (npriv << 16) | ((d & 1u) ? kMetaParity : 0u) | (nsh ? kMetaShared : 0u);
  return tb;
}

__device__ __forceinline__ void store_task(unsigned char* stream, size_t blk0, int t, const uint4& hot, uint32_t meta) {
  unsigned char* blk = stream + (blk0 + (size_t)(t >> 3)) * kBlkBytes;
  reinterpret_cast<uint4*>(blk)[t & 7] = hot;
  reinterpret_cast<uint32_t*>(blk + kBlkTasks * 16)[t & 7] = meta;
}

// `staged`: the tasks' stream bits are made in task order (coalesced reads) into shared memory (20 bytes per task) and
// each order then gathers them. Without (DAGs beyond ~10k tasks): every thread makes the bits of the task at its
// position, once per order.
__global__ void __launch_bounds__(kBuildThreads) stream_build_kernel(dls_batch_t b, Workspace w, int kind_mask, int staged) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int inst = blockIdx.x;
  Inst v;
  bind_instance(b, inst, v);
  if (w.topo_status[v.topo] != DLS_ST_OK) return;
  const int n = v.n;
  if (n == 0) return;
  const int nsh_topo = w.nshared[v.topo];
  const uint16_t* depth = w.depth + v.t_topo;
  const uint16_t* sid = w.par_sid + b.pid_off[v.topo];
  const uint16_t* g_order[3] = {w.order_dg + v.t_topo, w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo),
                                w.order_mru + v.t_topo};
  static_assert(kOrderDG == 0 && kOrderCP == 1 && kOrderMRU == 2, "order table follows the order kinds");
  bool use[3];
#pragma unroll
  for (int kind = 0; kind < 3; ++kind)  // DAGs the thread-per-DAG kernels leave to the general ones get no stream
    use[kind] = ((kind_mask >> kind) & 1) && (kind == kOrderMRU ? nsh_topo == 0 : nsh_topo <= kSharedCap);
  if (!use[0] && !use[1] && !use[2]) return;
  const size_t blk0 = stream_block(inst, v);
  int unstreamable = 0;
  if (staged) {
    uint4* s_hot = reinterpret_cast<uint4*>(smem_raw);            // [n]
    uint32_t* s_meta = reinterpret_cast<uint32_t*>(s_hot + n);    // [n]
    strided_for<4>(n, [&](int i) {
      const TaskBits tb = make_task(v, i, depth, nsh_topo, sid, unstreamable);
      s_hot[i] = tb.hot;
      s_meta[i] = tb.meta;
    });
    __syncthreads();
    // position t of all the orders in one iteration: three independent gathers per thread
    strided_for<DLS_BUILD_U>(n, [&](int t) {
#pragma unroll
      for (int kind = 0; kind < 3; ++kind) {
        if (!use[kind]) continue;
        const int i = __ldg(g_order[kind] + t);
        store_task(w.stream[kind], blk0, t, s_hot[i], s_meta[i]);
      }
    });
  } else {
    for (int kind = 0; kind < 3; ++kind) {
      if (!use[kind]) continue;
      strided_for<2>(n, [&](int t) {
        const TaskBits tb = make_task(v, __ldg(g_order[kind] + t), depth, nsh_topo, sid, unstreamable);
        store_task(w.stream[kind], blk0, t, tb.hot, tb.meta);
      });
    }
  }
  // a count no field can hold: the general kernels take this DAG (nshared is what routes it there)
  if (unstreamable) w.nshared[v.topo] = kUnstreamable;
}

// -------------------------------------------------------------------------------------------------------------------
// Shared pieces of the thread-per-DAG kernels
// -------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ RecRegs load32(const void* p) {
  RecRegs r;
  // One 256-bit load: two hot pairs or eight meta words. L2::128B: a miss in L2 brings in the whole 128-byte line, i.e.
  // most of this DAG's block. L1::no_allocate: the stream is read once; kept out of L1 the MRU kernel ran 2.19 instead
  // of 2.87 ms (round 2b).
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w)
               : "l"(p));
  return r;
}
#ifdef DLS_NO_PREFETCH
__device__ __forceinline__ void prefetch_l2(const void*) {}
#else
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#endif

// A thread's walk over its stream (blocks of eight tasks: 8 x 16 hot bytes, then 8 meta words). Sixteen registers
// hold the current pair of tasks and the current block's meta words; each is reloaded the moment its last value has
// been handed out, one step before the next one is needed (the L2 prefetch keeps that a hit). (With the next pair and
// the next block's words in registers of their own the kernels needed ~30 registers more and lost two CTAs per SM:
// DFS 1.38 -> 2.31 ms, round 2f. Issuing each load TWO steps ahead, with the odd task of the pair and the block's last
// meta word parked in five more registers, was slower too: DFS 1.64 -> 1.91 ms, MRU 2.29 -> 2.53 ms.)
struct StreamReader {
  const unsigned char* base;  // first block of the instance
  RecRegs hot;                // the current pair of tasks (lo: even position, hi: odd)
  RecRegs meta;               // the current block's eight meta words
  int n;
  __device__ __forceinline__ void init(const unsigned char* stream, size_t blk0, int n_) {
    base = stream + blk0 * kBlkBytes;
    n = n_;
    hot.lo = hot.hi = make_uint4(0, 0, 0, 0);
    meta = hot;
    if (n > 0) { hot = load32(base); meta = load32(base + kBlkTasks * 16); }
    if (n > 8) {
      prefetch_l2(base + kBlkBytes);
      if (((uintptr_t)(base + kBlkBytes) & 127u) == 96u) prefetch_l2(base + kBlkBytes + 32);
    }
  }
  // The task at position t; to be called for t = 0, 1, 2, ... in order (t is the same for all lanes of a warp).
  __device__ __forceinline__ void take(int t, double& m, double& c, uint32_t& mw) {
    const uint4 h = (t & 1) ? hot.hi : hot.lo;
    m = __hiloint2double((int)h.y, (int)h.x);
    c = __hiloint2double((int)h.w, (int)h.z);
    const uint4 q = (t & 4) ? meta.hi : meta.lo;
    const uint32_t qa = (t & 2) ? q.z : q.x, qb = (t & 2) ? q.w : q.y;
    mw = (t & 1) ? qb : qa;
    const int tn = t + 1;
    if ((t & 1) && tn < n) {
      const unsigned char* nb = base + (size_t)(tn >> 3) * kBlkBytes;
      hot = load32(nb + (tn & 7) * 16);
      if ((tn & 7) == 0) {
        meta = load32(nb + kBlkTasks * 16);
        if (tn + 8 < n) {  // the block after this one on its way into L2: blocks are 160 bytes, so the lines holding
          prefetch_l2(nb + kBlkBytes);  // block starts are four out of five; the fifth follows a start at byte 96
          if (((uintptr_t)(nb + kBlkBytes) & 127u) == 96u) prefetch_l2(nb + kBlkBytes + 32);
        }
      }
    }
  }
};

// A thread's verdict bytes (node or DLS_TASK_*) by order position. They leave in aligned words: 32 lanes storing one
// byte each to 32 different lines cost the L1-to-L2 request path — what bounds these kernels — as much as a record
// load.
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
};

// Per-warp shared-memory slab: word `wd` of the calling lane.
struct Slab {
  uint32_t* u;  // [words][32]
  double* d;    // [dwords][32]
  int lane;
  __device__ __forceinline__ uint32_t& w(int wd) const { return u[wd * 32 + lane]; }
  __device__ __forceinline__ double& f(int k) const { return d[k * 32 + lane]; }
};

// is_task_ready (ref schedulers.py:50-53) without looking at dependencies: a task is NOT ready exactly when one of its
// dependencies failed or was left pending, and that is known when the dependency is visited (it lies in an earlier
// round). So a task that ends without a node marks its dependents (reverse CSR of kernel P), and every task tests
// its own mark. While nothing has failed nothing is marked or tested. One bit per task in the thread's column of the
// slab; for DAGs too large for that (bm_words == 0) one byte per task in the node_of output, which the launcher
// cleared and unpermute_kernel rewrites afterwards.
struct BlockedSet {
  Slab sl;
  int bm;            // first mark word in the slab
  bool in_smem;
  int8_t* bytes;     // task-space mark bytes (node_of of this slot and instance)
  __device__ __forceinline__ bool test(int i) const {
    if (in_smem) return (sl.w(bm + (i >> 5)) >> (i & 31)) & 1u;
    return __ldcg(bytes + i) != 0;
  }
  __device__ __forceinline__ void mark(int s) const {
    if (in_smem) sl.w(bm + (s >> 5)) |= 1u << (s & 31);
    else bytes[s] = 1;
  }
  // (the reverse CSR's addresses are worked out here, on the rare path, instead of living in registers)
  __device__ __noinline__ void mark_dependents(const dls_batch_t& b, const Workspace& w, const Inst& v, int i) const {
    const uint32_t* succ_ptr = w.succ_ptr + v.t_topo + v.topo;
    const uint16_t* succ_idx = w.succ_idx + b.pred_off[v.topo];
    const uint32_t sb = succ_ptr[i], se = succ_ptr[i + 1];
    for (uint32_t e = sb; e < se; ++e) mark(succ_idx[e]);
  }
};

// mark_dependents spread over three steps of the walk, for the kernels whose tasks fail often (in the 1k-task sweep a
// tenth of the tasks of DFS / Greedy / Critical end without a node): the reverse CSR is per instance, so its two
// dependent loads (range, then successors) are two trips to DRAM, and done on the spot they stall the warp for the one
// lane that failed. Here the step that fails a task issues the range loads, the next step issues the loads of the first
// four successors, the one after sets the marks; each load has a step's worth of the other warps' work to land in.
// A dependent lies in a later round than the task (greater depth), so the marks are in time if everything in flight
// is drained when a round starts.
#ifndef DLS_PIPE_FAN
#define DLS_PIPE_FAN 4
#endif
constexpr int kPipeFan = DLS_PIPE_FAN;  // successors fetched ahead per failed task; the rest on the spot
struct MarkPipe {
  const uint32_t* succ_ptr;
  const uint16_t* succ_idx;
  uint32_t sb, se;      // stage 1: successor range of the task that just ended without a node (loads in flight)
  uint32_t rb, re;      // stage 2: range whose first successors are in q (loads in flight)
  uint32_t q[kPipeFan];
  __device__ __forceinline__ void init(const dls_batch_t& b, const Workspace& w, const Inst& v, bool have) {
    succ_ptr = w.succ_ptr + v.t_topo + v.topo;
    succ_idx = w.succ_idx + (have ? b.pred_off[v.topo] : 0);
    sb = se = rb = re = 0;
  }
  __device__ __forceinline__ void issue(int i) { sb = succ_ptr[i]; se = succ_ptr[i + 1]; }
  __device__ __forceinline__ void land(const BlockedSet& bs) {  // stage 2 -> marks
    if (re > rb) {
#pragma unroll
      for (int j = 0; j < kPipeFan; ++j)
        if (rb + j < re) bs.mark((int)q[j]);
      for (uint32_t e = rb + kPipeFan; e < re; ++e) bs.mark(succ_idx[e]);
      rb = re = 0;
    }
  }
  __device__ __forceinline__ void fetch() {  // stage 1 -> stage 2
    if (se > sb) {
#pragma unroll
      for (int j = 0; j < kPipeFan; ++j)
        if (sb + j < se) q[j] = succ_idx[sb + j];
      rb = sb; re = se;
      sb = se = 0;
    }
  }
  __device__ __forceinline__ void advance(const BlockedSet& bs) { land(bs); fetch(); }
  __device__ __forceinline__ void drain(const BlockedSet& bs) { land(bs); fetch(); land(bs); }
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

// amin_base >= 0: the slab holds, per node, the smallest available memory seen right after a task's own memory was
// taken (between ref schedulers.py:95 and :126). Rounding is monotone, so the largest fl(cap - avail) over those
// moments — what the warp-per-DAG kernels and the oracle accumulate — is fl(cap - smallest).
template <int KMAX>
__device__ __forceinline__ void finish_thread(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int slot,
                                              int inst, const Inst& v, const double (&avail)[KMAX], const Slab& sl,
                                              int fin_base, int cnt_base, int amin_base, const ThreadOut& to, int route_or = 0) {
  const int64_t ko = (int64_t)slot * b.nk_inst + v.k_out;
  double mk = 0.0, pk = 0.0;
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
      if (amin_base >= 0) {
        const double p = fmax(0.0, __dsub_rn(v.cap[k], sl.f(amin_base + k)));
        r.peak[ko + k] = p;
        pk = fmax(pk, p);
      }
    }
  }
  dls_summary_t s;
  s.makespan = any ? mk : 0.0;
  s.peak = pk;
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
template <int ALGO, int KMAX, bool PEAK>
__global__ void DLS_S_BOUNDS(kTPB) sched_stream_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot,
                                                       int twin_mode, int twin_slot, int bm_words) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  // slab layout per warp: doubles fin[KMAX], spd[KMAX] (+ amin[KMAX]) (+ avail[KMAX]); words smask[kSharedCap/4], cnt[KMAX],
  // mark-bits[bm_words]
  constexpr bool kAvSlab = ALGO == DLS_ALGO_CRITICAL;  // available memory in the slab (see `avail` below)
  constexpr int kD = ((PEAK ? 3 : 2) + (kAvSlab ? 1 : 0)) * KMAX;
  const int kW = kSharedCap / 4 + KMAX + bm_words;
  const size_t warp_bytes = (size_t)kD * 32 * 8 + (size_t)kW * 32 * 4;
  Slab sl;
  sl.d = reinterpret_cast<double*>(smem_raw + wib * warp_bytes);
  sl.u = reinterpret_cast<uint32_t*>(smem_raw + wib * warp_bytes + (size_t)kD * 32 * 8);
  sl.lane = lane;
  constexpr int FIN = 0, SPD = KMAX, AMIN = 2 * KMAX, SMASK = 0, CNT = kSharedCap / 4, BM = kSharedCap / 4 + KMAX;
  constexpr int AV = (PEAK ? 3 : 2) * KMAX;  // available memory per node

  const int inst = blockIdx.x * kTPB + threadIdx.x;
  const bool have = inst < b.B;
  Inst v;
  v.n = 0; v.K = 0; v.topo = 0; v.t_inst = 0; v.t_topo = 0;
  // DFS and Greedy walk the same order, and on a DAG without shared params they are the same function: every param
  // must be loaded whichever node is taken, so Greedy's (fewest loads, then most memory) is DFS's (most memory), see
  // the node choice below. The DFS launch (twin_mode 1) then stores its result in Greedy's slot as well, and the Greedy
  // launch (twin_mode 2) leaves those instances alone. The one case where the two differ even so (no node has more
  // than -1 available: DFS starts from max_memory = -1, Greedy from the first fitting node) withdraws the offer.
  bool elig = false, twin_ok = false;
  int nsh_topo = 0;
  if (have) {
    bind_instance(b, inst, v);
    bool taken = false;
    if (twin_mode == 2) {
      const int rt = w.route[(int64_t)slot * b.B + inst];
      taken = rt != kRouteOpen && (rt & kRouteTwin);
    }
    nsh_topo = w.nshared[v.topo];
    elig = !taken && w.topo_status[v.topo] == DLS_ST_OK && v.n > 0 && v.K <= KMAX && v.K >= 0 && nsh_topo <= kSharedCap;
    if (!elig && !taken) w.route[(int64_t)slot * b.B + inst] = kRouteGeneral;
    twin_ok = ALGO == DLS_ALGO_DFS && twin_mode == 1 && elig && nsh_topo == 0;
  }
  const int n = elig ? v.n : 0, K = v.K;
  const int n_warp = __reduce_max_sync(DLS_FULL_MASK, n);
  if (n_warp == 0) {
    if (have && twin_mode == 1) w.route[(int64_t)twin_slot * b.B + inst] = kRouteOpen;
    return;
  }
  const unsigned char* stream = w.stream[ALGO == DLS_ALGO_CRITICAL ? kOrderCP : kOrderDG];
  const uint16_t* sid = w.par_sid + (have ? b.pid_off[v.topo] : 0);
  VerdictOut vout;
  vout.init(w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst);
  BlockedSet blocked;
  blocked.sl = sl;
  blocked.bm = BM;
  blocked.in_smem = bm_words > 0;
  blocked.bytes = r.node_of + (int64_t)slot * b.nt_inst + v.t_inst;

  // Node state. avail[k] of a node that does not exist is -inf: it never fits and never wins a max. For Critical it
  // lives in the slab and is copied into registers for the duration of a step: the chosen node — found by its speed
  // rank, not by a scan over avail — is then ONE load and ONE store with a run-time index instead of two
  // compare-and-select chains over the nodes (1.57 -> 1.49 ms). DFS / Greedy find the node BY that scan and keep avail
  // in registers (in the slab: 1.64 -> 1.63 ms with peak tracking, 1.52 -> 1.63 ms without, where it costs a CTA).
  double avail[KMAX];
  // Critical: nodes ranked by (speed desc, index asc) once; key = rank*KMAX + k, nodes with speed <= 0 never chosen
  // ("node.compute_speed > best_speed" with best_speed = 0, ref schedulers.py:353-358).
  int ckey[ALGO == DLS_ALGO_CRITICAL ? KMAX : 1];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const bool on = elig && k < K;
    avail[k] = on ? v.cap[k] : -INFINITY;
    if constexpr (kAvSlab) sl.f(AV + k) = avail[k];
    const double sp = on ? v.spd[k] : 1.0;
    sl.f(FIN + k) = 0.0;
    sl.f(SPD + k) = sp;
    sl.w(CNT + k) = 0;
    if constexpr (PEAK) sl.f(AMIN + k) = INFINITY;
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
  for (int x = 0; x < bm_words; ++x) sl.w(BM + x) = 0;

  ThreadOut to;
  to.completed = to.failed = to.hits = to.misses = to.rounds = 0;
  to.status = DLS_ST_OK;
  to.fail_rest = false;
  bool running = elig, any_ready = false, any_assigned = false;
  bool clean = true;  // no task so far was failed or left pending: every dependency of the next task is done
  int prev_par = 0;
  StreamReader sr;
  sr.init(stream, stream_block(inst, v), n);
  MarkPipe mp;
  mp.init(b, w, v, have);
  for (int t = 0; t < n_warp; ++t) {
    if (t < n) {
      double m, c;
      uint32_t mw;
      sr.take(t, m, c, mw);
      const int i = (int)(mw & 0xffffu), npriv = (int)((mw >> 16) & 0xffu);
      const bool has_shared = (mw & kMetaShared) != 0;
      const int par = (mw & kMetaParity) ? 1 : 0;
      if (!clean) {
        if (par != prev_par) mp.drain(blocked);
        else mp.advance(blocked);
      }
      if (t == 0 || par != prev_par) {  // first task of a round
        if (t > 0 && running) {
          if (!any_ready) running = false;                       // ready list empty (ref schedulers.py:172-174)
          else if (!any_assigned) { running = false; to.fail_rest = true; }  // no progress (ref :202-206)
        }
        if (running) { ++to.rounds; any_ready = any_assigned = false; }
      }
      prev_par = par;
      int verdict = DLS_TASK_PENDING;
      int nsh = 0;  // shared params of the task (counted where they are walked)
      // the shared params of the task, through the input CSR and the param classes (LLM DAGs: a few tasks have any)
      auto for_each_shared = [&](auto f) {
        for (uint32_t e = v.par_ptr[i]; e < v.par_ptr[i + 1]; ++e) {
          const int sd = sid[v.par_idx[e]];
          if (sd != 0xFFFF) f(sd);
        }
      };
      if (running) {
        const bool ready = clean || !blocked.test(i);
        if (ready) {
          any_ready = true;
          int best = -1, lbest = npriv;
          double a = 0.0;  // available memory of the chosen node
          if constexpr (kAvSlab) {
#pragma unroll
            for (int k = 0; k < KMAX; ++k) avail[k] = sl.f(AV + k);
          }
          if (!has_shared) {
            // Every param of the task is private: params_to_load and hence the requirement are the same on every
            // node, so "fits" is avail >= need and the best fitting node is the first maximum, if that one fits.
            const double need = memory_requirement(m, npriv);
            if constexpr (ALGO == DLS_ALGO_CRITICAL) {
              int bk = 0x7fffffff;
#pragma unroll
              for (int k = 0; k < KMAX; ++k) bk = min(bk, need <= avail[k] ? ckey[k] : 0x7fffffff);
              if (bk != 0x7fffffff) { best = bk & (KMAX - 1); a = sl.f(AV + best); }
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
            for_each_shared([&](int sd) {
              ++nsh;
              const uint32_t mask = (sl.w(SMASK + (sd >> 2)) >> ((sd & 3) * 8)) & 0xffu;
              const unsigned long long spread = ((unsigned long long)(mask ^ 0xffu) * 0x0101010101010101ull) & 0x8040201008040201ull;
              miss += ((spread + 0x7f7f7f7f7f7f7f7full) >> 7) & 0x0101010101010101ull;
            });
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
            if constexpr (ALGO == DLS_ALGO_CRITICAL) { if (best >= 0) a = sl.f(AV + best); }
          }
          if (best >= 0) {
            if (lbest > 0) a = __dsub_rn(a, 0.5);  // node.available_memory -= 0.5 per loaded param (ref :89)
            if (lbest > 1) a = __dsub_rn(a, 0.5);
            for (int j = 2; j < lbest; ++j) a = __dsub_rn(a, 0.5);
            a = __dsub_rn(a, m);                   // ref :95
            if constexpr (PEAK) sl.f(AMIN + best) = fmin(sl.f(AMIN + best), a);
            a = __dadd_rn(a, m);                   // complete_task frees it again (ref :126)
            if constexpr (kAvSlab) {
              sl.f(AV + best) = a;
            } else {
#pragma unroll
              for (int k = 0; k < KMAX; ++k) avail[k] = (k == best) ? a : avail[k];
            }
            sl.f(FIN + best) = __dadd_rn(sl.f(FIN + best), __ddiv_rn(c, sl.f(SPD + best)));  // ref simulation.py:261-264
            sl.w(CNT + best) += 1;
            to.misses += lbest;
            to.hits += npriv + nsh - lbest;
            if (has_shared)
              for_each_shared([&](int sd) { sl.w(SMASK + (sd >> 2)) |= (1u << best) << ((sd & 3) * 8); });
            verdict = best;
            ++to.completed;
            any_assigned = true;
          } else {
            verdict = DLS_TASK_FAILED;  // fail_task (ref :128-131)
            ++to.failed;
          }
        }
        if (verdict < 0) {  // failed, or not ready: its dependents will not be ready either
          clean = false;
          mp.issue(i);
        }
      }
      vout.put(t, verdict, t == n - 1);
    }
  }
  if (elig) {
    if (running) {
      if (any_ready && !any_assigned) to.fail_rest = true;
      else if (any_ready && to.completed + to.failed < n) ++to.rounds;  // one more iteration finds nothing ready
    }
    if constexpr (kAvSlab) {
#pragma unroll
      for (int k = 0; k < KMAX; ++k) avail[k] = sl.f(AV + k);
    }
    finish_thread<KMAX>(b, r, w, slot, inst, v, avail, sl, FIN, CNT, PEAK ? AMIN : -1, to);
    if (twin_ok) finish_thread<KMAX>(b, r, w, twin_slot, inst, v, avail, sl, FIN, CNT, PEAK ? AMIN : -1, to, kRouteTwin);
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
// (Available memory in the slab, as in the kernel above, was measured here too: the slab grows to 16 KB per warp, six
// CTAs per SM instead of seven, and the eviction path writes every node back: 2.22 -> 2.78 ms.)
// -------------------------------------------------------------------------------------------------------------------

// evict_params_for_task (ref schedulers.py:404-442) on one node, as arithmetic. Each of the node's `ncached` params
// frees exactly 0.5 and the reference walks them until `freed >= shortage`: it needs j = ceil(shortage / 0.5) of
// them and succeeds iff j <= ncached. On success the node's available memory goes up by 0.5, j times (ref :429); on
// failure all ncached are evicted and rolled back, which leaves only the +0.5 / -0.5 rounding trace (ref :438-441).
//
// x + 0.5 + 0.5 + ... (j additions, one rounding each, as the reference does them) equals fl(x + 0.5 j) whenever that
// one sum is exact (TwoSum's error term is zero): every partial sum v = x + 0.5 i lies between x and the result s, is a
// multiple of ulp(x) counted from x and of ulp(s) counted back from s (0.5 is a multiple of both below 2^51), hence
// of ulp(v) <= max(ulp(x), ulp(s)): representable, so no addition of the chain rounds. Anything else (a sum that
// crosses into a binade where x's last bits do not fit, NaN, infinities) takes the chain itself. The chain was 10 % of
// this kernel's stall samples (ncu r2f): lanes evict different numbers of params and the warp waits for the longest.
__device__ __forceinline__ bool add_halves_exactly(double x, uint32_t j, double& s) {
  const double h = __dmul_rn(0.5, (double)j);  // exact
  s = __dadd_rn(x, h);
  const double bb = __dsub_rn(s, x);
  const double err = __dadd_rn(__dsub_rn(x, __dsub_rn(s, bb)), __dsub_rn(h, bb));
  return err == 0.0 && fabs(x) < 0x1p51 && fabs(s) < 0x1p51;
}

__device__ __forceinline__ bool plan_evict(double need, double& avail_k, uint32_t& ncached) {
  const double shortage = __dsub_rn(need, avail_k);
  if (shortage <= 0.0) return true;
  const double want = ceil(__dmul_rn(shortage, 2.0));  // exact: power-of-two scaling, then ceil
  double s;
  if (want <= (double)ncached) {
    const uint32_t j = (uint32_t)want;
    if (add_halves_exactly(avail_k, j, s)) avail_k = s;
    else
      for (uint32_t q = 0; q < j; ++q) avail_k = __dadd_rn(avail_k, 0.5);
    ncached -= j;
    return true;
  }
  if (!add_halves_exactly(avail_k, ncached, s)) {  // (exact on the way up: the way down retraces the same values)
    for (uint32_t q = 0; q < ncached; ++q) avail_k = __dadd_rn(avail_k, 0.5);
    for (uint32_t q = 0; q < ncached; ++q) avail_k = __dsub_rn(avail_k, 0.5);
  }
  return false;
}

template <int KMAX, bool PEAK>
__global__ void DLS_S_BOUNDS(kTPBM) sched_stream_mru_kernel(dls_batch_t b, dls_result_t r, Workspace w, int slot,
                                                                 int bm_words) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  // slab per warp: doubles fin[KMAX], spd[KMAX], hcnt[KMAX] (+ amin[KMAX]); words cnt[KMAX], ncached[KMAX],
  // mark-bits[bm_words]
  constexpr int kD = (PEAK ? 4 : 3) * KMAX;
  const int kW = 2 * KMAX + bm_words;
  const size_t warp_bytes = (size_t)kD * 32 * 8 + (size_t)kW * 32 * 4;
  Slab sl;
  sl.d = reinterpret_cast<double*>(smem_raw + wib * warp_bytes);
  sl.u = reinterpret_cast<uint32_t*>(smem_raw + wib * warp_bytes + (size_t)kD * 32 * 8);
  sl.lane = lane;
  // HCNT[k] = len(node.completed_tasks) * 0.5 kept as a double: every += 0.5 is exact (ref schedulers.py:498)
  constexpr int FIN = 0, SPD = KMAX, HCNT = 2 * KMAX, AMIN = 3 * KMAX, CNT = 0, NC = KMAX, BM = 2 * KMAX;

  const int inst = blockIdx.x * kTPBM + threadIdx.x;
  const bool have = inst < b.B;
  Inst v;
  v.n = 0; v.K = 0; v.topo = 0; v.P = 0; v.t_inst = 0; v.t_topo = 0;
  bool elig = false;
  if (have) {
    bind_instance(b, inst, v);
    elig = w.topo_status[v.topo] == DLS_ST_OK && v.n > 0 && v.K <= KMAX && v.K >= 0 && w.nshared[v.topo] == 0;
    if (!elig) w.route[(int64_t)slot * b.B + inst] = kRouteGeneral;
  }
  const int n = elig ? v.n : 0, K = v.K;
  const int n_warp = __reduce_max_sync(DLS_FULL_MASK, n);
  if (n_warp == 0) return;
  const unsigned char* stream = w.stream[kOrderMRU];
  VerdictOut vout;
  vout.init(w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst);
  BlockedSet blocked;
  blocked.sl = sl;
  blocked.bm = BM;
  blocked.in_smem = bm_words > 0;
  blocked.bytes = r.node_of + (int64_t)slot * b.nt_inst + v.t_inst;

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
    if constexpr (PEAK) sl.f(AMIN + k) = INFINITY;
  }
  for (int x = 0; x < bm_words; ++x) sl.w(BM + x) = 0;

  ThreadOut to;
  to.completed = to.failed = to.hits = to.misses = to.rounds = 0;
  to.status = DLS_ST_OK;
  to.fail_rest = false;
  int carried = 0;
  bool running = elig, any_ready = false, any_assigned = false, clean = true;
  int prev_par = 0;
  StreamReader sr;
  sr.init(stream, stream_block(inst, v), n);
  for (int t = 0; t < n_warp; ++t) {
    if (t < n) {
      double m, c;
      uint32_t mw;
      sr.take(t, m, c, mw);
      const int i = (int)(mw & 0xffffu), npar = (int)((mw >> 16) & 0xffu);
      const int par = (mw & kMetaParity) ? 1 : 0;
      if (t == 0 || par != prev_par) {
        if (t > 0 && running) {
          if (!any_ready) running = false;
          else if (!any_assigned) { running = false; to.fail_rest = true; }
        }
        if (running) { ++to.rounds; any_ready = any_assigned = false; }  // self.time_step += 1 (ref :453)
      }
      prev_par = par;
      int verdict = DLS_TASK_PENDING;
      if (running) {
        const bool ready = clean || !blocked.test(i);  // is_task_ready (ref :50-53), as in sched_stream_kernel
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
              if constexpr (PEAK) sl.f(AMIN + best) = fmin(sl.f(AMIN + best), a);
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
        if (verdict < 0) {  // failed, or not ready: its dependents will not be ready either
          clean = false;
          blocked.mark_dependents(b, w, v, i);
        }
      }
      vout.put(t, verdict, t == n - 1);
    }
  }
  if (elig) {
    if (running) {
      if (any_ready && !any_assigned) to.fail_rest = true;
      else if (any_ready && to.completed + to.failed + carried < n) ++to.rounds;
    }
    finish_thread<KMAX>(b, r, w, slot, inst, v, avail, sl, FIN, CNT, PEAK ? AMIN : -1, to);
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
  const uint16_t* order = kind == kOrderDG ? w.order_dg + v.t_topo
                        : kind == kOrderMRU ? w.order_mru + v.t_topo
                                            : w.order_cp + (w.cp_per_instance ? v.t_inst : v.t_topo);
  const int8_t* npos = w.node_pos + (int64_t)slot * b.nt_inst + v.t_inst;
  int8_t* node_of = r.node_of + (int64_t)slot * b.nt_inst + v.t_inst;
  uint16_t* seq_pos = r.seq_pos + (int64_t)slot * b.nt_inst + v.t_inst;
  int8_t* node_of2 = r.node_of + (int64_t)(dup ? fol : slot) * b.nt_inst + v.t_inst;
  uint16_t* seq_pos2 = r.seq_pos + (int64_t)(dup ? fol : slot) * b.nt_inst + v.t_inst;
  int8_t* s_node = reinterpret_cast<int8_t*>(smem_raw) + (size_t)(threadIdx.x >> 5) * 3 * n_al;
  uint16_t* s_seq = reinterpret_cast<uint16_t*>(s_node + n_al);
  int base = 0;
  // Fast path (everything 4-aligned: any batch of DAGs whose sizes are multiples of four): a lane takes four
  // consecutive positions — one 32-bit load of verdicts, one 64-bit load of task ids — and the prefix count of
  // assignments over the warp comes from three ballots (a lane's count is 0..4: three bits). A third of the
  // instructions per position of the one-position-per-lane loop below (0.85 -> see DESIGN.md).
  const bool quad = staged && (v.n & 3) == 0 && ((reinterpret_cast<uintptr_t>(npos) & 3) == 0) &&
                    ((reinterpret_cast<uintptr_t>(order) & 7) == 0) && ((reinterpret_cast<uintptr_t>(node_of) & 3) == 0) &&
                    ((reinterpret_cast<uintptr_t>(seq_pos) & 7) == 0) && ((reinterpret_cast<uintptr_t>(node_of2) & 3) == 0) &&
                    ((reinterpret_cast<uintptr_t>(seq_pos2) & 7) == 0);
  if (quad) {
    const unsigned lt = (1u << lane) - 1u;
    constexpr int kQ = 4;  // chunks of 128 positions whose (independent) loads are issued together: with one chunk per
                           // trip the wait for its two loads was 42 % of this kernel's stall samples (ncu r2f)
    for (int t0 = 0; t0 < v.n; t0 += 128 * kQ) {
      uint32_t vws[kQ];
      uint2 ows[kQ];
#pragma unroll
      for (int u = 0; u < kQ; ++u) {
        const int t = t0 + 128 * u + 4 * lane;
        vws[u] = 0xfefefefeu;  // four DLS_TASK_PENDING
        ows[u] = make_uint2(0, 0);
        if (t < v.n) {
          vws[u] = *reinterpret_cast<const uint32_t*>(npos + t);
          ows[u] = *reinterpret_cast<const uint2*>(order + t);
        }
      }
#pragma unroll
      for (int u = 0; u < kQ; ++u) {
        if (t0 + 128 * u >= v.n) break;  // (same for the whole warp)
        const int t = t0 + 128 * u + 4 * lane;
        const uint32_t vw = vws[u];
        const uint2 ow = ows[u];
        int vd[4], tk[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) vd[j] = (int)(int8_t)(vw >> (8 * j));
        tk[0] = ow.x & 0xffff; tk[1] = ow.x >> 16; tk[2] = ow.y & 0xffff; tk[3] = ow.y >> 16;
        const int c = (vd[0] >= 0) + (vd[1] >= 0) + (vd[2] >= 0) + (vd[3] >= 0);
        const unsigned b0 = __ballot_sync(DLS_FULL_MASK, c & 1), b1 = __ballot_sync(DLS_FULL_MASK, c & 2),
                       b2 = __ballot_sync(DLS_FULL_MASK, c & 4);
        int sq = base + __popc(b0 & lt) + 2 * __popc(b1 & lt) + 4 * __popc(b2 & lt);
        base += __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
        if (t < v.n) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            int verdict = vd[j];
            const bool asg = verdict >= 0;
            if ((route & kRouteFailRest) && verdict == DLS_TASK_PENDING) verdict = DLS_TASK_FAILED;  // "fail all pending" (ref :202-206)
            s_node[tk[j]] = (int8_t)verdict;
            s_seq[tk[j]] = asg ? (uint16_t)sq : (uint16_t)DLS_SEQ_NONE;
            sq += asg;
          }
        }
      }
    }
    __syncwarp();
    for (int i = 4 * lane; i < v.n; i += 128) {
      const uint32_t nd = *reinterpret_cast<const uint32_t*>(s_node + i);
      const uint2 sq = *reinterpret_cast<const uint2*>(s_seq + i);
      *reinterpret_cast<uint32_t*>(node_of + i) = nd;
      *reinterpret_cast<uint2*>(seq_pos + i) = sq;
      if (dup) {
        *reinterpret_cast<uint32_t*>(node_of2 + i) = nd;
        *reinterpret_cast<uint2*>(seq_pos2 + i) = sq;
      }
    }
    return;
  }
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

int bitmap_words(int n_max) {
  const int wds = (n_max + 31) >> 5;
  return wds <= kBitmapWordsMax ? wds : 0;
}
template <int KMAX>
size_t smem_simple(bool peak, int bm_words, bool critical = false) {
  return (size_t)(kTPB / 32) * ((size_t)((peak ? 3 : 2) + (critical ? 1 : 0)) * KMAX * 32 * 8 + (size_t)(kSharedCap / 4 + KMAX + bm_words) * 32 * 4);
}
template <int KMAX>
size_t smem_mru(bool peak, int bm_words) {
  return (size_t)(kTPBM / 32) * ((size_t)(peak ? 4 : 3) * KMAX * 32 * 8 + (size_t)(2 * KMAX + bm_words) * 32 * 4);
}

}  // namespace

// The streamed path needs K <= 8 nodes and per-thread state small enough for a useful number of resident warps.
bool stream_path_possible(const dls_batch_t& b, const dls_result_t* r, int smem_limit) {
  if (b.K_max > 8) return false;
  if (r && (r->cached || r->tfin)) return false;  // these extension outputs are produced by the general kernels only
  return smem_mru<8>(true, kBitmapWordsMax) <= (size_t)smem_limit / 2 && b.n_max >= 1;
}

template <class Kern, class... Args>
static cudaError_t launch_fast(Kern kern, int kernel_id, dim3 grid, int threads, size_t smem, LaunchCtx& ctx, Args... args) {
  ctx.why = "cudaFuncSetAttribute(stream kernel)";
  cudaError_t e = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), (size_t)smem);
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
  if (algo_mask & 3) kinds |= 1 << kOrderDG;
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) kinds |= 1 << kOrderCP;
  if (algo_mask & (1 << DLS_ALGO_MRU)) kinds |= 1 << kOrderMRU;
  const size_t smem = (size_t)b.n_max * 20;
  const int staged = smem <= 200 * 1024;
  return launch_fast(stream_build_kernel, DLS_KERNEL_STREAM_BUILD, dim3(b.B), kBuildThreads, staged ? smem : 0, ctx, b, w, kinds,
                     staged);
}

template <int KMAX, bool PEAK>
static cudaError_t launch_stream_sched_k(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                                         LaunchCtx& ctx) {
  const dim3 grid((b.B + kTPB - 1) / kTPB);
  const int bmw = bitmap_words(b.n_max);
  if (bmw == 0) {  // DAGs too large for mark bits in shared memory keep mark bytes in the node_of output: start from zero
    int A = 0;
    for (int a = 0; a < DLS_NUM_ALGOS; ++a) A += (algo_mask >> a) & 1;
    ctx.why = "cudaMemsetAsync(node_of)";
    cudaError_t e0 = cudaMemsetAsync(r.node_of, 0, (size_t)A * (size_t)b.nt_inst, ctx.st);
    if (e0 != cudaSuccess) return e0;
  }
  int slot = 0;
  SlotKinds sk = {{0, 0, 0, 0}, {-1, -1, -1, -1}};
  cudaError_t e = cudaSuccess;
  const bool twins = (algo_mask & 3) == 3;  // DFS and Greedy both asked for: slots 0 and 1
  if (algo_mask & (1 << DLS_ALGO_DFS)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_DFS, KMAX, PEAK>, DLS_KERNEL_STREAM_DFS, grid, kTPB,
                    smem_simple<KMAX>(PEAK, bmw), ctx, b, r, w, slot, twins ? 1 : 0, slot + 1, bmw);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kOrderDG;
  }
  if (algo_mask & (1 << DLS_ALGO_GREEDY)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_GREEDY, KMAX, PEAK>, DLS_KERNEL_STREAM_GREEDY, grid, kTPB,
                    smem_simple<KMAX>(PEAK, bmw), ctx, b, r, w, slot, twins ? 2 : 0, slot - 1, bmw);
    if (e != cudaSuccess) return e;
    if (twins) sk.follower[slot - 1] = slot;
    sk.kind[slot++] = kOrderDG;
  }
  if (algo_mask & (1 << DLS_ALGO_CRITICAL)) {
    e = launch_fast(sched_stream_kernel<DLS_ALGO_CRITICAL, KMAX, PEAK>, DLS_KERNEL_STREAM_CRITICAL, grid, kTPB,
                    smem_simple<KMAX>(PEAK, bmw, true), ctx, b, r, w, slot, 0, 0, bmw);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kOrderCP;
  }
  if (algo_mask & (1 << DLS_ALGO_MRU)) {
    e = launch_fast(sched_stream_mru_kernel<KMAX, PEAK>, DLS_KERNEL_STREAM_MRU, dim3((b.B + kTPBM - 1) / kTPBM), kTPBM,
                    smem_mru<KMAX>(PEAK, bmw), ctx, b, r, w, slot, bmw);
    if (e != cudaSuccess) return e;
    sk.kind[slot++] = kOrderMRU;
  }
  // verdicts of all slots back to task ids in one launch
  const int n_al = (b.n_max + 7) & ~7;  // (8: the staged seq_pos rows are read back eight bytes at a time)
  const size_t usm = (size_t)4 * 3 * n_al;
  const int staged = usm <= 160 * 1024;
  return launch_fast(unpermute_kernel, DLS_KERNEL_UNPERMUTE, dim3((b.B + 3) / 4, slot), 128, staged ? usm : 0, ctx, b, r, w, sk,
                     n_al, staged);
}

cudaError_t launch_stream_sched(const dls_batch_t& b, const dls_result_t& r, const Workspace& w, int algo_mask,
                                LaunchCtx& ctx) {
  if (r.peak) {
    if (b.K_max <= 2) return launch_stream_sched_k<2, true>(b, r, w, algo_mask, ctx);
    if (b.K_max <= 4) return launch_stream_sched_k<4, true>(b, r, w, algo_mask, ctx);
    return launch_stream_sched_k<8, true>(b, r, w, algo_mask, ctx);
  }
  if (b.K_max <= 2) return launch_stream_sched_k<2, false>(b, r, w, algo_mask, ctx);
  if (b.K_max <= 4) return launch_stream_sched_k<4, false>(b, r, w, algo_mask, ctx);
  return launch_stream_sched_k<8, false>(b, r, w, algo_mask, ctx);
}

}  // namespace dls
