// dls_abi.cu — the extern "C" entry points declared in include/dls.h.
//
// The library never allocates device memory and never synchronises the stream: scratch comes from the
// caller (dls_workspace_bytes), results go to caller buffers, kernels are enqueued on the caller's stream (one part
// of the replay path on an auxiliary stream forked from it and joined back into it: see ForkJoin).
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>

#include "dls_common.cuh"

namespace {

thread_local std::string g_err;
std::atomic<int64_t> g_launches{0};
dls::Profiler g_prof;
const char* const kKernelNames[DLS_NUM_KERNELS] = {"priority_topology_kernel", "priority_critical_kernel",
                                                   "sched_simple_kernel<DFS>", "sched_simple_kernel<Greedy>",
                                                   "sched_simple_kernel<Critical>", "sched_mru_kernel",
                                                   "param_class_kernel", "stream_build_kernel",
                                                   "sched_stream_kernel<DFS>", "sched_stream_kernel<Greedy>",
                                                   "sched_stream_kernel<Critical>", "sched_stream_mru_kernel",
                                                   "unpermute_kernel", "replay_kernel"};

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

constexpr size_t kAlign = 256;
size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

struct Layout {
  size_t off_depth, off_outdeg, off_order_dg, off_order_mru, off_order_cp, off_lvl_end, off_up, off_succ_ptr, off_succ_idx,
      off_status, total;
  // streamed path (present when `stream` is set)
  size_t off_sid, off_nshared, off_stream[3], off_node_pos, off_route;
  bool stream;
  int cp_per_instance;
};

int count_bits(int m) { int c = 0; for (; m; m &= m - 1) ++c; return c; }

Layout layout_of(const dls_batch_t& b, int algo_mask = DLS_MASK_ALL, bool stream = false) {
  Layout L;
  L.stream = stream;
  L.cp_per_instance = !(b.T == 1 && !(b.flags & DLS_F_DUR_PER_INSTANCE));
  const size_t nt = (size_t)b.nt_topo, ncp = (size_t)(L.cp_per_instance ? b.nt_inst : b.nt_topo);
  size_t o = 0;
  L.off_depth = o;     o = align_up(o + nt * sizeof(uint16_t));
  L.off_outdeg = o;    o = align_up(o + nt * sizeof(uint16_t));
  L.off_order_dg = o;  o = align_up(o + nt * sizeof(uint16_t));
  L.off_order_mru = o; o = align_up(o + nt * sizeof(uint16_t));
  L.off_order_cp = o;  o = align_up(o + ncp * sizeof(uint16_t));
  L.off_lvl_end = o;   o = align_up(o + nt * sizeof(uint16_t));
  L.off_up = o;        o = align_up(o + ncp * sizeof(double));
  L.off_succ_ptr = o;  o = align_up(o + (nt + (size_t)b.T) * sizeof(uint32_t));
  L.off_succ_idx = o;  o = align_up(o + (size_t)b.e_total * sizeof(uint16_t));
  L.off_status = o;    o = align_up(o + (size_t)b.T * sizeof(int32_t));
  L.off_sid = L.off_nshared = L.off_node_pos = L.off_route = 0;
  for (int k = 0; k < 3; ++k) L.off_stream[k] = 0;
  if (stream) {
    const size_t np = (size_t)(b.T == 1 ? b.P_max : b.np_inst);
    const int A = count_bits(algo_mask & DLS_MASK_ALL);
    L.off_sid = o;     o = align_up(o + np * sizeof(uint16_t));
    L.off_nshared = o; o = align_up(o + (size_t)b.T * sizeof(int32_t));
    // per processing order in use, one stream of 160-byte blocks (eight tasks each)
    const bool need[3] = {(algo_mask & 3) != 0, (algo_mask & (1 << DLS_ALGO_CRITICAL)) != 0,
                          (algo_mask & (1 << DLS_ALGO_MRU)) != 0};
    for (int k = 0; k < 3; ++k) {
      if (!need[k]) continue;
      // blocks of eight tasks; an instance starts at block (its first task >> 3) + (its index): see stream_block()
      L.off_stream[k] = o; o = align_up(o + ((size_t)b.nt_inst / dls::kBlkTasks + (size_t)b.B + 4) * dls::kBlkBytes);
    }
    L.off_node_pos = o; o = align_up(o + (size_t)A * (size_t)b.nt_inst);
    L.off_route = o;    o = align_up(o + (size_t)A * (size_t)b.B);
  }
  L.total = o;
  return L;
}

dls::Workspace carve(void* base, const Layout& L) {
  char* p = static_cast<char*>(base);
  dls::Workspace w;
  w.depth = reinterpret_cast<uint16_t*>(p + L.off_depth);
  w.outdeg = reinterpret_cast<uint16_t*>(p + L.off_outdeg);
  w.order_dg = reinterpret_cast<uint16_t*>(p + L.off_order_dg);
  w.order_mru = reinterpret_cast<uint16_t*>(p + L.off_order_mru);
  w.order_cp = reinterpret_cast<uint16_t*>(p + L.off_order_cp);
  w.lvl_end = reinterpret_cast<uint16_t*>(p + L.off_lvl_end);
  w.up = reinterpret_cast<double*>(p + L.off_up);
  w.succ_ptr = reinterpret_cast<uint32_t*>(p + L.off_succ_ptr);
  w.succ_idx = reinterpret_cast<uint16_t*>(p + L.off_succ_idx);
  w.topo_status = reinterpret_cast<int32_t*>(p + L.off_status);
  w.cp_per_instance = L.cp_per_instance;
  w.keep_up = 0;
  w.par_sid = nullptr; w.nshared = nullptr; w.node_pos = nullptr; w.route = nullptr;
  for (int k = 0; k < 3; ++k) w.stream[k] = nullptr;
  if (L.stream) {
    w.par_sid = reinterpret_cast<uint16_t*>(p + L.off_sid);
    w.nshared = reinterpret_cast<int32_t*>(p + L.off_nshared);
    for (int k = 0; k < 3; ++k)
      if (L.off_stream[k]) w.stream[k] = reinterpret_cast<unsigned char*>(p + L.off_stream[k]);
    w.node_pos = reinterpret_cast<int8_t*>(p + L.off_node_pos);
    w.route = reinterpret_cast<uint8_t*>(p + L.off_route);
  }
  return w;
}

int smem_limit() {
  int dev = 0, lim = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&lim, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess) return -1;
  return lim;
}

int validate(const dls_batch_t* b, int algo_mask, bool need_device) {
  if (!b) return fail(DLS_ERR_BAD_ARG, "batch is NULL");
  if (b->B <= 0) return fail(DLS_ERR_BAD_ARG, "B must be positive");
  if (b->T != b->B && b->T != 1) return fail(DLS_ERR_BAD_ARG, "T must be B or 1");
  if ((algo_mask & ~DLS_MASK_ALL) || algo_mask == 0) return fail(DLS_ERR_BAD_ARG, "algo_mask must select 1..4 of the algorithms");
  if (b->n_max < 0 || b->K_max <= 0 || b->P_max < 0 || b->E_max < 0 || b->R_max < 0 || b->nt_topo < 0 || b->nt_inst < 0 ||
      b->e_total < 0 || b->r_total < 0 || b->nk_inst < 0 || b->np_inst < 0)
    return fail(DLS_ERR_BAD_ARG, "negative size in descriptor");
  if (b->n_max > DLS_MAX_TASKS) return fail(DLS_ERR_TOO_LARGE, "n_max > 65535 (task ids are u16)");
  if (b->P_max > DLS_MAX_PARAMS) return fail(DLS_ERR_TOO_LARGE, "P_max > 65535 (param ids are u16)");
  if (b->K_max > DLS_MAX_NODES) return fail(DLS_ERR_TOO_LARGE, "K_max > 32 (one warp lane per node)");
  if (!b->task_off || !b->pred_off || !b->par_off || !b->pid_off || !b->pred_ptr || !b->par_ptr || !b->rank ||
      !b->mem || !b->dur || !b->cap || !b->spd)
    return fail(DLS_ERR_BAD_ARG, "NULL array in batch descriptor");
  if ((b->e_total > 0 && !b->pred_idx) || (b->r_total > 0 && !b->par_idx))
    return fail(DLS_ERR_BAD_ARG, "NULL index array in batch descriptor");
  if (need_device) {
    const int lim = smem_limit();
    if (lim < 0) return fail(DLS_ERR_CUDA, std::string("no usable CUDA device: ") + cudaGetErrorString(cudaGetLastError()));
    size_t need = dls::priority_smem_topo(b->n_max, -1);
    if (algo_mask & (1 << DLS_ALGO_CRITICAL)) need = std::max(need, dls::priority_smem_inst(b->n_max, false));
    if (algo_mask & 7) need = std::max(need, dls::sched_smem_simple(b->n_max, b->P_max));
    if (algo_mask & (1 << DLS_ALGO_MRU)) need = std::max(need, dls::sched_smem_mru(b->n_max, b->P_max, b->K_max));
    if (need > (size_t)lim) {
      char buf[160];
      snprintf(buf, sizeof buf, "per-DAG state needs %zu B of shared memory, device offers %d B (n_max=%d P_max=%d)", need,
               lim, b->n_max, b->P_max);
      return fail(DLS_ERR_TOO_LARGE, buf);
    }
  }
  return DLS_OK;
}

std::atomic<bool> g_stream_enabled{true};
std::atomic<int> g_replay_mode{1};  // 0 off, 1 on (template run beside the Critical run), 2 on, one stream

int cuda_fail(cudaError_t e, const char* why) {
  return fail(DLS_ERR_CUDA, std::string(why ? why : "cuda") + ": " + cudaGetErrorString(e));
}

// ---- built-in scratch for dls_schedule_batch_static --------------------------------------------------------------
// The array is part of this library's device image: each GPU that loads the library carries its own copy from
// load until exit. In HOST code the name below is only a handle for the runtime, not a device address; the
// address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[DLS_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {};

// Device address of g_static_scratch on the current device (looked up on the first call per device).
int static_scratch_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return fail(DLS_ERR_BAD_ARG, "device ordinal beyond the built-in scratch table");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(built-in scratch)");
    if (!p) return fail(DLS_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the built-in scratch");
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return DLS_OK;
}

// Enqueue everything for one batch on `ctx`. `ws_bytes` is what the caller provided; the streamed (thread-per-DAG) path
// is used when the batch qualifies and its larger scratch fits, otherwise the warp-per-DAG kernels do all the work.
int enqueue_schedule(const dls_batch_t* batch, const dls_result_t* result, int algo_mask, void* workspace, size_t ws_bytes,
                     dls::LaunchCtx& ctx, bool clear_status) {
  const int lim = smem_limit();
  if (result->tfin) {  // the timeline extension keeps one finish time per task next to the rest of the per-DAG state
    size_t need = 0;
    if (algo_mask & 7) need = std::max(need, dls::sched_smem_simple(batch->n_max, batch->P_max));
    if (algo_mask & (1 << DLS_ALGO_MRU)) need = std::max(need, dls::sched_smem_mru(batch->n_max, batch->P_max, batch->K_max));
    need = ((need + 15) & ~(size_t)15) + (((size_t)batch->n_max * sizeof(double) + 15) & ~(size_t)15);
    if (need > (size_t)lim) return fail(DLS_ERR_TOO_LARGE, "per-DAG state plus the timeline's finish times exceed shared memory");
  }
  bool stream = g_stream_enabled.load() && dls::stream_path_possible(*batch, result, lim);
  Layout L = layout_of(*batch, algo_mask, stream);
  if (stream && ws_bytes < L.total) {
    stream = false;
    L = layout_of(*batch, algo_mask, false);
  }
  if (!workspace || ws_bytes < L.total) {
    char buf[200];
    snprintf(buf, sizeof buf, "scratch of %zu bytes is smaller than the %zu bytes this batch needs", ws_bytes, L.total);
    return fail(DLS_ERR_WORKSPACE, buf);
  }
  if (reinterpret_cast<uintptr_t>(workspace) % kAlign) return fail(DLS_ERR_WORKSPACE, "workspace must be 256-byte aligned");
  const dls::Workspace w = carve(workspace, L);
  cudaError_t e = cudaSuccess;
  if (clear_status) {
    e = cudaMemsetAsync(w.topo_status, 0, (size_t)batch->T * sizeof(int32_t), ctx.st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(built-in scratch)");
  }
  e = dls::launch_priority(*batch, w, algo_mask, ctx);
  if (e == cudaSuccess && stream) e = dls::launch_stream_build(*batch, w, algo_mask, ctx);
  if (e == cudaSuccess && stream) e = dls::launch_stream_sched(*batch, *result, w, algo_mask, ctx);
  if (e == cudaSuccess) e = dls::launch_schedule(*batch, *result, w, algo_mask, ctx);
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  return DLS_OK;
}

// ---- one schedule for all instances (see dls_replay.cu) ------------------------------------------------------------
constexpr int kDurationFree = (1 << DLS_ALGO_DFS) | (1 << DLS_ALGO_GREEDY) | (1 << DLS_ALGO_MRU);

bool replay_applies(const dls_batch_t& b, const dls_result_t* r, int algo_mask) {
  return g_replay_mode.load() != 0 && b.T == 1 && b.B > 1 && b.flags == DLS_F_DUR_PER_INSTANCE && !b.node_off &&
         (algo_mask & kDurationFree) && !(r && r->tfin);  // (the timeline extension is made of durations)
}

dls_batch_t template_batch(const dls_batch_t& b) {
  dls_batch_t t = b;
  t.B = 1;
  t.nt_inst = b.nt_topo;
  t.nk_inst = b.K_max;
  t.np_inst = b.np_inst / b.B;
  return t;
}

// Scratch of the template run behind the caller-visible layout: its own workspace, its result arrays, the sequence.
struct TemplateLayout {
  size_t off_ws, ws_bytes, off_node_of, off_seq_pos, off_summary, off_fin, off_cnt, off_avail, off_peak, off_cached,
      off_seq_task, off_seq_node, total;
};

TemplateLayout template_layout(const dls_batch_t& b, int tmask, size_t base) {
  const dls_batch_t t = template_batch(b);
  const size_t A = (size_t)count_bits(tmask), n = (size_t)t.nt_inst, K = (size_t)t.K_max, P = (size_t)t.np_inst;
  TemplateLayout T;
  size_t o = align_up(base);
  T.off_ws = o;
  T.ws_bytes = layout_of(t, tmask, dls::stream_path_possible(t, nullptr, 227 * 1024)).total;
  o = align_up(o + T.ws_bytes);
  T.off_node_of = o;  o = align_up(o + A * n);
  T.off_seq_pos = o;  o = align_up(o + A * n * sizeof(uint16_t));
  T.off_summary = o;  o = align_up(o + A * sizeof(dls_summary_t));
  T.off_fin = o;      o = align_up(o + A * K * sizeof(double));
  T.off_cnt = o;      o = align_up(o + A * K * sizeof(int32_t));
  T.off_avail = o;    o = align_up(o + A * K * sizeof(double));
  T.off_peak = o;     o = align_up(o + A * K * sizeof(double));
  T.off_cached = o;   o = align_up(o + A * P * sizeof(uint32_t));
  T.off_seq_task = o; o = align_up(o + A * n * sizeof(uint16_t));
  T.off_seq_node = o; o = align_up(o + A * n);
  T.total = o;
  return T;
}

// The replay path's two halves — the one-instance template run of DFS / Greedy / MRU and the per-instance Critical
// run — touch disjoint scratch and disjoint outputs, and the first is a handful of single-warp kernels whose time is
// pure latency (GPT-2: 0.36 ms of a 2.2 ms step; the 9,986-task DAG: 10.6 of 22 ms). They are therefore enqueued side
// by side: the template run on an auxiliary stream forked from the caller's stream by an event and joined back by
// another before the replay kernel, so that everything is still ordered on the caller's stream (and capturable).
// One auxiliary stream and two events per device, created on first use and kept; calls that use them are serialised
// while they ENQUEUE (the events are re-recorded per call). dls_schedule_batch_static never comes here.
struct ForkJoin {
  std::mutex mu;
  cudaStream_t aux = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool tried = false, ok = false;
};

ForkJoin* fork_join_for_current_device() {
  static ForkJoin table[kMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  ForkJoin& f = table[dev];
  std::lock_guard<std::mutex> lock(f.mu);
  if (!f.tried) {
    f.tried = true;
    f.ok = cudaStreamCreateWithFlags(&f.aux, cudaStreamNonBlocking) == cudaSuccess &&
           cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming) == cudaSuccess;
    if (!f.ok) (void)cudaGetLastError();  // (the calls then stay on one stream)
  }
  return f.ok ? &f : nullptr;
}

int run_schedule(const dls_batch_t* batch, dls_result_t* result, int algo_mask, void* workspace, size_t ws_bytes,
                 void* cuda_stream, bool clear_status, bool allow_replay) {
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  ctx.prof = g_prof.on ? &g_prof : nullptr;
  int rc;
  if (allow_replay && replay_applies(*batch, result, algo_mask)) {
    const int tmask = algo_mask & kDurationFree, cmask = algo_mask & ~kDurationFree;
    const size_t main_bytes = cmask ? layout_of(*batch, cmask, dls::stream_path_possible(*batch, result, 227 * 1024)).total : 0;
    const TemplateLayout T = template_layout(*batch, tmask, main_bytes);
    if (workspace && ws_bytes >= T.total && reinterpret_cast<uintptr_t>(workspace) % kAlign == 0) {
      char* p = static_cast<char*>(workspace);
      const dls_batch_t tb = template_batch(*batch);
      dls_result_t tr;
      tr.node_of = reinterpret_cast<int8_t*>(p + T.off_node_of);
      tr.seq_pos = reinterpret_cast<uint16_t*>(p + T.off_seq_pos);
      tr.summary = reinterpret_cast<dls_summary_t*>(p + T.off_summary);
      tr.fin = reinterpret_cast<double*>(p + T.off_fin);
      tr.cnt = reinterpret_cast<int32_t*>(p + T.off_cnt);
      tr.avail = reinterpret_cast<double*>(p + T.off_avail);
      tr.peak = result->peak ? reinterpret_cast<double*>(p + T.off_peak) : nullptr;
      tr.cached = result->cached ? reinterpret_cast<uint32_t*>(p + T.off_cached) : nullptr;
      tr.tfin = nullptr;
      ForkJoin* fj = (cmask && g_replay_mode.load() == 1) ? fork_join_for_current_device() : nullptr;
      std::unique_lock<std::mutex> fj_lock;
      dls::LaunchCtx tctx = ctx;  // the template run's launches
      if (fj) {
        fj_lock = std::unique_lock<std::mutex>(fj->mu);
        if (cudaEventRecord(fj->fork, ctx.st) == cudaSuccess && cudaStreamWaitEvent(fj->aux, fj->fork, 0) == cudaSuccess) {
          tctx.st = fj->aux;
        } else {
          (void)cudaGetLastError();
          fj = nullptr;
          fj_lock.unlock();
        }
      }
      rc = enqueue_schedule(&tb, &tr, tmask, p + T.off_ws, T.ws_bytes, tctx, true);
      ctx.launches = tctx.launches;
      if (rc != DLS_OK) ctx.why = tctx.why;
      if (rc == DLS_OK && cmask) {
        // Critical for every instance, written to its own slot of the caller's arrays
        int cslot = 0;
        for (int a = 0; a < DLS_ALGO_CRITICAL; ++a) cslot += (algo_mask >> a) & 1;
        dls_result_t cr = *result;
        cr.node_of += (int64_t)cslot * batch->nt_inst;
        cr.seq_pos += (int64_t)cslot * batch->nt_inst;
        cr.summary += (int64_t)cslot * batch->B;
        cr.fin += (int64_t)cslot * batch->nk_inst;
        cr.cnt += (int64_t)cslot * batch->nk_inst;
        cr.avail += (int64_t)cslot * batch->nk_inst;
        if (cr.peak) cr.peak += (int64_t)cslot * batch->nk_inst;
        if (cr.cached) cr.cached += (int64_t)cslot * batch->np_inst;
        rc = enqueue_schedule(batch, &cr, cmask, workspace, main_bytes, ctx, true);
      }
      if (fj) {  // join, whatever happened above: the caller's stream must not run ahead of the auxiliary one
        cudaError_t e = cudaEventRecord(fj->join, fj->aux);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx.st, fj->join, 0);
        if (e != cudaSuccess && rc == DLS_OK) rc = cuda_fail(e, "joining the template run's stream");
        fj_lock.unlock();
      }
      if (rc == DLS_OK) {
        dls::ReplayArgs a;
        a.nslots = count_bits(tmask);
        a.n = (int)tb.nt_inst;
        int ts = 0, fs = 0;
        for (int al = 0; al < DLS_NUM_ALGOS; ++al) {
          if (!((algo_mask >> al) & 1)) continue;
          if ((tmask >> al) & 1) a.full_slot[ts++] = fs;
          ++fs;
        }
        for (; ts < 4; ++ts) a.full_slot[ts] = 0;
        a.t_node_of = tr.node_of; a.t_seq_pos = tr.seq_pos; a.t_summary = tr.summary; a.t_cnt = tr.cnt;
        a.t_avail = tr.avail; a.t_peak = tr.peak; a.t_cached = tr.cached;
        a.seq_task = reinterpret_cast<uint16_t*>(p + T.off_seq_task);
        a.seq_node = reinterpret_cast<int8_t*>(p + T.off_seq_node);
        cudaError_t e = dls::launch_replay(*batch, *result, a, ctx);
        if (e != cudaSuccess) rc = cuda_fail(e, ctx.why);
      }
      g_launches += ctx.launches;
      if (rc == DLS_OK) g_err.clear();
      return rc;
    }
    // (scratch too small for the template's extra arrays: every instance is scheduled the ordinary way)
  }
  rc = enqueue_schedule(batch, result, algo_mask, workspace, ws_bytes, ctx, clear_status);
  g_launches += ctx.launches;
  if (rc == DLS_OK) g_err.clear();
  return rc;
}

}  // namespace

namespace dls {

cudaError_t ensure_dynamic_smem(const void* kernel, size_t bytes) {
  static std::mutex mu;
  static std::unordered_map<const void*, size_t> granted[kMaxDevices];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  std::lock_guard<std::mutex> lock(mu);
  size_t& have = granted[dev][kernel];
  if (bytes <= have) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) have = bytes;
  return e;
}

}  // namespace dls

extern "C" {

int dls_version(void) { return DLS_VERSION; }

const char* dls_last_error(void) { return g_err.c_str(); }

int64_t dls_launch_count(void) { return g_launches.load(); }

size_t dls_workspace_bytes(const dls_batch_t* batch, int algo_mask) {
  if (!batch) return 0;
  // Size for the larger of the two paths; which one runs is decided per call (see run_schedule).
  const bool stream = dls::stream_path_possible(*batch, nullptr, 227 * 1024);
  const size_t plain = layout_of(*batch, algo_mask, stream).total;
  if (!replay_applies(*batch, nullptr, algo_mask)) return plain;
  // room for the one-instance template run behind the ordinary layout (whichever path a call then takes)
  return std::max(plain, template_layout(*batch, algo_mask & kDurationFree, plain).total);
}

int dls_schedule_batch(const dls_batch_t* batch, dls_result_t* result, int algo_mask, void* workspace,
                       size_t workspace_bytes, void* cuda_stream) {
  int rc = validate(batch, algo_mask, true);
  if (rc != DLS_OK) return rc;
  if (!result || !result->node_of || !result->seq_pos || !result->summary || !result->fin || !result->cnt || !result->avail)
    return fail(DLS_ERR_BAD_ARG, "NULL required output in result");
  return run_schedule(batch, result, algo_mask, workspace, workspace_bytes, cuda_stream, false, true);
}

// See include/dls.h. Same work as dls_schedule_batch; the only difference is where the scratch lives.
int dls_schedule_batch_static(const dls_batch_t* batch, dls_result_t* result, int algo_mask, void* cuda_stream) {
  int rc = validate(batch, algo_mask, true);
  if (rc != DLS_OK) return rc;
  if (!result || !result->node_of || !result->seq_pos || !result->summary || !result->fin || !result->cnt || !result->avail)
    return fail(DLS_ERR_BAD_ARG, "NULL required output in result");
  const Layout L = layout_of(*batch, algo_mask, false);  // smallest scratch this batch can run with
  if (L.total > (size_t)DLS_STATIC_SCRATCH_BYTES) {
    char buf[200];
    snprintf(buf, sizeof buf,
             "batch needs %zu bytes of scratch but the built-in scratch holds %u bytes; split the batch or call "
             "dls_schedule_batch with your own workspace", L.total, (unsigned)DLS_STATIC_SCRATCH_BYTES);
    return fail(DLS_ERR_TOO_LARGE, buf);
  }
  void* base = nullptr;
  rc = static_scratch_address(&base);  // first call on this device: resolves the array's device address
  if (rc != DLS_OK) return rc;
  // The array still holds whatever the previous call left. The per-topology status words are the one region the
  // scheduling kernels trust before looking at anything else, so run_schedule clears them on the stream first.
  return run_schedule(batch, result, algo_mask, base, (size_t)DLS_STATIC_SCRATCH_BYTES, cuda_stream, true, false);
}

static int check_gen(const dls_gen_t* g) {
  if (!g || !g->rank || !g->param_pos) return fail(DLS_ERR_BAD_ARG, "NULL in dls_gen_t");
  if (g->B <= 0 || g->n <= 0 || g->K <= 0) return fail(DLS_ERR_BAD_ARG, "B, n, K must be positive");
  if (g->n > 20000 || g->K > DLS_MAX_NODES || g->max_deps < 0 || g->max_deps > 8)
    return fail(DLS_ERR_TOO_LARGE, "generator limits: n <= 20000, K <= 32, max_deps <= 8");
  return DLS_OK;
}

int dls_generate_counts(const dls_gen_t* gen, int32_t* e_count, int32_t* r_count, void* cuda_stream) {
  int rc = check_gen(gen);
  if (rc != DLS_OK) return rc;
  if (!e_count || !r_count) return fail(DLS_ERR_BAD_ARG, "NULL count output");
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e = dls::launch_gen_count(*gen, e_count, r_count, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  g_err.clear();
  return DLS_OK;
}

int dls_generate_fill(const dls_gen_t* gen, const dls_batch_t* out, void* cuda_stream) {
  int rc = check_gen(gen);
  if (rc != DLS_OK) return rc;
  if (!out || out->B != gen->B || out->T != gen->B || !out->node_off) return fail(DLS_ERR_BAD_ARG, "output batch does not match the generator");
  rc = validate(out, DLS_MASK_ALL, false);
  if (rc != DLS_OK) return rc;
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e = dls::launch_gen_fill(*gen, *out, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  g_err.clear();
  return DLS_OK;
}

int dls_aggregate(const dls_batch_t* batch, const dls_result_t* result, int algo_mask, const int32_t* group_of,
                  int32_t n_groups, double* out, void* cuda_stream) {
  int rc = validate(batch, algo_mask, false);
  if (rc != DLS_OK) return rc;
  if (!result || !result->summary || !result->fin || !result->cnt || !out || n_groups <= 0)
    return fail(DLS_ERR_BAD_ARG, "NULL or empty argument in dls_aggregate");
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e = dls::launch_aggregate(*batch, *result, count_bits(algo_mask & DLS_MASK_ALL), group_of, n_groups, out, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  g_err.clear();
  return DLS_OK;
}

int dls_rows_from_degrees(const uint8_t* deg, const int64_t* task_off, int32_t T, uint32_t* ptr, void* cuda_stream) {
  if (!deg || !task_off || !ptr || T <= 0) return fail(DLS_ERR_BAD_ARG, "NULL or empty argument in dls_rows_from_degrees");
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e = dls::launch_rows_from_degrees(deg, task_off, T, ptr, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  g_err.clear();
  return DLS_OK;
}

int dls_repeat_rank_row(const uint16_t* row, int32_t n, int32_t T, uint16_t* out, void* cuda_stream) {
  if (!row || !out || n <= 0 || T <= 0) return fail(DLS_ERR_BAD_ARG, "NULL or empty argument in dls_repeat_rank_row");
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  cudaError_t e = dls::launch_repeat_row(row, n, T, out, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  g_err.clear();
  return DLS_OK;
}

/* Test/measurement switch (no reference counterpart): 0 forces the general warp-per-DAG kernels for every
 * instance, 1 (default) lets qualifying batches take the thread-per-DAG streamed path. Returns the old value. */
int dls_set_stream_path(int enabled) { return g_stream_enabled.exchange(enabled != 0) ? 1 : 0; }

int dls_set_replay_path(int mode) { return g_replay_mode.exchange(mode == 0 ? 0 : mode == 2 ? 2 : 1); }

int dls_priorities(const dls_batch_t* batch, dls_prio_t* prio, void* workspace, size_t workspace_bytes, void* cuda_stream) {
  int rc = validate(batch, DLS_MASK_ALL, true);
  if (rc != DLS_OK) return rc;
  if (!prio) return fail(DLS_ERR_BAD_ARG, "prio is NULL");
  const Layout L = layout_of(*batch);
  if (!workspace || workspace_bytes < L.total) return fail(DLS_ERR_WORKSPACE, "workspace smaller than dls_workspace_bytes()");
  if (reinterpret_cast<uintptr_t>(workspace) % kAlign) return fail(DLS_ERR_WORKSPACE, "workspace must be 256-byte aligned");
  dls::Workspace w = carve(workspace, L);
  w.keep_up = 1;
  dls::LaunchCtx ctx;
  ctx.st = static_cast<cudaStream_t>(cuda_stream);
  ctx.prof = g_prof.on ? &g_prof : nullptr;
  cudaStream_t st = ctx.st;
  cudaError_t e = dls::launch_priority(*batch, w, DLS_MASK_ALL, ctx);
  g_launches += ctx.launches;
  if (e != cudaSuccess) return cuda_fail(e, ctx.why);
  const size_t nt = (size_t)batch->nt_topo, ncp = (size_t)(L.cp_per_instance ? batch->nt_inst : batch->nt_topo);
  struct { void* dst; const void* src; size_t bytes; } copies[] = {
      {prio->depth, w.depth, nt * sizeof(uint16_t)},         {prio->outdeg, w.outdeg, nt * sizeof(uint16_t)},
      {prio->order_dg, w.order_dg, nt * sizeof(uint16_t)},   {prio->order_mru, w.order_mru, nt * sizeof(uint16_t)},
      {prio->order_cp, w.order_cp, ncp * sizeof(uint16_t)},  {prio->up, w.up, ncp * sizeof(double)},
  };
  for (auto& c : copies) {
    if (!c.dst || c.bytes == 0) continue;
    e = cudaMemcpyAsync(c.dst, c.src, c.bytes, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(priorities)");
  }
  g_err.clear();
  return DLS_OK;
}

int dls_profile_enable(int on) {
  g_prof.on = on != 0;
  return DLS_OK;
}

int dls_profile_read(double* ms_per_kernel, int64_t* launches_per_kernel) {
  if (!ms_per_kernel || !launches_per_kernel) return fail(DLS_ERR_BAD_ARG, "NULL output in dls_profile_read");
  for (int i = 0; i < g_prof.n; ++i) {
    cudaError_t e = cudaEventSynchronize(g_prof.ev[i][1]);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, g_prof.ev[i][0], g_prof.ev[i][1]);
    if (e != cudaSuccess) { g_prof.n = 0; return cuda_fail(e, "dls_profile_read"); }
    ms_per_kernel[g_prof.kid[i]] += (double)ms;
    launches_per_kernel[g_prof.kid[i]] += 1;
  }
  g_prof.n = 0;
  return DLS_OK;
}

const char* dls_kernel_name(int kernel_id) {
  return (kernel_id >= 0 && kernel_id < DLS_NUM_KERNELS) ? kKernelNames[kernel_id] : "";
}

}  // extern "C"
