// dls_replay.cu — shared-topology batches whose instances differ in their durations only (config 3: one DAG,
// 100,000 perturbed-duration instances on one node set).
//
// DFS, Greedy and MRU never look at a duration: the processing order (depth, set order, out-degree), every fit test
// and every node choice depend on memory, params and the nodes alone (ref schedulers.py:154-208, 244-296, 444-525;
// SURVEY.md section 8a'). For such a batch their schedule is therefore ONE schedule: the library runs the three
// algorithms on instance 0 only (the "template"), and every instance then needs just what simulate_execution adds
// (ref simulation.py:261-271): the node finish times — fin[k] += dur/spd along the fixed assignment sequence, in that
// order, one rounding per task like the reference — and their maximum. Critical keeps its per-instance run: its order
// sorts by the upward rank, which is made of durations.
#include "dls_common.cuh"

namespace dls {

namespace {

constexpr int kReplayWarps = 4;

// seq_task[ts][q] = task assigned q-th, seq_node[ts][q] = its node, for each template slot; one CTA per slot.
__global__ void __launch_bounds__(256) replay_sequence_kernel(ReplayArgs a) {
  const int ts = blockIdx.x;
  const int8_t* node_of = a.t_node_of + (size_t)ts * a.n;
  const uint16_t* seq_pos = a.t_seq_pos + (size_t)ts * a.n;
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) {
    const int nd = node_of[i];
    if (nd >= 0) {
      const int q = seq_pos[i];
      a.seq_task[(size_t)ts * a.n + q] = (uint16_t)i;
      a.seq_node[(size_t)ts * a.n + q] = (int8_t)nd;
    }
  }
}

// Lane = (template slot, node): with K nodes a warp holds 32 / K slots of one instance, so that at K = 8 ONE warp
// replays all three algorithms of an instance. A lane walks its slot's assignment sequence (the lanes of a slot read
// the same entries) and adds the tasks that ran on its node; the per-task outputs are the template's, verbatim.
__global__ void __launch_bounds__(kReplayWarps * 32) replay_kernel(dls_batch_t b, dls_result_t r, ReplayArgs a, int per_warp,
                                                                    int warps_per_inst) {
  const int lane = threadIdx.x & 31;
  const long long wid = (long long)blockIdx.x * kReplayWarps + (threadIdx.x >> 5);
  if (wid >= (long long)warps_per_inst * b.B) return;
  const int inst = (int)(wid / warps_per_inst), part = (int)(wid % warps_per_inst);
  const int n = a.n, K = b.K_max;
  const int ts0 = part * per_warp, ts1 = min(a.nslots, ts0 + per_warp);
  for (int ts = ts0; ts < ts1; ++ts) {
    const int fs = a.full_slot[ts];
    const int8_t* src_n = a.t_node_of + (size_t)ts * n;
    const uint16_t* src_s = a.t_seq_pos + (size_t)ts * n;
    int8_t* dst_n = r.node_of + (int64_t)fs * b.nt_inst + (int64_t)inst * n;
    uint16_t* dst_s = r.seq_pos + (int64_t)fs * b.nt_inst + (int64_t)inst * n;
    for (int i = lane; i < n; i += 32) {
      dst_n[i] = src_n[i];
      dst_s[i] = src_s[i];
    }
    if (r.cached) {
      const int P = (int)(b.np_inst / b.B);
      const uint32_t* src_c = a.t_cached + (size_t)ts * P;
      uint32_t* dst_c = r.cached + (int64_t)fs * b.np_inst + (int64_t)inst * P;
      for (int p = lane; p < P; p += 32) dst_c[p] = src_c[p];
    }
  }
  // node finish times along the sequence (ref simulation.py:261-265)
  const int ts = ts0 + lane / K, k = lane % K;
  const bool on = lane < per_warp * K && ts < ts1;
  const int tsx = on ? ts : ts0;
  const dls_summary_t tsum = a.t_summary[tsx];
  const double* dur = b.dur + (int64_t)inst * n;
  const double spd = b.spd[k];
  const uint16_t* st = a.seq_task + (size_t)tsx * n;
  const int8_t* sn = a.seq_node + (size_t)tsx * n;
  double fin = 0.0;
  const int nseq = (on && tsum.status == DLS_ST_OK) ? tsum.completed : 0;
  int q = 0;
  for (; q + 4 <= nseq; q += 4) {  // loads of four steps in flight, additions in order
    const int t0 = st[q], t1 = st[q + 1], t2 = st[q + 2], t3 = st[q + 3];
    const int k0 = sn[q], k1 = sn[q + 1], k2 = sn[q + 2], k3 = sn[q + 3];
    const double c0 = __ldg(dur + t0), c1 = __ldg(dur + t1), c2 = __ldg(dur + t2), c3 = __ldg(dur + t3);
    if (k0 == k) fin = __dadd_rn(fin, __ddiv_rn(c0, spd));
    if (k1 == k) fin = __dadd_rn(fin, __ddiv_rn(c1, spd));
    if (k2 == k) fin = __dadd_rn(fin, __ddiv_rn(c2, spd));
    if (k3 == k) fin = __dadd_rn(fin, __ddiv_rn(c3, spd));
  }
  for (; q < nseq; ++q)
    if (sn[q] == k) fin = __dadd_rn(fin, __ddiv_rn(__ldg(dur + st[q]), spd));
  const int cnt = on ? a.t_cnt[(size_t)ts * K + k] : 0;
  double mk = (on && cnt > 0) ? fin : -INFINITY;  // makespan over the nodes that ran something (ref :269-271)
  for (int off = 1; off < K; off <<= 1) {  // max over the K lanes of a slot (K need not be a power of two)
    const double o = __shfl_down_sync(DLS_FULL_MASK, mk, off);
    if (k + off < K) mk = fmax(mk, o);
  }
  if (on) {
    const int fs = a.full_slot[ts];
    const int64_t ko = (int64_t)fs * b.nk_inst + (int64_t)inst * K;
    r.fin[ko + k] = fin;
    r.cnt[ko + k] = cnt;
    r.avail[ko + k] = a.t_avail[(size_t)ts * K + k];
    if (r.peak) r.peak[ko + k] = a.t_peak[(size_t)ts * K + k];
    if (k == 0) {
      dls_summary_t s = tsum;
      s.makespan = (tsum.status == DLS_ST_OK && tsum.completed > 0) ? mk : 0.0;
      r.summary[(int64_t)fs * b.B + inst] = s;
    }
  }
}

}  // namespace

cudaError_t launch_replay(const dls_batch_t& b, const dls_result_t& r, const ReplayArgs& a, LaunchCtx& ctx) {
  int rec = ctx.prof ? ctx.prof->begin(DLS_KERNEL_REPLAY, ctx.st) : -1;
  replay_sequence_kernel<<<a.nslots, 256, 0, ctx.st>>>(a);
  ++ctx.launches;
  ctx.why = "replay_sequence_kernel launch";
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    const int per_warp = 32 / b.K_max;  // (K_max <= 32)
    const int warps_per_inst = (a.nslots + per_warp - 1) / per_warp;
    const long long warps = (long long)warps_per_inst * b.B;
    replay_kernel<<<(unsigned)((warps + kReplayWarps - 1) / kReplayWarps), kReplayWarps * 32, 0, ctx.st>>>(b, r, a, per_warp,
                                                                                                        warps_per_inst);
    ++ctx.launches;
    ctx.why = "replay_kernel launch";
    e = cudaGetLastError();
  }
  if (ctx.prof) ctx.prof->end(rec, ctx.st);
  return e;
}

}  // namespace dls
