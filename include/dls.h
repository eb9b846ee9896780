/*
 * dls.h — C ABI of the B200-native batched DAG list scheduler.
 *
 * Drop-in boundary for ONE hot path of 2alaaa/distributed-llm-scheduler: the four
 * memory-constrained list-scheduling algorithms of schedulers.py run over a batch of
 * DAG instances, plus the makespan/cache statistics of simulation.py:simulate_execution.
 * (All file:line citations below are into the reference repository.)
 *
 * Everything crossing this boundary is plain pointers and sizes. The library owns no
 * memory between calls: inputs, outputs and scratch are caller-allocated DEVICE buffers,
 * borrowed for the duration of the call; work is enqueued on the caller's CUDA stream.
 * There is no CPU fallback: every entry point that computes needs a CUDA device.
 *
 * Lowered representation (what a DAG builder emits; SURVEY.md section 8(a')):
 *   tasks    0..n-1 in add_task order                         (schedulers.py:43-48)
 *   pred CSR dependencies of each task, in list order          (schedulers.py:14)
 *   par  CSR distinct params of each task; param id = lexicographic rank of the param
 *            name inside its DAG (needed by the (score, name) sort, schedulers.py:420)
 *   rank[i]  position of task i in list(scheduler.pending_tasks) taken after the last
 *            add_task — CPython set order, the reference's tie-break (schedulers.py:41,58)
 *   mem,dur  Task.memory_required / Task.compute_time, fp64   (schedulers.py:12-13)
 *   cap,spd  Node.total_memory / Node.compute_speed, fp64, in the order of the `nodes`
 *            list given to the scheduler = node tie-break order (schedulers.py:23-25,34)
 */
#ifndef DLS_H_
#define DLS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLS_VERSION 200 /* major*100 + minor; 2.0: dls_summary_t carries the instance's peak memory (48 bytes) */

/* algorithm ids (slot order inside results = increasing id among those requested) */
#define DLS_ALGO_DFS 0      /* DFSScheduler           schedulers.py:138-208 */
#define DLS_ALGO_GREEDY 1   /* GreedyScheduler        schedulers.py:211-296 */
#define DLS_ALGO_CRITICAL 2 /* CriticalPathScheduler  schedulers.py:299-372 */
#define DLS_ALGO_MRU 3      /* MRUScheduler           schedulers.py:375-525 */
#define DLS_NUM_ALGOS 4
#define DLS_MASK_ALL 0xF

/* limits */
#define DLS_MAX_TASKS 65535  /* task ids are u16; 0xFFFF is the "none" marker */
#define DLS_MAX_PARAMS 65535 /* param ids are u16 */
#define DLS_MAX_NODES 32     /* one warp lane per node */

/* error codes (negative); dls_last_error() gives the text for the calling thread */
#define DLS_OK 0
#define DLS_ERR_BAD_ARG (-1)   /* null pointer, negative size, inconsistent descriptor */
#define DLS_ERR_TOO_LARGE (-2) /* n > 65535, K > 32, P > 65535, or per-DAG state exceeds shared memory */
#define DLS_ERR_WORKSPACE (-3) /* workspace smaller than dls_workspace_bytes() */
#define DLS_ERR_CUDA (-4)      /* a CUDA runtime call failed (text in dls_last_error) */

/* dls_batch_t.flags (only meaningful when T == 1, i.e. topology shared by all instances) */
#define DLS_F_DUR_PER_INSTANCE 1 /* dur is [B][n] instead of [n]  (config 3: perturbed durations) */
#define DLS_F_MEM_PER_INSTANCE 2 /* mem is [B][n] instead of [n] */

/* per-task verdict in node_of[] */
#define DLS_TASK_FAILED (-1)  /* in scheduler.failed_tasks            (schedulers.py:128-131) */
#define DLS_TASK_PENDING (-2) /* left in scheduler.pending_tasks      (schedulers.py:172-174) */
#define DLS_SEQ_NONE 0xFFFFu

/* dls_summary_t.status bits */
#define DLS_ST_OK 0
#define DLS_ST_BAD_TOPOLOGY 1 /* dependency cycle or index out of range: instance not scheduled */
#define DLS_ST_CARRY 2        /* MRU: assign_task_to_node returned False after a node was chosen
                                 (schedulers.py:82-83,509); the task would be retried next round.
                                 Not replayed on the device: results of this instance are invalid. */

/*
 * A batch of B DAG instances over T topologies. T == B: every instance has its own
 * topology (config 5, the simulation.py sweep). T == 1: one topology shared by all
 * instances, values optionally per instance (config 3).
 *
 * Topology t owns tasks [task_off[t], task_off[t+1]) of the per-task arrays, entries
 * [pred_off[t], pred_off[t+1]) of pred_idx, [par_off[t], par_off[t+1]) of par_idx and
 * param ids [pid_off[t], pid_off[t+1]) of the per-param outputs. CSR row pointers are
 * local to the topology's slice and have n_t+1 entries starting at task_off[t]+t.
 */
typedef struct dls_batch {
  int32_t B;     /* DAG instances */
  int32_t T;     /* topologies: B or 1 */
  int32_t n_max; /* max tasks of any topology   (<= DLS_MAX_TASKS) */
  int32_t K_max; /* max nodes of any instance   (<= DLS_MAX_NODES) */
  int32_t P_max; /* max distinct params of any topology (<= DLS_MAX_PARAMS) */
  int32_t E_max; /* max dependency entries of any topology */
  int32_t R_max; /* max param references of any topology */
  int32_t flags; /* DLS_F_* */
  /* totals, known to the caller on the host (the library never reads device memory on the host) */
  int64_t nt_topo; /* task_off[T]: tasks over all topologies */
  int64_t nt_inst; /* NT: instance-tasks = T==1 ? B*n : task_off[T] */
  int64_t e_total; /* pred_off[T] */
  int64_t r_total; /* par_off[T] */
  int64_t nk_inst; /* NK: instance-nodes = node_off ? node_off[B] : B*K_max */
  int64_t np_inst; /* NP: instance-params = T==1 ? B*pid_off[1] : pid_off[T] */
  const int64_t* task_off; /* [T+1] */
  const int64_t* pred_off; /* [T+1] */
  const int64_t* par_off;  /* [T+1] */
  const int64_t* pid_off;  /* [T+1] */
  const uint32_t* pred_ptr; /* [task_off[T]+T] */
  const uint16_t* pred_idx; /* [pred_off[T]]   Task.dependencies             schedulers.py:14 */
  const uint32_t* par_ptr;  /* [task_off[T]+T] */
  const uint16_t* par_idx;  /* [par_off[T]]    Task.params_needed            schedulers.py:15 */
  const uint16_t* rank;     /* [task_off[T]]   pi, set order of pending_tasks schedulers.py:41 */
  const double* mem;        /* [task_off[T]] or [B*n] with DLS_F_MEM_PER_INSTANCE */
  const double* dur;        /* [task_off[T]] or [B*n] with DLS_F_DUR_PER_INSTANCE */
  const int64_t* node_off;  /* [B+1], or NULL: all instances use nodes 0..K_max-1 */
  const double* cap;        /* Node.total_memory   schedulers.py:23 */
  const double* spd;        /* Node.compute_speed  schedulers.py:25 */
} dls_batch_t;

/* Per (algorithm, instance) record. Mirrors what run_single_test reads after schedule()
 * (simulation.py:336-346) plus simulate_execution's outputs (simulation.py:216-278). */
typedef struct dls_summary {
  double makespan;   /* max over nodes of the sequential sum of dur/spd   simulation.py:261-271 */
  double peak;       /* extension, no reference counterpart: largest per-node peak memory of the instance (max over
                        nodes of dls_result_t.peak); 0.0 unless the peak output was requested */
  int32_t completed; /* len(scheduler.completed_tasks)                    simulation.py:339 */
  int32_t failed;    /* len(scheduler.failed_tasks)                       simulation.py:340 */
  int32_t pending;   /* len(scheduler.pending_tasks) after schedule() */
  int32_t hits;      /* stats['param_cache_hits']                         simulation.py:253-258 */
  int32_t misses;    /* stats['param_cache_misses'] */
  int32_t rounds;    /* iterations of the while loop                      schedulers.py:167 */
  int32_t status;    /* DLS_ST_* */
  int32_t reserved;
} dls_summary_t;

/*
 * Caller-allocated outputs. Leading dimension A = number of algorithms in algo_mask
 * (slot s = s-th set bit). NT = batch->nt_inst, NK = batch->nk_inst, NP = batch->np_inst.
 * Nullable: peak, cached, tfin.
 */
typedef struct dls_result {
  int8_t* node_of;        /* [A][NT] node index, DLS_TASK_FAILED or DLS_TASK_PENDING   Task.assigned_node schedulers.py:93 */
  uint16_t* seq_pos;      /* [A][NT] position in the global assignment sequence, DLS_SEQ_NONE if unassigned */
  dls_summary_t* summary; /* [A][B] */
  double* fin;            /* [A][NK] node finish time                       simulation.py:269 */
  int32_t* cnt;           /* [A][NK] len(node.completed_tasks)              schedulers.py:123 */
  double* avail;          /* [A][NK] final node.available_memory (bit-exact fp64 replay) schedulers.py:89,95,126 */
  double* peak;           /* [A][NK] extension, no reference counterpart: max of total-available sampled
                             between schedulers.py:95 and :126; NULL to skip. Produced by every kernel path. */
  uint32_t* cached;       /* [A][NP] bit k set = param cached on node k at the end (Node.cached_params,
                             schedulers.py:26,88,428); NULL to skip */
  double* tfin;           /* [A][NK] extension, no reference counterpart: node finish
                             time on a DEPENDENCY-AWARE timeline. `fin` follows simulate_execution, which lets a node
                             run its tasks back to back whatever they depend on (simulation.py:238-271); here a task
                             starts when its node is free AND all its dependencies have finished:
                             t_i = max(node free, max finish of deps) + compute_time / compute_speed, tasks taken in
                             assignment order. The dependency-aware makespan is the maximum over nodes with cnt > 0.
                             NULL to skip; like cached it is produced by the warp-per-DAG kernels only. */
} dls_result_t;

/* Outputs of the priority pass alone (kernel P). All nullable. Per topology-task arrays have
 * nt_topo entries; per instance-task arrays have nt_inst entries. */
typedef struct dls_prio {
  uint16_t* depth;     /* [task_off[T]] DFSScheduler.compute_depth                 schedulers.py:140-152 */
  uint16_t* outdeg;    /* [task_off[T]] len(task_dependencies[id]) (MRU urgency)   schedulers.py:468-469 */
  double* up;          /* [NT]          CriticalPathScheduler.compute_critical_path schedulers.py:301-321 */
  uint16_t* order_dg;  /* [task_off[T]] static processing order for DFS/Greedy: (depth asc, rank asc) */
  uint16_t* order_mru; /* [task_off[T]] for MRU: (depth asc, outdeg desc, rank asc)  schedulers.py:475 */
  uint16_t* order_cp;  /* [NT]          for Critical: (depth asc, up desc, rank asc) schedulers.py:343 */
} dls_prio_t;

/* Library version (DLS_VERSION of the build). */
int dls_version(void);

/* Text of the last error on the calling thread ("" if none). */
const char* dls_last_error(void);

/* Bytes of device scratch dls_schedule_batch / dls_priorities need for this batch
 * (only the scalar fields of the descriptor are read). */
size_t dls_workspace_bytes(const dls_batch_t* batch, int algo_mask);

/*
 * Replaces, for every instance b and every algorithm a in algo_mask:
 *   scheduler = XScheduler(nodes); for t in tasks: scheduler.add_task(t); schedule = scheduler.schedule()
 *   makespan, stats = evaluator.simulate_execution(scheduler, schedule)
 * (schedulers.py:154,244,323,444; simulation.py:320-336). All pointers inside batch/result and
 * `workspace` are device pointers; the structs themselves are host memory. Asynchronous on `cuda_stream`
 * (a cudaStream_t, NULL = default stream). Returns DLS_OK or a negative DLS_ERR_*; never throws.
 * Infeasible tasks are data (node_of = -1), not errors, like the reference (schedulers.py:128-131).
 */
int dls_schedule_batch(const dls_batch_t* batch, dls_result_t* result, int algo_mask, void* workspace,
                       size_t workspace_bytes, void* cuda_stream);

/* Kernel P only: depth / out-degree / upward rank / static processing orders
 * (compute_depth schedulers.py:140-152, compute_critical_path :301-321, the ready-list sorts :177,:343,:475). */
int dls_priorities(const dls_batch_t* batch, dls_prio_t* prio, void* workspace, size_t workspace_bytes,
                   void* cuda_stream);

/*
 * dls_schedule_batch without a caller-supplied workspace. (No counterpart in the reference.)
 *
 * Some hosts of this library cannot hand it scratch memory at all: a plugin that is only ever given its input
 * and output buffers by an application that has already claimed the whole GPU for its own pools; a control loop
 * that must not call cudaMalloc because that call may synchronise the device and stall everything queued on it;
 * a binding in a runtime that has no device allocator to call in the first place. For them this entry point
 * keeps the scratch (depth, orders, upward rank, reverse adjacency, status words) in a fixed array that lives
 * inside the library's device image: DLS_STATIC_SCRATCH_BYTES bytes, resident on every GPU that loads the
 * library until the process exits. It never calls cudaMalloc, cudaFree or any other allocating routine; inputs and
 * outputs are still the caller's device buffers, exactly as for dls_schedule_batch, and results are identical.
 *
 * - A batch whose smallest workable scratch exceeds DLS_STATIC_SCRATCH_BYTES is refused with DLS_ERR_TOO_LARGE and
 *   a message naming both numbers; nothing is launched. (About 1,300 DAGs of 1,000 tasks. Within the limit, the
 *   faster streamed kernels are used when their larger scratch also fits, else the general ones.)
 * - The first call on each device looks up the array's device address (one cudaGetSymbolAddress, made with that
 *   device current). Make that first call outside any stream capture.
 * - There is ONE array per device and every call uses it. It still holds the previous call's contents when the
 *   next call starts; the words that must start at zero are cleared on the stream before the kernels run. Two calls
 *   whose GPU work can overlap (different streams, or different host threads on one device) race on it: serialise
 *   them, or use dls_schedule_batch with separate workspaces instead.
 */
#define DLS_STATIC_SCRATCH_BYTES (32u << 20)
int dls_schedule_batch_static(const dls_batch_t* batch, dls_result_t* result, int algo_mask, void* cuda_stream);

/*
 * Device-side DAG builder: emits B random DAGs of n tasks straight into the CSR arrays of a
 * dls_batch_t, with the distribution of DAGGenerator.generate_random_dag (simulation.py:91-114) and node sets from
 * calculate_total_memory_needed / create_nodes_with_memory_regime (simulation.py:161-214). The random stream is a
 * counter-based hash of (seed, instance, task, draw), so instances are reproducible but are not the reference's
 * Mersenne-Twister instances. Two steps because CSR sizes are data dependent:
 *   1. dls_generate_counts: e_count[b], r_count[b] = dependency / param-reference entries of instance b;
 *   2. the caller prefix-sums them into pred_off/par_off/pid_off (pid_off == par_off: params are private), sets
 *      task_off[b] = b*n and node_off[b] = b*K, allocates the arrays, and calls dls_generate_fill.
 */
typedef struct dls_gen {
  int32_t B, n, K, max_deps; /* max_deps <= 8 (reference default 3) */
  double regime;             /* memory regime, e.g. 0.9 */
  uint64_t seed;
  const uint16_t* rank;      /* [n]  device: pi of the task names 'task_0'..'task_{n-1}' (host hash order) */
  const uint16_t* param_pos; /* [2n] device: position of 'param_{i}_{j}' (slot 2i+j) among all 2n names sorted */
} dls_gen_t;
int dls_generate_counts(const dls_gen_t* gen, int32_t* e_count, int32_t* r_count, void* cuda_stream);
int dls_generate_fill(const dls_gen_t* gen, const dls_batch_t* out, void* cuda_stream);

/*
 * Sweep-level aggregation on the device (replaces the groupby(...).mean() tables of
 * analyze_results, simulation.py:452,518-524,552-563). group_of[b] in [0, n_groups) assigns every instance to a group
 * (e.g. its memory regime); NULL puts all instances in group 0. out[slot][group][DLS_AGG_FIELDS] receives SUMS:
 *   0 count, 1 completion_rate, 2 makespan, 3 avg_node_utilization, 4 load_balance_score, 5 param_cache_hits,
 *   6 param_cache_misses, 7 instances with completed_tasks > 0 (the reference averages makespan over those only).
 * `result` is the device result of a finished or enqueued dls_schedule_batch on the same stream.
 */
#define DLS_AGG_FIELDS 8
int dls_aggregate(const dls_batch_t* batch, const dls_result_t* result, int algo_mask, const int32_t* group_of,
                  int32_t n_groups, double* out, void* cuda_stream);

/* Host-to-device traffic savers for pipelined HOST batches (no reference counterpart; the reference never leaves the
 * host). A CSR row-pointer array is the prefix sum of its row lengths, and a homogeneous sweep repeats one rank row:
 * the host ships one byte per task instead of a u32 (possible while no task has more than 255 dependencies /
 * params) and one rank row instead of T, and these calls rebuild the arrays dls_schedule_batch reads, in device
 * memory, on the caller's stream. All pointers are device pointers.
 *   dls_rows_from_degrees: ptr[task_off[t] + t + i] = deg[task_off[t]] + ... + deg[task_off[t] + i - 1], i = 0..n_t
 *     (i.e. the layout of dls_batch_t.pred_ptr / par_ptr); deg has task_off[T] entries, ptr task_off[T] + T.
 *   dls_repeat_rank_row: out[t * n + i] = row[i] for t < T (dls_batch_t.rank of T topologies of n tasks each). */
int dls_rows_from_degrees(const uint8_t* deg, const int64_t* task_off, int32_t T, uint32_t* ptr, void* cuda_stream);
int dls_repeat_rank_row(const uint16_t* row, int32_t n, int32_t T, uint16_t* out, void* cuda_stream);

/* Test/measurement switch (no reference counterpart). Qualifying batches (K <= 8 nodes, small per-DAG state, no
 * extension outputs requested) normally run on the thread-per-DAG streamed kernels, fed by the priority and
 * stream-build kernels. 0 forces the general warp-per-DAG kernels for everything; any other value restores the
 * default. Returns the previous setting (0 or 1). */
int dls_set_stream_path(int enabled);

/* Test/measurement switch (no reference counterpart). A batch with ONE topology, one node set (node_off == NULL) and
 * per-instance durations only (flags == DLS_F_DUR_PER_INSTANCE) normally has DFS, Greedy and MRU scheduled once, on
 * instance 0 — their decisions never read a duration (schedulers.py:154-208,244-296,444-525) — and every instance's
 * node finish times and makespan replayed along that one assignment sequence (simulation.py:261-271); Critical still
 * runs per instance. The one-instance run is enqueued on an auxiliary stream of the library (one per device, created
 * on first use), forked from the caller's stream by an event and joined back into it before the replay, so that it
 * runs beside the Critical kernels; all of the call's work is still ordered on the caller's stream, and the call can
 * be captured into a graph. mode 0 schedules every instance with every algorithm instead; mode 2 keeps the replay
 * but enqueues everything on the caller's stream alone; any other value restores the default (1).
 * Returns the previous mode. dls_schedule_batch_static does not use this path. */
int dls_set_replay_path(int mode);

/* Number of kernel launches the library has enqueued since load (for bench accounting). */
int64_t dls_launch_count(void);

/* Measurement support (no reference counterpart): when enabled, every kernel the library enqueues is bracketed by
 * CUDA events on the caller's stream. dls_profile_read waits for the recorded events, ADDS the elapsed
 * milliseconds / launch counts per kernel id into the caller's arrays (DLS_NUM_KERNELS entries each) and clears
 * the record list. At most 4096 launches are kept between reads. The switch and the record list are process-wide and
 * not synchronised: measurement code for one host thread. While it is on, events are created on demand
 * (cudaEventCreate), also under dls_schedule_batch_static — leave it off where that entry point's "allocates nothing"
 * matters. */
#define DLS_KERNEL_PRIO_TOPOLOGY 0
#define DLS_KERNEL_PRIO_CRITICAL 1
#define DLS_KERNEL_SCHED_DFS 2
#define DLS_KERNEL_SCHED_GREEDY 3
#define DLS_KERNEL_SCHED_CRITICAL 4
#define DLS_KERNEL_SCHED_MRU 5
#define DLS_KERNEL_PARAM_CLASS 6
#define DLS_KERNEL_STREAM_BUILD 7
#define DLS_KERNEL_STREAM_DFS 8
#define DLS_KERNEL_STREAM_GREEDY 9
#define DLS_KERNEL_STREAM_CRITICAL 10
#define DLS_KERNEL_STREAM_MRU 11
#define DLS_KERNEL_UNPERMUTE 12
#define DLS_KERNEL_REPLAY 13
#define DLS_NUM_KERNELS 14
int dls_profile_enable(int on);
int dls_profile_read(double* ms_per_kernel, int64_t* launches_per_kernel);
const char* dls_kernel_name(int kernel_id);

#ifdef __cplusplus
}
#endif
#endif /* DLS_H_ */
