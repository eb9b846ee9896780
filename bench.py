#!/usr/bin/env python
"""bench.py — DAG schedules/sec on 1k-task DAGs (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--instances B_per_gpu] [--impl reference]
                    [--workload random1k|cfg3|cfg4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A step = one pass of the hot path (priority kernels, stream build, the four list-scheduling kernels, unpermute) over
one batch of synthetic DAG instances per GPU. Default workload `random1k` = config-5 shape (n=1000, K=8 nodes, memory
regime 0.9, all four algorithms, peak memory on); `cfg3` = GPT-2 DAG x 100,000 perturbed-duration instances on 8
capped devices (one shared topology); `cfg4` = 48-layer x 205-head LLM DAG (9,986 tasks) x 512 instances. Instances
are sharded across ranks with no data-path collective; the only collective is ONE all-gather of the fixed-size
per-instance result records (makespan, peak memory, counters) after the last step of the sweep.

value  = schedules/s with inputs resident in HBM (CUDA events, max over ranks).
e2e    = the same through the public API with HOST inputs: pinned host->device copy of the batch, the C-ABI calls,
         device->host copy of the results, each step (random1k: PipelinedScheduler, chunks overlapped on 3 streams).
roofline = dominant kernel (largest share of the timed region, CUDA events recorded by the library on the launch
         stream): algorithmic bytes per launch (29n + 4(E+R) + 28K + 32 per schedule) / duration,
         against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
parity = the CPU oracle's results for the first instances of the step (the cpu_baseline sample) compared field by
         field, bit for bit, with what the GPU produced for the same instances (outside every timed region).
cpu_baseline / --impl reference = the CPU oracle (C port of the reference's algorithm, oracle/dls_oracle.c) on the
         host's threads over a bounded sample of THE SAME instances (same generator, seed and tie-break order; built
         on the host by the numpy port of the generator kernels). The reference itself is CPython and does not exist
         on the GPU box; its measured numbers are in BASELINE.md section 2.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_TASKS, K_NODES, REGIME = 1000, 8, 0.9
SEED = 1000
METRIC = "DAG schedules/sec (1k-task DAGs)"
UNIT = "schedules/s"
ALGO_LABEL = ("DFS", "Greedy", "Critical", "MRU_spec")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--instances", type=int, default=0, help="DAG instances per GPU per step (0 = the workload's default)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="random1k", choices=["random1k", "cfg3", "cfg4"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-peak", action="store_true", help="do not track peak memory (for the with/without comparison)")
    ap.add_argument("--results", default="full", choices=["full", "summary"],
                    help="what the end-to-end leg brings back to the host: every array, or the per-instance records, "
                         "node finish times and task counts only")
    return ap.parse_args()


def tie_break_rank(n):
    """pi is an input (the reference takes it from CPython's set order, which changes with the process's hash seed):
    the bench fixes one permutation so that every rank, both arms and every run schedule the same instances."""
    return np.random.default_rng(12345).permutation(n).astype(np.uint16)


def algorithmic_bytes(batch):
    """B(n,E,R,K) = 29n + 4(E+R) + 28K + 32 bytes per (instance, algorithm), summed over the batch.
    Shared-topology batches (T == 1): per-instance unique bytes only — the perturbed durations 8n (+ 8n memory) and the
    outputs 3n + 12K + 24 — plus the topology once."""
    if batch.T == batch.B:
        return 29 * batch.nt_inst + 4 * (batch.e_total + batch.r_total) + 28 * batch.nk_inst + 32 * batch.B
    n = int(batch.task_off[1])
    per_inst = 3 * n + 12 * batch.K_max + 24
    per_inst += 8 * n * (bool(batch.flags & 1) + bool(batch.flags & 2))
    once = 26 * n + 4 * (batch.e_total + batch.r_total) + 16 * batch.K_max + 8
    return per_inst * batch.B + once


def priority_bytes(batch):
    """SURVEY.md 8(d) kernel P alone: 4(n+1)+4E+8n in, 2n+8n+2n out."""
    return 4 * (batch.nt_inst + batch.B) + 4 * batch.e_total + 8 * batch.nt_inst + 12 * batch.nt_inst


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel_name, instances):
    """dram bytes per launch from the committed ncu capture, scaled per instance (profiles/ncu_traffic.json)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    try:
        per_inst = json.load(open(p)).get(kernel_name, {}).get("dram_bytes_per_instance")
        return None if per_inst is None else float(per_inst) * instances
    except Exception:
        return None


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.gpu = gpu_index
        self.skip = 0
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def mark(self):
        """The timed region starts here: samples written before are dropped. (The sampler is started BEFORE the region:
        nvidia-smi attaching to the driver stalled rank 0's first launches by 35-80 ms on an 8-GPU box, which the other
        ranks then waited for in the sweep's collective — r2f, 22.6 instead of 15.0 ms per step.)"""
        try:
            self.skip = os.path.getsize(self.f.name)
        except OSError:
            self.skip = 0

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(self.skip)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


# ---- workloads ----------------------------------------------------------------------------------------------------

class Random1k:
    """config 5 shape: every instance its own random 1k-task DAG (generate_random_dag's distribution)."""
    name, default_B, mask = "random1k", 131072, 0xF

    def describe(self):
        return (f"random-1k (config-5 shape): n={N_TASKS} tasks, K={K_NODES} nodes, memory regime {REGIME}, algorithms "
                f"DFS+Greedy+Critical+MRU_spec, counter-hash generator seed {SEED}+rank, fixed tie-break permutation")

    def device_batch(self, B, rank_id, device):
        from dls_b200.gen import random_dag_batch_device
        return random_dag_batch_device(B, N_TASKS, K_NODES, REGIME, seed=SEED + rank_id, device=device, rank=tie_break_rank(N_TASKS))

    def host_batch(self, B, rank_id, first=0):
        from dls_b200.gen import random_dag_batch_counter
        return random_dag_batch_counter(B, N_TASKS, K_NODES, REGIME, seed=SEED + rank_id, rank=tie_break_rank(N_TASKS), first=first)


class SharedTopology:
    """configs 3 and 4: one DAG, B instances with durations perturbed by +-10 % (T == 1, DLS_F_DUR_PER_INSTANCE)."""
    mask = 0xF

    def __init__(self, name):
        from dls_b200 import gen
        self.name = name
        if name == "cfg3":  # GPT-2 small DAG of test_gpt2.py on the four laptops listed twice
            self.dag, self.default_B = gen.gpt2_dag(), 100_000
            self.cap = np.array([8.0, 8.0, 6.0, 6.0] * 2)
            self.spd = np.array([1.0, 1.2, 0.8, 0.9] * 2)
            self.text = "cfg3: GPT-2 small DAG (99 tasks, 1 shared param) on 8 capped devices"
        else:               # 48 layers x 205 heads, create_nodes_with_memory_regime(total, 0.9, 8)
            self.dag, self.default_B = gen.llm_dag(48, 205, 205), 512
            npar = np.diff(self.dag.par_ptr.astype(np.int64))
            total = np.array([(self.dag.mem + npar * 0.5).max() + self.dag.P * 0.5])
            cap, spd = gen.node_sets(total, 0.9, 8, np.random.default_rng(SEED))
            self.cap, self.spd = cap[0], spd[0]
            self.text = "cfg4: 48-layer x 205-head LLM DAG (9,986 tasks) on 8 nodes at memory regime 0.9"
        self.dag.rank = tie_break_rank(self.dag.n)

    def describe(self):
        return (f"{self.text}, durations perturbed +-10 % per instance (seed {SEED}+rank), one shared topology, algorithms "
                "DFS+Greedy+Critical+MRU_spec, fixed tie-break permutation")

    def host_batch(self, B, rank_id, first=0):
        from dls_b200.gen import perturbed_durations
        from dls_b200.lowering import DagBatch
        dur = perturbed_durations(self.dag, first + B, 0.1, seed=SEED + rank_id)[first:]
        return DagBatch.shared_topology(self.dag, self.cap, self.spd, B, dur=dur)

    def device_batch(self, B, rank_id, device):
        from dls_b200.batch import DeviceBatch
        return DeviceBatch(self.host_batch(B, rank_id), device)


def make_workload(name):
    return Random1k() if name == "random1k" else SharedTopology(name)


# ---- the CPU arm ----------------------------------------------------------------------------------------------------

def oracle_time(oracle, sample, mask, threads, reps=1):
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        res = oracle.schedule_batch(sample, mask, nthreads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, res


def cpu_sample_size(oracle, wl, rank_id, threads, target_s, cap):
    """How many of the step's first instances ~target_s seconds of all-thread oracle time cover (all four algorithms)."""
    pilot = wl.host_batch(min(cap, max(threads, 8)), rank_id)
    dt, _ = oracle_time(oracle, pilot, wl.mask, threads)
    per_inst = max(dt, 1e-4) / pilot.B
    m = int(min(cap, max(pilot.B, target_s / per_inst)))
    return max(threads, m // threads * threads) if cap >= threads else cap


def cpu_baseline(oracle, wl, rank_id, threads, target_s, cap):
    """The oracle over the first instances of this rank's step: all algorithms on all host threads (the blended
    figure), each algorithm on its own (MRU is the slow one: per-eviction rescans of the pending set,
    ref schedulers.py:396-400), and one thread alone. Returns (record, sample batch, oracle results of the blended run)."""
    m = cpu_sample_size(oracle, wl, rank_id, threads, target_s * 0.55, cap)
    sample = wl.host_batch(m, rank_id)
    dt_all, res = oracle_time(oracle, sample, wl.mask, threads)
    per_algo, per_algo_1 = {}, {}
    small = sample.shard(0, max(1, min(m, max(threads, m // 4))))
    for a in range(4):
        dt, _ = oracle_time(oracle, small, 1 << a, threads)
        per_algo[ALGO_LABEL[a]] = small.B / dt
    one = sample.shard(0, max(1, min(m, max(2, m // (4 * threads)))))
    for a in range(4):
        dt, _ = oracle_time(oracle, one, 1 << a, 1)
        per_algo_1[ALGO_LABEL[a]] = one.B / dt
    rec = {"value": 4 * sample.B / dt_all, "unit": UNIT, "cores": threads, "kind": "port",
           "sample": f"first {sample.B} instances of the step x 4 algorithms, {dt_all:.1f} s on {threads} threads "
                     "(oracle/dls_oracle.c, C port of the reference's algorithm)",
           "host_threads": os.cpu_count(),
           "per_algorithm": {k: {"all_threads": v, "one_thread": per_algo_1[k]} for k, v in per_algo.items()},
           "per_algorithm_sample": f"{small.B} instances on {threads} threads, {one.B} on one thread"}
    return rec, sample, res


def parity_check(res, want, m, n, K):
    """Device results (BatchResult of the resident run) against the oracle's for the first m instances: node_of, seq_pos,
    the whole per-instance record, cnt and the fp64 bits of fin / avail / peak. Returns the `parity` record."""
    A = len(res.algos)
    bits = lambda a: np.ascontiguousarray(a, np.float64).view(np.uint64)
    bad = np.zeros((A, m), bool)
    got_nodes = res.node_of[:, :m * n].cpu().numpy().reshape(A, m, n)
    got_seq = res.seq_pos[:, :m * n].cpu().numpy().view(np.uint16).reshape(A, m, n)
    bad |= (got_nodes != want.node_of.reshape(A, m, n)).any(2)
    bad |= (got_seq != want.seq_pos.reshape(A, m, n)).any(2)
    got_summ = res.summary[:, :m].cpu().numpy().view(want.summary.dtype).reshape(A, m)
    fields = [f for f in want.summary.dtype.names if f != "peak" or res.peak is not None]
    for f in fields:
        g, w = got_summ[f], want.summary[f]
        bad |= (bits(g) != bits(w)) if g.dtype.kind == "f" else (g != w)
    bad |= (res.cnt[:, :m * K].cpu().numpy().reshape(A, m, K) != want.cnt.reshape(A, m, K)).any(2)
    per_node = [("fin", want.fin), ("avail", want.avail)] + ([("peak", want.peak)] if res.peak is not None else [])
    for f, w in per_node:
        g = getattr(res, f)[:, :m * K].cpu().numpy()
        bad |= (bits(g).reshape(A, m, K) != bits(w).reshape(A, m, K)).any(2)
    return {"instances": int(m), "schedules": int(A * m), "mismatches": int(bad.sum()),
            "compared": "node_of, seq_pos, summary record" + (" incl. peak" if res.peak is not None else "") +
                        ", cnt, fin/avail" + ("/peak" if res.peak is not None else "") + " bits vs oracle/dls_oracle.c"}


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle port on all host threads, each step a bounded sample of the GPU arm's own
    instances (same generator, seed and tie-break order, built on the host). Rank 0 only."""
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    wl = make_workload(args.workload)
    threads = os.cpu_count() or 1
    B = args.instances or wl.default_B
    m = cpu_sample_size(oracle, wl, 0, threads, 4.0, min(B, 4096))   # one step ~4 s of wall time
    sample = wl.host_batch(m, 0)
    for _ in range(args.warmup):
        oracle.schedule_batch(sample, wl.mask, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.schedule_batch(sample, wl.mask, nthreads=threads)
    dt = time.perf_counter() - t0
    val = 4 * sample.B * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl.describe(), "sample": f"first {sample.B} instances of rank 0's step, per step, on the host CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"first {sample.B} instances x 4 algorithms per step, oracle/dls_oracle.c (C port of the "
                                   "reference's algorithm; the CPython reference cannot run on this box)"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from dls_b200 import _native as N
    from dls_b200.batch import BatchResult, DeviceBatch, PipelinedScheduler, schedule_batch
    from dls_b200.dist import gather_summaries

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scheduling path has no CPU implementation")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = N.load()
    wl = make_workload(args.workload)
    B = args.instances or wl.default_B
    peak = not args.no_peak

    # this rank's shard of the sweep (random1k: built on the device by the library's generator kernels);
    # the host copy feeds the end-to-end leg (pinned host inputs)
    dev = wl.device_batch(B, rank, device)
    host = dev.to_host() if wl.name == "random1k" else dev.host
    n, K = int(host.task_off[1] - host.task_off[0]), host.K_max
    res = BatchResult(host, wl.mask, device, peak=peak, cached=False)
    ws = torch.empty(int(lib.dls_workspace_bytes(C.byref(dev.struct), wl.mask)), dtype=torch.uint8, device=device)
    gathered = None

    def step_resident():
        schedule_batch(dev, wl.mask, workspace=ws, result=res)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- resident-input timing -----------------------------------------------------------------------------------
    clocks = ClockSampler(local) if rank == 0 else None  # (running before the region starts: see ClockSampler.mark)
    for _ in range(args.warmup):
        step_resident()
    if world > 1:
        gathered = gather_summaries(res.summary, world)  # (warm-up of the collective as well)
    barrier()
    if clocks:
        clocks.mark()
    lib.dls_profile_enable(1)
    ms_k = (C.c_double * N.NUM_KERNELS)()
    cnt_k = (C.c_int64 * N.NUM_KERNELS)()
    lib.dls_profile_read(ms_k, cnt_k)  # drop anything recorded earlier
    for i in range(N.NUM_KERNELS):
        ms_k[i] = 0.0
        cnt_k[i] = 0
    launches0 = lib.dls_launch_count()
    gathered = None  # (the warm-up gather's 200 MB go back to torch's allocator: the timed one reuses them)
    ev0, ev1, evm = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    ev0.record()
    for _ in range(args.steps):
        step_resident()
    evm.record()
    if world > 1:  # the sweep's one collective: every rank's per-instance records (makespan, peak memory, counters)
        gathered = gather_summaries(res.summary, world)
    ev1.record()
    barrier()
    launches = int(lib.dls_launch_count() - launches0)
    N.check(lib.dls_profile_read(ms_k, cnt_k))
    lib.dls_profile_enable(0)
    # max over ranks of the whole region; beside it, where a multi-rank region's time went: this rank's steps (max and
    # min over ranks) and the collective including the wait for the slowest rank
    elapsed = torch.tensor([ev0.elapsed_time(ev1), ev0.elapsed_time(evm), -ev0.elapsed_time(evm), evm.elapsed_time(ev1)],
                           dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    ms_total = float(elapsed[0].item())
    region = {"steps_ms_max_over_ranks": float(elapsed[1].item()), "steps_ms_min_over_ranks": -float(elapsed[2].item()),
              "collective_and_wait_ms_max_over_ranks": float(elapsed[3].item())}

    # ---- end-to-end timing: host inputs in, host results out, every step -------------------------------------------
    if wl.name == "random1k":
        # Public API for host batches: PipelinedScheduler cuts the batch into chunks and overlaps the pinned
        # host->device copy, the kernels and the device->host copy of the results on three streams.
        pipe = PipelinedScheduler(host, wl.mask, device, chunks=24, nstreams=3, results=args.results)

        def step_e2e():
            pipe.run().synchronize()

        e2e_bytes = lambda: (pipe.h2d_bytes, pipe.d2h_bytes)
        e2e_summary = lambda: pipe.result().summary
        e2e_api = f"PipelinedScheduler(chunks=24, nstreams=3, results='{args.results}')"
    else:
        # one shared topology: a single schedule_batch call on a pinned host batch, every result array back
        from dls_b200.batch import pin_batch
        pinned = pin_batch(host)
        state = {}

        def step_e2e():
            d = DeviceBatch(host, device, pinned=pinned)
            r = schedule_batch(d, wl.mask, workspace=ws, peak=peak)
            state["h2d"] = d.h2d_bytes
            state["res"] = r.host() if args.results == "full" else r.summary_host()
            state["d2h"] = r.d2h_bytes(full=args.results == "full")

        e2e_bytes = lambda: (state["h2d"], state["d2h"])
        e2e_summary = lambda: state["res"].summary if args.results == "full" else state["res"]
        e2e_api = f"schedule_batch(pinned host DagBatch) + {'BatchResult.host()' if args.results == 'full' else 'summary_host()'}"
    step_e2e()
    barrier()
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    t_wall = (time.perf_counter() - t_wall0) * 1e3
    e2e_ms = torch.tensor([t_wall], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    clk = clocks.stop() if clocks else None
    h2d, d2h = e2e_bytes()

    # ---- the sweep as run_experiments consumes it (ref simulation.py:365-416,518-524): build the instances, schedule them,
    # average per (algorithm, group) — all on the device; only the group table crosses the bus ---------------------------
    sweep = None
    if wl.name == "random1k":
        from dls_b200.batch import aggregate
        from dls_b200.gen import random_dag_batch_device
        groups = torch.arange(B, dtype=torch.int32, device=device) % 3   # three "regime" groups, like the reference's tables
        res_sweep = BatchResult(host, wl.mask, device, peak=False, cached=False)   # (`res` keeps the resident run's results)

        def step_sweep(k):
            d = random_dag_batch_device(B, N_TASKS, K_NODES, REGIME, seed=SEED + 7919 * (k + 1) + rank, device=device,
                                        rank=tie_break_rank(N_TASKS))
            r = schedule_batch(d, wl.mask, workspace=ws, result=res_sweep)
            return aggregate(r, groups, 3)   # returns host arrays: the D2H of the [4, 3, 8] table is the synchronisation

        step_sweep(-1)
        barrier()
        t0 = time.perf_counter()
        for k in range(args.steps):
            table = step_sweep(k)
        barrier()
        sweep_ms = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(sweep_ms, op=dist.ReduceOp.MAX)
        sweep = {"value": 4 * B * world * args.steps / (float(sweep_ms.item()) * 1e-3), "unit": UNIT,
                 "ms_per_step": float(sweep_ms.item()) / args.steps, "d2h_bytes_per_step": int(4 * 3 * 8 * 8), "h2d_bytes_per_step": 0,
                 "what": "dls_generate_* -> dls_schedule_batch -> dls_aggregate -> group table to the host, every step a fresh batch",
                 "completion_rate_mean_last_step": [float(x) for x in table["completion_rate"].mean(1)]}

    # sanity: every instance scheduled, statuses clean (not part of any timed region)
    summ = res.summary_host()
    assert (summ["status"] == 0).all() and (summ["completed"] + summ["failed"] + summ["pending"] == n).all()
    e2e_summ = e2e_summary()
    for f in summ.dtype.names:
        if f != "peak" or wl.name != "random1k":  # (the pipelined leg does not ask for peaks)
            assert (e2e_summ[f] == summ[f]).all(), f"end-to-end results differ from the resident run in {f}"
    if world > 1:
        assert gathered is not None and gathered.shape[0] == world

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_sched = 4 * B * world * args.steps
    value = total_sched / (ms_total * 1e-3)
    e2e_value = total_sched / (float(e2e_ms.item()) * 1e-3)
    # dominant kernel of the timed region
    shares = {lib.dls_kernel_name(k).decode(): (ms_k[k], int(cnt_k[k])) for k in range(N.NUM_KERNELS) if cnt_k[k]}
    ksum = sum(v[0] for v in shares.values())
    dom = max(shares, key=lambda k: shares[k][0])
    dom_ms = shares[dom][0] / shares[dom][1]
    alg_of = lambda k: priority_bytes(host) if k.startswith("priority") else algorithmic_bytes(host)
    peak_bw, peak_src = measured_peak()
    achieved = alg_of(dom) / (dom_ms * 1e-3) / 1e9
    step_alg = 4 * algorithmic_bytes(host)
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak_bw, "unit": "GB/s", "frac": achieved / peak_bw,
        "traffic": ncu_traffic(dom, B) if wl.name == "random1k" else None, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": alg_of(dom), "kernel_ms": dom_ms, "share_of_step": shares[dom][0] / ksum,
        # the step as a whole: four schedules per instance, every kernel of the step, against the same peak
        "step": {"algorithmic_bytes": step_alg, "achieved": step_alg / (ms_total / args.steps * 1e-3) / 1e9,
                 "frac": step_alg / (ms_total / args.steps * 1e-3) / 1e9 / peak_bw},
        # per kernel: share of the step and, for kernels doing real work, algorithmic GB/s over the measured peak
        # (sched_* kernels of the general path only pick up routed instances here: a few tens of microseconds;
        # param_class and unpermute touch a fraction of a schedule's bytes: no figure)
        "kernels": {k: {"ms_per_launch": v[0] / v[1], "share": v[0] / ksum,
                        "frac": ((alg_of(k) / (v[0] / v[1] * 1e-3) / 1e9) / peak_bw)
                        if v[0] / ksum > 0.01 and k.startswith(("priority", "sched_", "stream_build")) else None}
                    for k, v in shares.items()},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl.describe(), "instances_per_gpu": B, "schedules_per_step": 4 * B * world,
                   "peak_memory_tracked": peak,
                   "l2": f"inputs {host.input_bytes() / 1e6:.0f} MB per GPU > 126 MB L2 (no flush needed)"
                         if host.input_bytes() > 126e6 else
                         f"inputs {host.input_bytes() / 1e6:.0f} MB + {ws.numel() / 1e6:.0f} MB of scratch rewritten per step",
                   "collective": "one all_gather of the 48-byte per-instance records after the last step" if world > 1 else "none (1 GPU)"},
        "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": float(e2e_ms.item()) / args.steps, "api": e2e_api},
        "gpu_launches": launches,
        "clocks": clk, "timed_region": region,
    }
    if sweep is not None:
        line["sweep"] = sweep
    if world == 1:
        from oracle import oracle
        oracle.build()
        threads = os.cpu_count() or 1
        rec, sample, want = cpu_baseline(oracle, wl, rank, threads, args.cpu_seconds, min(B, 8192))
        line["cpu_baseline"] = rec
        line["parity"] = parity_check(res, want, sample.B, n, K)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if line.get("parity", {}).get("mismatches", 0):
        raise SystemExit("parity gate: the GPU results differ from the oracle's; the timings above do not count")


if __name__ == "__main__":
    main()
