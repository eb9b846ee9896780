#!/usr/bin/env python
"""bench.py — DAG schedules/sec on 1k-task DAGs (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--instances B_per_gpu] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A step = one pass of the hot path (priority kernels + the four list-scheduling kernels) over one batch of
synthetic random 1k-task DAG instances per GPU (config-5 shape: n=1000, K=8 nodes, memory regime 0.9, all four
algorithms). Instances are sharded across ranks with no data-path collective; the only collective is the
all-gather of the fixed-size per-instance result records, done inside every step when N>1.

value  = schedules/s with inputs resident in HBM (CUDA events, max over ranks).
e2e    = the same through the public API with HOST inputs (PipelinedScheduler): pinned host->device copy of the
         batch, the C-ABI calls, device->host copy of every result array, each step; chunks overlap on 3 streams.
roofline = dominant kernel (largest share of the timed region, CUDA events recorded by the library on the launch
         stream): algorithmic bytes per launch (29n + 4(E+R) + 28K + 32 per schedule) / duration,
         against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
cpu_baseline / --impl reference = the CPU oracle (C port of the reference's algorithm, oracle/dls_oracle.c) on all
         host threads over a bounded sample of the same workload. The reference itself is CPython and does not
         exist on the GPU box; its measured single-core numbers are in BASELINE.md section 2.
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
METRIC = "DAG schedules/sec (1k-task DAGs)"
UNIT = "schedules/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--instances", type=int, default=131072, help="DAG instances per GPU per step")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


def algorithmic_bytes(batch, per_schedule=True):
    """B(n,E,R,K) = 29n + 4(E+R) + 28K + 32 bytes per (instance, algorithm), summed over the batch."""
    return 29 * batch.nt_inst + 4 * (batch.e_total + batch.r_total) + 28 * batch.nk_inst + 32 * batch.B


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
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

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
        self.f.seek(0)
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


def cpu_sample(oracle, batch, threads, target_s):
    """Oracle (C port) on all host threads over a bounded sample; returns (schedules/s, sample description)."""
    from dls_b200 import _native as N
    pilot = batch.shard(0, min(batch.B, max(threads, 8)))
    t0 = time.perf_counter()
    oracle.schedule_batch(pilot, N.MASK_ALL, nthreads=threads)
    dt = max(time.perf_counter() - t0, 1e-4)
    per_inst = dt / pilot.B
    m = int(min(batch.B, max(pilot.B, target_s / per_inst)))
    m = max(threads, m // threads * threads)
    sample = batch.shard(0, min(m, batch.B))
    t0 = time.perf_counter()
    oracle.schedule_batch(sample, N.MASK_ALL, nthreads=threads)
    dt = time.perf_counter() - t0
    return 4 * sample.B / dt, f"{sample.B} of the step's instances x 4 algorithms, {dt:.1f} s on {threads} threads", sample.B, dt


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle port, all host threads, bounded sample per step. Rank 0 only."""
    if rank != 0:
        return
    from dls_b200.gen import random_dag_batch
    from oracle import oracle
    from dls_b200 import _native as N
    oracle.build()
    threads = os.cpu_count() or 1
    batch = random_dag_batch(max(threads * 8, 256), N_TASKS, K_NODES, REGIME, seed=0)
    # size one step to ~4 s of wall time
    t0 = time.perf_counter()
    oracle.schedule_batch(batch.shard(0, threads), N.MASK_ALL, nthreads=threads)
    per_inst = (time.perf_counter() - t0) / threads
    m = int(min(batch.B, max(threads, 4.0 / max(per_inst, 1e-6))))
    m = max(threads, m // threads * threads)
    sample = batch.shard(0, m)
    for _ in range(args.warmup):
        oracle.schedule_batch(sample, N.MASK_ALL, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle.schedule_batch(sample, N.MASK_ALL, nthreads=threads)
    dt = time.perf_counter() - t0
    val = 4 * sample.B * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"random-1k (config-5 shape): n={N_TASKS}, K={K_NODES}, regime {REGIME}, DFS+Greedy+Critical+MRU; "
                               f"bounded sample of {sample.B} instances per step on the host CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample.B} instances x 4 algorithms per step, oracle/dls_oracle.c (C port of the "
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
    from dls_b200.batch import BatchResult, DeviceBatch, pin_batch, schedule_batch
    from dls_b200.dist import gather_summaries
    from dls_b200.gen import random_dag_batch_device

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scheduling path has no CPU implementation")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = N.load()

    B = args.instances
    # this rank's shard of the sweep, built on the device by the library's generator kernels;
    # the host copy feeds the end-to-end leg (pinned host inputs) and the CPU baseline sample
    dev = random_dag_batch_device(B, N_TASKS, K_NODES, REGIME, seed=1000 + rank, device=device)
    host = dev.to_host()
    res = BatchResult(host, N.MASK_ALL, device, peak=False, cached=False)
    ws = torch.empty(int(lib.dls_workspace_bytes(C.byref(dev.struct), N.MASK_ALL)), dtype=torch.uint8, device=device)
    gathered = None

    def step_resident():
        nonlocal gathered
        schedule_batch(dev, N.MASK_ALL, workspace=ws, result=res)
        if world > 1:
            gathered = gather_summaries(res.summary, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- resident-input timing -----------------------------------------------------------------------------------
    for _ in range(args.warmup):
        step_resident()
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    lib.dls_profile_enable(1)
    ms_k = (C.c_double * N.NUM_KERNELS)()
    cnt_k = (C.c_int64 * N.NUM_KERNELS)()
    lib.dls_profile_read(ms_k, cnt_k)  # drop anything recorded earlier
    for i in range(N.NUM_KERNELS):
        ms_k[i] = 0.0
        cnt_k[i] = 0
    launches0 = lib.dls_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step_resident()
    ev1.record()
    barrier()
    launches = int(lib.dls_launch_count() - launches0)
    N.check(lib.dls_profile_read(ms_k, cnt_k))
    lib.dls_profile_enable(0)
    elapsed = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    ms_total = float(elapsed.item())

    # ---- end-to-end timing: host inputs in, host results out, every step -------------------------------------------
    # Public API for host batches: PipelinedScheduler cuts the batch into chunks and overlaps the pinned
    # host->device copy, the kernels and the device->host copy of the results on three streams.
    from dls_b200.batch import PipelinedScheduler
    pipe = PipelinedScheduler(host, N.MASK_ALL, device, chunks=24, nstreams=3)

    def step_e2e():
        nonlocal gathered
        pipe.run().synchronize()
        if world > 1:
            gathered = gather_summaries(pipe.out["summary"].to(device, non_blocking=True), world)

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
    h2d = pipe.h2d_bytes
    d2h = pipe.d2h_bytes
    e2e_summ = pipe.result().summary

    # sanity: every instance scheduled, statuses clean (not part of any timed region)
    summ = res.summary_host()
    assert (summ["status"] == 0).all() and (summ["completed"] + summ["failed"] + summ["pending"] == N_TASKS).all()
    assert (e2e_summ == summ).all(), "pipelined end-to-end results differ from the resident run"
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
    alg = priority_bytes(host) if dom.startswith("priority") else algorithmic_bytes(host)
    peak, peak_src = measured_peak()
    achieved = alg / (dom_ms * 1e-3) / 1e9
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": ncu_traffic(dom, B), "peak_source": peak_src, "algorithmic_bytes_per_launch": alg,
        "kernel_ms": dom_ms, "share_of_step": shares[dom][0] / ksum,
        # the step as a whole: four schedules per instance, every kernel of the step, against the same peak
        "step": {"algorithmic_bytes": 4 * algorithmic_bytes(host),
                 "achieved": 4 * algorithmic_bytes(host) / (ms_total / args.steps * 1e-3) / 1e9,
                 "frac": 4 * algorithmic_bytes(host) / (ms_total / args.steps * 1e-3) / 1e9 / peak},
        # per kernel: share of the step and, for kernels doing real work, algorithmic GB/s over the measured peak
        # (sched_* kernels of the general path only pick up routed instances here: a few tens of microseconds;
        # param_class and unpermute touch a fraction of a schedule's bytes: no figure)
        "kernels": {k: {"ms_per_launch": v[0] / v[1], "share": v[0] / ksum,
                        "frac": (((priority_bytes(host) if k.startswith("priority") else algorithmic_bytes(host))
                                  / (v[0] / v[1] * 1e-3) / 1e9) / peak)
                        if v[0] / ksum > 0.01 and k.startswith(("priority", "sched_", "stream_build")) else None}
                    for k, v in shares.items()},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"random-1k (config-5 shape): {B} DAG instances per GPU per step, n={N_TASKS} tasks, "
                               f"K={K_NODES} nodes, memory regime {REGIME}, algorithms DFS+Greedy+Critical+MRU_spec",
                   "instances_per_gpu": B, "schedules_per_step": 4 * B * world,
                   "l2": f"inputs {host.input_bytes() / 1e6:.0f} MB per GPU > 126 MB L2 (no flush needed)",
                   "collective": "all_gather of 40-byte per-instance records each step" if world > 1 else "none (1 GPU)"},
        "roofline": roofline,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": float(e2e_ms.item()) / args.steps},
        "gpu_launches": launches,
        "clocks": clk,
    }
    if world == 1:
        from oracle import oracle
        oracle.build()
        threads = os.cpu_count() or 1
        v, desc, _, _ = cpu_sample(oracle, host, threads, args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
