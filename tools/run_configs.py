#!/usr/bin/env python
"""Run the five BASELINE.json configs on the GPU (timing + an oracle spot check each) and print one JSON object.

    python tools/run_configs.py > profiles/<round>_configs.json
"""
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch

import golden_io
from common import assert_results_equal
from dls_b200 import _native as N
from dls_b200.batch import DeviceBatch, schedule_batch
from dls_b200.gen import perturbed_durations, random_dag_batch_device
from dls_b200.lowering import DagBatch
from dls_b200.schedulers import SCHEDULERS, Node, Task
from dls_b200.simulation import ImprovedSchedulerEvaluator
from oracle import oracle


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps, r


out = {}
# config 1: built-in 4-task example through the reference-shaped classes
tasks = [("t1", [], {"p1"}), ("t2", ["t1"], {"p2"}), ("t3", ["t1"], {"p3"}), ("t4", ["t2", "t3"], {"p1", "p2"})]
c1 = {}
for name, cls in SCHEDULERS.items():
    s = cls([Node("n1", 3.0), Node("n2", 2.5)])
    for tid, deps, pars in tasks:
        s.add_task(Task(tid, 1.0, 0.1, list(deps), set(pars)))
    sched = s.schedule()
    c1[name] = {"completed": len(s.completed_tasks), "failed": len(s.failed_tasks), "schedule": sched}
out["config1_example"] = c1

# config 2: the default sweep, batched
random.seed(0)
ev = ImprovedSchedulerEvaluator(SCHEDULERS)
inst = ev.sweep_instances(3)
dt, res = timed(lambda: ev.run_batch(inst), reps=3)
comp = {}
for r in res:
    comp.setdefault((r.scheduler_name, r.memory_regime), []).append(r.completion_rate)
out["config2_sweep"] = {"schedules": len(res), "seconds_per_sweep_incl_lowering": dt,
                        "schedules_per_s": len(res) / dt,
                        "avg_completion": {f"{k[0]}@{k[1]}": float(np.mean(v)) for k, v in sorted(comp.items())}}

# config 3: GPT-2 DAG on 8 capped devices, 100k perturbed-duration instances, shared topology
base = [c for c in golden_io.load("cfg3_gpt2.npz") if c.label == "gpt2|8laptops"][0]
B3 = 100_000
dur = perturbed_durations(base.dag, B3, 0.1, seed=0)
b3 = DagBatch.shared_topology(base.dag, base.cap, base.spd, B3, dur=dur)
d3 = DeviceBatch(b3, "cuda:0")
dt, r3 = timed(lambda: schedule_batch(d3, N.MASK_ALL))
sub = b3.shard(0, 2000)
assert_results_equal(schedule_batch(sub, N.MASK_ALL).host(), oracle.schedule_batch(sub, N.MASK_ALL, nthreads=16), "cfg3")
s3 = r3.summary_host()
out["config3_gpt2_100k"] = {"instances": B3, "seconds": dt, "schedules_per_s": 4 * B3 / dt,
                            "completed_mean": s3["completed"].mean(1).tolist(), "makespan_mean": s3["makespan"].mean(1).tolist(),
                            "parity_subset": "2000 instances x 4 algorithms bit-exact vs oracle"}

# config 4: 48-layer x 205-head DAG (9,986 tasks), K=8, regime 0.9, durations perturbed per instance
big = [c for c in golden_io.load("cfg4_llm10k.npz") if c.label.endswith("r0.9")][0]
B4 = 512
b4 = DagBatch.shared_topology(big.dag, big.cap, big.spd, B4, dur=perturbed_durations(big.dag, B4, 0.1, seed=1))
d4 = DeviceBatch(b4, "cuda:0")
dt, r4 = timed(lambda: schedule_batch(d4, 0b0111), reps=2)
dtm, r4m = timed(lambda: schedule_batch(d4, 0b1000), reps=1)
sub = b4.shard(0, 4)
assert_results_equal(schedule_batch(sub, N.MASK_ALL).host(), oracle.schedule_batch(sub, N.MASK_ALL, nthreads=4), "cfg4")
out["config4_llm10k"] = {"instances": B4, "tasks": big.n, "seconds_dfs_greedy_critical": dt,
                         "schedules_per_s_dfs_greedy_critical": 3 * B4 / dt, "seconds_mru": dtm,
                         "schedules_per_s_mru": B4 / dtm, "completed_mean_mru": float(r4m.summary_host()["completed"].mean()),
                         "parity_subset": "4 instances x 4 algorithms bit-exact vs oracle"}

# config 5 (one GPU's share of the 1M-instance sweep): see bench.py; here a parity spot check on device-generated DAGs
d5 = random_dag_batch_device(2048, 1000, 8, 0.9, seed=77)
h5 = d5.to_host()
assert_results_equal(schedule_batch(d5, N.MASK_ALL).host(), oracle.schedule_batch(h5, N.MASK_ALL, nthreads=16), "cfg5")
out["config5_random1k"] = {"parity_subset": "2048 device-generated instances x 4 algorithms bit-exact vs oracle",
                           "throughput": "bench.py (131,072 instances per GPU per step)"}
print(json.dumps(out, indent=1))
