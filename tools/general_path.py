#!/usr/bin/env python
"""Rate of the GENERAL (warp-per-DAG) kernels, the path taken by K > 8 nodes, more than 32 shared params, MRU with any
shared param, or the `cached` / `tfin` outputs: (a) the bench's random 1k-task DAGs with the thread-per-DAG path
switched off (same instances, same results), (b) random 1k-task DAGs on 16 nodes, which only this path takes.
    python tools/general_path.py [instances]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from dls_b200 import _native as N
from dls_b200.batch import DeviceBatch, schedule_batch
from dls_b200.gen import random_dag_batch, random_dag_batch_device

lib = N.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384


def timed(dev, label, steps=3):
    res = schedule_batch(dev, 0xF)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        schedule_batch(dev, 0xF, workspace=res.workspace, result=res)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    s = res.summary_host()
    print(f"{label}: {dev.host.B} instances x 4 algorithms in {ms:.2f} ms = {4 * dev.host.B / ms * 1e3 / 1e6:.2f} M schedules/s "
          f"(completed per algorithm {s['completed'].mean(1).round(1).tolist()}, status any {bool(s['status'].any())})")
    return res


rank = np.random.default_rng(12345).permutation(1000).astype(np.uint16)
dev = random_dag_batch_device(B, 1000, 8, 0.9, seed=1000, rank=rank)
fast = timed(dev, "K=8, thread-per-DAG path")
lib.dls_set_stream_path(0)
slow = timed(dev, "K=8, general path (thread-per-DAG path switched off)")
lib.dls_set_stream_path(1)
same = all(torch.equal(getattr(fast, f), getattr(slow, f)) for f in ("node_of", "seq_pos", "fin", "cnt", "avail"))
print("results of the two paths identical:", same)
host = random_dag_batch(min(B, 4096), 1000, 16, 0.9, seed=7)
timed(DeviceBatch(host, torch.device("cuda", 0)), "K=16 (general path only)")
