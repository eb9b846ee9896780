#!/usr/bin/env python
"""Whole-step time of the scheduling call on device-generated 1k-task DAGs, per-kernel events OFF (so that launches
which overlap their predecessor's tail really do).  python tools/step_time.py [instances]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dls_b200 import _native as N
from dls_b200.batch import schedule_batch
from dls_b200.gen import random_dag_batch_device
lib = N.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
dev = random_dag_batch_device(B, 1000, 8, 0.9, seed=1000)
res = schedule_batch(dev, 0xF)
for _ in range(3):
    res = schedule_batch(dev, 0xF, workspace=res.workspace, result=res)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 5
e0.record()
for _ in range(K):
    res = schedule_batch(dev, 0xF, workspace=res.workspace, result=res)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print(f"step {ms:.3f} ms  -> {4 * B / ms * 1e3 / 1e6:.2f} M schedules/s")
