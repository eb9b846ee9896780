#!/usr/bin/env python
"""Wall and device time of building one 131,072-instance batch with the device generator (dls_generate_*), four times.
    python tools/gen_time.py"""
import sys, time, os
sys.path.insert(0, os.getcwd())
import torch
from dls_b200.gen import random_dag_batch_device
for it in range(4):
    torch.cuda.synchronize(); t0=time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d = random_dag_batch_device(131072, 1000, 8, 0.9, seed=1000+it)
    e1.record(); torch.cuda.synchronize()
    print(f"wall {1e3*(time.perf_counter()-t0):.2f} ms, device span {e0.elapsed_time(e1):.2f} ms")
    del d
