import sys, time
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dls_b200 import _native as N
from dls_b200.batch import PipelinedScheduler
from dls_b200.gen import random_dag_batch_device
B = 131072
dev = random_dag_batch_device(B, 1000, 8, 0.9, seed=1000)
host = dev.to_host()
del dev
torch.cuda.empty_cache()
for chunks, ns in ((16, 3), (32, 3), (32, 4), (64, 4), (24, 3)):
    pipe = PipelinedScheduler(host, N.MASK_ALL, chunks=chunks, nstreams=ns)
    pipe.run().synchronize()
    ts = []
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        pipe.run().synchronize()
        ts.append(time.perf_counter() - t0)
    print(f"chunks={chunks} streams={ns}: {min(ts)*1e3:.1f} ms  -> {4*B/min(ts)/1e6:.2f} M/s  h2d {pipe.h2d_bytes/1e9:.2f} GB")
    del pipe
    torch.cuda.empty_cache()
