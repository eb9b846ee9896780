#!/usr/bin/env python
"""Per-kernel CUDA-event times of the scheduling step on device-generated 1k-task DAGs (no e2e leg, no CPU baseline):
the quick loop used between kernel changes.  python tools/kernel_times.py [instances]"""
import os
import sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from dls_b200 import _native as N
from dls_b200.batch import schedule_batch
from dls_b200.gen import random_dag_batch_device
lib = N.load()
if os.environ.get("DLS_L2_FETCH"):   # cudaLimitMaxL2FetchGranularity (0x05): 32, 64 or 128 bytes
    cu = C.CDLL("libcuda.so.1")
    torch.cuda.init(); torch.zeros(1, device="cuda")
    old = C.c_size_t(0)
    cu.cuCtxGetLimit(C.byref(old), 5)
    rc = cu.cuCtxSetLimit(5, C.c_size_t(int(os.environ["DLS_L2_FETCH"])))
    new = C.c_size_t(0)
    cu.cuCtxGetLimit(C.byref(new), 5)
    print("L2 fetch granularity", old.value, "->", new.value, "rc", rc)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
PEAK = bool(int(os.environ.get("DLS_PEAK", "0")))
rank = None
if os.environ.get("DLS_RANK_PERM"):
    rank = np.random.default_rng(12345).permutation(1000).astype(np.uint16)
dev = random_dag_batch_device(B, 1000, 8, 0.9, seed=1000, rank=rank)
res = schedule_batch(dev, 0xF, peak=PEAK)
torch.cuda.synchronize()
lib.dls_profile_enable(1)
ms = (C.c_double * N.NUM_KERNELS)(); cnt = (C.c_int64 * N.NUM_KERNELS)()
lib.dls_profile_read(ms, cnt)
for i in range(N.NUM_KERNELS): ms[i]=0; cnt[i]=0
for it in range(3):
    res = schedule_batch(dev, 0xF, workspace=res.workspace, result=res)
torch.cuda.synchronize()
lib.dls_profile_read(ms, cnt)
for k in range(N.NUM_KERNELS):
    if cnt[k] and ms[k]/cnt[k] > 0.02:
        print(f"{lib.dls_kernel_name(k).decode():34s} avg ms {ms[k]/cnt[k]:9.3f}  -> inst/s {B/(ms[k]/cnt[k])*1e3:14.0f}")
tot = sum(ms) / 3
print("total ms/step", tot, "schedules/s", 4 * B / tot * 1e3)
s = res.summary_host()
print("completed mean per algo", s["completed"].mean(1), "status any", s["status"].any())
