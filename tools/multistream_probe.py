#!/usr/bin/env python
"""Does running the step as S concurrent calls on S streams (each over B/S instances) beat one call over B?
python tools/multistream_probe.py [instances]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dls_b200 import _native as N
from dls_b200.batch import schedule_batch
from dls_b200.gen import random_dag_batch_device
B = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
STAGGER = os.environ.get("DLS_STAGGER")  # stream s starts s/S of a step late, so that different kernels overlap
for S in (1, 2, 4):
    devs = [random_dag_batch_device(B // S, 1000, 8, 0.9, seed=1000 + s) for s in range(S)]
    streams = [torch.cuda.Stream() for _ in range(S)]
    res = []
    for s in range(S):
        with torch.cuda.stream(streams[s]):
            res.append(schedule_batch(devs[s], 0xF))
    torch.cuda.synchronize()

    def step():
        for s in range(S):
            with torch.cuda.stream(streams[s]):
                schedule_batch(devs[s], 0xF, workspace=res[s].workspace, result=res[s])
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 10
    e0.record()
    for s, st in enumerate(streams):
        st.wait_event(e0)
        if STAGGER and s:
            with torch.cuda.stream(st):
                torch.cuda._sleep(int(float(STAGGER) * 1.9e6 * s / S))  # DLS_STAGGER = ms of a step, in cycles at 1.9 GHz
    for _ in range(K):
        step()
    for st in streams:
        torch.cuda.current_stream().wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    print(f"{S} stream(s) x {B // S} instances: {ms:.3f} ms per step -> {4 * B / ms * 1e3 / 1e6:.2f} M schedules/s")
    del devs, res
    torch.cuda.empty_cache()
