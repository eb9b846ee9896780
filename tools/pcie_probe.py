#!/usr/bin/env python
"""Host<->device copy rates with pinned memory: the roofline of bench.py's end-to-end leg.

    python tools/pcie_probe.py                                      one GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_probe.py
                                                                    N ranks copying AT THE SAME TIME (one box: they share the
                                                                    host's memory controllers and PCIe root complexes)
"""
import os
import time

import torch
import torch.distributed as dist

rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
GB = 1 << 30
x = torch.empty(GB, dtype=torch.uint8).pin_memory()
h = torch.empty(GB, dtype=torch.uint8).pin_memory()
d = torch.empty(GB, dtype=torch.uint8, device="cuda")
d2 = torch.empty(GB, dtype=torch.uint8, device="cuda")


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def timed(fn, reps):
    fn()
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    barrier()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item())


s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    with torch.cuda.stream(s1):
        d.copy_(x, non_blocking=True)
    with torch.cuda.stream(s2):
        h.copy_(d2, non_blocking=True)


rows = []
for name, fn, nbytes in (("H2D 1 GiB", lambda: d.copy_(x, non_blocking=True), GB),
                         ("H2D 16 MiB x64", lambda: [d[k << 24:(k + 1) << 24].copy_(x[k << 24:(k + 1) << 24], non_blocking=True) for k in range(64)], GB),
                         ("D2H 1 GiB", lambda: h.copy_(d, non_blocking=True), GB),
                         ("H2D + D2H at once, 1 GiB each way", both, GB)):
    dt = timed(fn, 4)
    rows.append(f"{name:36s} per rank {4 * nbytes / dt / 1e9:6.1f} GB/s" + (" each way" if "at once" in name else "") +
                f"   all {world} rank(s) {world * 4 * nbytes / dt / 1e9:7.1f} GB/s")
if rank == 0:
    print(f"# {world} rank(s) copying at the same time, pinned host memory, max time over ranks")
    print("\n".join(rows))
if world > 1:
    dist.destroy_process_group()
