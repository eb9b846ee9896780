#!/usr/bin/env python
"""Time the sweep's one collective on its own: dist.gather_summaries of [4, 131072, 48]-byte records per rank, ten
times, CUDA events on every rank, max over ranks printed by rank 0 (what it costs inside bench.py's timed region).
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/gather_probe.py [instances]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from dls_b200.dist import gather_summaries

B = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rec = torch.zeros((4, B, 48), dtype=torch.uint8, device=dev)
times = []
for it in range(12):
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = gather_summaries(rec, world)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    times.append(float(t.item()))
    del out
if rank == 0:
    mb = rec.numel() * world / 1e6
    print(f"ranks {world}  {mb:.0f} MB gathered per rank  ms per call (max over ranks): " + " ".join(f"{t:.2f}" for t in times))
dist.destroy_process_group()
