"""N>1 host logic on CPU: world_size-2 gloo processes shard a batch, each computes its shard's records (with the
oracle standing in for the GPU), all-gather them with the product's gather code, and both ranks must see the
records of a single-process run in global instance order."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from common import ROOT
from dls_b200 import _native as N
from dls_b200.dist import gather_summaries, merge_gathered, shard_bounds
from dls_b200.gen import random_dag_batch


def test_shard_bounds_partition():
    for B in (1, 7, 8, 1000, 1_000_000):
        for world in (1, 2, 4, 8):
            spans = [shard_bounds(B, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_shard_is_selfcontained(oracle_mod):
    batch = random_dag_batch(11, 40, 8, 0.8, seed=2)
    full = oracle_mod.schedule_batch(batch, N.MASK_ALL)
    for world in (2, 3):
        for r in range(world):
            lo, hi = shard_bounds(batch.B, world, r)
            part = oracle_mod.schedule_batch(batch.shard(lo, hi), N.MASK_ALL)
            np.testing.assert_array_equal(part.summary, full.summary[:, lo:hi])
            t0, t1 = int(batch.task_off[lo]), int(batch.task_off[hi])
            np.testing.assert_array_equal(part.node_of, full.node_of[:, t0:t1])
            np.testing.assert_array_equal(part.seq_pos, full.seq_pos[:, t0:t1])


def _batch(B):
    # the tie-break order pi is an INPUT (it depends on the hash seed of the process that built the task set):
    # fix it explicitly so parent and spawned workers schedule the same instances
    return random_dag_batch(B, 30, 4, 0.7, seed=9, rank=np.arange(30, dtype=np.uint16)[::-1].copy())


def _worker(rank, world, port, sizes, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        out = []
        for B in sizes:  # several gathers in ONE process group: even shards, then uneven ones, then even again
            batch = _batch(B)
            lo, hi = shard_bounds(B, world, rank)
            local = oracle.schedule_batch(batch.shard(lo, hi), N.MASK_ALL)
            rec = torch.from_numpy(np.ascontiguousarray(local.summary).view(np.uint8).reshape(4, hi - lo, N.SUMMARY_BYTES))
            out.append(merge_gathered(gather_summaries(rec, world), B, world).tobytes())
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("sizes", [(8,), (9,), (8, 9, 8, 11)])
def test_two_rank_gather_matches_single_process(oracle_mod, sizes):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sizes, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for k, B in enumerate(sizes):
        want = oracle_mod.schedule_batch(_batch(B), N.MASK_ALL).summary
        for r in range(2):
            np.testing.assert_array_equal(np.frombuffer(got[r][k], dtype=want.dtype).reshape(want.shape), want)
