"""Array-level DAG builders: emit the CSR batch of include/dls.h directly, without Task objects.

`random_dag_batch` follows the distribution of the reference's DAGGenerator.generate_random_dag
(ref simulation.py:91-114: deps = sample(previous tasks, randint(0, min(max_deps, i))), 1-2 private params,
mem ~ U(0.1, 0.5), time ~ U(0.05, 0.15)) and node sets follow create_nodes_with_memory_regime /
calculate_total_memory_needed (ref simulation.py:161-214). The random stream is numpy's, not CPython's
Mersenne Twister calls, so instances are distributed like the reference's but are not the same instances;
parity is always checked against an oracle run on these very arrays.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from .lowering import DagBatch, LoweredDag, set_order


def task_rank(n: int, fmt: str = "task_{}") -> np.ndarray:
    """rank[] for tasks named fmt.format(0..n-1) added in index order (this process's hash seed)."""
    names = [fmt.format(i) for i in range(n)]
    pos = {nm: r for r, nm in enumerate(set_order(names))}
    return np.array([pos[nm] for nm in names], np.uint16)


def _param_name_order(n: int) -> np.ndarray:
    """Position of 'param_{i}_{j}' (slot 2*i+j) in the lexicographic order of all 2n possible names."""
    names = [f"param_{i}_{j}" for i in range(n) for j in range(2)]
    order = np.argsort(np.array(names))  # numpy sorts unicode strings by code point, like Python
    pos = np.empty(2 * n, np.int64)
    pos[order] = np.arange(2 * n)
    return pos


def node_sets(total: np.ndarray, regime: float, K: int, rng: np.random.Generator):
    """create_nodes_with_memory_regime (ref simulation.py:161-192) for a vector of totals -> cap[B,K], spd[B,K]."""
    B = len(total)
    available = total * regime
    if K == 2:
        cap = np.stack([available * 0.6, available * 0.4], 1)
        spd = np.tile(np.array([1.2, 1.0]), (B, 1))
    elif K == 4:
        cap = np.stack([available * f for f in (0.35, 0.25, 0.25, 0.15)], 1)
        spd = np.tile(np.array([1.2, 1.0, 1.0, 0.8]), (B, 1))
    else:
        cap = np.repeat((available / K)[:, None], K, 1)
        spd = rng.uniform(0.7, 1.3, (B, K))
    return np.ascontiguousarray(cap, np.float64), np.ascontiguousarray(spd, np.float64)


def random_dag_batch(B: int, n: int, K: int = 8, regime: float = 0.9, seed: int = 0, max_deps: int = 3,
                     rank: Optional[np.ndarray] = None) -> DagBatch:
    """B independent random DAGs of n tasks each (config 5 shape), one topology per instance."""
    rng = np.random.default_rng(seed)
    idx = np.arange(n)
    # dependencies: an ordered sample without replacement of nd in 0..min(max_deps, i) earlier tasks
    hi = np.minimum(max_deps, idx)
    nd = (rng.random((B, n)) * (hi + 1)).astype(np.int64)
    nd = np.minimum(nd, hi)
    picks = []
    for j in range(max_deps):
        span = np.maximum(idx - j, 1)
        r = (rng.random((B, n)) * span).astype(np.int64)
        r = np.minimum(r, span - 1)
        prev = np.sort(np.stack(picks, 0), 0) if picks else None
        if prev is not None:
            for q in range(prev.shape[0]):
                r = r + (r >= prev[q])
        picks.append(r)
    picks = np.stack(picks, 2)  # [B, n, max_deps]
    keep = np.arange(max_deps)[None, None, :] < nd[:, :, None]
    pred_idx = picks[keep].astype(np.uint16)  # row-major: instance, task, slot -> CSR order
    pred_ptr = np.zeros((B, n + 1), np.uint32)
    pred_ptr[:, 1:] = np.cumsum(nd, 1)
    e_per = pred_ptr[:, -1].astype(np.int64)
    # params: 1-2 private params per task, ids = lexicographic rank of 'param_{i}_{j}' among those present
    npar = rng.integers(1, 3, (B, n))
    present = np.zeros((B, n, 2), bool)
    present[:, :, 0] = True
    present[:, :, 1] = npar == 2
    pos = _param_name_order(n)  # slot -> position in name order
    by_name = np.zeros((B, 2 * n), bool)
    by_name[:, pos] = present.reshape(B, 2 * n)
    ids_by_name = np.cumsum(by_name, 1) - 1
    ids = ids_by_name[:, pos].reshape(B, n, 2)
    lo = np.where(present[:, :, 1], np.minimum(ids[:, :, 0], ids[:, :, 1]), ids[:, :, 0])
    hi2 = np.maximum(ids[:, :, 0], ids[:, :, 1])
    pair = np.stack([lo, hi2], 2)
    par_idx = pair[present].astype(np.uint16)
    par_ptr = np.zeros((B, n + 1), np.uint32)
    par_ptr[:, 1:] = np.cumsum(npar, 1)
    r_per = par_ptr[:, -1].astype(np.int64)
    mem = rng.uniform(0.1, 0.5, (B, n))
    dur = rng.uniform(0.05, 0.15, (B, n))
    # calculate_total_memory_needed (ref simulation.py:194-214)
    total = (mem + npar * 0.5).max(1) + r_per * 0.5
    cap, spd = node_sets(total, regime, K, rng)
    if rank is None:
        rank = task_rank(n)
    off = lambda per: np.concatenate([[0], np.cumsum(per)]).astype(np.int64)
    ones = np.full(B, n, np.int64)
    return DagBatch(
        B=B, T=B, n_max=n, K_max=K, P_max=int(r_per.max()), E_max=int(e_per.max()), R_max=int(r_per.max()), flags=0,
        task_off=off(ones), pred_off=off(e_per), par_off=off(r_per), pid_off=off(r_per),
        pred_ptr=np.ascontiguousarray(pred_ptr.reshape(-1)), pred_idx=pred_idx,
        par_ptr=np.ascontiguousarray(par_ptr.reshape(-1)), par_idx=par_idx,
        rank=np.tile(rank.astype(np.uint16), B), mem=mem.reshape(-1), dur=dur.reshape(-1),
        node_off=off(np.full(B, K, np.int64)), cap=cap.reshape(-1), spd=spd.reshape(-1), dags=None)


def perturbed_durations(dag: LoweredDag, B: int, eps: float = 0.1, seed: int = 0) -> np.ndarray:
    """Config 3: dur_i * (1 + e), e ~ U(-eps, eps), one row per instance."""
    rng = np.random.default_rng(seed)
    return dag.dur[None, :] * (1.0 + rng.uniform(-eps, eps, (B, dag.n)))
