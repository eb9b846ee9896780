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


_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on uint64 arrays (wrap-around arithmetic), as csrc/dls_gen.cu::mix64."""
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def _u01(seed: int, inst: np.ndarray, task: np.ndarray, draw: int) -> np.ndarray:
    """uniform double in [0, 1) for (instance, task, draw): csrc/dls_gen.cu::u01."""
    with np.errstate(over="ignore"):
        key = np.uint64(seed) ^ ((inst.astype(np.uint64) << np.uint64(32)) | task.astype(np.uint64))
        h = _mix64(_mix64(key) + np.uint64(draw))
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def random_dag_batch_counter(B: int, n: int, K: int = 8, regime: float = 0.9, seed: int = 0, max_deps: int = 3,
                             rank: Optional[np.ndarray] = None, first: int = 0) -> DagBatch:
    """Instances first .. first+B-1 of the DEVICE generator's stream (dls_generate_counts / dls_generate_fill), built
    on the host with numpy: the same counter-based hash of (seed, instance, task, draw), the same arithmetic, bit for
    bit (tests/test_gpu_parity.py::test_device_generator_equals_its_host_port). What the CPU arm of bench.py schedules,
    so that both arms time the same instances without the CPU arm touching the GPU."""
    inst = (np.arange(B, dtype=np.uint64) + np.uint64(first))[:, None]
    task = np.arange(n, dtype=np.uint64)[None, :]
    inst_b, task_b = np.broadcast_arrays(inst, task)
    idx = np.arange(n)
    hi = np.minimum(max_deps, idx)
    nd = np.minimum((_u01(seed, inst_b, task_b, 0) * (hi + 1)).astype(np.int64), hi)
    npar = np.where(_u01(seed, inst_b, task_b, 1) < 0.5, 1, 2)
    mem = 0.1 + (0.5 - 0.1) * _u01(seed, inst_b, task_b, 2)
    dur = 0.05 + (0.15 - 0.05) * _u01(seed, inst_b, task_b, 3)
    picks = []
    for j in range(max_deps):
        span = np.maximum(idx - j, 1)
        r = np.minimum((_u01(seed, inst_b, task_b, 8 + j) * span).astype(np.int64), span - 1)
        if picks:
            prev = np.sort(np.stack(picks, 0), 0)
            for q in range(prev.shape[0]):
                r = r + (r >= prev[q])
        picks.append(r)
    picks = np.stack(picks, 2) if picks else np.zeros((B, n, 0), np.int64)
    keep = np.arange(max_deps)[None, None, :] < nd[:, :, None]
    pred_idx = picks[keep].astype(np.uint16)
    pred_ptr = np.zeros((B, n + 1), np.uint32)
    pred_ptr[:, 1:] = np.cumsum(nd, 1)
    e_per = pred_ptr[:, -1].astype(np.int64)
    present = np.zeros((B, n, 2), bool)
    present[:, :, 0] = True
    present[:, :, 1] = npar == 2
    pos = _param_name_order(n)
    by_name = np.zeros((B, 2 * n), bool)
    by_name[:, pos] = present.reshape(B, 2 * n)
    ids = (np.cumsum(by_name, 1) - by_name)[:, pos].reshape(B, n, 2)  # exclusive count = id in name order
    lo = np.where(present[:, :, 1], np.minimum(ids[:, :, 0], ids[:, :, 1]), ids[:, :, 0])
    pair = np.stack([lo, np.maximum(ids[:, :, 0], ids[:, :, 1])], 2)
    par_idx = pair[present].astype(np.uint16)
    par_ptr = np.zeros((B, n + 1), np.uint32)
    par_ptr[:, 1:] = np.cumsum(npar, 1)
    r_per = par_ptr[:, -1].astype(np.int64)
    total = (mem + npar * 0.5).max(1) + r_per * 0.5
    available = total * regime
    if K == 2:
        cap = np.stack([available * 0.6, available * 0.4], 1)
        spd = np.tile(np.array([1.2, 1.0]), (B, 1))
    elif K == 4:
        cap = np.stack([available * f for f in (0.35, 0.25, 0.25, 0.15)], 1)
        spd = np.tile(np.array([1.2, 1.0, 1.0, 0.8]), (B, 1))
    else:
        cap = np.repeat((available / K)[:, None], K, 1)
        kk = np.arange(K, dtype=np.uint64)[None, :]
        inst_k, _ = np.broadcast_arrays(inst, kk)
        spd = np.stack([0.7 + (1.3 - 0.7) * _u01(seed, inst_k[:, k], np.full(B, 0xFFFFFFFF, np.uint64), k) for k in range(K)], 1)
    if rank is None:
        rank = task_rank(n)
    off = lambda per: np.concatenate([[0], np.cumsum(per)]).astype(np.int64)
    return DagBatch(
        B=B, T=B, n_max=n, K_max=K, P_max=int(r_per.max()), E_max=int(e_per.max()), R_max=int(r_per.max()), flags=0,
        task_off=off(np.full(B, n, np.int64)), pred_off=off(e_per), par_off=off(r_per), pid_off=off(r_per),
        pred_ptr=np.ascontiguousarray(pred_ptr.reshape(-1)), pred_idx=pred_idx,
        par_ptr=np.ascontiguousarray(par_ptr.reshape(-1)), par_idx=par_idx,
        rank=np.tile(np.asarray(rank, np.uint16), B), mem=np.ascontiguousarray(mem.reshape(-1)),
        dur=np.ascontiguousarray(dur.reshape(-1)), node_off=off(np.full(B, K, np.int64)),
        cap=np.ascontiguousarray(cap.reshape(-1), np.float64), spd=np.ascontiguousarray(spd.reshape(-1), np.float64), dags=None)


def random_shared_param_batch(B: int, n: int, K: int, pool: int, regime: float = 0.7, seed: int = 0,
                              max_deps: int = 4, max_params: int = 3) -> DagBatch:
    """Random DAGs whose tasks draw 0..max_params params from a pool of `pool` names, so params are SHARED between
    tasks (pipeline-like pressure: cache hits, evictions with live readiness, rollbacks). Not a reference generator —
    a fuzzing input for the parity tests; instances are built task by task (small sizes)."""
    rng = np.random.default_rng(seed)
    dags, nodes = [], []
    for _ in range(B):
        names = [f"task_{i}" for i in range(n)]
        deps, params = [], []
        for i in range(n):
            nd = int(rng.integers(0, min(max_deps, i) + 1))
            d = rng.choice(i, size=nd, replace=False).tolist() if nd else []
            if d and rng.random() < 0.15:
                d.append(d[0])  # duplicate dependency entry (kept with multiplicity, ref schedulers.py:47-48)
            deps.append(d)
            npar = int(rng.integers(0, max_params + 1))
            params.append({f"w{int(x):03d}" for x in rng.choice(pool, size=min(npar, pool), replace=False)})
        mem = rng.uniform(0.05, 0.6, n)
        mem[rng.random(n) < 0.05] = 0.0
        dur = rng.uniform(0.01, 0.2, n)
        dag = lower_arrays(names, deps, params, mem, dur)
        total = float((mem + np.array([len(p) for p in params]) * 0.5).max() + dag.P * 0.5)
        cap = rng.uniform(0.3, 1.7, K) * total * regime / K
        spd = rng.choice([0.5, 0.8, 1.0, 1.0, 1.2, 1.3], K)  # repeated speeds: ties between nodes
        dags.append(dag)
        nodes.append((cap.astype(np.float64), spd.astype(np.float64)))
    return DagBatch.from_instances(dags, nodes)


def perturbed_durations(dag: LoweredDag, B: int, eps: float = 0.1, seed: int = 0) -> np.ndarray:
    """Config 3: dur_i * (1 + e), e ~ U(-eps, eps), one row per instance."""
    rng = np.random.default_rng(seed)
    return dag.dur[None, :] * (1.0 + rng.uniform(-eps, eps, (B, dag.n)))


def lower_arrays(names, deps, params, mem, dur) -> LoweredDag:
    """Array-level lowering: task names, dependency index lists, param-name sets, memory and time -> LoweredDag.
    The tie-break order pi is this process's set order of `names` (as add_task would produce it)."""
    n = len(names)
    pos = {nm: r for r, nm in enumerate(set_order(names))}
    pnames = sorted({p for ps in params for p in ps})
    pid = {p: i for i, p in enumerate(pnames)}
    pred_ptr = np.zeros(n + 1, np.uint32)
    par_ptr = np.zeros(n + 1, np.uint32)
    pred_ptr[1:] = np.cumsum([len(d) for d in deps])
    par_ptr[1:] = np.cumsum([len(p) for p in params])
    return LoweredDag(
        names=list(names), param_names=pnames, pred_ptr=pred_ptr,
        pred_idx=np.asarray([j for d in deps for j in d], np.uint16), par_ptr=par_ptr,
        par_idx=np.asarray([i for ps in params for i in sorted(pid[p] for p in ps)], np.uint16),
        rank=np.asarray([pos[nm] for nm in names], np.uint16), mem=np.asarray(mem, np.float64),
        dur=np.asarray(dur, np.float64))


def llm_dag(num_layers: int, attention_heads: int = 8, head_cap: int = 4) -> LoweredDag:
    """generate_llm_dag (ref simulation.py:36-88) straight to arrays; head_cap=4 is the reference's fixed cap (:52),
    head_cap=205 with 48 layers gives config 4's 9,986-task DAG."""
    names, deps, params, mem, dur = ["embedding"], [[]], [{"embedding_weights"}], [0.5], [0.1]
    prev = 0
    for layer in range(num_layers):
        heads = []
        for h in range(min(attention_heads, head_cap)):
            nm = f"layer_{layer}_attention_head_{h}"
            heads.append(len(names))
            names.append(nm); deps.append([prev]); params.append({f"{nm}_weights"}); mem.append(0.2); dur.append(0.05)
        names.append(f"layer_{layer}_attention_output"); deps.append(heads)
        params.append({f"layer_{layer}_attention_output_weights"}); mem.append(0.3); dur.append(0.05)
        names.append(f"layer_{layer}_ffn"); deps.append([len(names) - 2])
        params.append({f"layer_{layer}_ffn_weights"}); mem.append(0.5); dur.append(0.1)
        names.append(f"layer_{layer}_output"); deps.append([len(names) - 2]); params.append(set()); mem.append(0.1); dur.append(0.02)
        prev = len(names) - 1
    names.append("output"); deps.append([prev]); params.append({"output_weights"}); mem.append(0.3); dur.append(0.05)
    return lower_arrays(names, deps, params, mem, dur)


def pipeline_dag(num_stages: int, width: int = 3) -> LoweredDag:
    """generate_pipeline_dag (ref simulation.py:117-151) straight to arrays."""
    names, deps, params, mem, dur = [], [], [], [], []
    for stage in range(num_stages):
        prev = [] if stage == 0 else list(range((stage - 1) * width, stage * width))
        for wk in range(width):
            names.append(f"stage_{stage}_worker_{wk}"); deps.append(list(prev)); params.append({f"stage_{stage}_params"})
            mem.append(0.3); dur.append(0.1)
    names.append("final_output"); deps.append(list(range((num_stages - 1) * width, num_stages * width)))
    params.append({"output_params"}); mem.append(0.2); dur.append(0.05)
    return lower_arrays(names, deps, params, mem, dur)


def gpt2_dag(n_layer: int = 12, n_embd: int = 768, vocab_size: int = 50257) -> LoweredDag:
    """The GPT-2 DAG of LLMDAGExtractor.extract_gpt2_dag (ref test_gpt2.py:45-168) from the model's SHAPES alone (no
    torch module is instantiated): estimate_memory_gb (ref :18-31) is parameter bytes / 1e9 plus, for a module with a
    `.weight`, prod(weight.shape) * batch * 4 / 1e9, else 0.1. Defaults = gpt2-small: 99 tasks, 75 params."""
    emb_numel = vocab_size * n_embd
    embedding_memory = emb_numel * 4 / 1e9 + float(np.prod((vocab_size, n_embd))) * 1 * 4 / 1e9          # model.wte
    attn_numel = (n_embd * 3 * n_embd + 3 * n_embd) + (n_embd * n_embd + n_embd)                          # c_attn + c_proj
    attn_memory = attn_numel * 4 / 1e9 + 0.1                                                              # no .weight
    fc_numel = n_embd * 4 * n_embd + 4 * n_embd
    ffn_memory = fc_numel * 4 / 1e9 + float(np.prod((n_embd, 4 * n_embd))) * 1 * 4 / 1e9                 # mlp.c_fc (Conv1D)
    names, deps, params, mem, dur = ["embedding"], [[]], [{"embedding_weights", "position_weights"}], [embedding_memory], [0.1]
    prev = 0
    for L in range(n_layer):
        base = len(names)
        rows = [
            (f"layer_{L}_ln1", [prev], {f"layer_{L}_ln1_weights"}, 0.01, 0.01),
            (f"layer_{L}_attention", [base], {f"layer_{L}_attn_qkv_weights", f"layer_{L}_attn_proj_weights"}, attn_memory, 0.05),
            (f"layer_{L}_attn_residual", [base + 1, prev], set(), 0.01, 0.01),
            (f"layer_{L}_ln2", [base + 2], {f"layer_{L}_ln2_weights"}, 0.01, 0.01),
            (f"layer_{L}_ffn_expand", [base + 3], {f"layer_{L}_ffn_expand_weights"}, ffn_memory, 0.08),
            (f"layer_{L}_ffn_activation", [base + 4], set(), 0.01, 0.01),
            (f"layer_{L}_ffn_contract", [base + 5], {f"layer_{L}_ffn_contract_weights"}, ffn_memory, 0.08),
            (f"layer_{L}_output", [base + 6, base + 2], set(), 0.01, 0.01),
        ]
        for nm, d, p, m, c in rows:
            names.append(nm); deps.append(d); params.append(p); mem.append(m); dur.append(c)
        prev = base + 7
    names.append("final_ln"); deps.append([prev]); params.append({"final_ln_weights"}); mem.append(0.01); dur.append(0.01)
    names.append("output_projection"); deps.append([len(names) - 2]); params.append({"embedding_weights"})
    mem.append(embedding_memory); dur.append(0.1)
    return lower_arrays(names, deps, params, mem, dur)


def random_dag_batch_device(B: int, n: int, K: int = 8, regime: float = 0.9, seed: int = 0, max_deps: int = 3,
                            device=None, rank: Optional[np.ndarray] = None):
    """Config-5 sweep inputs built ON the GPU by the library's generator kernels (dls_generate_counts/_fill).
    Returns a DeviceBatch; `.to_host()` copies it back for oracle checks. torch is used for the buffers and for the
    two prefix sums over per-instance counts."""
    import ctypes as C

    import torch

    from . import _native as N
    from .batch import DeviceBatch
    lib = N.load()
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device: the generator kernels have no CPU implementation")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    with torch.cuda.device(dev):
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        rank_h = task_rank(n) if rank is None else rank.astype(np.uint16)
        rank_d = torch.from_numpy(rank_h.view(np.int16)).to(dev)
        ppos_d = torch.from_numpy(_param_name_order(n).astype(np.uint16).view(np.int16)).to(dev)
        g = N.DlsGen(B=B, n=n, K=K, max_deps=max_deps, regime=regime, seed=seed, rank=rank_d.data_ptr(),
                     param_pos=ppos_d.data_ptr())
        e_cnt = torch.empty(B, dtype=torch.int32, device=dev)
        r_cnt = torch.empty(B, dtype=torch.int32, device=dev)
        N.check(lib.dls_generate_counts(C.byref(g), e_cnt.data_ptr(), r_cnt.data_ptr(), stream))
        zero = torch.zeros(1, dtype=torch.int64, device=dev)
        pred_off = torch.cat([zero, torch.cumsum(e_cnt.to(torch.int64), 0)])
        par_off = torch.cat([zero, torch.cumsum(r_cnt.to(torch.int64), 0)])
        task_off_h = np.arange(B + 1, dtype=np.int64) * n
        node_off_h = np.arange(B + 1, dtype=np.int64) * K
        task_off = torch.from_numpy(task_off_h).to(dev, non_blocking=True)
        node_off = torch.from_numpy(node_off_h).to(dev, non_blocking=True)
        # ONE device-to-host read for everything the host needs to size the arrays: both offset tables and the two maxima
        head = torch.cat([pred_off, par_off, e_cnt.max().to(torch.int64).reshape(1), r_cnt.max().to(torch.int64).reshape(1)]).cpu().numpy()
        pred_off_h, par_off_h = head[:B + 1].copy(), head[B + 1:2 * B + 2].copy()
        e_total, r_total, e_max, r_max = int(pred_off_h[-1]), int(par_off_h[-1]), int(head[-2]), int(head[-1])
        t = {
            "task_off": task_off, "pred_off": pred_off, "par_off": par_off, "pid_off": par_off, "node_off": node_off,
            "pred_ptr": torch.empty(B * (n + 1), dtype=torch.int32, device=dev),
            "pred_idx": torch.empty(max(e_total, 1), dtype=torch.int16, device=dev),
            "par_ptr": torch.empty(B * (n + 1), dtype=torch.int32, device=dev),
            "par_idx": torch.empty(max(r_total, 1), dtype=torch.int16, device=dev),
            "rank": torch.empty(B * n, dtype=torch.int16, device=dev),
            "mem": torch.empty(B * n, dtype=torch.float64, device=dev),
            "dur": torch.empty(B * n, dtype=torch.float64, device=dev),
            "cap": torch.empty(B * K, dtype=torch.float64, device=dev),
            "spd": torch.empty(B * K, dtype=torch.float64, device=dev),
        }
        meta = DagBatch(
            B=B, T=B, n_max=n, K_max=K, P_max=r_max, E_max=e_max, R_max=r_max, flags=0,
            task_off=task_off_h, pred_off=pred_off_h, par_off=par_off_h,
            pid_off=par_off_h.copy(), pred_ptr=None, pred_idx=None, par_ptr=None, par_idx=None, rank=None,
            mem=None, dur=None, node_off=node_off_h, cap=None, spd=None, dags=None)
        db = DeviceBatch.from_tensors(meta, t, dev)
        N.check(lib.dls_generate_fill(C.byref(g), C.byref(db.struct), stream))
        db._keep = (rank_d, ppos_d)
    return db
