"""Harness that imports the UNMODIFIED reference from /root/reference (this container only) and runs it on
seeded inputs, returning results in the lowered array form used everywhere else.

Nothing here is used on the GPU box: tests that need it skip when /root/reference is absent, and
tests/golden/make_golden.py freezes its outputs into fixtures that travel.
"""
from __future__ import annotations

import os
import random
import sys
import types

import numpy as np

REF_DIR = os.environ.get("DLS_REFERENCE_DIR", "/root/reference")


def available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "schedulers.py"))


_mods = {}


def ref_modules():
    """(schedulers, simulation) modules of the reference; matplotlib is stubbed (not installed; only
    analyze_results touches it, ref simulation.py:418-563)."""
    if _mods:
        return _mods["schedulers"], _mods["simulation"]
    sys.dont_write_bytecode = True  # the reference directory is read-only
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    import importlib.util

    def load(name, alias):
        spec = importlib.util.spec_from_file_location(alias, os.path.join(REF_DIR, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[alias] = mod
        spec.loader.exec_module(mod)
        return mod

    # simulation.py does `from schedulers import *`: register the reference module under that bare name while
    # it is imported, then restore whatever was there.
    saved = sys.modules.get("schedulers")
    sch = load("schedulers", "schedulers")
    sim = load("simulation", "ref_simulation")
    if saved is not None:
        sys.modules["schedulers"] = saved
    else:
        sys.modules["ref_schedulers"] = sys.modules.pop("schedulers")
    _mods["schedulers"], _mods["simulation"] = sch, sim
    return sch, sim


ALGO_CLASS = {0: "DFSScheduler", 1: "GreedyScheduler", 2: "CriticalPathScheduler", 3: "MRUScheduler"}


def gpt2_tasks():
    """The 99-task GPT-2 DAG of ref test_gpt2.py:45-168, with the hub lookup patched to library defaults
    (= gpt2-small)."""
    sch, _ = ref_modules()
    from transformers import GPT2Config
    orig = GPT2Config.from_pretrained
    GPT2Config.from_pretrained = classmethod(lambda cls, *a, **k: cls())
    saved = sys.modules.get("schedulers")
    sys.modules["schedulers"] = sch
    try:
        import importlib.util
        spec = importlib.util.spec_from_file_location("ref_test_gpt2", os.path.join(REF_DIR, "test_gpt2.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        tasks = mod.LLMDAGExtractor("gpt2").extract_gpt2_dag()
    finally:
        GPT2Config.from_pretrained = orig
        if saved is not None:
            sys.modules["schedulers"] = saved
        else:
            sys.modules.pop("schedulers", None)
    return tasks


class PeakMixin:
    """Pins the peak-memory extension: sample total-available between ref schedulers.py:95 and :126."""

    def complete_task(self, task_id):
        task = self.tasks.get(task_id)
        if task is not None and task.assigned_node:
            node = self.nodes[task.assigned_node]
            pk = getattr(node, "_peak", 0.0)
            node._peak = max(pk, node.total_memory - node.available_memory)
            self.__dict__.setdefault("_order_log", []).append(task_id)
        return super().complete_task(task_id)

    def assign_task_to_node(self, task_id, node_id):
        """Counts the assignments the reference refuses AFTER a node was chosen (ref schedulers.py:82-83; reached from
        MRU's :509 when an eviction's rounding leaves the node a hair short and nothing is left to evict)."""
        ok = super().assign_task_to_node(task_id, node_id)
        if not ok:
            self.__dict__["_refused"] = self.__dict__.get("_refused", 0) + 1
        return ok


def run_reference(algo: int, tasks, nodes):
    """Run one reference scheduler on copies of (tasks, nodes) exactly as run_single_test does
    (ref simulation.py:308-336). Returns a dict with the live scheduler, pi, schedule and stats."""
    sch, sim = ref_modules()
    base = getattr(sch, ALGO_CLASS[algo])
    cls = type("Peak" + base.__name__, (PeakMixin, base), {})
    task_copies = [sch.Task(t.id, t.memory_required, t.compute_time, list(t.dependencies), set(t.params_needed))
                   for t in tasks]
    node_copies = [sch.Node(n.id, n.total_memory, n.compute_speed) for n in nodes]
    s = cls(node_copies)
    for t in task_copies:
        s.add_task(t)
    pi = list(s.pending_tasks)
    lim = sys.getrecursionlimit()
    sys.setrecursionlimit(max(lim, 4 * len(tasks) + 1000))
    try:
        schedule = s.schedule()
    finally:
        sys.setrecursionlimit(lim)
    ev = sim.ImprovedSchedulerEvaluator({})
    makespan, stats = ev.simulate_execution(s, schedule)
    util = list(stats["node_utilization"].values())
    return {"scheduler": s, "pi": pi, "schedule": schedule, "makespan": float(makespan), "stats": stats,
            "tasks": task_copies, "nodes": node_copies, "evaluator": ev,
            "avg_node_utilization": float(np.mean(util)) if util else 0.0,              # ref simulation.py:344-345
            "load_balance_score": float(ev.calculate_load_balance(s, schedule)),       # ref simulation.py:347
            "completion_rate": (len(s.completed_tasks) / len(tasks) * 100) if tasks else 0.0}


def expected_arrays(run, dag):
    """Reference outcome in the array form of dls_result_t for one (instance, algo)."""
    s = run["scheduler"]
    idx = {nm: i for i, nm in enumerate(dag.names)}
    nodes = run["nodes"]
    kidx = {nd.id: k for k, nd in enumerate(nodes)}
    n, K = dag.n, len(nodes)
    node_of = np.full(n, -2, np.int8)
    for tid in s.failed_tasks:
        node_of[idx[tid]] = -1
    # global assignment order: node.completed_tasks keeps per-node order; the global order is recovered from
    # the schedule dict only per node, so take it from the per-task assignment log kept by `order_log`.
    order_log = [idx[t] for t in getattr(s, "_order_log", [])]
    for t in run["tasks"]:
        if t.completed:
            node_of[idx[t.id]] = kidx[t.assigned_node]
    fin = np.zeros(K)
    cnt = np.zeros(K, np.int32)
    for nd in nodes:
        k = kidx[nd.id]
        cnt[k] = len(nd.completed_tasks)
        tt = 0
        for tid in run["schedule"].get(nd.id, []):
            tt += s.tasks[tid].compute_time / nd.compute_speed
        fin[k] = tt
    avail = np.array([nd.available_memory for nd in nodes], np.float64)
    peak = np.array([getattr(nd, "_peak", 0.0) for nd in nodes], np.float64)
    cached = np.zeros(dag.P, np.uint32)
    pidx = {p: i for i, p in enumerate(dag.param_names)}
    for nd in nodes:
        for p in nd.cached_params:
            cached[pidx[p]] |= np.uint32(1 << kidx[nd.id])
    per_node_seq = [[idx[t] for t in run["schedule"].get(nd.id, [])] for nd in nodes]
    return {"order": order_log, "node_of": node_of, "fin": fin, "cnt": cnt, "avail": avail, "peak": peak, "cached": cached,
            "per_node_seq": per_node_seq, "completed": len(s.completed_tasks), "failed": len(s.failed_tasks),
            "pending": len(s.pending_tasks), "makespan": run["makespan"],
            "hits": run["stats"]["param_cache_hits"], "misses": run["stats"]["param_cache_misses"],
            "key_order": [kidx[k] for k in run["schedule"].keys()], "refused": int(getattr(s, "_refused", 0)),
            "avg_node_utilization": run["avg_node_utilization"], "load_balance_score": run["load_balance_score"],
            "completion_rate": run["completion_rate"]}
