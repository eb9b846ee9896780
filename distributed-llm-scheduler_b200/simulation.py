"""Host-side mirror of the reference's evaluation harness around the hot path (ref simulation.py:15-416).

Kept: `TestResult` (13 fields + the dynamic `num_nodes`), `DAGGenerator` (same task literals and the same sequence of
`random` calls, so a seeded run builds the same DAGs as the reference), the node/memory-regime factory,
`run_single_test`, and `run_experiments`. Changed: scheduling and `simulate_execution` run on the GPU. `run_experiments`
builds every (DAG, node set) of the sweep first and schedules them all, for all four algorithms, in ONE batched device
call (`evaluate_batch`), then packs the results in the reference's order. Plots (`analyze_results`' figures) are out of
scope — matplotlib output is unavailable on this box; the CSV schema is kept (`results_frame`, `save_csv`).
"""
from __future__ import annotations

import random
import time
from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np

from . import _native as N
from .batch import ResultArrays, schedule_batch
from .lowering import DagBatch, lower_nodes, lower_tasks
from .schedulers import (BaseScheduler, CriticalPathScheduler, DFSScheduler, GreedyScheduler, MRUScheduler, Node,  # noqa: F401
                         Task)

CSV_COLUMNS = ["scheduler_name", "dag_type", "memory_regime", "total_tasks", "completed_tasks", "failed_tasks", "makespan",
               "avg_node_utilization", "param_cache_hits", "param_cache_misses", "load_balance_score", "execution_time",
               "completion_rate", "num_nodes"]  # ref simulation.py:424-439


@dataclass
class TestResult:
    """ref simulation.py:15-30"""
    __test__ = False  # not a pytest class
    scheduler_name: str
    dag_type: str
    memory_regime: float
    total_tasks: int
    completed_tasks: int
    failed_tasks: int
    makespan: float
    avg_node_utilization: float
    param_cache_hits: int
    param_cache_misses: int
    load_balance_score: float
    execution_time: float
    completion_rate: float


class DAGGenerator:
    """ref simulation.py:33-151. Same Task literals; same order of `random` calls."""

    @staticmethod
    def generate_llm_dag(num_layers: int, layer_width: int = 1, attention_heads: int = 8, ffn_multiplier: int = 4,
                         head_cap: int = 4) -> List[Task]:
        """ref simulation.py:36-88. `head_cap` (reference: fixed 4, :52) is exposed for config 4's wide DAG."""
        tasks = [Task("embedding", 0.5, 0.1, [], {"embedding_weights"})]
        for layer in range(num_layers):
            prev = ["embedding"] if layer == 0 else [f"layer_{layer-1}_output"]
            heads = []
            for head in range(min(attention_heads, head_cap)):
                name = f"layer_{layer}_attention_head_{head}"
                tasks.append(Task(name, 0.2, 0.05, list(prev), {f"{name}_weights"}))
                heads.append(name)
            tasks.append(Task(f"layer_{layer}_attention_output", 0.3, 0.05, heads, {f"layer_{layer}_attention_output_weights"}))
            tasks.append(Task(f"layer_{layer}_ffn", 0.5, 0.1, [f"layer_{layer}_attention_output"], {f"layer_{layer}_ffn_weights"}))
            tasks.append(Task(f"layer_{layer}_output", 0.1, 0.02, [f"layer_{layer}_ffn"], set()))
        tasks.append(Task("output", 0.3, 0.05, [f"layer_{num_layers-1}_output"], {"output_weights"}))
        return tasks

    @staticmethod
    def generate_random_dag(num_tasks: int, max_deps: int = 3) -> List[Task]:
        """ref simulation.py:91-114"""
        tasks = []
        for i in range(num_tasks):
            deps = []
            if i > 0:
                num_deps = min(random.randint(0, min(max_deps, i)), i)
                if num_deps > 0:
                    # sampling indices of a length-i population draws the same random numbers as sampling its names
                    deps = [f"task_{j}" for j in random.sample(range(i), num_deps)]
            num_params = random.randint(1, 2)
            params = {f"param_{i}_{j}" for j in range(num_params)}
            mem = random.uniform(0.1, 0.5)
            dur = random.uniform(0.05, 0.15)
            tasks.append(Task(f"task_{i}", mem, dur, deps, params))
        return tasks

    @staticmethod
    def generate_pipeline_dag(num_stages: int, width: int = 3) -> List[Task]:
        """ref simulation.py:117-151"""
        tasks = []
        for stage in range(num_stages):
            deps = [] if stage == 0 else [f"stage_{stage-1}_worker_{i}" for i in range(width)]
            for wk in range(width):
                tasks.append(Task(f"stage_{stage}_worker_{wk}", 0.3, 0.1, list(deps), {f"stage_{stage}_params"}))
        tasks.append(Task("final_output", 0.2, 0.05, [f"stage_{num_stages-1}_worker_{i}" for i in range(width)],
                          {"output_params"}))
        return tasks


def evaluate_batch(batch: DagBatch, res: ResultArrays, slot: int) -> Dict[str, np.ndarray]:
    """TestResult's derived metrics for every instance of a scheduled batch, from the device records
    (ref simulation.py:271-302,339-347), with no per-instance Python: segment reductions over the per-node arrays.
    fin[] already holds simulate_execution's per-node sums; nodes with cnt == 0 are absent from the schedule dict."""
    B = batch.B
    summ = res.summary[slot]
    makespan = summ["makespan"].astype(np.float64)
    fin, cnt = res.fin[slot], res.cnt[slot]
    if batch.node_off is None:
        seg = np.repeat(np.arange(B), batch.K_max)
        fin, cnt = fin[:B * batch.K_max], cnt[:B * batch.K_max]
    else:
        seg = np.repeat(np.arange(B), np.diff(batch.node_off))
    used = cnt > 0
    loads = np.where(used, fin, 0.0)
    k_used = np.bincount(seg, weights=used, minlength=B)
    total = np.bincount(seg, weights=loads, minlength=B)
    have = k_used > 0
    mean = np.divide(total, k_used, out=np.zeros(B), where=have)
    # np.mean(loads / makespan) (ref :274-276,344)
    ms_safe = np.where(makespan > 0, makespan, 1.0)
    util = np.divide(np.bincount(seg, weights=loads / ms_safe[seg], minlength=B), k_used, out=np.zeros(B),
                     where=have & (makespan > 0))
    # 1 / (1 + np.std(loads) / np.mean(loads)), ddof 0, over the used nodes (ref :295-301)
    dev2 = np.where(used, (fin - mean[seg]) ** 2, 0.0)
    std = np.sqrt(np.divide(np.bincount(seg, weights=dev2, minlength=B), k_used, out=np.zeros(B), where=have))
    lmax = np.full(B, -np.inf)
    np.maximum.at(lmax, seg[used], fin[used])
    ok = have & (lmax != 0) & (mean > 0)
    balance = np.divide(1.0, 1.0 + np.divide(std, mean, out=np.zeros(B), where=ok), out=np.zeros(B), where=ok)
    n_tasks = np.diff(batch.task_off) if batch.T == batch.B else np.full(B, int(batch.task_off[1]))
    completed = summ["completed"].astype(np.int64)
    rate = np.where(n_tasks > 0, completed / np.maximum(n_tasks, 1) * 100, 0.0)
    return {"total_tasks": n_tasks, "completed_tasks": completed, "failed_tasks": summ["failed"].astype(np.int64),
            "makespan": makespan, "avg_node_utilization": util, "param_cache_hits": summ["hits"].astype(np.int64),
            "param_cache_misses": summ["misses"].astype(np.int64), "load_balance_score": balance, "completion_rate": rate}


class ImprovedSchedulerEvaluator:
    """ref simulation.py:154-416"""

    ALGO_OF = {DFSScheduler: N.ALGO_DFS, GreedyScheduler: N.ALGO_GREEDY, CriticalPathScheduler: N.ALGO_CRITICAL,
               MRUScheduler: N.ALGO_MRU}

    def __init__(self, schedulers: Dict[str, type]):
        self.schedulers = schedulers
        self.results: List[TestResult] = []

    def create_nodes_with_memory_regime(self, total_memory_needed: float, memory_regime: float, num_nodes: int = 4) -> List[Node]:
        """ref simulation.py:161-192"""
        available = total_memory_needed * memory_regime
        if num_nodes == 2:
            return [Node("node_0", available * 0.6, 1.2), Node("node_1", available * 0.4, 1.0)]
        if num_nodes == 4:
            return [Node(f"node_{i}", available * f, s)
                    for i, (f, s) in enumerate(zip([0.35, 0.25, 0.25, 0.15], [1.2, 1.0, 1.0, 0.8]))]
        per_node = available / num_nodes
        return [Node(f"node_{i}", per_node, random.uniform(0.7, 1.3)) for i in range(num_nodes)]

    def calculate_total_memory_needed(self, tasks: List[Task]) -> float:
        """ref simulation.py:194-214"""
        max_concurrent = 0
        all_params = set()
        for task in tasks:
            task_total = task.memory_required
            task_total += len(task.params_needed) * 0.5
            max_concurrent = max(max_concurrent, task_total)
            all_params.update(task.params_needed)
        return max_concurrent + len(all_params) * 0.5

    def simulate_execution(self, scheduler: BaseScheduler, schedule: Dict[str, List[str]]) -> Tuple[float, Dict]:
        """ref simulation.py:216-278. For the schedule that `scheduler.schedule()` just returned the device already
        produced these numbers in the scheduling kernel's epilogue; any other (scheduler, schedule) pair — the reference
        accepts whatever dict it is handed — is replayed here the way the reference does it."""
        if not schedule:
            return 0.0, {"param_cache_hits": 0, "param_cache_misses": 0, "node_utilization": {}}
        rec = getattr(scheduler, "last_device_result", None)
        if rec is not None and rec["schedule"] == schedule:
            makespan = float(rec["makespan"])
            util = {}
            for nid in schedule:
                util[nid] = rec["fin"][nid] / makespan if makespan > 0 else rec["fin"][nid]
            return makespan, {"param_cache_hits": int(rec["hits"]), "param_cache_misses": int(rec["misses"]),
                              "node_utilization": util}
        hits = misses = 0
        finish, util = {}, {}
        for node_id, task_ids in schedule.items():
            if node_id not in scheduler.nodes:
                continue
            node = scheduler.nodes[node_id]
            now, seen = 0, set()
            for task_id in task_ids:
                if task_id not in scheduler.tasks:
                    continue
                task = scheduler.tasks[task_id]
                for param in task.params_needed:  # per-node seen-set replay, evictions ignored (ref :253-258)
                    if param in seen:
                        hits += 1
                    else:
                        misses += 1
                        seen.add(param)
                dur = task.compute_time / node.compute_speed
                now += dur
                util[node_id] = util.get(node_id, 0.0) + dur
            finish[node_id] = now
        makespan = max(finish.values()) if finish else 0
        if makespan > 0:
            for node_id in util:
                util[node_id] /= makespan
        return makespan, {"param_cache_hits": hits, "param_cache_misses": misses, "node_utilization": util}

    def calculate_load_balance(self, scheduler: BaseScheduler, schedule: Dict[str, List[str]]) -> float:
        """ref simulation.py:280-302 (a derived metric of the lifted schedule; Python's sum() is kept because its
        compensated summation can differ from a running += by an ulp)."""
        node_loads = []
        for node_id, task_ids in schedule.items():
            if node_id not in scheduler.nodes:
                continue
            node = scheduler.nodes[node_id]
            node_loads.append(sum(scheduler.tasks[t].compute_time / node.compute_speed for t in task_ids if t in scheduler.tasks))
        if not node_loads or max(node_loads) == 0:
            return 0
        avg_load, std_load = np.mean(node_loads), np.std(node_loads)
        return 1 / (1 + std_load / avg_load) if avg_load > 0 else 0

    def run_single_test(self, scheduler_class: type, scheduler_name: str, tasks: List[Task], nodes: List[Node],
                        dag_type: str, memory_regime: float) -> TestResult:
        """ref simulation.py:304-363 — same steps, the scheduler class is GPU-backed."""
        task_copies = [Task(t.id, t.memory_required, t.compute_time, t.dependencies.copy(), t.params_needed.copy())
                       for t in tasks]
        node_copies = [Node(nd.id, nd.total_memory, nd.compute_speed) for nd in nodes]
        scheduler = scheduler_class(node_copies)
        for t in task_copies:
            scheduler.add_task(t)
        start = time.time()
        try:
            schedule = scheduler.schedule()
        except Exception as e:  # the reference swallows scheduler errors the same way (:328-332)
            print(f"Error in {scheduler_name}: {e}")
            schedule = {}
        execution_time = time.time() - start
        makespan, stats = self.simulate_execution(scheduler, schedule)
        completed, failed, total = len(scheduler.completed_tasks), len(scheduler.failed_tasks), len(tasks)
        util_vals = list(stats["node_utilization"].values())
        return TestResult(
            scheduler_name=scheduler_name, dag_type=dag_type, memory_regime=memory_regime, total_tasks=total,
            completed_tasks=completed, failed_tasks=failed, makespan=makespan,
            avg_node_utilization=np.mean(util_vals) if util_vals else 0, param_cache_hits=stats["param_cache_hits"],
            param_cache_misses=stats["param_cache_misses"], load_balance_score=self.calculate_load_balance(scheduler, schedule),
            execution_time=execution_time, completion_rate=(completed / total * 100) if total > 0 else 0)

    # ---- the sweep, batched ------------------------------------------------------------------------------------
    def sweep_instances(self, num_runs: int = 5) -> List[Tuple[str, int, float, List[Task], List[Node]]]:
        """The (dag_type, num_nodes, regime, tasks, nodes) list run_experiments iterates, in its order and with its
        sequence of `random` calls (ref simulation.py:366-401)."""
        dag_configs = [("LLM-Small", lambda: DAGGenerator.generate_llm_dag(4, attention_heads=4)),
                       ("LLM-Medium", lambda: DAGGenerator.generate_llm_dag(8, attention_heads=4)),
                       ("LLM-Large", lambda: DAGGenerator.generate_llm_dag(12, attention_heads=4)),
                       ("Random-Small", lambda: DAGGenerator.generate_random_dag(30)),
                       ("Random-Medium", lambda: DAGGenerator.generate_random_dag(60)),
                       ("Pipeline", lambda: DAGGenerator.generate_pipeline_dag(5, width=3))]
        out = []
        for dag_name, gen in dag_configs:
            for num_nodes in (2, 4, 8):
                for regime in (1.0, 0.9, 0.8):
                    for _ in range(num_runs):
                        tasks = gen()
                        total = self.calculate_total_memory_needed(tasks)
                        out.append((dag_name, num_nodes, regime, tasks, self.create_nodes_with_memory_regime(total, regime, num_nodes)))
        return out

    def run_batch(self, instances: Sequence[Tuple[str, int, float, List[Task], List[Node]]]) -> List[TestResult]:
        """Schedule all instances x all configured schedulers in one device call and pack TestResults in the
        reference's order (instance-major, scheduler-minor; ref simulation.py:404-410)."""
        algos = {name: self.ALGO_OF[next(c for c in cls.__mro__ if c in self.ALGO_OF)] for name, cls in self.schedulers.items()}
        mask = 0
        for a in algos.values():
            mask |= 1 << a
        dags = [lower_tasks(tasks) for _, _, _, tasks, _ in instances]
        batch = DagBatch.from_instances(dags, [lower_nodes(nodes) for *_, nodes in instances])
        t0 = time.time()
        res = schedule_batch(batch, mask).host()
        per_call = (time.time() - t0) / max(1, len(instances) * len(algos))
        metrics = {a: evaluate_batch(batch, res, res.slot(a)) for a in set(algos.values())}
        out = []
        for b, (dag_name, num_nodes, regime, _, _) in enumerate(instances):
            for name, a in algos.items():
                m = metrics[a]
                r = TestResult(
                    scheduler_name=name, dag_type=dag_name, memory_regime=regime, total_tasks=int(m["total_tasks"][b]),
                    completed_tasks=int(m["completed_tasks"][b]), failed_tasks=int(m["failed_tasks"][b]),
                    makespan=float(m["makespan"][b]), avg_node_utilization=float(m["avg_node_utilization"][b]),
                    param_cache_hits=int(m["param_cache_hits"][b]), param_cache_misses=int(m["param_cache_misses"][b]),
                    load_balance_score=float(m["load_balance_score"][b]), execution_time=per_call,
                    completion_rate=float(m["completion_rate"][b]))
                r.num_nodes = num_nodes  # ref :409
                out.append(r)
        return out

    def run_experiments(self, num_runs: int = 5):
        """ref simulation.py:365-416, batched."""
        instances = self.sweep_instances(num_runs)
        self.results.extend(self.run_batch(instances))
        print(f"\nCompleted {len(instances)} test configurations")

    def results_frame(self):
        """The DataFrame analyze_results builds (ref simulation.py:424-439), same columns in the same order."""
        import pandas as pd
        return pd.DataFrame([{c: getattr(r, c, 4) for c in CSV_COLUMNS} for r in self.results], columns=CSV_COLUMNS)

    def save_csv(self, path: str = "evaluation_results/raw_results.csv"):
        import os
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        self.results_frame().to_csv(path, index=False)  # ref :445
