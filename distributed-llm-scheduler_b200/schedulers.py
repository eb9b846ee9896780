"""Host-side mirror of the reference's scheduler classes (ref schedulers.py:7-525).

Same names, constructor signatures, `add_task` / `schedule()` contract and post-call state as the reference,
so callers such as `run_single_test` (ref simulation.py:304-363), `test_schedulers` (ref schedulers.py:529-568)
and `test_with_your_scheduler` (ref test_gpt2.py:274-299) work unchanged. `schedule()` does no scheduling in
Python: it lowers the registered tasks to CSR arrays and runs a batch of one through the CUDA library
(`batch.schedule_batch`), then lifts the device records back into the reference's objects. There is no CPU
fallback — without the CUDA library or a GPU, `schedule()` raises.
"""
from __future__ import annotations

from collections import defaultdict, deque
from typing import Dict, List, Optional, Set

import numpy as np

from . import _native as N
from .batch import ResultArrays, schedule_batch, schedule_dict, sequence_of
from .lowering import DagBatch, lower_nodes, lower_tasks


class Task:
    """ref schedulers.py:7-17"""

    def __init__(self, task_id: str, memory_required: float, compute_time: float,
                 dependencies: List[str] = None, params_needed: Set[str] = None):
        self.id = task_id
        self.memory_required = memory_required  # GB
        self.compute_time = compute_time  # seconds
        self.dependencies = dependencies or []
        self.params_needed = params_needed or set()
        self.completed = False
        self.assigned_node = None


class Node:
    """ref schedulers.py:19-29"""

    def __init__(self, node_id: str, total_memory: float, compute_speed: float = 1.0):
        self.id = node_id
        self.total_memory = total_memory
        self.available_memory = total_memory
        self.compute_speed = compute_speed
        self.cached_params = set()
        self.running_tasks = []
        self.completed_tasks = []
        self.last_used_params = deque(maxlen=10)


class BaseScheduler:
    """ref schedulers.py:31-135. Registry + the lowering/lifting glue around the device call."""

    ALGO: Optional[int] = None

    def __init__(self, nodes: List[Node]):
        self.nodes = {node.id: node for node in nodes}
        self.tasks = {}
        self.ready_queue = []
        self.completed_tasks = set()
        self.failed_tasks = set()
        self.task_dependencies = defaultdict(list)
        self.param_locations = defaultdict(set)
        # A real set filled by add(): its iteration order IS the reference's tie-break among ready tasks.
        self.pending_tasks = set()
        self._added: List[Task] = []
        self.last_summary = None  # dls_summary_t of the last schedule() as a numpy record
        self.last_device_result = None

    def add_task(self, task: Task):
        """ref schedulers.py:43-48"""
        self.tasks[task.id] = task
        self.pending_tasks.add(task.id)
        self._added.append(task)
        for dep in task.dependencies:
            self.task_dependencies[dep].append(task.id)

    # The small predicates are kept because callers and subclasses of the reference may use them; they read
    # the lifted state and do not take part in scheduling.
    def is_task_ready(self, task_id: str) -> bool:
        """ref schedulers.py:50-53"""
        return all(dep in self.completed_tasks for dep in self.tasks[task_id].dependencies)

    def get_ready_tasks(self) -> List[Task]:
        """ref schedulers.py:55-61"""
        return [self.tasks[t] for t in self.pending_tasks if self.is_task_ready(t)]

    def calculate_memory_requirement(self, task: Task, node: Node) -> float:
        """ref schedulers.py:63-72"""
        memory_needed = task.memory_required
        memory_needed += len(task.params_needed - node.cached_params) * 0.5
        return memory_needed

    def can_fit_on_node(self, task: Task, node: Node) -> bool:
        """ref schedulers.py:74-76"""
        return self.calculate_memory_requirement(task, node) <= node.available_memory

    def schedule(self) -> Dict[str, List[str]]:
        """ref schedulers.py:133,154,244,323,444 — runs on the GPU as a batch of one."""
        if self.ALGO is None:
            raise NotImplementedError
        tasks = self._added
        if len(self.tasks) != len(tasks):
            raise ValueError("duplicate task ids are not supported")
        node_list = list(self.nodes.values())
        if not tasks:
            return {}
        if self.completed_tasks or self.failed_tasks:
            # A second call after a run: whatever is still pending waits for a task that failed, so the reference's
            # loop finds nothing ready and returns an empty schedule (schedulers.py:167-174); same here, without a launch.
            return {}
        if any(nd.cached_params or nd.completed_tasks for nd in node_list):
            # (the device starts every node with an empty cache and no completed tasks; available_memory is taken as is)
            raise ValueError("nodes that already hold cached params or completed tasks are not supported: "
                             "pass fresh Node objects (their available_memory is honoured)")
        if not node_list:
            # ref: every ready task fails, then nothing was scheduled -> all pending fail (schedulers.py:198-206)
            for tid in list(self.pending_tasks):
                self.failed_tasks.add(tid)
                self.pending_tasks.discard(tid)
            return {}
        dag = lower_tasks(tasks, pending=self._pending_order())
        cap = np.asarray([float(x.available_memory) for x in node_list], np.float64)
        _, spd = lower_nodes(node_list)
        batch = DagBatch.from_instances([dag], [(cap, spd)])
        res = schedule_batch(batch, 1 << self.ALGO, peak=False, cached=True).host()
        self._lift(batch, res, dag, node_list)
        sched = schedule_dict(batch, res, 0, 0, dag.names, [x.id for x in node_list])
        # simulate_execution's outputs (ref simulation.py:216-278) came out of the same kernel; keep them for callers
        self.last_device_result = {
            "schedule": sched, "makespan": float(res.summary[0, 0]["makespan"]), "hits": int(res.summary[0, 0]["hits"]),
            "misses": int(res.summary[0, 0]["misses"]),
            "fin": {nd.id: float(res.fin[0, k]) for k, nd in enumerate(node_list)}}
        return sched

    def _pending_order(self) -> List[str]:
        """Iteration order of the pending set = the reference's tie-break among ready tasks (ref schedulers.py:58)."""
        return list(self.pending_tasks)

    def _lift(self, batch: DagBatch, res: ResultArrays, dag, node_list: List[Node]):
        summ = res.summary[0, 0]
        self.last_summary = summ
        if summ["status"] & N.ST_BAD_TOPOLOGY:
            raise RecursionError("dependency cycle in the task graph")  # ref: compute_depth recursion overflows
        if summ["status"] & N.ST_CARRY:
            raise RuntimeError("MRU: assignment refused after node choice (fp rounding corner, "
                               "ref schedulers.py:82-83); not replayed on the device")
        node_of = res.node_of[0]
        order, nodes = sequence_of(batch, res, 0, 0)
        for i, k in zip(order.tolist(), nodes.tolist()):
            t, nd = self._added[i], node_list[k]
            t.assigned_node = nd.id
            t.completed = True
            nd.completed_tasks.append(t.id)
            nd.last_used_params.extend(t.params_needed)
            self.completed_tasks.add(t.id)
            self.pending_tasks.discard(t.id)
        for i in np.nonzero(node_of == N.TASK_FAILED)[0].tolist():
            tid = self._added[i].id
            self.failed_tasks.add(tid)
            self.pending_tasks.discard(tid)
        for k, nd in enumerate(node_list):
            nd.available_memory = float(res.avail[0, k])
            if res.cached is not None:
                bit = 1 << k
                nd.cached_params = {dag.param_names[p] for p in np.nonzero(res.cached[0] & bit)[0].tolist()}
                for p in nd.cached_params:
                    self.param_locations[p].add(nd.id)


class DFSScheduler(BaseScheduler):
    """ref schedulers.py:138-208"""
    ALGO = N.ALGO_DFS


class GreedyScheduler(BaseScheduler):
    """ref schedulers.py:211-296 (identify_sequential_chains is dead code in the reference and is not mirrored)"""
    ALGO = N.ALGO_GREEDY


class CriticalPathScheduler(BaseScheduler):
    """ref schedulers.py:299-372"""
    ALGO = N.ALGO_CRITICAL


class MRUScheduler(BaseScheduler):
    """ref schedulers.py:375-525"""
    ALGO = N.ALGO_MRU

    def __init__(self, nodes: List[Node]):
        super().__init__(nodes)
        self.param_usage_count = defaultdict(int)
        self.param_last_used = {}
        self.time_step = 0

    def _lift(self, batch, res, dag, node_list):
        step0 = self.time_step
        super()._lift(batch, res, dag, node_list)
        self.time_step += int(res.summary[0, 0]["rounds"])
        # usage counts and last-used steps (ref schedulers.py:514-516): a task of depth d is assigned in round d + 1
        pp, pi = dag.pred_ptr, dag.pred_idx
        depth = np.zeros(len(self._added), np.int64)
        changed = True  # (tasks may be listed in any order: relax until stable)
        while changed:
            changed = False
            for i in range(len(self._added)):
                if pp[i + 1] > pp[i]:
                    d = int(depth[pi[pp[i]:pp[i + 1]]].max()) + 1
                    if d != depth[i]:
                        depth[i] = d
                        changed = True
        order, _ = sequence_of(batch, res, 0, 0)
        for i in order.tolist():
            for p in self._added[i].params_needed:
                self.param_usage_count[p] += 1
                self.param_last_used[p] = step0 + int(depth[i]) + 1


SCHEDULERS = {"DFS": DFSScheduler, "Greedy": GreedyScheduler, "Critical": CriticalPathScheduler,
              "MRU_spec": MRUScheduler}
