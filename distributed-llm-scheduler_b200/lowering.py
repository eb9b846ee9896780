"""Lowering: reference-style `Task` / `Node` objects -> the CSR + value arrays of include/dls.h.

This is the "DAG builder emits CSR adjacency plus per-task duration and memory arrays" step of the north
star. The arrays follow SURVEY.md section 8(a'):

  idx      task name -> 0..n-1 in add_task order                      (ref schedulers.py:43-48)
  pred     dependencies in list order                                  (ref schedulers.py:14)
  par      distinct params; param id = lexicographic rank of its name  (ref schedulers.py:420 sorts (score, name))
  rank     position in list(scheduler.pending_tasks) after the last add_task — CPython set iteration order,
           which is what breaks ties among ready tasks in the reference (ref schedulers.py:41,58)
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import _native as N


def set_order(names: Sequence[str]) -> List[str]:
    """Iteration order of a Python set filled by `.add(name)` in the given order (ref schedulers.py:41,46).

    Must run in the SAME process as whatever it is compared with: str hashes depend on PYTHONHASHSEED.
    """
    s = set()
    for nm in names:
        s.add(nm)
    return list(s)


@dataclass
class LoweredDag:
    names: List[str]
    param_names: List[str]
    pred_ptr: np.ndarray  # u32 [n+1]
    pred_idx: np.ndarray  # u16 [E]
    par_ptr: np.ndarray   # u32 [n+1]
    par_idx: np.ndarray   # u16 [R]
    rank: np.ndarray      # u16 [n]
    mem: np.ndarray       # f64 [n]
    dur: np.ndarray       # f64 [n]

    @property
    def n(self) -> int:
        return len(self.rank)

    @property
    def E(self) -> int:
        return len(self.pred_idx)

    @property
    def R(self) -> int:
        return len(self.par_idx)

    @property
    def P(self) -> int:
        return len(self.param_names)


def lower_tasks(tasks: Sequence, pending: Optional[Iterable[str]] = None) -> LoweredDag:
    """Lower a list of Task-like objects (attributes id, memory_required, compute_time, dependencies,
    params_needed — ref schedulers.py:9-17), given in add_task order.

    `pending` is the live `scheduler.pending_tasks` set (or its list()); if omitted, the set order is
    reproduced by filling a fresh set in add order, which is what add_task does.
    """
    names = [t.id for t in tasks]
    n = len(names)
    if n > N.MAX_TASKS:
        raise ValueError(f"{n} tasks > {N.MAX_TASKS}")
    idx = {}
    for i, nm in enumerate(names):
        if nm in idx:
            raise ValueError(f"duplicate task id {nm!r}")
        idx[nm] = i
    order = list(pending) if pending is not None else set_order(names)
    if len(order) != n or set(order) != set(names):
        raise ValueError("pending set does not match the task list")
    rank = np.empty(n, np.uint16)
    for r, nm in enumerate(order):
        rank[idx[nm]] = r
    all_params = set()
    for t in tasks:
        all_params.update(t.params_needed)
    param_names = sorted(all_params)
    if len(param_names) > N.MAX_PARAMS:
        raise ValueError(f"{len(param_names)} params > {N.MAX_PARAMS}")
    pid = {p: i for i, p in enumerate(param_names)}
    pred_ptr = np.zeros(n + 1, np.uint32)
    par_ptr = np.zeros(n + 1, np.uint32)
    preds: List[int] = []
    pars: List[int] = []
    for i, t in enumerate(tasks):
        for d in t.dependencies:
            if d not in idx:
                # ref: DFS raises KeyError in compute_depth (schedulers.py:145); the others never find the
                # task ready. Neither is a schedule; reject at the boundary.
                raise KeyError(d)
            preds.append(idx[d])
        pred_ptr[i + 1] = len(preds)
        pars.extend(sorted(pid[p] for p in set(t.params_needed)))
        par_ptr[i + 1] = len(pars)
    return LoweredDag(
        names=names, param_names=param_names, pred_ptr=pred_ptr,
        pred_idx=np.asarray(preds, np.uint16), par_ptr=par_ptr, par_idx=np.asarray(pars, np.uint16), rank=rank,
        mem=np.asarray([float(t.memory_required) for t in tasks], np.float64),
        dur=np.asarray([float(t.compute_time) for t in tasks], np.float64))


def lower_nodes(nodes: Sequence) -> Tuple[np.ndarray, np.ndarray]:
    """Node list (ref schedulers.py:21-25,34) -> (cap, spd) fp64 arrays in list order."""
    if len(nodes) > N.MAX_NODES:
        raise ValueError(f"{len(nodes)} nodes > {N.MAX_NODES}")
    return (np.asarray([float(x.total_memory) for x in nodes], np.float64),
            np.asarray([float(x.compute_speed) for x in nodes], np.float64))


_FIELDS_TOPO = ("task_off", "pred_off", "par_off", "pid_off", "pred_ptr", "pred_idx", "par_ptr", "par_idx", "rank")
_FIELDS_ALL = _FIELDS_TOPO + ("mem", "dur", "node_off", "cap", "spd")


@dataclass
class DagBatch:
    """Host (numpy) image of dls_batch_t."""
    B: int
    T: int
    n_max: int
    K_max: int
    P_max: int
    E_max: int
    R_max: int
    flags: int
    task_off: np.ndarray
    pred_off: np.ndarray
    par_off: np.ndarray
    pid_off: np.ndarray
    pred_ptr: np.ndarray
    pred_idx: np.ndarray
    par_ptr: np.ndarray
    par_idx: np.ndarray
    rank: np.ndarray
    mem: np.ndarray
    dur: np.ndarray
    node_off: Optional[np.ndarray]
    cap: np.ndarray
    spd: np.ndarray
    dags: Optional[List[LoweredDag]] = field(default=None, repr=False)  # names, for lifting

    # ---- totals -------------------------------------------------------------------------------------
    @property
    def nt_topo(self) -> int:
        return int(self.task_off[-1])

    @property
    def nt_inst(self) -> int:
        return self.B * int(self.task_off[1]) if self.T == 1 else int(self.task_off[-1])

    @property
    def e_total(self) -> int:
        return int(self.pred_off[-1])

    @property
    def r_total(self) -> int:
        return int(self.par_off[-1])

    @property
    def nk_inst(self) -> int:
        return int(self.node_off[-1]) if self.node_off is not None else self.B * self.K_max

    @property
    def np_inst(self) -> int:
        return self.B * int(self.pid_off[1]) if self.T == 1 else int(self.pid_off[-1])

    # ---- per-instance slicing helpers -----------------------------------------------------------------
    def n_of(self, b: int) -> int:
        t = 0 if self.T == 1 else b
        return int(self.task_off[t + 1] - self.task_off[t])

    def task_slice(self, b: int) -> slice:
        if self.T == 1:
            n = int(self.task_off[1])
            return slice(b * n, (b + 1) * n)
        return slice(int(self.task_off[b]), int(self.task_off[b + 1]))

    def node_slice(self, b: int) -> slice:
        if self.node_off is not None:
            return slice(int(self.node_off[b]), int(self.node_off[b + 1]))
        return slice(b * self.K_max, (b + 1) * self.K_max)

    def node_in_slice(self, b: int) -> slice:
        if self.node_off is not None:
            return slice(int(self.node_off[b]), int(self.node_off[b + 1]))
        return slice(0, self.K_max)

    def param_slice(self, b: int) -> slice:
        if self.T == 1:
            p = int(self.pid_off[1])
            return slice(b * p, (b + 1) * p)
        return slice(int(self.pid_off[b]), int(self.pid_off[b + 1]))

    def input_bytes(self) -> int:
        """Bytes of the input arrays as laid out (what a host->device copy moves)."""
        tot = 0
        for f in _FIELDS_ALL:
            a = getattr(self, f)
            if a is not None:
                tot += a.nbytes
        return tot

    # ---- constructors -------------------------------------------------------------------------------------
    @classmethod
    def from_instances(cls, dags: Sequence[LoweredDag], nodes: Sequence[Tuple[np.ndarray, np.ndarray]]) -> "DagBatch":
        """One topology per instance (T == B)."""
        B = len(dags)
        assert B == len(nodes) and B > 0
        ns = np.array([d.n for d in dags], np.int64)
        es = np.array([d.E for d in dags], np.int64)
        rs = np.array([d.R for d in dags], np.int64)
        ps = np.array([d.P for d in dags], np.int64)
        ks = np.array([len(c) for c, _ in nodes], np.int64)
        off = lambda a: np.concatenate([[0], np.cumsum(a)]).astype(np.int64)
        return cls(
            B=B, T=B, n_max=int(ns.max()), K_max=int(ks.max()), P_max=int(ps.max()), E_max=int(es.max()),
            R_max=int(rs.max()), flags=0, task_off=off(ns), pred_off=off(es), par_off=off(rs), pid_off=off(ps),
            pred_ptr=np.concatenate([d.pred_ptr for d in dags]).astype(np.uint32),
            pred_idx=np.concatenate([d.pred_idx for d in dags]).astype(np.uint16),
            par_ptr=np.concatenate([d.par_ptr for d in dags]).astype(np.uint32),
            par_idx=np.concatenate([d.par_idx for d in dags]).astype(np.uint16),
            rank=np.concatenate([d.rank for d in dags]).astype(np.uint16),
            mem=np.concatenate([d.mem for d in dags]).astype(np.float64),
            dur=np.concatenate([d.dur for d in dags]).astype(np.float64),
            node_off=off(ks), cap=np.concatenate([c for c, _ in nodes]).astype(np.float64),
            spd=np.concatenate([s for _, s in nodes]).astype(np.float64), dags=list(dags))

    @classmethod
    def shared_topology(cls, dag: LoweredDag, cap: np.ndarray, spd: np.ndarray, B: int,
                        dur: Optional[np.ndarray] = None, mem: Optional[np.ndarray] = None) -> "DagBatch":
        """One topology shared by B instances (T == 1); `dur` / `mem` optionally [B, n] per instance."""
        n = dag.n
        flags = 0
        if dur is not None:
            dur = np.ascontiguousarray(dur, np.float64).reshape(B * n)
            flags |= N.F_DUR_PER_INSTANCE
        if mem is not None:
            mem = np.ascontiguousarray(mem, np.float64).reshape(B * n)
            flags |= N.F_MEM_PER_INSTANCE
        one = lambda v: np.array([0, v], np.int64)
        return cls(
            B=B, T=1, n_max=n, K_max=len(cap), P_max=dag.P, E_max=dag.E, R_max=dag.R, flags=flags,
            task_off=one(n), pred_off=one(dag.E), par_off=one(dag.R), pid_off=one(dag.P),
            pred_ptr=dag.pred_ptr.astype(np.uint32), pred_idx=dag.pred_idx.astype(np.uint16),
            par_ptr=dag.par_ptr.astype(np.uint32), par_idx=dag.par_idx.astype(np.uint16),
            rank=dag.rank.astype(np.uint16), mem=dag.mem if mem is None else mem,
            dur=dag.dur if dur is None else dur, node_off=None,
            cap=np.ascontiguousarray(cap, np.float64), spd=np.ascontiguousarray(spd, np.float64), dags=[dag])

    def shard(self, lo: int, hi: int) -> "DagBatch":
        """Instances [lo, hi) as a batch of their own (multi-GPU partitioning)."""
        assert 0 <= lo < hi <= self.B
        if self.T == 1:
            n = int(self.task_off[1])
            per = lambda a, on: a[lo * n:hi * n] if on else a
            return DagBatch(
                B=hi - lo, T=1, n_max=self.n_max, K_max=self.K_max, P_max=self.P_max, E_max=self.E_max,
                R_max=self.R_max, flags=self.flags, task_off=self.task_off, pred_off=self.pred_off,
                par_off=self.par_off, pid_off=self.pid_off, pred_ptr=self.pred_ptr, pred_idx=self.pred_idx,
                par_ptr=self.par_ptr, par_idx=self.par_idx, rank=self.rank,
                mem=per(self.mem, self.flags & N.F_MEM_PER_INSTANCE),
                dur=per(self.dur, self.flags & N.F_DUR_PER_INSTANCE),
                node_off=None if self.node_off is None else self.node_off[lo:hi + 1] - self.node_off[lo],
                cap=self.cap if self.node_off is None else self.cap[self.node_off[lo]:self.node_off[hi]],
                spd=self.spd if self.node_off is None else self.spd[self.node_off[lo]:self.node_off[hi]],
                dags=self.dags)
        t0, t1 = int(self.task_off[lo]), int(self.task_off[hi])
        e0, e1 = int(self.pred_off[lo]), int(self.pred_off[hi])
        r0, r1 = int(self.par_off[lo]), int(self.par_off[hi])
        reb = lambda a: (a[lo:hi + 1] - a[lo]).astype(np.int64)
        ns = np.diff(self.task_off[lo:hi + 1])
        if self.node_off is not None:
            k0, k1 = int(self.node_off[lo]), int(self.node_off[hi])
            node_off, cap, spd = reb(self.node_off), self.cap[k0:k1], self.spd[k0:k1]
            kmax = int(np.diff(node_off).max())
        else:
            node_off, cap, spd, kmax = None, self.cap, self.spd, self.K_max
        return DagBatch(
            B=hi - lo, T=hi - lo, n_max=int(ns.max()), K_max=kmax,
            P_max=int(np.diff(self.pid_off[lo:hi + 1]).max()), E_max=int(np.diff(self.pred_off[lo:hi + 1]).max()),
            R_max=int(np.diff(self.par_off[lo:hi + 1]).max()), flags=0,
            task_off=reb(self.task_off), pred_off=reb(self.pred_off), par_off=reb(self.par_off),
            pid_off=reb(self.pid_off), pred_ptr=self.pred_ptr[t0 + lo:t1 + hi], pred_idx=self.pred_idx[e0:e1],
            par_ptr=self.par_ptr[t0 + lo:t1 + hi], par_idx=self.par_idx[r0:r1], rank=self.rank[t0:t1],
            mem=self.mem[t0:t1], dur=self.dur[t0:t1], node_off=node_off, cap=cap, spd=spd,
            dags=None if self.dags is None else self.dags[lo:hi])

    def repeat(self, k: int) -> "DagBatch":
        """The batch followed by k-1 copies of itself (T == B only): cheap way to build large synthetic sweeps."""
        assert self.T == self.B and k >= 1
        if k == 1:
            return self
        til = lambda a: np.tile(a, k)
        offs = lambda a: np.concatenate([[0], np.cumsum(np.tile(np.diff(a), k))]).astype(np.int64)
        return DagBatch(
            B=self.B * k, T=self.B * k, n_max=self.n_max, K_max=self.K_max, P_max=self.P_max, E_max=self.E_max,
            R_max=self.R_max, flags=0, task_off=offs(self.task_off), pred_off=offs(self.pred_off),
            par_off=offs(self.par_off), pid_off=offs(self.pid_off), pred_ptr=til(self.pred_ptr),
            pred_idx=til(self.pred_idx), par_ptr=til(self.par_ptr), par_idx=til(self.par_idx), rank=til(self.rank),
            mem=til(self.mem), dur=til(self.dur), node_off=None if self.node_off is None else offs(self.node_off),
            cap=self.cap if self.node_off is None else til(self.cap),
            spd=self.spd if self.node_off is None else til(self.spd), dags=None)

    def validate(self):
        if self.K_max > N.MAX_NODES:
            raise ValueError(f"K_max={self.K_max} > {N.MAX_NODES}")
        if self.n_max > N.MAX_TASKS or self.P_max > N.MAX_PARAMS:
            raise ValueError("n_max / P_max beyond u16 ids")
        if self.T not in (1, self.B):
            raise ValueError("T must be 1 or B")
        return self

    # ---- ctypes struct over arbitrary pointers ------------------------------------------------------------
    def as_struct(self, ptr_of) -> N.DlsBatch:
        """Build dls_batch_t; `ptr_of(field_name, array)` returns the address to store (host or device)."""
        s = N.DlsBatch()
        s.B, s.T, s.n_max, s.K_max = self.B, self.T, self.n_max, self.K_max
        s.P_max, s.E_max, s.R_max, s.flags = self.P_max, self.E_max, self.R_max, self.flags
        s.nt_topo, s.nt_inst, s.e_total, s.r_total = self.nt_topo, self.nt_inst, self.e_total, self.r_total
        s.nk_inst, s.np_inst = self.nk_inst, self.np_inst
        for f in _FIELDS_ALL:
            a = getattr(self, f, None)
            setattr(s, f, None if a is None else ptr_of(f, a))
        return s

    def as_struct_present(self, ptr_of_name) -> N.DlsBatch:
        """dls_batch_t from addresses looked up by field name (arrays need not exist on the host)."""
        s = self.as_struct(lambda f, a: None)
        for f in _FIELDS_ALL:
            setattr(s, f, ptr_of_name(f))
        return s

    def host_struct(self):
        """dls_batch_t over the numpy arrays themselves (for the CPU oracle). Keeps the arrays alive."""
        keep = {}

        def ptr_of(f, a):
            a = np.ascontiguousarray(a)
            keep[f] = a
            return a.ctypes.data

        s = self.as_struct(ptr_of)
        s._keep = keep
        return s
