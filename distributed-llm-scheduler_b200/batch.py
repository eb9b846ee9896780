"""Batched entry point `schedule_batch` and the lifting of device records back to reference-shaped results.

`schedule_batch(batch, algo_mask)` replaces, for every DAG instance and algorithm, the reference sequence
  scheduler = XScheduler(nodes); [scheduler.add_task(t) ...]; schedule = scheduler.schedule()
  makespan, stats = evaluator.simulate_execution(scheduler, schedule)
(ref simulation.py:320-336) with one call into the CUDA library. PyTorch is used only for device buffers,
pinned staging and streams.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _native as N
from .lowering import DagBatch, _FIELDS_ALL


def algos_of(mask: int) -> List[int]:
    return [a for a in range(N.NUM_ALGOS) if mask & (1 << a)]


@dataclass
class ResultArrays:
    """Host (numpy) image of dls_result_t for `algos` (slot order)."""
    algos: List[int]
    node_of: Optional[np.ndarray]   # i8  [A, NT]   (None in a summary-only result of PipelinedScheduler)
    seq_pos: Optional[np.ndarray]   # u16 [A, NT]   (likewise)
    summary: np.ndarray   # structured [A, B]
    fin: np.ndarray       # f64 [A, NK]
    cnt: np.ndarray       # i32 [A, NK]
    avail: Optional[np.ndarray]     # f64 [A, NK]   (likewise)
    peak: Optional[np.ndarray]    # f64 [A, NK]
    cached: Optional[np.ndarray]  # u32 [A, NP]
    tfin: Optional[np.ndarray] = None  # f64 [A, NK] extension: node finish times on the dependency-aware timeline

    @classmethod
    def empty(cls, batch: DagBatch, mask: int, peak: bool = True, cached: bool = True, timeline: bool = False) -> "ResultArrays":
        al = algos_of(mask)
        A = len(al)
        return cls(
            algos=al, node_of=np.full((A, batch.nt_inst), N.TASK_PENDING, np.int8),
            seq_pos=np.full((A, batch.nt_inst), N.SEQ_NONE, np.uint16),
            summary=np.zeros((A, batch.B), np.dtype(N.SUMMARY_DTYPE)), fin=np.zeros((A, batch.nk_inst), np.float64),
            cnt=np.zeros((A, batch.nk_inst), np.int32), avail=np.zeros((A, batch.nk_inst), np.float64),
            peak=np.zeros((A, batch.nk_inst), np.float64) if peak else None,
            cached=np.zeros((A, batch.np_inst), np.uint32) if cached else None,
            tfin=np.zeros((A, batch.nk_inst), np.float64) if timeline else None)

    def slot(self, algo: int) -> int:
        return self.algos.index(algo)

    def host_struct(self) -> N.DlsResult:
        r = N.DlsResult()
        for f in ("node_of", "seq_pos", "summary", "fin", "cnt", "avail", "peak", "cached", "tfin"):
            a = getattr(self, f)
            setattr(r, f, None if a is None else a.ctypes.data)
        return r

    def dependency_aware_makespan(self, batch: DagBatch) -> np.ndarray:
        """[A, B] makespan on the dependency-aware timeline (extension, needs `timeline=True`): max of `tfin` over the
        nodes of the instance that ran at least one task, 0.0 if none did."""
        if self.tfin is None:
            raise ValueError("no timeline in this result: ask for it with timeline=True")
        out = np.zeros((len(self.algos), batch.B))
        for b in range(batch.B):
            ks = batch.node_slice(b)
            used = self.cnt[:, ks] > 0
            out[:, b] = np.where(used, self.tfin[:, ks], 0.0).max(axis=1) if used.shape[1] else 0.0
        return out


# ---- lifting (SURVEY.md section 8a' "RESULT LIFT") ---------------------------------------------------------

def sequence_of(batch: DagBatch, res: ResultArrays, slot: int, b: int):
    """(task indices in assignment order, their node indices) of instance b."""
    ts = batch.task_slice(b)
    node_of = res.node_of[slot, ts]
    seq_pos = res.seq_pos[slot, ts]
    done = np.nonzero(node_of >= 0)[0]
    order = done[np.argsort(seq_pos[done], kind="stable")]
    return order, node_of[order]


def schedule_dict(batch: DagBatch, res: ResultArrays, slot: int, b: int, task_names: Sequence[str],
                  node_ids: Sequence[str]) -> Dict[str, List[str]]:
    """The dict `schedule()` returns: node id -> task ids in assignment order, keys in order of first
    assignment, nodes without tasks absent (ref schedulers.py:156,196,208)."""
    order, nodes = sequence_of(batch, res, slot, b)
    out: Dict[str, List[str]] = {}
    for i, k in zip(order.tolist(), nodes.tolist()):
        out.setdefault(node_ids[k], []).append(task_names[i])
    return out


# ---- device path -----------------------------------------------------------------------------------------

def _torch():
    import torch
    return torch


_NP_VIEW = {np.dtype(np.uint16): np.int16, np.dtype(np.uint32): np.int32}


def _to_torch_host(a: np.ndarray, pin: bool):
    torch = _torch()
    a = np.ascontiguousarray(a)
    v = a.view(_NP_VIEW.get(a.dtype, a.dtype))
    t = torch.from_numpy(v)
    if pin and t.numel() > 0:
        t = t.pin_memory()
    return t


class DeviceBatch:
    """dls_batch_t whose arrays live in HBM (torch tensors keep them alive)."""

    def __init__(self, host: DagBatch, device, non_blocking: bool = True, pinned: Optional[dict] = None):
        torch = _torch()
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: the scheduling path has no CPU implementation")
        self.host = host.validate()
        self.device = torch.device(device)
        self.tensors = {}
        self.h2d_bytes = 0
        for f in _FIELDS_ALL:
            a = getattr(host, f)
            if a is None:
                continue
            src = pinned[f] if pinned is not None else _to_torch_host(a, pin=False)
            self.tensors[f] = src.to(self.device, non_blocking=non_blocking)
            self.h2d_bytes += src.numel() * src.element_size()
        self.struct = host.as_struct(lambda f, a: self.tensors[f].data_ptr())


    @classmethod
    def from_tensors(cls, meta: DagBatch, tensors: dict, device) -> "DeviceBatch":
        """Wrap arrays that already live on the device (e.g. built by gen.random_dag_batch_device). `meta` carries
        the scalar fields and the (small) offset tables on the host; its large arrays may be None."""
        self = cls.__new__(cls)
        self.host = meta.validate()
        self.device = _torch().device(device)
        self.tensors = dict(tensors)
        self.h2d_bytes = 0
        self.struct = meta.as_struct_present(lambda f: self.tensors[f].data_ptr() if f in self.tensors else None)
        return self

    def to_host(self) -> DagBatch:
        """Copy every array back (device-generated batches: to feed the CPU oracle in tests)."""
        import dataclasses
        vals = {}
        for f, t in self.tensors.items():
            a = t.cpu().numpy()
            want = {"pred_ptr": np.uint32, "par_ptr": np.uint32, "pred_idx": np.uint16, "par_idx": np.uint16,
                    "rank": np.uint16}.get(f)
            vals[f] = a.view(want) if want is not None else a
        return dataclasses.replace(self.host, **vals)


def _row_lengths(ptr: np.ndarray, task_off: np.ndarray) -> Optional[np.ndarray]:
    """u8 row lengths of a concatenation of per-topology row-pointer arrays ([n_t + 1] each), or None if a row is
    longer than 255."""
    T = len(task_off) - 1
    d = np.diff(ptr.astype(np.int64))
    if T > 1:  # drop the differences that straddle two topologies
        d = np.delete(d, (task_off[1:-1] + np.arange(T - 1)).astype(np.int64))
    if d.size and (d.max() > 255 or d.min() < 0):
        return None
    return d.astype(np.uint8)


def pin_batch(host: DagBatch) -> dict:
    """Pinned host staging copies of a batch's arrays (for timed host->device transfers)."""
    return {f: _to_torch_host(getattr(host, f), pin=True) for f in _FIELDS_ALL if getattr(host, f) is not None}


class BatchResult:
    """Device-resident outputs of one `schedule_batch` call; `.host()` brings them back as ResultArrays."""

    def __init__(self, batch: DagBatch, mask: int, device, peak: bool, cached: bool, timeline: bool = False):
        torch = _torch()
        self.batch, self.mask, self.algos = batch, mask, algos_of(mask)
        A = len(self.algos)
        dev = device
        self.node_of = torch.empty((A, batch.nt_inst), dtype=torch.int8, device=dev)
        self.seq_pos = torch.empty((A, batch.nt_inst), dtype=torch.int16, device=dev)
        self.summary = torch.empty((A, batch.B, N.SUMMARY_BYTES), dtype=torch.uint8, device=dev)
        self.fin = torch.empty((A, batch.nk_inst), dtype=torch.float64, device=dev)
        self.cnt = torch.empty((A, batch.nk_inst), dtype=torch.int32, device=dev)
        self.avail = torch.empty((A, batch.nk_inst), dtype=torch.float64, device=dev)
        self.peak = torch.empty((A, batch.nk_inst), dtype=torch.float64, device=dev) if peak else None
        self.cached = torch.empty((A, batch.np_inst), dtype=torch.int32, device=dev) if cached else None
        self.tfin = torch.empty((A, batch.nk_inst), dtype=torch.float64, device=dev) if timeline else None
        r = N.DlsResult()
        for f in ("node_of", "seq_pos", "summary", "fin", "cnt", "avail", "peak", "cached", "tfin"):
            t = getattr(self, f)
            setattr(r, f, None if t is None else t.data_ptr())
        self.struct = r

    def summary_host(self) -> np.ndarray:
        """Only the fixed-size per-instance records (what multi-GPU runs all-gather)."""
        return self.summary.cpu().numpy().view(np.dtype(N.SUMMARY_DTYPE)).reshape(len(self.algos), self.batch.B)

    def d2h_bytes(self, full: bool = True) -> int:
        ts = [self.summary] + ([self.node_of, self.seq_pos, self.fin, self.cnt, self.avail] if full else [])
        if full:
            ts += [t for t in (self.peak, self.cached, self.tfin) if t is not None]
        return sum(t.numel() * t.element_size() for t in ts)

    def host(self) -> ResultArrays:
        A = len(self.algos)
        return ResultArrays(
            algos=self.algos, node_of=self.node_of.cpu().numpy(),
            seq_pos=self.seq_pos.cpu().numpy().view(np.uint16),
            summary=self.summary_host(), fin=self.fin.cpu().numpy(), cnt=self.cnt.cpu().numpy(),
            avail=self.avail.cpu().numpy(), peak=None if self.peak is None else self.peak.cpu().numpy(),
            cached=None if self.cached is None else self.cached.cpu().numpy().view(np.uint32),
            tfin=None if self.tfin is None else self.tfin.cpu().numpy())


def workspace_bytes(dev_batch: DeviceBatch, mask: int) -> int:
    return int(N.load().dls_workspace_bytes(C.byref(dev_batch.struct), mask))


def schedule_batch(batch, algo_mask: int = N.MASK_ALL, device=None, peak: bool = False, cached: bool = False,
                   workspace=None, result: Optional[BatchResult] = None, timeline: bool = False) -> BatchResult:
    """Run the algorithms in `algo_mask` over every instance of `batch` on the GPU.

    `batch` is a host `DagBatch` (copied to HBM here) or an already resident `DeviceBatch`. Work is enqueued
    on torch's current stream of `device`; the returned BatchResult holds device tensors (synchronise / call
    `.host()` to read them). Raises DlsError on a library error; infeasible tasks are data, not errors.

    `peak` (extension: per-node peak memory) and `cached` (final Node.cached_params as bitmasks) are produced by
    the general warp-per-DAG kernels only; asking for either keeps the whole batch off the faster thread-per-DAG
    path, so leave them off for sweeps. The same holds for `timeline` (extension: node finish times on a
    dependency-aware timeline, see dls_result_t.tfin in include/dls.h and ResultArrays.dependency_aware_makespan).
    """
    torch = _torch()
    lib = N.load()
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device: the scheduling path has no CPU implementation")
    if isinstance(batch, DagBatch):
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        batch = DeviceBatch(batch, device)
    device = batch.device
    with torch.cuda.device(device):
        if result is None:
            result = BatchResult(batch.host, algo_mask, device, peak, cached, timeline)
        need = int(lib.dls_workspace_bytes(C.byref(batch.struct), algo_mask))
        if workspace is None or workspace.numel() < need:
            workspace = torch.empty(max(need, 8), dtype=torch.uint8, device=device)
        stream = torch.cuda.current_stream(device).cuda_stream
        rc = lib.dls_schedule_batch(C.byref(batch.struct), C.byref(result.struct), algo_mask, workspace.data_ptr(),
                                    workspace.numel(), C.c_void_p(stream))
        N.check(rc)
    result.workspace = workspace  # keep alive until the stream has run
    result.dev_batch = batch
    return result


def aggregate(result: BatchResult, group_of=None, n_groups: int = 1) -> Dict[str, np.ndarray]:
    """Group means of the sweep on the device (ref analyze_results' groupby tables, simulation.py:518-524): returns
    {field: [A, G] array}. `group_of` is an int array [B] (host or device); makespan and load balance are averaged over
    instances that completed at least one task, like the reference (`df[df.completed_tasks > 0]`)."""
    torch = _torch()
    lib = N.load()
    dev_batch = result.dev_batch
    device = dev_batch.device
    A = len(result.algos)
    with torch.cuda.device(device):
        g = None
        if group_of is not None:
            g = torch.as_tensor(np.asarray(group_of, np.int32) if not torch.is_tensor(group_of) else group_of,
                                dtype=torch.int32, device=device)
        out = torch.empty((A, n_groups, N.AGG_FIELDS), dtype=torch.float64, device=device)
        stream = torch.cuda.current_stream(device).cuda_stream
        N.check(lib.dls_aggregate(C.byref(dev_batch.struct), C.byref(result.struct), result.mask,
                                  None if g is None else g.data_ptr(), n_groups, out.data_ptr(), C.c_void_p(stream)))
        sums = out.cpu().numpy()
    cnt = np.maximum(sums[..., 0], 1.0)
    with_c = np.maximum(sums[..., 7], 1.0)
    return {"count": sums[..., 0], "completion_rate": sums[..., 1] / cnt, "makespan_all": sums[..., 2] / cnt,
            "makespan": sums[..., 2] / with_c, "avg_node_utilization": sums[..., 3] / cnt,
            "load_balance_score": sums[..., 4] / with_c, "param_cache_hits": sums[..., 5],
            "param_cache_misses": sums[..., 6]}


def priorities(batch, device=None):
    """Kernel P alone: returns dict of numpy arrays depth/outdeg/up/order_dg/order_mru/order_cp."""
    torch = _torch()
    lib = N.load()
    if isinstance(batch, DagBatch):
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        batch = DeviceBatch(batch, device)
    hb = batch.host
    dev = batch.device
    with torch.cuda.device(dev):
        out = {
            "depth": torch.empty(hb.nt_topo, dtype=torch.int16, device=dev),
            "outdeg": torch.empty(hb.nt_topo, dtype=torch.int16, device=dev),
            "up": torch.empty(hb.nt_inst, dtype=torch.float64, device=dev),
            "order_dg": torch.empty(hb.nt_topo, dtype=torch.int16, device=dev),
            "order_mru": torch.empty(hb.nt_topo, dtype=torch.int16, device=dev),
            "order_cp": torch.empty(hb.nt_inst, dtype=torch.int16, device=dev),
        }
        pr = N.DlsPrio()
        for k, t in out.items():
            setattr(pr, k, t.data_ptr())
        need = int(lib.dls_workspace_bytes(C.byref(batch.struct), N.MASK_ALL))
        ws = torch.empty(max(need, 8), dtype=torch.uint8, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        N.check(lib.dls_priorities(C.byref(batch.struct), C.byref(pr), ws.data_ptr(), ws.numel(), C.c_void_p(stream)))
        torch.cuda.synchronize(dev)
    res = {}
    for k, t in out.items():
        a = t.cpu().numpy()
        res[k] = a.view(np.uint16) if a.dtype == np.int16 else a
    return res


class PipelinedScheduler:
    """End-to-end path for HOST batches: instances are cut into chunks and pushed through `nstreams` CUDA streams so
    that the host->device copy of chunk i+1, the kernels of chunk i and the device->host copy of chunk i-1 overlap
    (PCIe is full duplex; the kernels need neither direction). Inputs are staged in pinned memory once per batch
    object; results land in pinned host arrays (`.out`, same layout as ResultArrays). Everything a pass needs —
    device buffers per stream, the descriptor of every chunk, the list of copies — is built here, once: `run()` only
    enqueues copies and library calls (no allocation, no lowering), which is what lets the host keep three streams fed.
    """

    def __init__(self, host: DagBatch, algo_mask: int = N.MASK_ALL, device=None, chunks: int = 8, nstreams: int = 3,
                 compact: bool = True, results: str = "full"):
        """`results`: "full" brings every result array back to the host; "summary" only what a sweep's result table
        is made of — the per-instance records, node finish times and task counts (what evaluate_batch reads; 544 bytes
        per 8-node instance and algorithm instead of 3.2 KB for a 1k-task DAG) — per-task outputs stay on the device
        ("per-task outputs stay sharded unless requested")."""
        torch = _torch()
        lib = N.load()
        if host.T != host.B:
            raise ValueError("pipelining shards by instance: needs one topology per instance (T == B)")
        if results not in ("full", "summary"):
            raise ValueError("results must be 'full' or 'summary'")
        self.results = results
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.mask, self.algos = algo_mask, algos_of(algo_mask)
        A = len(self.algos)
        # The copy to the device is what bounds this path (PCIe), so ship fewer bytes where the device can rebuild
        # them: row pointers as one-byte row lengths (dls_rows_from_degrees), one rank row for a homogeneous batch
        # (dls_repeat_rank_row). 32 KB -> 24 KB per 1k-task instance.
        deg, rank_row, skip = {}, None, set()
        if compact:
            for f in ("pred_ptr", "par_ptr"):
                d = _row_lengths(getattr(host, f), host.task_off)
                if d is not None:
                    deg[f] = torch.from_numpy(d).pin_memory()
                    skip.add(f)
            ns = np.diff(host.task_off)
            if host.B > 1 and (ns == ns[0]).all() and ns[0] > 0:
                rows = host.rank.reshape(host.B, int(ns[0]))
                if (rows == rows[0]).all():
                    rank_row = torch.from_numpy(np.ascontiguousarray(rows[0]).view(np.int16)).pin_memory()
                    skip.add("rank")
        self._skip = skip
        # pinned copies of the inputs that travel as they are
        pinned = {f: _to_torch_host(getattr(host, f), pin=True) for f in _FIELDS_ALL
                  if getattr(host, f) is not None and f not in skip}
        self._pinned, self._deg, self._rank_row = pinned, deg, rank_row
        self.host = host
        chunks = max(1, min(chunks, host.B))
        bounds = [(c * host.B) // chunks for c in range(chunks + 1)]
        parts = [(lo, hi, host.shard(lo, hi)) for lo, hi in zip(bounds[:-1], bounds[1:]) if hi > lo]
        self.parts = parts
        self.streams = [torch.cuda.Stream(self.device) for _ in range(min(nstreams, len(parts)))]
        S = len(self.streams)

        # where the arrays of chunk [lo, hi) start in the batch's arrays (offset tables are rebased copies, see below)
        def span(f, lo, hi):
            t0, t1 = int(host.task_off[lo]), int(host.task_off[hi])
            if f in ("pred_ptr", "par_ptr"):
                return t0 + lo, t1 + hi
            if f in ("rank", "mem", "dur"):
                return t0, t1
            if f == "pred_idx":
                return int(host.pred_off[lo]), int(host.pred_off[hi])
            if f == "par_idx":
                return int(host.par_off[lo]), int(host.par_off[hi])
            if f in ("cap", "spd"):
                return (int(host.node_off[lo]), int(host.node_off[hi])) if host.node_off is not None else (0, len(host.cap))
            raise KeyError(f)

        tables = ("task_off", "pred_off", "par_off", "pid_off", "node_off")
        size = {}  # largest chunk, per array
        for lo, hi, part in parts:
            for f in _FIELDS_ALL:
                a = getattr(part, f)
                if a is not None:
                    size[f] = max(size.get(f, 0), int(a.size))
        pin = lambda shape, dt: torch.empty(shape, dtype=dt).pin_memory()
        self.out = {"summary": pin((A, host.B, N.SUMMARY_BYTES), torch.uint8), "fin": pin((A, host.nk_inst), torch.float64),
                    "cnt": pin((A, host.nk_inst), torch.int32)}
        if results == "full":
            self.out.update({"node_of": pin((A, host.nt_inst), torch.int8), "seq_pos": pin((A, host.nt_inst), torch.int16),
                             "avail": pin((A, host.nk_inst), torch.float64)})
        need = 0
        for _, _, part in parts:
            need = max(need, int(lib.dls_workspace_bytes(C.byref(part.as_struct(lambda f, a: None)), algo_mask)))
        nt_max, b_max, nk_max = (max(p[2].nt_inst for p in parts), max(p[2].B for p in parts), max(p[2].nk_inst for p in parts))
        with torch.cuda.device(self.device):
            self._slot = []
            for _ in range(S):
                buf = {f: torch.empty(n, dtype=_to_torch_host(np.empty(0, getattr(host, f).dtype), False).dtype,
                                      device=self.device) for f, n in size.items()}
                if deg:
                    buf["_deg"] = {f: torch.empty(size["rank"], dtype=torch.uint8, device=self.device) for f in deg}
                if rank_row is not None:
                    buf["_rank_row"] = torch.empty(rank_row.numel(), dtype=torch.int16, device=self.device)
                buf["_ws"] = torch.empty(max(need, 8), dtype=torch.uint8, device=self.device)
                buf["_res"] = {"node_of": torch.empty(A * nt_max, dtype=torch.int8, device=self.device),
                               "seq_pos": torch.empty(A * nt_max, dtype=torch.int16, device=self.device),
                               "summary": torch.empty(A * b_max * N.SUMMARY_BYTES, dtype=torch.uint8, device=self.device),
                               "fin": torch.empty(A * nk_max, dtype=torch.float64, device=self.device),
                               "cnt": torch.empty(A * nk_max, dtype=torch.int32, device=self.device),
                               "avail": torch.empty(A * nk_max, dtype=torch.float64, device=self.device)}
                r = N.DlsResult()
                for f, t in buf["_res"].items():
                    setattr(r, f, t.data_ptr())
                buf["_rstruct"] = r
                self._slot.append(buf)
        # the plan: per chunk, what to copy where and the descriptor to hand to the library
        self._plan, self._tables = [], []
        h2d = 0
        for c, (lo, hi, part) in enumerate(parts):
            buf = self._slot[c % S]
            copies, expand = [], []
            for f in _FIELDS_ALL:
                a = getattr(part, f)
                if a is None:
                    continue
                if f in tables:  # rebased by shard(): small arrays of their own
                    src = _to_torch_host(a, pin=True)
                    self._tables.append(src)
                elif f in skip:
                    continue
                else:
                    a0, a1 = span(f, lo, hi)
                    src = pinned[f][a0:a1]
                copies.append((buf[f][:src.numel()], src))
                h2d += src.numel() * src.element_size()
            t0, t1 = int(host.task_off[lo]), int(host.task_off[hi])
            for f, d in deg.items():
                copies.append((buf["_deg"][f][:t1 - t0], d[t0:t1]))
                h2d += t1 - t0
                expand.append(("rows", buf["_deg"][f].data_ptr(), buf["task_off"].data_ptr(), hi - lo, buf[f].data_ptr()))
            if rank_row is not None:
                copies.append((buf["_rank_row"], rank_row))
                h2d += 2 * rank_row.numel()
                expand.append(("rank", buf["_rank_row"].data_ptr(), rank_row.numel(), hi - lo, buf["rank"].data_ptr()))
            struct = part.as_struct_present(lambda f: buf[f].data_ptr() if f in buf else None)
            k0, k1 = (int(host.node_off[lo]), int(host.node_off[hi])) if host.node_off is not None else (
                lo * host.K_max, hi * host.K_max)
            back = []
            for a in range(A):  # one contiguous async copy per array and algorithm slot (strided copies would be staged)
                res = buf["_res"]
                nt, nb, nk = part.nt_inst, part.B, part.nk_inst
                if results == "full":
                    back.append((self.out["node_of"][a, t0:t1], res["node_of"][a * nt:(a + 1) * nt]))
                    back.append((self.out["seq_pos"][a, t0:t1], res["seq_pos"][a * nt:(a + 1) * nt]))
                    back.append((self.out["avail"][a, k0:k1], res["avail"][a * nk:(a + 1) * nk]))
                back.append((self.out["summary"][a, lo:hi].view(-1),
                             res["summary"][a * nb * N.SUMMARY_BYTES:(a + 1) * nb * N.SUMMARY_BYTES]))
                back.append((self.out["fin"][a, k0:k1], res["fin"][a * nk:(a + 1) * nk]))
                back.append((self.out["cnt"][a, k0:k1], res["cnt"][a * nk:(a + 1) * nk]))
            self._plan.append((c % S, copies, expand, struct, back))
        self.h2d_bytes = h2d
        self.d2h_bytes = sum(t.numel() * t.element_size() for t in self.out.values())

    def run(self):
        """One pass over the whole batch. Returns after everything has been ENQUEUED; call `.synchronize()`."""
        torch = _torch()
        lib = N.load()
        for slot, copies, expand, struct, back in self._plan:
            st = self.streams[slot]
            buf = self._slot[slot]
            stream = C.c_void_p(st.cuda_stream)
            with torch.cuda.stream(st):
                for dst, src in copies:
                    dst.copy_(src, non_blocking=True)
                for e in expand:
                    if e[0] == "rows":
                        N.check(lib.dls_rows_from_degrees(e[1], e[2], e[3], e[4], stream))
                    else:
                        N.check(lib.dls_repeat_rank_row(e[1], e[2], e[3], e[4], stream))
                N.check(lib.dls_schedule_batch(C.byref(struct), C.byref(buf["_rstruct"]), self.mask, buf["_ws"].data_ptr(),
                                               buf["_ws"].numel(), stream))
                for dst, src in back:
                    dst.copy_(src, non_blocking=True)
        return self

    def synchronize(self):
        for st in self.streams:
            st.synchronize()
        return self

    def result(self) -> ResultArrays:
        self.synchronize()
        o = self.out
        full = self.results == "full"
        return ResultArrays(
            algos=self.algos, node_of=o["node_of"].numpy() if full else None,
            seq_pos=o["seq_pos"].numpy().view(np.uint16) if full else None,
            summary=o["summary"].numpy().view(np.dtype(N.SUMMARY_DTYPE)).reshape(len(self.algos), self.host.B),
            fin=o["fin"].numpy(), cnt=o["cnt"].numpy(), avail=o["avail"].numpy() if full else None, peak=None, cached=None)
