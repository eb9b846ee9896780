"""No-GPU checks of the boundary: the library builds for sm_100a, loads, exports every symbol include/dls.h
declares, validates arguments without a device, and the lowering reproduces the reference's array contract."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from common import ROOT
from dls_b200 import _native as N
from dls_b200.gen import random_dag_batch
from dls_b200.lowering import DagBatch, lower_nodes, lower_tasks, set_order
from dls_b200.schedulers import Node, Task


@pytest.fixture(scope="module")
def lib():
    N.build()
    return N.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "dls.h")).read()
    declared = set(re.findall(r"^\s*(?:const\s+)?[a-z_0-9]+\s*\*?\s*(dls_[a-z_]+)\(", hdr, re.M))
    assert declared == set(N.EXPORTED_SYMBOLS), declared ^ set(N.EXPORTED_SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.dls_version() == 200


def test_struct_layout_matches_header():
    assert C.sizeof(N.DlsSummary) == 48
    assert C.sizeof(N.DlsBatch) == 8 * 4 + 6 * 8 + 14 * 8
    assert C.sizeof(N.DlsResult) == 9 * 8 and C.sizeof(N.DlsPrio) == 6 * 8


def test_argument_validation_needs_no_device(lib):
    assert lib.dls_schedule_batch(None, None, 1, None, 0, None) == N.ERR_BAD_ARG
    b = random_dag_batch(2, 8, 4, 0.9, seed=0)
    s = b.host_struct()
    assert lib.dls_workspace_bytes(C.byref(s), N.MASK_ALL) > 0
    r = N.DlsResult()
    assert lib.dls_schedule_batch(C.byref(s), C.byref(r), 0, None, 0, None) == N.ERR_BAD_ARG
    s.K_max = 40
    assert lib.dls_schedule_batch(C.byref(s), C.byref(r), 1, None, 0, None) == N.ERR_TOO_LARGE
    assert b"K_max" in lib.dls_last_error()


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from dls_b200.batch import schedule_batch
    from dls_b200.schedulers import DFSScheduler
    with pytest.raises(RuntimeError, match="no CPU"):
        schedule_batch(random_dag_batch(2, 8, 4, 0.9, seed=0))
    s = DFSScheduler([Node("n", 1.0)])
    s.add_task(Task("a", 0.1, 0.1))
    with pytest.raises(RuntimeError, match="no CPU"):
        s.schedule()


def test_lowering_contract():
    tasks = [Task("t1", 1.0, 0.1, [], {"p1"}), Task("t2", 1.0, 0.1, ["t1"], {"p2"}),
             Task("t3", 1.0, 0.1, ["t1"], {"p3"}), Task("t4", 1.0, 0.1, ["t2", "t3", "t2"], {"p2", "p1"})]
    d = lower_tasks(tasks)
    assert d.names == ["t1", "t2", "t3", "t4"] and d.param_names == ["p1", "p2", "p3"]
    assert d.pred_ptr.tolist() == [0, 0, 1, 2, 5] and d.pred_idx.tolist() == [0, 0, 1, 2, 1]
    assert d.par_ptr.tolist() == [0, 1, 2, 3, 5] and d.par_idx.tolist() == [0, 1, 2, 0, 1]
    order = set_order(d.names)
    assert [d.names[i] for i in np.argsort(d.rank)] == order
    with pytest.raises(KeyError):
        lower_tasks([Task("a", 1, 1, ["missing"])])
    with pytest.raises(ValueError):
        lower_tasks([Task("a", 1, 1), Task("a", 1, 1)])
    cap, spd = lower_nodes([Node("n1", 3.0), Node("n2", 2.5, 1.5)])
    assert cap.tolist() == [3.0, 2.5] and spd.tolist() == [1.0, 1.5]
    b = DagBatch.from_instances([d, d], [(cap, spd), (cap[:1], spd[:1])])
    assert (b.B, b.T, b.n_max, b.K_max, b.nt_inst, b.nk_inst, b.np_inst) == (2, 2, 4, 2, 8, 3, 6)
    sh = b.shard(1, 2)
    assert sh.B == 1 and sh.K_max == 1 and sh.pred_ptr.tolist() == d.pred_ptr.tolist()


def check_random_dag_structure(b, n):
    """Invariants of generate_random_dag's shape (ref simulation.py:91-114): deps only on earlier tasks, <= 3 distinct
    deps, 1-2 private params, value ranges; param ids are lexicographic ranks of the names 'param_{i}_{j}'; node
    memory follows calculate_total_memory_needed / create_nodes_with_memory_regime (ref simulation.py:161-214)."""
    for i in range(b.B):
        pp = b.pred_ptr[i * (n + 1):(i + 1) * (n + 1)]
        pi = b.pred_idx[b.pred_off[i]:b.pred_off[i + 1]]
        deg = np.diff(pp)
        assert pp[0] == 0 and pp[-1] == len(pi)
        assert deg.max() <= 3 and deg[0] == 0 and deg[1] <= 1
        child = np.repeat(np.arange(n), deg)
        assert (pi < child).all()
        for t in range(n):
            assert len(set(pi[pp[t]:pp[t + 1]])) == deg[t]
        qp = b.par_ptr[i * (n + 1):(i + 1) * (n + 1)]
        qi = b.par_idx[b.par_off[i]:b.par_off[i + 1]]
        npar = np.diff(qp)
        assert set(npar.tolist()) <= {1, 2}
        assert sorted(qi.tolist()) == list(range(len(qi)))  # private params, ids dense
        names = [f"param_{t}_{j}" for t in range(n) for j in range(npar[t])]
        rank = {nm: r for r, nm in enumerate(sorted(names))}
        want = [sorted(rank[f"param_{t}_{j}"] for j in range(npar[t])) for t in range(n)]
        assert [qi[qp[t]:qp[t + 1]].tolist() for t in range(n)] == want
        mem = b.mem[i * n:(i + 1) * n]
        total = (mem + npar * 0.5).max() + len(qi) * 0.5
        K = int(b.node_off[i + 1] - b.node_off[i])
        cap = b.cap[b.node_off[i]:b.node_off[i + 1]]
        if K not in (2, 4):
            assert (cap == total * 0.9 / K).all() or np.allclose(cap, cap[0])
    assert 0.1 <= b.mem.min() and b.mem.max() <= 0.5 and 0.05 <= b.dur.min() and b.dur.max() <= 0.15
    assert abs(np.diff(b.pred_ptr.reshape(b.B, n + 1), axis=1).mean() - 1.5) < 0.1


def test_generator_matches_reference_distribution():
    check_random_dag_structure(random_dag_batch(32, 300, 8, 0.9, seed=3), 300)


def test_chain_of_half_additions_has_a_closed_form_when_exact():
    """dls_stream.cu::add_halves_exactly replaces MRU's eviction chain x + 0.5 + 0.5 + ... (ref schedulers.py:429, one
    rounding per param) by ONE addition whenever TwoSum says that addition is exact; the claim is that the chain then
    never rounds either, and that the rollback chain (ref :438-441) returns to x. Checked here in numpy on magnitudes,
    binade edges and signs the kernels see, including many cases where the one-shot sum is NOT exact (those keep the
    chain on the device)."""
    rng = np.random.default_rng(1)
    n = 300_000
    exp = rng.integers(-8, 12, n)
    x = np.ldexp(1 + rng.random(n), exp) * rng.choice([1, -1], n)
    edge = rng.random(n) < 0.3
    x[edge] = np.ldexp(1.0, exp[edge]) - np.ldexp(rng.integers(0, 8, edge.sum()).astype(float), exp[edge] - 53) * rng.choice([1, -1], edge.sum())
    j = rng.integers(0, 70, n)
    h = 0.5 * j
    s = x + h
    bb = s - x
    err = (x - (s - bb)) + (h - bb)
    exact = (err == 0) & (np.abs(x) < 2.0 ** 51) & (np.abs(s) < 2.0 ** 51)
    up = x.copy()
    for q in range(70):
        up[q < j] += 0.5
    down = up.copy()
    for q in range(70):
        down[q < j] -= 0.5
    assert 0.2 < exact.mean() < 0.9 and (up != s).sum() > 1000          # both kinds of case are in the sample
    assert (up[exact] == s[exact]).all() and (down[exact] == x[exact]).all()


def test_mirror_scheduler_rejects_state_it_would_ignore():
    """The device starts every node with an empty cache: nodes that already carry cached params or completed tasks
    are refused instead of being scheduled as if they were fresh. A scheduler that has run returns the
    empty schedule on a second call, like the reference's loop that finds nothing ready (ref schedulers.py:167-174)."""
    from dls_b200.schedulers import GreedyScheduler
    nd = Node("n1", 4.0)
    nd.cached_params.add("p1")
    s = GreedyScheduler([nd])
    s.add_task(Task("t1", 1.0, 0.1, [], {"p1"}))
    with pytest.raises(ValueError, match="cached params"):
        s.schedule()
    s2 = GreedyScheduler([Node("n1", 4.0)])
    s2.add_task(Task("t1", 1.0, 0.1, [], {"p1"}))
    s2.add_task(Task("t2", 1.0, 0.1, ["t1"], {"p2"}))
    s2.failed_tasks.add("t1")          # as after a run in which t1 failed and t2 stayed pending
    s2.pending_tasks.discard("t1")
    assert s2.schedule() == {}
