"""The evaluation-harness mirror (dls_b200.simulation; ref simulation.py:15-416): generators draw the same DAGs as
the reference for a given seed, the batched sweep reproduces the reference's TestResult fields."""
import random

import numpy as np
import pytest

import golden_io
import refharness as rh
from dls_b200 import _native as N
from dls_b200.lowering import lower_nodes, lower_tasks
from dls_b200.simulation import (CSV_COLUMNS, DAGGenerator, ImprovedSchedulerEvaluator, TestResult, evaluate_batch)
from dls_b200.schedulers import SCHEDULERS


def test_sweep_instances_equal_the_reference_sweep_fixture():
    """random.seed(0) + sweep_instances(3) must build exactly the 162 (DAG, node set) instances the reference's
    run_experiments(num_runs=3) builds (fixture cfg2 was written from the reference's own generators): same
    dependencies, params, memory, durations and node sets, in the same order."""
    cases = golden_io.load("cfg2_sweep_seed0.npz")
    random.seed(0)
    inst = ImprovedSchedulerEvaluator(SCHEDULERS).sweep_instances(3)
    assert len(inst) == len(cases) == 162
    for (dag_name, k, regime, tasks, nodes), c in zip(inst, cases):
        assert c.label.startswith(f"{dag_name}|K{k}|r{regime}|")
        d = lower_tasks(tasks)
        for f in ("pred_ptr", "pred_idx", "par_ptr", "par_idx", "mem", "dur"):
            np.testing.assert_array_equal(getattr(d, f), getattr(c.dag, f), err_msg=f"{c.label} {f}")
        cap, spd = lower_nodes(nodes)
        np.testing.assert_array_equal(cap, c.cap)
        np.testing.assert_array_equal(spd, c.spd)


def test_array_level_builders_equal_object_lowering():
    """llm / pipeline / GPT-2 builders that emit arrays directly give exactly what lowering the
    reference-style Task objects gives; the analytic GPT-2 DAG equals the fixture extracted from the real model."""
    from dls_b200.gen import gpt2_dag, llm_dag, pipeline_dag
    pairs = [(llm_dag(6, attention_heads=8), lower_tasks(DAGGenerator.generate_llm_dag(6, attention_heads=8))),
             (llm_dag(3, attention_heads=7, head_cap=7), lower_tasks(DAGGenerator.generate_llm_dag(3, attention_heads=7, head_cap=7))),
             (pipeline_dag(5, 3), lower_tasks(DAGGenerator.generate_pipeline_dag(5, 3)))]
    for a, b in pairs:
        assert a.names == b.names and a.param_names == b.param_names
        for f in ("pred_ptr", "pred_idx", "par_ptr", "par_idx", "rank", "mem", "dur"):
            np.testing.assert_array_equal(getattr(a, f), getattr(b, f), err_msg=f)
    g = gpt2_dag()
    fx = [c for c in golden_io.load("cfg3_gpt2.npz") if c.label == "gpt2|4laptops"][0].dag
    assert g.n == 99 and g.P == 75 and g.names == fx.names and g.param_names == fx.param_names
    for f in ("pred_ptr", "pred_idx", "par_ptr", "par_idx"):
        np.testing.assert_array_equal(getattr(g, f), getattr(fx, f), err_msg=f)
    np.testing.assert_array_equal(g.dur, fx.dur)
    np.testing.assert_array_equal(g.mem.view(np.uint64), fx.mem.view(np.uint64))  # bit-identical memory estimates
    big = llm_dag(48, attention_heads=205, head_cap=205)
    assert big.n == 9986 and big.E == 19777


@pytest.mark.skipif(not rh.available(), reason="/root/reference not present")
def test_generators_match_live_reference():
    _, sim = rh.ref_modules()
    for seed in range(3):
        for mine, theirs in ((lambda: DAGGenerator.generate_random_dag(80, 4), lambda: sim.DAGGenerator.generate_random_dag(80, 4)),
                             (lambda: DAGGenerator.generate_llm_dag(6, attention_heads=8), lambda: sim.DAGGenerator.generate_llm_dag(6, attention_heads=8)),
                             (lambda: DAGGenerator.generate_pipeline_dag(4, 5), lambda: sim.DAGGenerator.generate_pipeline_dag(4, 5))):
            random.seed(seed)
            a = mine()
            sa = random.getstate()
            random.seed(seed)
            b = theirs()
            assert random.getstate() == sa  # consumed the same random numbers
            assert [(t.id, t.memory_required, t.compute_time, t.dependencies, t.params_needed) for t in a] == \
                   [(t.id, t.memory_required, t.compute_time, t.dependencies, t.params_needed) for t in b]


def _check_metrics(batch, res, cases):
    for a in range(4):
        m = evaluate_batch(batch, res, res.slot(a))
        for b, c in enumerate(cases):
            e = c.exp[a]
            assert int(m["total_tasks"][b]) == c.n
            assert (int(m["completed_tasks"][b]), int(m["failed_tasks"][b])) == (e["completed"], e["failed"])
            assert (int(m["param_cache_hits"][b]), int(m["param_cache_misses"][b])) == (e["hits"], e["misses"])
            assert m["makespan"][b] == e["makespan"]
            for f in ("avg_node_utilization", "load_balance_score", "completion_rate"):
                np.testing.assert_allclose(m[f][b], e[f], rtol=1e-6, atol=1e-12, err_msg=f"{c.label} algo {a} {f}")


def test_evaluate_batch_on_oracle_records(oracle_mod):
    """The result-format emitter alone (host logic, no GPU): the oracle's records of the default sweep, pushed through
    evaluate_batch's segment reductions, give the reference's 648 TestResults; and a ragged batch with unused nodes,
    empty schedules and zero makespans goes through the same code as a plain per-instance loop would."""
    cases = golden_io.load("cfg2_sweep_seed0.npz")
    batch = golden_io.batch_of(cases)
    res = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=4)
    _check_metrics(batch, res, cases)
    # against the definition, instance by instance (ref simulation.py:271-302,339-347)
    for slot in range(4):
        m = evaluate_batch(batch, res, slot)
        for b in range(0, batch.B, 7):
            ks = batch.node_slice(b)
            loads = res.fin[slot, ks][res.cnt[slot, ks] > 0]
            mk = res.summary[slot, b]["makespan"]
            util = np.mean(loads / mk) if loads.size and mk > 0 else 0.0
            bal = 1 / (1 + np.std(loads) / np.mean(loads)) if loads.size and loads.max() != 0 and np.mean(loads) > 0 else 0.0
            np.testing.assert_allclose([m["avg_node_utilization"][b], m["load_balance_score"][b]], [util, bal], rtol=1e-12)


@pytest.mark.skipif(not rh.available(), reason="needs the reference checkout (build container only)")
def test_simulate_execution_replays_any_schedule_like_the_reference():
    """simulate_execution accepts ANY (scheduler, schedule) pair in the reference (simulation.py:216-278): hand both
    implementations the same foreign schedule — shuffled, with an unknown node and an unknown task in it."""
    sch, sim = rh.ref_modules()
    random.seed(3)
    tasks = sim.DAGGenerator.generate_pipeline_dag(4, 3)
    nodes = [sch.Node("a", 8.0, 1.2), sch.Node("b", 6.0, 0.8), sch.Node("c", 6.0, 1.0)]
    s = sch.GreedyScheduler(nodes)
    for t in tasks:
        s.add_task(t)
    ids = [t.id for t in tasks]
    random.shuffle(ids)
    schedule = {"b": ids[:5] + ["ghost_task"], "nowhere": ids[5:7], "a": ids[7:], "c": []}
    want = sim.ImprovedSchedulerEvaluator({}).simulate_execution(s, schedule)
    got = ImprovedSchedulerEvaluator({}).simulate_execution(s, schedule)
    assert got[0] == want[0]
    assert (got[1]["param_cache_hits"], got[1]["param_cache_misses"]) == (want[1]["param_cache_hits"], want[1]["param_cache_misses"])
    assert dict(got[1]["node_utilization"]) == dict(want[1]["node_utilization"])
    assert ImprovedSchedulerEvaluator({}).simulate_execution(s, {}) == sim.ImprovedSchedulerEvaluator({}).simulate_execution(s, {})


@pytest.mark.gpu
def test_evaluate_batch_reproduces_reference_testresults():
    """All 648 TestResults of the default sweep: counts and cache statistics exact, makespan bit-equal,
    utilisation / load balance / completion rate within 1e-6 relative (north-star tolerance)."""
    from dls_b200.batch import schedule_batch
    cases = golden_io.load("cfg2_sweep_seed0.npz")
    batch = golden_io.batch_of(cases)
    res = schedule_batch(batch, N.MASK_ALL).host()
    _check_metrics(batch, res, cases)


@pytest.mark.gpu
def test_run_single_test_and_run_experiments_on_gpu():
    """Object-level path: run_single_test on the 4-task example and GPT-2 against the fixture (with the fixture's
    tie-break order imposed), then the whole batched sweep end to end with its CSV frame."""
    from test_gpu_host_api import _objects
    ev = ImprovedSchedulerEvaluator(SCHEDULERS)
    for fname, label in (("cfg1_example_and_edges.npz", "example4"), ("cfg3_gpt2.npz", "gpt2|4laptops")):
        case = [c for c in golden_io.load(fname) if c.label == label][0]
        tasks, nodes, pi = _objects(case)
        for name, cls in SCHEDULERS.items():
            fixed = type("Fixed" + cls.__name__, (cls,), {"_pending_order": lambda self, pi=pi: list(pi)})
            r = ev.run_single_test(fixed, name, tasks, nodes, label, 1.0)
            e = case.exp[cls.ALGO]
            assert isinstance(r, TestResult)
            assert (r.total_tasks, r.completed_tasks, r.failed_tasks) == (case.n, e["completed"], e["failed"])
            assert (r.param_cache_hits, r.param_cache_misses) == (e["hits"], e["misses"])
            assert r.makespan == e["makespan"]
            assert r.load_balance_score == pytest.approx(e["load_balance_score"], rel=1e-6)
            assert r.avg_node_utilization == pytest.approx(e["avg_node_utilization"], rel=1e-6)
            assert r.completion_rate == pytest.approx(e["completion_rate"], rel=1e-9)
    random.seed(0)
    ev.run_experiments(num_runs=3)
    assert len(ev.results) == 648
    df = ev.results_frame()
    assert list(df.columns) == CSV_COLUMNS and len(df) == 648
    assert set(df["scheduler_name"]) == {"DFS", "Greedy", "Critical", "MRU_spec"} and set(df["num_nodes"]) == {2, 4, 8}
    # order: instance-major, scheduler-minor, like the reference's nested loops
    assert df["scheduler_name"].tolist()[:4] == ["DFS", "Greedy", "Critical", "MRU_spec"]
    # MRU completes every LLM DAG at every regime (ref PDF 5.2.2; hash-order independent: level width <= 4 ties aside)
    llm_mru = df[(df["dag_type"].str.startswith("LLM")) & (df["scheduler_name"] == "MRU_spec")]
    assert (llm_mru["completion_rate"] == 100.0).all()


@pytest.mark.gpu
def test_device_aggregation_matches_groupby():
    """the per-(scheduler, regime) mean tables of analyze_results computed on the device equal a
    pandas groupby over the reference's own TestResult metrics (fixture), to 1e-9."""
    import pandas as pd
    from dls_b200.batch import aggregate, schedule_batch
    cases = golden_io.load("cfg2_sweep_seed0.npz")
    batch = golden_io.batch_of(cases)
    regimes = [float(c.label.split("|")[2][1:]) for c in cases]
    levels = sorted(set(regimes))
    group = np.array([levels.index(x) for x in regimes], np.int32)
    res = schedule_batch(batch, N.MASK_ALL)
    agg = aggregate(res, group, len(levels))
    rows = []
    for a in range(4):
        for c, g in zip(cases, group):
            e = c.exp[a]
            rows.append({"a": a, "g": int(g), "completion_rate": e["completion_rate"], "makespan": e["makespan"],
                         "avg_node_utilization": e["avg_node_utilization"], "load_balance_score": e["load_balance_score"],
                         "completed": e["completed"]})
    df = pd.DataFrame(rows)
    m_all = df.groupby(["a", "g"])[["completion_rate", "avg_node_utilization"]].mean()
    m_done = df[df["completed"] > 0].groupby(["a", "g"])[["makespan", "load_balance_score"]].mean()
    for a in range(4):
        for g in range(len(levels)):
            assert agg["count"][a, g] == (group == g).sum()
            for f in ("completion_rate", "avg_node_utilization"):
                np.testing.assert_allclose(agg[f][a, g], m_all.loc[(a, g), f], rtol=1e-9, err_msg=f)
            for f in ("makespan", "load_balance_score"):
                np.testing.assert_allclose(agg[f][a, g], m_done.loc[(a, g), f], rtol=1e-9, err_msg=f)
