"""Live differential test: the C oracle against the unmodified reference imported from /root/reference.
Runs only where the reference exists (this container); skipped on the GPU box."""
import random

import pytest

import refharness as rh
from common import assert_matches_expected
from dls_b200.lowering import DagBatch, lower_nodes, lower_tasks

pytestmark = pytest.mark.skipif(not rh.available(), reason="/root/reference not present")


def _gens(sim):
    G = sim.DAGGenerator
    return [("LLM-Small", lambda: G.generate_llm_dag(4, attention_heads=4)),
            ("LLM-Large", lambda: G.generate_llm_dag(12, attention_heads=4)),
            ("Random-Small", lambda: G.generate_random_dag(30)),
            ("Random-Medium", lambda: G.generate_random_dag(60)),
            ("Random-150", lambda: G.generate_random_dag(150, max_deps=5)),
            ("Pipeline", lambda: G.generate_pipeline_dag(5, width=3)),
            ("Pipeline-wide", lambda: G.generate_pipeline_dag(4, width=6))]


@pytest.mark.parametrize("seed", range(4))
def test_oracle_matches_live_reference(oracle_mod, seed):
    sch, sim = rh.ref_modules()
    ev = sim.ImprovedSchedulerEvaluator({})
    n = 0
    for name, g in _gens(sim):
        for K in (2, 4, 8):
            for regime in (1.0, 0.9, 0.8, 0.6, 0.45):
                random.seed(seed * 7919 + K * 10 + int(regime * 100))
                tasks = g()
                nodes = ev.create_nodes_with_memory_regime(ev.calculate_total_memory_needed(tasks), regime, K)
                for algo in range(4):
                    run = rh.run_reference(algo, tasks, nodes)
                    dag = lower_tasks(tasks, pending=run["pi"])
                    batch = DagBatch.from_instances([dag], [lower_nodes(nodes)])
                    res = oracle_mod.schedule_batch(batch, 1 << algo)
                    assert_matches_expected(rh.expected_arrays(run, dag), batch, res, 0, 0,
                                            f"{name} K={K} r={regime} seed={seed} algo={algo}")
                    n += 1
    assert n == 7 * 3 * 5 * 4


def test_four_task_example_matches_live_reference(oracle_mod):
    """Config 1 (ref schedulers.py:529-568). The printed schedules depend on the hash seed, so
    compare with the live reference in this process rather than a frozen string."""
    sch, _ = rh.ref_modules()
    tasks = [sch.Task("t1", 1.0, 0.1, [], {"p1"}), sch.Task("t2", 1.0, 0.1, ["t1"], {"p2"}),
             sch.Task("t3", 1.0, 0.1, ["t1"], {"p3"}), sch.Task("t4", 1.0, 0.1, ["t2", "t3"], {"p1", "p2"})]
    nodes = [sch.Node("n1", 3.0), sch.Node("n2", 2.5)]
    for algo in range(4):
        run = rh.run_reference(algo, tasks, nodes)
        dag = lower_tasks(tasks, pending=run["pi"])
        batch = DagBatch.from_instances([dag], [lower_nodes(nodes)])
        res = oracle_mod.schedule_batch(batch, 1 << algo)
        assert_matches_expected(rh.expected_arrays(run, dag), batch, res, 0, 0, f"example algo={algo}")
