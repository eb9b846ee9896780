#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) in this container.

    PYTHONHASHSEED=0 python tests/golden/make_golden.py [--big]

Each fixture holds, per case, the lowered inputs (include/dls.h arrays, INCLUDING the rank array pi taken
from the live scheduler's pending set, so the fixture does not depend on the hash seed of whoever reads it)
and, per algorithm, the reference's outcome: per-task verdict/node, global assignment order, per-node
finish time / count / final available memory (fp64, bit-exact), cached sets, completed/failed/pending,
makespan, cache hits/misses, plus the peak-memory extension sampled by tests/refharness.PeakMixin.

The reference cannot travel to the GPU box; these files can.
"""
from __future__ import annotations

import argparse
import os
import random
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import refharness as rh  # noqa: E402
from dls_b200.lowering import lower_nodes, lower_tasks  # noqa: E402

IN_FIELDS = ("pred_ptr", "pred_idx", "par_ptr", "par_idx", "rank", "mem", "dur")
EXP_ARR = ("node_of", "order", "fin", "cnt", "avail", "peak", "cached")
EXP_SCALAR = ("completed", "failed", "pending", "makespan", "hits", "misses", "avg_node_utilization",
              "load_balance_score", "completion_rate")


def llm_dag_wide(sch, num_layers: int, heads: int):
    """Config 4's ~10k-task shape: ref generate_llm_dag (simulation.py:36-88) with the 4-head cap (:52) lifted.
    Same Task format, memory/time literals and naming as the reference generator."""
    tasks = [sch.Task("embedding", 0.5, 0.1, [], {"embedding_weights"})]
    for layer in range(num_layers):
        layer_tasks = []
        for head in range(heads):
            deps = ["embedding"] if layer == 0 else [f"layer_{layer-1}_output"]
            t = sch.Task(f"layer_{layer}_attention_head_{head}", 0.2, 0.05, deps,
                         {f"layer_{layer}_attention_head_{head}_weights"})
            tasks.append(t)
            layer_tasks.append(t.id)
        tasks.append(sch.Task(f"layer_{layer}_attention_output", 0.3, 0.05, layer_tasks,
                              {f"layer_{layer}_attention_output_weights"}))
        tasks.append(sch.Task(f"layer_{layer}_ffn", 0.5, 0.1, [f"layer_{layer}_attention_output"],
                              {f"layer_{layer}_ffn_weights"}))
        tasks.append(sch.Task(f"layer_{layer}_output", 0.1, 0.02, [f"layer_{layer}_ffn"], set()))
    tasks.append(sch.Task("output", 0.3, 0.05, [f"layer_{num_layers-1}_output"], {"output_weights"}))
    return tasks


class Writer:
    def __init__(self):
        self.arrays = {}
        self.ncases = 0

    def add(self, label: str, tasks, nodes, algos=(0, 1, 2, 3), names=False):
        c = self.ncases
        pi = None
        algmask = 0
        for algo in algos:
            t0 = time.time()
            run = rh.run_reference(algo, tasks, nodes)
            if pi is None:
                pi = run["pi"]
                dag = lower_tasks(tasks, pending=pi)
            assert run["pi"] == pi
            exp = rh.expected_arrays(run, dag)
            for f in EXP_ARR:
                self.arrays[f"c{c}_a{algo}_{f}"] = np.asarray(exp[f])
            self.arrays[f"c{c}_a{algo}_scalars"] = np.array([float(exp[f]) for f in EXP_SCALAR], np.float64)
            self.arrays[f"c{c}_a{algo}_pernode_len"] = np.array([len(x) for x in exp["per_node_seq"]], np.int32)
            self.arrays[f"c{c}_a{algo}_pernode"] = np.array([i for x in exp["per_node_seq"] for i in x], np.int32)
            self.arrays[f"c{c}_a{algo}_refused"] = np.array(exp["refused"], np.int64)
            algmask |= 1 << algo
            dt = time.time() - t0
            if dt > 5:
                print(f"   {label} algo {algo}: {dt:.1f}s", flush=True)
        for f in IN_FIELDS:
            self.arrays[f"c{c}_{f}"] = getattr(dag, f)
        cap, spd = lower_nodes(nodes)
        self.arrays[f"c{c}_cap"], self.arrays[f"c{c}_spd"] = cap, spd
        self.arrays[f"c{c}_meta"] = np.array([dag.n, dag.P, len(nodes), algmask], np.int64)
        self.arrays[f"c{c}_label"] = np.array(label)
        if names:
            self.arrays[f"c{c}_names"] = np.array(dag.names)
            self.arrays[f"c{c}_param_names"] = np.array(dag.param_names)
            self.arrays[f"c{c}_node_ids"] = np.array([n.id for n in nodes])
        self.ncases += 1

    def save(self, path):
        self.arrays["ncases"] = np.array(self.ncases)
        np.savez_compressed(path, **self.arrays)
        print(f"{path}: {self.ncases} cases, {os.path.getsize(path)/1024:.0f} KiB")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true", help="also regenerate the slow 1k/10k-task fixtures")
    ap.add_argument("--only-refused", action="store_true", help="write cfg1b_mru_refused.npz only")
    args = ap.parse_args()
    sch, sim = rh.ref_modules()
    ev = sim.ImprovedSchedulerEvaluator({})
    G = sim.DAGGenerator

    # ---- MRU's refused assignment (ref schedulers.py:82-83 reached from :509) -----------------------------------------
    # assign_task_to_node re-checks the fit after MRU has chosen a node. It can only say no when the eviction that made
    # the node a candidate freed "enough" by its own arithmetic (freed >= need - avail, ref :425) while avail + 0.5 * j
    # still rounds below need, and nothing is left to evict for a second try (:506-507). With 0.5-steps that takes an
    # available memory in (-0.25, 0): the numbers below were found by search (negative task memory is outside the
    # generators' domain but is data like any other to the scheduler). Task "late" is refused in round 2, stays pending,
    # is looked at again in round 3 and fails there.
    w = Writer()
    for C, m0, m1 in ((0.3909436547441535, -0.3765207077218265, -0.10905634525584645),
                      (0.3238587717729853, -0.4348473194366214, -0.17614122822701467),
                      (0.36289824318069075, -0.5366170053406539, -0.1371017568193092)):
        tasks = [sch.Task("first", m0, 0.1, [], {"a"}), sch.Task("late", m1, 0.2, ["first"], {"b"}),
                 sch.Task("side", 0.0, 0.05, ["first"], set()), sch.Task("tail", 0.0, 0.05, ["late", "side"], set())]
        w.add(f"mru_refused|C{C:.3f}", tasks, [sch.Node("only", C, 1.0)], names=True)
    w.save(os.path.join(HERE, "cfg1b_mru_refused.npz"))
    if args.only_refused:
        return

    # ---- config 1: the built-in 4-task example (ref schedulers.py:534-548) + hand-made edge cases -----------------
    w = Writer()
    ex_tasks = [sch.Task("t1", 1.0, 0.1, [], {"p1"}), sch.Task("t2", 1.0, 0.1, ["t1"], {"p2"}),
                sch.Task("t3", 1.0, 0.1, ["t1"], {"p3"}), sch.Task("t4", 1.0, 0.1, ["t2", "t3"], {"p1", "p2"})]
    ex_nodes = [sch.Node("n1", 3.0), sch.Node("n2", 2.5)]
    w.add("example4", ex_tasks, ex_nodes, names=True)
    w.add("single", [sch.Task("a", 0.3, 0.2, [], {"w"})], [sch.Node("n", 1.0)])
    w.add("single_nofit", [sch.Task("a", 2.0, 0.2, [], {"w"})], [sch.Node("n", 1.0), sch.Node("m", 2.4, 2.0)])
    w.add("noparams_chain", [sch.Task(f"c{i}", 0.25, 0.1 * (i + 1), [f"c{i-1}"] if i else []) for i in range(7)],
          [sch.Node("x", 0.3, 0.5), sch.Node("y", 0.25, 2.0), sch.Node("z", 0.2, 3.0)])
    w.add("dup_deps", [sch.Task("a", 0.1, 0.1, [], {"p"}), sch.Task("b", 0.1, 0.2, ["a", "a"], {"p", "q"}),
                       sch.Task("c", 0.1, 0.3, ["a", "b", "a"], {"q"})], [sch.Node("n0", 1.2), sch.Node("n1", 1.2, 1.5)])
    w.add("all_fail", [sch.Task(f"t{i}", 1.0, 0.1, [f"t{i-1}"] if i else [], {f"p{i}"}) for i in range(5)],
          [sch.Node("n0", 1.0), sch.Node("n1", 1.2)])
    w.add("fail_midway_leaves_pending",
          [sch.Task("r", 0.2, 0.1, [], {"a"}), sch.Task("big", 5.0, 0.1, ["r"], {"b"}),
           sch.Task("ok", 0.2, 0.1, ["r"], {"c"}), sch.Task("after_big", 0.1, 0.1, ["big"], set()),
           sch.Task("after_ok", 0.1, 0.1, ["ok"], {"a"}), sch.Task("join", 0.1, 0.1, ["after_big", "after_ok"], set())],
          [sch.Node("n0", 2.0, 1.0), sch.Node("n1", 1.5, 1.3)])
    w.add("zero_mem_equal_speed", [sch.Task(f"z{i}", 0.0, 0.05, [], {"s"}) for i in range(9)],
          [sch.Node(f"n{k}", 1.0, 1.0) for k in range(4)])
    # shared params under pressure: MRU evictions incl. rollback
    for seed, (stages, width, regime, K) in enumerate([(5, 3, 0.45, 2), (6, 4, 0.5, 4), (8, 2, 0.6, 2), (4, 5, 0.4, 8)]):
        random.seed(100 + seed)
        tasks = G.generate_pipeline_dag(stages, width)
        nodes = ev.create_nodes_with_memory_regime(ev.calculate_total_memory_needed(tasks), regime, K)
        w.add(f"pipeline_{stages}x{width}_r{regime}_K{K}", tasks, nodes)
    w.save(os.path.join(HERE, "cfg1_example_and_edges.npz"))

    # ---- config 2: run_experiments(num_runs=3) replayed with random.seed(0) (ref simulation.py:365-416) ---------
    w = Writer()
    random.seed(0)
    dag_configs = [("LLM-Small", lambda: G.generate_llm_dag(4, attention_heads=4)),
                   ("LLM-Medium", lambda: G.generate_llm_dag(8, attention_heads=4)),
                   ("LLM-Large", lambda: G.generate_llm_dag(12, attention_heads=4)),
                   ("Random-Small", lambda: G.generate_random_dag(30)),
                   ("Random-Medium", lambda: G.generate_random_dag(60)),
                   ("Pipeline", lambda: G.generate_pipeline_dag(5, width=3))]
    for dag_name, gen in dag_configs:
        for num_nodes in (2, 4, 8):
            for regime in (1.0, 0.9, 0.8):
                for run in range(3):
                    tasks = gen()
                    total = ev.calculate_total_memory_needed(tasks)
                    nodes = ev.create_nodes_with_memory_regime(total, regime, num_nodes)
                    w.add(f"{dag_name}|K{num_nodes}|r{regime}|run{run}", tasks, nodes)
    w.save(os.path.join(HERE, "cfg2_sweep_seed0.npz"))

    # ---- config 3: GPT-2 small DAG (ref test_gpt2.py:45-168) on the 4 laptops (:278-283) and on 8 ---------------
    w = Writer()
    gpt2 = rh.gpt2_tasks()
    laptops4 = [sch.Node("laptop_0", 8.0, 1.0), sch.Node("laptop_1", 8.0, 1.2), sch.Node("laptop_2", 6.0, 0.8),
                sch.Node("laptop_3", 6.0, 0.9)]
    laptops8 = [sch.Node(f"laptop_{i}", [8.0, 8.0, 6.0, 6.0][i % 4], [1.0, 1.2, 0.8, 0.9][i % 4]) for i in range(8)]
    w.add("gpt2|4laptops", gpt2, laptops4, names=True)
    w.add("gpt2|8laptops", gpt2, laptops8)
    rng = np.random.default_rng(0)
    for j in range(6):  # perturbed durations, as config 3 defines them: dur*(1+eps), eps~U(-0.1,0.1)
        eps = rng.uniform(-0.1, 0.1, len(gpt2))
        pert = [sch.Task(t.id, t.memory_required, float(t.compute_time * (1.0 + e)), list(t.dependencies),
                         set(t.params_needed)) for t, e in zip(gpt2, eps)]
        w.add(f"gpt2|8laptops|perturbed{j}", pert, laptops8)
    w.save(os.path.join(HERE, "cfg3_gpt2.npz"))

    if not args.big:
        return
    # ---- config 5 parity subset: reference generate_random_dag(1000) -----------------------------------------
    w = Writer()
    for seed, (K, regime) in enumerate([(8, 0.9), (8, 1.0), (4, 0.8), (8, 0.8), (2, 0.9), (8, 0.9)]):
        random.seed(seed)
        tasks = G.generate_random_dag(1000)
        nodes = ev.create_nodes_with_memory_regime(ev.calculate_total_memory_needed(tasks), regime, K)
        w.add(f"random1k|seed{seed}|K{K}|r{regime}", tasks, nodes)
    w.save(os.path.join(HERE, "cfg5_random1k.npz"))

    # ---- config 4: 48 layers x 205 heads = 9,986 tasks ------------------------------------------------------------
    w = Writer()
    random.seed(4)
    tasks = llm_dag_wide(sch, 48, 205)
    total = ev.calculate_total_memory_needed(tasks)
    nodes = ev.create_nodes_with_memory_regime(total, 0.9, 8)
    w.add("llm48x205|K8|r0.9", tasks, nodes, algos=(0, 1, 2))
    nodes = ev.create_nodes_with_memory_regime(total, 1.0, 8)
    w.add("llm48x205|K8|r1.0", tasks, nodes, algos=(0, 1, 2, 3))
    w.save(os.path.join(HERE, "cfg4_llm10k.npz"))


if __name__ == "__main__":
    main()
