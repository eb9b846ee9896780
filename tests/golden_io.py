"""Reader for the fixtures written by tests/golden/make_golden.py."""
from __future__ import annotations

import os

import numpy as np

from common import ROOT  # noqa: F401  (puts the repo root on sys.path)
from dls_b200.lowering import DagBatch, LoweredDag

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCALARS = ("completed", "failed", "pending", "makespan", "hits", "misses", "avg_node_utilization",
           "load_balance_score", "completion_rate")
FLOAT_SCALARS = ("makespan", "avg_node_utilization", "load_balance_score", "completion_rate")


class Case:
    def __init__(self, z, c):
        n, P, K, mask = (int(v) for v in z[f"c{c}_meta"])
        self.label = str(z[f"c{c}_label"])
        self.n, self.P, self.K, self.mask = n, P, K, mask
        names = [str(x) for x in z[f"c{c}_names"]] if f"c{c}_names" in z else [f"t{i}" for i in range(n)]
        pnames = [str(x) for x in z[f"c{c}_param_names"]] if f"c{c}_param_names" in z else [f"p{i:06d}" for i in range(P)]
        self.node_ids = [str(x) for x in z[f"c{c}_node_ids"]] if f"c{c}_node_ids" in z else [f"n{k}" for k in range(K)]
        self.dag = LoweredDag(names=names, param_names=pnames, pred_ptr=z[f"c{c}_pred_ptr"], pred_idx=z[f"c{c}_pred_idx"],
                              par_ptr=z[f"c{c}_par_ptr"], par_idx=z[f"c{c}_par_idx"], rank=z[f"c{c}_rank"],
                              mem=z[f"c{c}_mem"], dur=z[f"c{c}_dur"])
        self.cap, self.spd = z[f"c{c}_cap"], z[f"c{c}_spd"]
        self.exp = {}
        for a in range(4):
            if not mask & (1 << a):
                continue
            e = {f: z[f"c{c}_a{a}_{f}"] for f in ("node_of", "order", "fin", "cnt", "avail", "peak", "cached")}
            sc = z[f"c{c}_a{a}_scalars"]
            for f, v in zip(SCALARS, sc):
                e[f] = float(v) if f in FLOAT_SCALARS else int(v)
            lens = z[f"c{c}_a{a}_pernode_len"]
            flat = z[f"c{c}_a{a}_pernode"]
            offs = np.concatenate([[0], np.cumsum(lens)])
            e["per_node_seq"] = [flat[offs[k]:offs[k + 1]].tolist() for k in range(K)]
            e["order"] = e["order"].tolist()
            e["refused"] = int(z[f"c{c}_a{a}_refused"]) if f"c{c}_a{a}_refused" in z else 0
            self.exp[a] = e


def load(name: str):
    path = os.path.join(GOLDEN_DIR, name)
    if not os.path.exists(path):
        return []
    with np.load(path) as z:
        return [Case(z, c) for c in range(int(z["ncases"]))]


def batch_of(cases) -> DagBatch:
    return DagBatch.from_instances([c.dag for c in cases], [(c.cap, c.spd) for c in cases])


ALL_FILES = ("cfg1_example_and_edges.npz", "cfg2_sweep_seed0.npz", "cfg3_gpt2.npz", "cfg5_random1k.npz",
             "cfg4_llm10k.npz")
