#!/usr/bin/env python
"""Freeze the MRU scheduler's post-call bookkeeping (param_usage_count, param_last_used, time_step;
ref schedulers.py:377-381,453,514-516) as the UNMODIFIED reference leaves it, for the two object-level fixtures, with
each fixture's own tie-break order imposed on the pending set. Output: tests/golden/mru_state.json.

    python tests/golden/make_mru_state.py        (needs /root/reference; run in the build container only)
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import golden_io  # noqa: E402
import refharness as rh  # noqa: E402
from test_gpu_host_api import _OrderedPending, _objects  # noqa: E402


def main():
    sch, _ = rh.ref_modules()
    out = {}
    for fname, label in (("cfg1_example_and_edges.npz", "example4"), ("cfg3_gpt2.npz", "gpt2|4laptops")):
        case = [c for c in golden_io.load(fname) if c.label == label][0]
        tasks, nodes, pi = _objects(case)
        s = sch.MRUScheduler([sch.Node(n.id, n.total_memory, n.compute_speed) for n in nodes])
        for t in tasks:
            s.add_task(sch.Task(t.id, t.memory_required, t.compute_time, list(t.dependencies), set(t.params_needed)))
        s.pending_tasks = _OrderedPending(pi)
        s.schedule()
        out[label] = {"time_step": s.time_step, "param_usage_count": dict(s.param_usage_count),
                      "param_last_used": dict(s.param_last_used), "completed": len(s.completed_tasks)}
    with open(os.path.join(HERE, "mru_state.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print({k: (v["time_step"], len(v["param_last_used"])) for k, v in out.items()})


if __name__ == "__main__":
    main()
