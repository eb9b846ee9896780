"""Shared helpers for the parity tests."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from dls_b200 import _native as N  # noqa: E402
from dls_b200.batch import ResultArrays, sequence_of  # noqa: E402
from dls_b200.lowering import DagBatch  # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def assert_matches_expected(exp: dict, batch: DagBatch, res: ResultArrays, slot: int, b: int, what: str = "", status: int = 0):
    """Compare one (algo slot, instance) of a ResultArrays with a reference outcome from
    refharness.expected_arrays / a golden fixture. Integer/index data and fp64 state: bit-exact."""
    ts, ks, ps = batch.task_slice(b), batch.node_slice(b), batch.param_slice(b)
    K = ks.stop - ks.start
    tag = f"[{what} slot={slot} inst={b}] "
    summ = res.summary[slot, b]
    assert summ["status"] == status, tag + f"status {summ['status']}"
    np.testing.assert_array_equal(res.node_of[slot, ts], exp["node_of"], err_msg=tag + "node_of")
    order, nodes = sequence_of(batch, res, slot, b)
    assert order.tolist() == list(exp["order"]), tag + "global assignment order"
    for k in range(K):
        assert order[nodes == k].tolist() == list(exp["per_node_seq"][k]), tag + f"sequence on node {k}"
    assert (int(summ["completed"]), int(summ["failed"]), int(summ["pending"])) == (
        int(exp["completed"]), int(exp["failed"]), int(exp["pending"])), tag + "completed/failed/pending"
    assert (int(summ["hits"]), int(summ["misses"])) == (int(exp["hits"]), int(exp["misses"])), tag + "hits/misses"
    np.testing.assert_array_equal(res.cnt[slot, ks][:K], exp["cnt"], err_msg=tag + "cnt")
    np.testing.assert_array_equal(bits(res.avail[slot, ks][:K]), bits(exp["avail"]), err_msg=tag + "avail bits")
    np.testing.assert_array_equal(bits(res.fin[slot, ks][:K]), bits(exp["fin"]), err_msg=tag + "fin bits")
    # north star: makespan within 1e-6 relative; in practice bit-equal
    np.testing.assert_allclose(summ["makespan"], exp["makespan"], rtol=1e-6, atol=0, err_msg=tag + "makespan")
    assert bits([summ["makespan"]])[0] == bits([exp["makespan"]])[0], tag + "makespan bits"
    if res.peak is not None and "peak" in exp:
        np.testing.assert_allclose(res.peak[slot, ks][:K], exp["peak"], rtol=1e-6, atol=0, err_msg=tag + "peak")
    if res.cached is not None and "cached" in exp:
        np.testing.assert_array_equal(res.cached[slot, ps], exp["cached"], err_msg=tag + "cached")


def assert_results_equal(a: ResultArrays, b: ResultArrays, what: str = ""):
    """Two ResultArrays (e.g. CUDA vs oracle) must agree: exact on everything, fp64 bit-for-bit."""
    assert a.algos == b.algos
    np.testing.assert_array_equal(a.summary["status"], b.summary["status"], err_msg=what + " status")
    np.testing.assert_array_equal(a.node_of, b.node_of, err_msg=what + " node_of")
    np.testing.assert_array_equal(a.seq_pos, b.seq_pos, err_msg=what + " seq_pos")
    for f in ("completed", "failed", "pending", "hits", "misses", "rounds"):
        np.testing.assert_array_equal(a.summary[f], b.summary[f], err_msg=what + " " + f)
    np.testing.assert_allclose(a.summary["makespan"], b.summary["makespan"], rtol=1e-6, atol=0)
    np.testing.assert_array_equal(bits(a.summary["makespan"]), bits(b.summary["makespan"]), err_msg=what + " makespan bits")
    np.testing.assert_array_equal(a.cnt, b.cnt, err_msg=what + " cnt")
    np.testing.assert_array_equal(bits(a.fin), bits(b.fin), err_msg=what + " fin bits")
    np.testing.assert_array_equal(bits(a.avail), bits(b.avail), err_msg=what + " avail bits")
    if a.peak is not None and b.peak is not None:
        np.testing.assert_array_equal(bits(a.peak), bits(b.peak), err_msg=what + " peak bits")
        np.testing.assert_array_equal(bits(a.summary["peak"]), bits(b.summary["peak"]), err_msg=what + " summary peak bits")
    if a.cached is not None and b.cached is not None:
        np.testing.assert_array_equal(a.cached, b.cached, err_msg=what + " cached")


def run_both_paths(batch, mask):
    """Yield (label, ResultArrays) for the device implementations: the general warp-per-DAG kernels (which also
    produce the cached / timeline extensions) and the thread-per-DAG streamed kernels (taken when the batch qualifies;
    otherwise the library falls back to the general ones by itself), the latter with and without peak memory."""
    from dls_b200.batch import schedule_batch
    from dls_b200 import _native as N
    yield "general", schedule_batch(batch, mask, peak=True, cached=True).host()
    yield "streamed", schedule_batch(batch, mask, peak=False, cached=False).host()
    yield "streamed, peak", schedule_batch(batch, mask, peak=True, cached=False).host()
    old = N.load().dls_set_stream_path(0)  # the general kernels again, this time without the extension outputs
    try:
        yield "general, no extensions", schedule_batch(batch, mask, peak=False, cached=False).host()
    finally:
        N.load().dls_set_stream_path(old)
