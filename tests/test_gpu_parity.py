"""Parity tests proper: the CUDA path, called through the C ABI (dls_schedule_batch / dls_priorities via
dls_b200.batch), against (1) the golden fixtures produced by the reference and (2) the CPU oracle on the same
seeded inputs. Bar: bit-exact task order, node assignment, verdicts, fp64 node state and makespan bits
(north star asks 1e-6 relative on makespan; assert_* checks both)."""
import numpy as np
import pytest
import torch

import golden_io
from common import assert_matches_expected, assert_results_equal, run_both_paths
from dls_b200 import _native as N
from dls_b200.batch import priorities, schedule_batch
from dls_b200.gen import perturbed_durations, random_dag_batch
from dls_b200.lowering import DagBatch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fname", golden_io.ALL_FILES)
def test_cuda_reproduces_reference_fixtures(fname):
    cases = golden_io.load(fname)
    assert cases, f"{fname} missing"
    for mask in sorted({c.mask for c in cases}):
        sub = [c for c in cases if c.mask == mask]
        batch = golden_io.batch_of(sub)
        for path, res in run_both_paths(batch, mask):
            for b, c in enumerate(sub):
                for slot, a in enumerate(res.algos):
                    assert_matches_expected(c.exp[a], batch, res, slot, b, f"{path} {fname}:{c.label} algo={a}")


@pytest.mark.parametrize("n,K,regime,B", [(1000, 8, 0.9, 96), (1000, 8, 0.8, 48), (1000, 4, 1.0, 48), (1000, 2, 0.9, 32),
                                          (257, 8, 0.6, 64), (64, 8, 0.45, 128), (33, 4, 0.7, 128), (1, 2, 1.0, 8)])
def test_cuda_matches_oracle_random_dags(oracle_mod, n, K, regime, B):
    batch = random_dag_batch(B, n, K, regime, seed=n * 131 + K)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} random n={n} K={K} r={regime}")


@pytest.mark.parametrize("n,K,pool,regime", [(40, 3, 6, 1.2), (60, 8, 12, 1.5), (30, 2, 3, 1.0), (80, 8, 40, 1.0),
                                             (50, 17, 10, 2.5), (24, 32, 8, 6.0), (120, 5, 16, 0.9), (60, 8, 12, 0.6)])
def test_cuda_matches_oracle_shared_params_fuzz(oracle_mod, n, K, pool, regime):
    """Shared params (cache hits, MRU eviction scores with live readiness, rollbacks), duplicate dependency entries,
    zero-memory tasks, tied node speeds, up to 32 nodes: every device path against the oracle, bit for bit."""
    from dls_b200.gen import random_shared_param_batch
    batch = random_shared_param_batch(96, n, K, pool, regime, seed=n * 7 + K)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    assert want.summary["failed"].sum() > 0 and want.summary["completed"].sum() > 0   # pressure, but not only failures
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} shared-param fuzz n={n} K={K} pool={pool}")


def test_dfs_result_shared_with_greedy_only_where_they_agree(oracle_mod):
    """On DAGs without shared params the DFS launch stores its result for Greedy too. The one input class where the
    two differ even there (no node above DFS's max_memory = -1 start, requirements negative enough to "fit":
    ref schedulers.py:187 vs :267-281) must still come out as the reference's, next to ordinary instances."""
    batch = random_dag_batch(64, 60, 4, 0.9, seed=3)
    odd = np.repeat(np.arange(64) % 2 == 1, 4)
    batch.cap[odd] = -2.0 - (np.arange(odd.sum()) % 3)
    odd_t = np.repeat(np.arange(64) % 2 == 1, 60)
    batch.mem[odd_t] -= 6.0
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=4)
    dfs, greedy = want.summary[N.ALGO_DFS], want.summary[N.ALGO_GREEDY]
    assert (dfs["completed"][1::2] != greedy["completed"][1::2]).any()      # they do differ here
    assert (dfs["completed"][0::2] == greedy["completed"][0::2]).all()
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} dfs/greedy twins")
    for mask in (3, 2, 1, 11):
        got = schedule_batch(batch, mask, peak=True).host()
        for slot, a in enumerate(got.algos):
            np.testing.assert_array_equal(got.node_of[slot], want.node_of[a])
            np.testing.assert_array_equal(got.summary[slot], want.summary[a])


def test_single_algorithm_masks_and_slots(oracle_mod):
    """Any subset of the algorithms gives, slot by slot, what the oracle gives for that algorithm."""
    batch = random_dag_batch(16, 200, 8, 0.8, seed=5)
    full = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=4)
    for mask in (1, 2, 4, 8, 5, 10, 12):
        got = schedule_batch(batch, mask, peak=True).host()
        for slot, a in enumerate(got.algos):
            np.testing.assert_array_equal(got.node_of[slot], full.node_of[a])
            np.testing.assert_array_equal(got.seq_pos[slot], full.seq_pos[a])
            np.testing.assert_array_equal(got.summary[slot], full.summary[a])


def test_priorities_match_oracle(oracle_mod):
    for batch in (random_dag_batch(32, 1000, 8, 0.9, seed=3), random_dag_batch(64, 77, 4, 0.9, seed=4),
                  golden_io.batch_of(golden_io.load("cfg2_sweep_seed0.npz")[::5]),
                  golden_io.batch_of(golden_io.load("cfg3_gpt2.npz"))):
        got = priorities(batch)
        want = oracle_mod.priorities(batch, nthreads=4)
        for k in ("depth", "outdeg", "order_dg", "order_mru", "order_cp"):
            np.testing.assert_array_equal(got[k], want[k], err_msg=k)
        np.testing.assert_array_equal(got["up"].view(np.uint64), want["up"].view(np.uint64), err_msg="up bits")


def test_shared_topology_perturbed_durations(oracle_mod):
    """Config 3 shape: GPT-2 DAG on 8 capped devices, durations perturbed per instance (T == 1)."""
    base = [c for c in golden_io.load("cfg3_gpt2.npz") if c.label == "gpt2|8laptops"][0]
    B = 512
    dur = perturbed_durations(base.dag, B, 0.1, seed=0)
    batch = DagBatch.shared_topology(base.dag, base.cap, base.spd, B, dur=dur)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} gpt2 shared topology")
    # instance 0 with unperturbed durations reproduces the fixture
    plain = DagBatch.shared_topology(base.dag, base.cap, base.spd, 3)
    for path, res in run_both_paths(plain, N.MASK_ALL):
        for b in range(3):
            for slot, a in enumerate(res.algos):
                assert_matches_expected(base.exp[a], plain, res, slot, b, f"{path} gpt2 shared b={b} algo={a}")
    # DFS/Greedy/MRU decisions do not depend on durations (SURVEY.md 8a'): only makespan moves
    for slot in (0, 1, 3):
        n = base.n
        assert (got.node_of[slot].reshape(B, n) == got.node_of[slot, :n][None, :]).all()


def test_one_schedule_for_duration_only_batches(oracle_mod):
    """One topology, one node set, durations per instance: DFS / Greedy / MRU are scheduled once and replayed
    (dls_replay.cu) — bit for bit what scheduling every instance gives and what the oracle gives, under memory
    pressure too (failed and pending tasks, rounds, cache statistics are the template's), for any algorithm mask,
    with the extension outputs; and the short cut steps aside when instances differ in more than their durations."""
    import ctypes as C
    lib = N.load()
    base = [c for c in golden_io.load("cfg3_gpt2.npz") if c.label == "gpt2|8laptops"][0]
    pipe = [c for c in golden_io.load("cfg2_sweep_seed0.npz") if c.label.startswith("Pipeline")][4]
    rnd = random_dag_batch(1, 300, 8, 0.8, seed=12)
    from dls_b200.lowering import LoweredDag
    rdag = LoweredDag(names=[f"task_{i}" for i in range(300)], param_names=[f"p{i}" for i in range(rnd.P_max)],
                      pred_ptr=rnd.pred_ptr, pred_idx=rnd.pred_idx, par_ptr=rnd.par_ptr, par_idx=rnd.par_idx, rank=rnd.rank,
                      mem=rnd.mem, dur=rnd.dur)
    cases = [(base.dag, base.cap, base.spd), (base.dag, base.cap * 0.5, base.spd), (pipe.dag, pipe.cap * 0.8, pipe.spd),
             (rdag, rnd.cap, rnd.spd)]
    ms = (C.c_double * N.NUM_KERNELS)()
    cnt = (C.c_int64 * N.NUM_KERNELS)()
    for dag, cap, spd in cases:
        B = 96
        batch = DagBatch.shared_topology(dag, cap, spd, B, dur=perturbed_durations(dag, B, 0.1, seed=3))
        want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
        lib.dls_profile_enable(1)
        lib.dls_profile_read(ms, cnt)
        for i in range(N.NUM_KERNELS):
            cnt[i] = 0
        got = schedule_batch(batch, N.MASK_ALL, peak=True, cached=True).host()
        lib.dls_profile_read(ms, cnt)
        lib.dls_profile_enable(0)
        names = {lib.dls_kernel_name(k).decode(): int(cnt[k]) for k in range(N.NUM_KERNELS)}
        assert names["replay_kernel"] == 1, names
        assert_results_equal(got, want, "replayed, all algorithms")
        for mask in (1, 2, 8, 3, 11, 5, 4):
            got = schedule_batch(batch, mask).host()
            for slot, a in enumerate(got.algos):
                np.testing.assert_array_equal(got.node_of[slot], want.node_of[a])
                np.testing.assert_array_equal(got.seq_pos[slot], want.seq_pos[a])
                np.testing.assert_array_equal(got.fin[slot].view(np.uint64), want.fin[a].view(np.uint64))
                _summary_equal(got.summary[slot], want.summary[a], False)
        old = lib.dls_set_replay_path(0)
        try:
            plain = schedule_batch(batch, N.MASK_ALL, peak=True, cached=True).host()
        finally:
            lib.dls_set_replay_path(old)
        assert_results_equal(got := schedule_batch(batch, N.MASK_ALL, peak=True, cached=True).host(), plain, "replay vs every instance")
        # the template run beside the Critical run (auxiliary stream, the default) and behind it (one stream)
        assert lib.dls_set_replay_path(2) == 1
        try:
            one_stream = schedule_batch(batch, N.MASK_ALL, peak=True, cached=True).host()
        finally:
            assert lib.dls_set_replay_path(1) == 2
        assert_results_equal(one_stream, want, "replayed on one stream")
        # ... and on a stream of the caller's, again and again into the same buffers (fork / join per call)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            res = schedule_batch(batch, N.MASK_ALL, peak=True, cached=True)
            for _ in range(4):
                res = schedule_batch(batch, N.MASK_ALL, workspace=res.workspace, result=res)
        st.synchronize()
        assert_results_equal(res.host(), want, "replayed repeatedly on a side stream")
    # memory per instance as well: decisions differ between instances, nothing is replayed
    dag, cap, spd = cases[1]
    rng = np.random.default_rng(1)
    batch = DagBatch.shared_topology(dag, cap, spd, 64, dur=perturbed_durations(dag, 64, 0.1, seed=3),
                                     mem=dag.mem[None, :] * rng.uniform(0.7, 1.3, (64, dag.n)))
    lib.dls_profile_enable(1)
    lib.dls_profile_read(ms, cnt)
    for i in range(N.NUM_KERNELS):
        cnt[i] = 0
    got = schedule_batch(batch, N.MASK_ALL).host()
    lib.dls_profile_read(ms, cnt)
    lib.dls_profile_enable(0)
    assert cnt[N.NUM_KERNELS - 1] == 0
    assert_results_equal(got, oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8), "per-instance memory: no replay")


def test_size_independent_properties_full_size():
    """At BASELINE.json's 1k-task size without an oracle run: structural invariants of any valid outcome."""
    batch = random_dag_batch(2048, 1000, 8, 0.9, seed=11)
    res = schedule_batch(batch, N.MASK_ALL).host()
    n = 1000
    for slot in range(4):
        s = res.summary[slot]
        assert (s["status"] == 0).all()
        assert (s["completed"] + s["failed"] + s["pending"] == n).all()
        node_of = res.node_of[slot].reshape(-1, n)
        seq = res.seq_pos[slot].reshape(-1, n)
        done = node_of >= 0
        assert (done.sum(1) == s["completed"]).all()
        assert ((node_of == -1).sum(1) == s["failed"]).all()
        # assignment positions are a permutation of 0..completed-1
        srt = np.sort(np.where(done, seq, 0xFFFF), 1)
        want = np.where(np.arange(n)[None, :] < s["completed"][:, None], np.arange(n)[None, :], 0xFFFF)
        assert (srt == want).all()
        cnt = res.cnt[slot].reshape(-1, 8)
        assert (cnt.sum(1) == s["completed"]).all()
        fin = res.fin[slot].reshape(-1, 8)
        np.testing.assert_array_equal(np.where(s["completed"] > 0, fin.max(1), 0.0), s["makespan"])
        assert (s["hits"] + s["misses"] >= 0).all()
    # dependencies respected: a completed task is sequenced after all its dependencies (spot check 64 instances)
    for b in range(0, 2048, 32):
        ts = batch.task_slice(b)
        pp = batch.pred_ptr[b * (n + 1):(b + 1) * (n + 1)]
        pi = batch.pred_idx[batch.pred_off[b]:batch.pred_off[b + 1]]
        for slot in range(4):
            node_of, seq = res.node_of[slot, ts], res.seq_pos[slot, ts].astype(np.int64)
            child = np.repeat(np.arange(n), np.diff(pp))
            ok = node_of[child] >= 0
            assert (node_of[pi][ok] >= 0).all() and (seq[pi][ok] < seq[child][ok]).all()


def test_streamed_path_is_taken_and_falls_back():
    """The thread-per-DAG kernels really run for qualifying batches (their launches are counted), instances they
    cannot take (shared params under MRU: pipeline DAGs) are routed to the general kernels, and forcing the general
    path gives identical results."""
    import ctypes as C
    lib = N.load()
    lib.dls_profile_enable(1)
    ms = (C.c_double * N.NUM_KERNELS)()
    cnt = (C.c_int64 * N.NUM_KERNELS)()
    lib.dls_profile_read(ms, cnt)
    for i in range(N.NUM_KERNELS):
        cnt[i] = 0
    batch = random_dag_batch(64, 100, 8, 0.7, seed=77)
    fast = schedule_batch(batch, N.MASK_ALL).host()
    lib.dls_profile_read(ms, cnt)
    lib.dls_profile_enable(0)
    names = {lib.dls_kernel_name(k).decode(): int(cnt[k]) for k in range(N.NUM_KERNELS)}
    assert names["sched_stream_kernel<DFS>"] == 1 and names["sched_stream_mru_kernel"] == 1 and names["unpermute_kernel"] == 1
    assert names["priority_topology_kernel"] == 1 and names["param_class_kernel"] == 1
    assert names["stream_build_kernel"] == 1
    old = lib.dls_set_stream_path(0)
    try:
        slow = schedule_batch(batch, N.MASK_ALL).host()
    finally:
        lib.dls_set_stream_path(old)
    assert_results_equal(fast, slow, "streamed vs forced general")
    # mixed batch: random DAGs (all private params) + pipeline DAGs (shared params)
    cases = golden_io.load("cfg2_sweep_seed0.npz")
    mixed = [c for c in cases if c.label.startswith(("Pipeline", "Random-Small"))][:40]
    mb = golden_io.batch_of(mixed)
    res = schedule_batch(mb, N.MASK_ALL).host()
    for b, c in enumerate(mixed):
        for slot, a in enumerate(res.algos):
            assert_matches_expected(c.exp[a], mb, res, slot, b, f"mixed {c.label} algo={a}")


def test_device_generator_parity_and_shape(oracle_mod):
    """the device-side DAG builder emits valid CSR batches with generate_random_dag's shape, is
    reproducible, and what it emits schedules identically on the GPU and in the oracle."""
    from dls_b200.gen import random_dag_batch_device
    from test_abi_cpu import check_random_dag_structure
    for n, K, regime in ((300, 8, 0.9), (64, 4, 0.7), (1000, 2, 0.8)):
        dev = random_dag_batch_device(96, n, K, regime, seed=5)
        host = dev.to_host()
        check_random_dag_structure(host, n)
        again = random_dag_batch_device(96, n, K, regime, seed=5).to_host()
        other = random_dag_batch_device(96, n, K, regime, seed=6).to_host()
        for f in ("pred_ptr", "pred_idx", "par_idx", "mem", "dur", "cap", "spd"):
            np.testing.assert_array_equal(getattr(host, f), getattr(again, f))
        assert not np.array_equal(host.mem, other.mem)
        want = oracle_mod.schedule_batch(host, N.MASK_ALL, nthreads=8)
        got = schedule_batch(dev, N.MASK_ALL).host()          # straight from the device-resident batch
        assert_results_equal(got, want, f"device-generated n={n} K={K}")
        for path, res in run_both_paths(host, N.MASK_ALL):   # and through the host round trip
            assert_results_equal(res, want, f"{path} device-generated n={n} K={K}")


def test_device_generator_equals_its_host_port():
    """gen.random_dag_batch_counter restates the generator kernels in numpy (same counter hash, same arithmetic): the
    CPU arm of bench.py schedules the instances the GPU arm schedules without touching the GPU. Every array, bit for bit,
    for K = 2, 4 (fixed node sets) and 8 (drawn speeds), and for a window of instances out of the middle."""
    from dls_b200.gen import random_dag_batch_counter, random_dag_batch_device
    for n, K, regime, seed in ((300, 8, 0.9, 5), (64, 4, 0.7, 1000), (1000, 2, 0.8, 7), (1000, 8, 0.9, 1000)):
        rank = np.random.default_rng(n).permutation(n).astype(np.uint16)
        dev = random_dag_batch_device(40, n, K, regime, seed=seed, rank=rank).to_host()
        host = random_dag_batch_counter(40, n, K, regime, seed=seed, rank=rank)
        for f in ("task_off", "pred_off", "par_off", "pid_off", "node_off", "pred_ptr", "pred_idx", "par_ptr", "par_idx", "rank"):
            np.testing.assert_array_equal(getattr(dev, f), getattr(host, f), err_msg=f)
        for f in ("mem", "dur", "cap", "spd"):
            np.testing.assert_array_equal(getattr(dev, f).view(np.uint64), getattr(host, f).view(np.uint64), err_msg=f)
        assert (dev.E_max, dev.R_max, dev.P_max) == (host.E_max, host.R_max, host.P_max)
        window = random_dag_batch_counter(7, n, K, regime, seed=seed, rank=rank, first=21)
        np.testing.assert_array_equal(window.mem, host.mem[21 * n: 28 * n])
        np.testing.assert_array_equal(window.pred_idx, host.pred_idx[host.pred_off[21]: host.pred_off[28]])
        np.testing.assert_array_equal(window.cap, host.cap[21 * K: 28 * K])


def test_pipelined_host_path_matches_single_call(oracle_mod):
    """The chunked, multi-stream end-to-end path (host in, host out) returns exactly what one call returns."""
    from dls_b200.batch import PipelinedScheduler
    batch = random_dag_batch(150, 120, 8, 0.8, seed=31)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    want.peak = want.cached = None
    full = PipelinedScheduler(batch, N.MASK_ALL, chunks=7, nstreams=3, compact=False)
    pipe = PipelinedScheduler(batch, N.MASK_ALL, chunks=7, nstreams=3)
    # row pointers travel as one-byte row lengths, the (shared) rank row once per chunk: 10 of 32 bytes per task saved
    assert set(pipe._skip) == {"pred_ptr", "par_ptr", "rank"} and pipe.h2d_bytes < 0.8 * full.h2d_bytes
    for p in (pipe, full):
        for _ in range(2):  # buffers are reused across runs
            assert_results_equal(p.run().result(), want, "pipelined")
    # summary mode: only the per-instance records, node finish times and task counts come back
    brief = PipelinedScheduler(batch, N.MASK_ALL, chunks=7, nstreams=3, results="summary")
    got = brief.run().result()
    assert got.node_of is None and got.seq_pos is None and brief.d2h_bytes < 0.3 * pipe.d2h_bytes
    for f in want.summary.dtype.names:
        if f != "peak":
            np.testing.assert_array_equal(got.summary[f], want.summary[f], err_msg=f)
    np.testing.assert_array_equal(got.fin, want.fin)
    np.testing.assert_array_equal(got.cnt, want.cnt)


def test_pipelined_compact_transfer_ragged_and_fallback(oracle_mod):
    """dls_rows_from_degrees / dls_repeat_rank_row on what they must not mangle: instances of different sizes and
    their own rank rows (row pointers compacted, ranks not), and a task with more than 255 dependencies (its row
    pointers travel as they are)."""
    import ctypes as C
    import torch
    from dls_b200.batch import PipelinedScheduler, _row_lengths
    from dls_b200.lowering import LoweredDag
    cases = golden_io.load("cfg2_sweep_seed0.npz")[:60]       # n in {16, 30, 58, 60, 86}, own pi per instance
    rb = golden_io.batch_of(cases)
    pipe = PipelinedScheduler(rb, N.MASK_ALL, chunks=5, nstreams=2)
    assert set(pipe._skip) == {"pred_ptr", "par_ptr"}
    res = pipe.run().result()
    for b, c in enumerate(cases):
        for slot, a in enumerate(res.algos):
            assert_matches_expected(c.exp[a], rb, res, slot, b, f"compact ragged {c.label} algo={a}")
    # the expansion kernel alone, against the arrays it replaces
    lib = N.load()
    for f in ("pred_ptr", "par_ptr"):
        deg = torch.from_numpy(_row_lengths(getattr(rb, f), rb.task_off)).cuda()
        off = torch.from_numpy(rb.task_off).cuda()
        out = torch.zeros(rb.nt_topo + rb.T, dtype=torch.int32, device="cuda")
        N.check(lib.dls_rows_from_degrees(deg.data_ptr(), off.data_ptr(), rb.T, out.data_ptr(), None))
        np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), getattr(rb, f))
    # 300 dependencies on one task: no byte holds that
    n = 320
    wide = LoweredDag(names=[f"t{i}" for i in range(n)], param_names=[f"p{i}" for i in range(n)],
                      pred_ptr=np.concatenate([np.zeros(n, np.uint32), [300]]).astype(np.uint32),
                      pred_idx=np.arange(300, dtype=np.uint16), par_ptr=np.arange(n + 1, dtype=np.uint32),
                      par_idx=np.arange(n, dtype=np.uint16), rank=np.arange(n, dtype=np.uint16),
                      mem=np.full(n, 0.25), dur=np.full(n, 0.1))
    wb = DagBatch.from_instances([wide, wide, wide], [(np.full(4, 30.0), np.array([1.0, 1.2, 0.8, 1.0]))] * 3)
    wp = PipelinedScheduler(wb, N.MASK_ALL, chunks=2, nstreams=2)
    assert set(wp._skip) == {"par_ptr", "rank"}
    want = oracle_mod.schedule_batch(wb, N.MASK_ALL, nthreads=2)
    want.peak = want.cached = None
    assert_results_equal(wp.run().result(), want, "compact fallback (wide row)")


def test_error_codes():
    import ctypes as C
    lib = N.load()
    batch = random_dag_batch(4, 16, 8, 0.9, seed=1)
    with pytest.raises(N.DlsError) as ei:
        schedule_batch(batch, 0)
    assert ei.value.code == N.ERR_BAD_ARG
    batch.K_max = 33
    with pytest.raises(ValueError):
        schedule_batch(batch, 1)
    assert lib.dls_schedule_batch(None, None, 1, None, 0, None) == N.ERR_BAD_ARG
    assert b"NULL" in lib.dls_last_error()


def test_cycle_is_reported_not_hung():
    from dls_b200.lowering import LoweredDag
    dag = LoweredDag(names=["a", "b", "c"], param_names=["p"], pred_ptr=np.array([0, 1, 2, 3], np.uint32),
                     pred_idx=np.array([2, 0, 1], np.uint16), par_ptr=np.array([0, 1, 1, 1], np.uint32),
                     par_idx=np.array([0], np.uint16), rank=np.array([0, 1, 2], np.uint16),
                     mem=np.ones(3), dur=np.ones(3))
    batch = DagBatch.from_instances([dag], [(np.array([4.0, 4.0]), np.array([1.0, 1.0]))])
    res = schedule_batch(batch, N.MASK_ALL).host()
    assert (res.summary["status"] == N.ST_BAD_TOPOLOGY).all()
    assert (res.node_of == N.TASK_PENDING).all()


def _summary_equal(got, want, with_peak):
    """dls_summary_t records field by field; `peak` only where the run asked for peaks (it is 0.0 otherwise)."""
    for f in got.dtype.names:
        if f == "peak" and not with_peak:
            continue
        np.testing.assert_array_equal(got[f], want[f], err_msg=f)


def test_garbage_inputs_are_contained(oracle_mod):
    """Inputs outside the reference's domain must not derail the kernels: a rank[] that is not a permutation is a bad
    topology (it would leave positions of every processing order unwritten), and NaN durations — for which the
    reference's own sort order is undefined — still give a processing order that is a permutation, every task a
    verdict, and no effect on the neighbouring instances of the batch."""
    good = random_dag_batch(8, 200, 8, 0.9, seed=21)
    want = oracle_mod.schedule_batch(good, N.MASK_ALL, nthreads=4)   # what the untouched neighbours must still give
    bad = random_dag_batch(8, 200, 8, 0.9, seed=21)
    bad.rank[3 * 200 + 17] = bad.rank[3 * 200 + 18]          # instance 3: repeated rank
    bad.dur[5 * 200: 6 * 200][::7] = np.nan                  # instance 5: NaN durations
    for path, got in run_both_paths(bad, N.MASK_ALL):
        st = got.summary["status"]
        assert (st[:, 3] == N.ST_BAD_TOPOLOGY).all() and (got.node_of[:, 3 * 200: 4 * 200] == N.TASK_PENDING).all(), path
        s5 = got.summary[:, 5]
        assert (s5["status"] == 0).all() and (s5["completed"] + s5["failed"] + s5["pending"] == 200).all(), path
        for b in (0, 1, 2, 4, 6, 7):
            np.testing.assert_array_equal(got.node_of[:, b * 200: (b + 1) * 200], want.node_of[:, b * 200: (b + 1) * 200])
            _summary_equal(got.summary[:, b], want.summary[:, b], got.peak is not None)
    prio = priorities(bad)
    for b in (5,):
        for name in ("order_dg", "order_cp", "order_mru"):
            o = prio[name][b * 200: (b + 1) * 200]
            assert sorted(o.tolist()) == list(range(200)), (name, b)


def test_refused_assignment_is_flagged(oracle_mod):
    """The one corner the device does not replay: MRU's assign_task_to_node refusing after the node was chosen
    (ref schedulers.py:82-83,509). On the three reference-generated instances where that happens
    (cfg1b_mru_refused.npz; the oracle replays them literally, tests/test_oracle_golden.py) every device path sets
    DLS_ST_CARRY for MRU — exactly there: the other algorithms of those instances and every algorithm of the ordinary
    instances batched next to them come out as the reference's, and the Python class raises instead of returning a
    schedule the reference would not have produced."""
    from dls_b200.schedulers import MRUScheduler, Node, Task
    odd = golden_io.load("cfg1b_mru_refused.npz")
    plain = [c for c in golden_io.load("cfg1_example_and_edges.npz") if c.mask == N.MASK_ALL][:5]
    cases = [plain[0], odd[0], plain[1], odd[1], odd[2]] + plain[2:]
    batch = golden_io.batch_of(cases)
    for path, res in run_both_paths(batch, N.MASK_ALL):
        for b, c in enumerate(cases):
            refused = c in odd
            for slot, a in enumerate(res.algos):
                if refused and a == N.ALGO_MRU:
                    assert res.summary[slot, b]["status"] == N.ST_CARRY, f"{path} {c.label}: refusal not flagged"
                else:
                    assert_matches_expected(c.exp[a], batch, res, slot, b, f"{path} {c.label} algo={a}")
    c = odd[0]
    d = c.dag
    s = MRUScheduler([Node("only", float(c.cap[0]), float(c.spd[0]))])
    for i in range(d.n):
        s.add_task(Task(d.names[i], float(d.mem[i]), float(d.dur[i]), [d.names[j] for j in d.pred_idx[d.pred_ptr[i]:d.pred_ptr[i + 1]]],
                        {d.param_names[p] for p in d.par_idx[d.par_ptr[i]:d.par_ptr[i + 1]]}))
    # (with "side" ahead of "late" in the set order the eviction happens for "side" and nothing is refused: impose the
    # fixture's order, as tests/test_gpu_host_api.py does)
    from test_gpu_host_api import _OrderedPending
    s.pending_tasks = _OrderedPending([d.names[i] for i in np.argsort(d.rank)])
    with pytest.raises(RuntimeError, match="refused"):
        s.schedule()


def test_timeline_extension_matches_oracle(oracle_mod):
    """dls_result_t.tfin (extension: node finish times on a dependency-aware timeline) — no reference output exists
    for it, so parity is pinned by the oracle only: bit for bit, on random DAGs under memory pressure, on DAGs with
    shared params and wide dependency lists, and without disturbing any other output."""
    from dls_b200.gen import random_shared_param_batch
    for batch in (random_dag_batch(48, 300, 8, 0.8, seed=9), random_shared_param_batch(64, 60, 8, 12, 1.2, seed=4),
                  random_dag_batch(4, 1000, 4, 1.0, seed=2)):
        want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=4, timeline=True)
        got = schedule_batch(batch, N.MASK_ALL, timeline=True).host()
        np.testing.assert_array_equal(got.tfin, want.tfin)
        assert (got.tfin >= got.fin).all() and (got.tfin > got.fin).any()
        plain = schedule_batch(batch, N.MASK_ALL).host()
        for f in ("node_of", "seq_pos", "summary", "fin", "cnt", "avail"):
            np.testing.assert_array_equal(getattr(got, f), getattr(plain, f))
        np.testing.assert_array_equal(got.dependency_aware_makespan(batch), want.dependency_aware_makespan(batch))


def _relabelled(batch, b, perm, K):
    """Instance b of `batch` with task i renamed perm[i] (same DAG, listed in another order)."""
    from dls_b200.lowering import LoweredDag
    ts = batch.task_slice(b)
    n = ts.stop - ts.start
    pp = batch.pred_ptr[ts.start + b: ts.stop + b + 1].astype(np.int64)
    qp = batch.par_ptr[ts.start + b: ts.stop + b + 1].astype(np.int64)
    pi = batch.pred_idx[batch.pred_off[b]: batch.pred_off[b + 1]]
    qi = batch.par_idx[batch.par_off[b]: batch.par_off[b + 1]]
    inv = np.argsort(perm)                      # new index j holds old task inv[j]
    preds = [perm[pi[pp[i] - pp[0]: pp[i + 1] - pp[0]]] for i in inv]
    pars = [qi[qp[i] - qp[0]: qp[i + 1] - qp[0]] for i in inv]
    ks = batch.node_slice(b)
    dag = LoweredDag(names=[f"t{j}" for j in range(n)], param_names=[f"p{j}" for j in range(int(qi.max()) + 1)],
                     pred_ptr=np.concatenate([[0], np.cumsum([len(p) for p in preds])]).astype(np.uint32),
                     pred_idx=np.concatenate(preds).astype(np.uint16) if n else np.zeros(0, np.uint16),
                     par_ptr=np.concatenate([[0], np.cumsum([len(p) for p in pars])]).astype(np.uint32),
                     par_idx=np.concatenate(pars).astype(np.uint16), rank=batch.rank[ts][inv].astype(np.uint16),
                     mem=batch.mem[ts][inv], dur=batch.dur[ts][inv])
    return dag, (batch.cap[ks].copy(), batch.spd[ks].copy())


def test_tasks_listed_in_any_order(oracle_mod):
    """The reference's generators list every task after its dependencies, and kernel P has a fast pass for that; a
    DAG listed in any other order (dependencies with larger indices) takes the level-synchronous relaxation instead
    and must give the reference's answer all the same — reversed, and randomly relabelled."""
    base = random_dag_batch(24, 180, 8, 0.85, seed=41)
    rng = np.random.default_rng(5)
    dags, nodes = [], []
    for b in range(base.B):
        perm = np.arange(180)[::-1].copy() if b % 2 == 0 else rng.permutation(180)
        d, nd = _relabelled(base, b, perm, 8)
        dags.append(d)
        nodes.append(nd)
    batch = DagBatch.from_instances(dags, nodes)
    assert (batch.pred_idx.astype(np.int64).max() > 0)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=4)
    assert (want.summary["status"] == 0).all() and want.summary["completed"].sum() > 0
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} relabelled tasks")


def _hub_dag(rng, n, fan, K):
    """A DAG in which one early task (the hub) cannot fit anywhere: it fails, and everything that depends on it,
    directly (`fan` tasks) or through other tasks, must stay pending. Other dependencies are random."""
    from dls_b200.lowering import LoweredDag
    hub = int(rng.integers(0, max(1, n // 4)))
    later = np.arange(hub + 1, n)
    direct = set(rng.choice(later, size=min(fan, len(later)), replace=False).tolist()) if len(later) else set()
    preds = []
    for i in range(n):
        p = set()
        if i in direct:
            p.add(hub)
        for _ in range(int(rng.integers(0, 3))):
            if i > 0:
                p.add(int(rng.integers(0, i)))
        preds.append(sorted(p))
    npar = rng.integers(1, 3, size=n)
    mem = rng.uniform(0.1, 0.5, size=n)
    cap = np.full(K, float((mem + 0.5 * npar).max() + 0.5 * npar.sum() * rng.uniform(0.3, 1.1) / K))
    mem[hub] = 10.0 * cap[0] + 5.0
    dag = LoweredDag(names=[f"t{j}" for j in range(n)], param_names=[f"p{j}" for j in range(int(npar.sum()))],
                     pred_ptr=np.concatenate([[0], np.cumsum([len(p) for p in preds])]).astype(np.uint32),
                     pred_idx=np.asarray([q for p in preds for q in p], np.uint16),
                     par_ptr=np.concatenate([[0], np.cumsum(npar)]).astype(np.uint32),
                     par_idx=np.arange(int(npar.sum()), dtype=np.uint16), rank=rng.permutation(n).astype(np.uint16),
                     mem=mem, dur=rng.uniform(0.05, 0.15, size=n))
    return dag, (cap, rng.uniform(0.5, 2.0, size=K))


@pytest.mark.parametrize("sizes,fans", [((5, 8, 9, 17, 40, 100, 300), (1, 3, 4, 5, 17, 10 ** 6)), ((2100, 2500), (4, 5, 600, 10 ** 6))])
def test_failed_tasks_block_their_dependents_at_any_fan_out(oracle_mod, sizes, fans):
    """is_task_ready on the thread-per-DAG path is a mark set by the dependency that ended without a node (reverse
    CSR, the first four dependents fetched ahead, the rest on the spot; marks drained when a round starts): hubs with
    fewer, exactly, and many more dependents than the pipeline's width, cascades over several rounds, instance lengths
    around the 8-task blocks of the streams, and DAGs beyond 2,048 tasks (mark bytes in HBM instead of bits in shared
    memory) — every device path against the oracle, bit for bit."""
    rng = np.random.default_rng(len(sizes) * 1000 + 7)
    dags, nodes = [], []
    for n in sizes:
        for fan in fans:
            for _ in range(2):
                d, nd = _hub_dag(rng, n, fan, 4)
                dags.append(d)
                nodes.append(nd)
    batch = DagBatch.from_instances(dags, nodes)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    assert (want.summary["status"] == 0).all()
    assert (want.summary["failed"] >= 1).all() and want.summary["pending"].sum() > 0 and want.summary["completed"].sum() > 0
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} hub DAGs n={sizes}")


def test_dense_dags_take_the_unstaged_paths(oracle_mod):
    """Many dependencies per task (E up to ~8n, in-degrees up to 12, out-degrees in the dozens): the dependency CSR no
    longer fits the shared-memory bytes the priority kernels stage it in (they then read it from global memory), the
    depth pass meets tasks with more than three dependencies, failed tasks have more dependents than the marks
    pipeline fetches ahead — under memory pressure, with private and shared params, against the oracle."""
    from dls_b200.lowering import LoweredDag
    rng = np.random.default_rng(77)
    dags, nodes = [], []
    for b in range(40):
        n = int(rng.choice([24, 90, 200, 640]))
        K = int(rng.choice([2, 5, 8]))
        preds = [sorted(set(rng.integers(0, i, size=int(rng.integers(0, 13))).tolist())) if i else [] for i in range(n)]
        pool = 0 if b % 2 == 0 else 6                      # every other instance shares a few params
        pars, P = [], pool
        for i in range(n):
            own = list(range(P, P + int(rng.integers(0, 3))))
            P += len(own)
            shared = sorted(set(rng.integers(0, pool, size=int(rng.integers(0, 3))).tolist())) if pool else []
            pars.append(shared + own)
        mem = rng.uniform(0.1, 0.5, size=n)
        npar = np.array([len(p) for p in pars])
        cap = np.full(K, float((mem + 0.5 * npar).max() + 0.5 * npar.sum() * rng.uniform(0.25, 0.9) / K))
        dags.append(LoweredDag(
            names=[f"t{j}" for j in range(n)], param_names=[f"p{j:05d}" for j in range(max(P, 1))],
            pred_ptr=np.concatenate([[0], np.cumsum([len(p) for p in preds])]).astype(np.uint32),
            pred_idx=np.asarray([q for p in preds for q in p], np.uint16),
            par_ptr=np.concatenate([[0], np.cumsum(npar)]).astype(np.uint32),
            par_idx=np.asarray([q for p in pars for q in p], np.uint16), rank=rng.permutation(n).astype(np.uint16),
            mem=mem, dur=rng.uniform(0.05, 0.15, size=n)))
        nodes.append((cap, rng.uniform(0.5, 2.0, size=K)))
    batch = DagBatch.from_instances(dags, nodes)
    assert batch.e_total > 4 * batch.nt_inst
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    assert (want.summary["status"] == 0).all()
    assert want.summary["failed"].sum() > 0 and want.summary["pending"].sum() > 0 and want.summary["completed"].sum() > 0
    for path, got in run_both_paths(batch, N.MASK_ALL):
        assert_results_equal(got, want, f"{path} dense DAGs")


def test_scheduling_call_can_be_captured_into_a_cuda_graph(oracle_mod):
    """dls_schedule_batch only enqueues work on the caller's stream (the replay path forks an auxiliary stream from it
    by an event and joins it back): after one ordinary call it can be captured into a CUDA graph and replayed, for an
    instance-per-topology batch (thread-per-DAG kernels) and for a duration-only batch (template run + replay). The
    result buffers are cleared between replays; what the graph writes must be the oracle's result every time."""
    from dls_b200.batch import DeviceBatch
    base = [c for c in golden_io.load("cfg3_gpt2.npz") if c.label == "gpt2|8laptops"][0]
    shared = DagBatch.shared_topology(base.dag, base.cap * 0.5, base.spd, 64, dur=perturbed_durations(base.dag, 64, 0.1, seed=9))
    for batch in (random_dag_batch(96, 200, 8, 0.8, seed=5), shared):
        want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
        dev = DeviceBatch(batch, torch.device("cuda", 0))
        res = schedule_batch(dev, N.MASK_ALL)           # ordinary call first: function attributes, auxiliary stream
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            schedule_batch(dev, N.MASK_ALL, workspace=res.workspace, result=res)
        for _ in range(3):
            for t in (res.node_of, res.seq_pos, res.summary, res.fin, res.cnt, res.avail):
                t.zero_()
            g.replay()
            torch.cuda.synchronize()
            assert_results_equal(res.host(), want, "graph replay")
