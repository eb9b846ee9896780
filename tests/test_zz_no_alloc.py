"""dls_schedule_batch_static: the entry point for hosts that cannot give the library scratch memory (a plugin
handed only its input/output buffers by an application that owns the rest of the GPU, a loop that must not risk
the device-wide stall a cudaMalloc can cause, a runtime with no device allocator). Its scratch is a fixed array
inside the library, shared by every call on a device and never cleared between calls, so the thing to prove is
that a second call, with a different batch, is not disturbed by what the first one left behind.

The file name keeps this test last in collection order: if it fails under `-x`, everything else has already run.
It must also pass on its own in a fresh process (the first call then does the per-device address look-up).
"""
import ctypes as C

import pytest
import torch

from common import assert_results_equal
from dls_b200 import _native as N
from dls_b200.batch import BatchResult, DeviceBatch, schedule_batch
from dls_b200.gen import random_dag_batch

pytestmark = pytest.mark.gpu


def _static_call(host_batch, mask, peak=True, cached=True):
    """Call the new entry point directly through the C ABI: device inputs, device outputs, no workspace argument."""
    lib = N.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    d = DeviceBatch(host_batch, dev)
    out = BatchResult(host_batch, mask, dev, peak=peak, cached=cached)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.dls_schedule_batch_static(C.byref(d.struct), C.byref(out.struct), mask, C.c_void_p(stream))
    N.check(rc)
    torch.cuda.synchronize(dev)
    return out.host()


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    # first call: many small DAGs with shared memory pressure; second call: fewer, larger DAGs, another node count,
    # so every scratch region is reused at different offsets and with stale contents from the first call
    first = random_dag_batch(192, 120, 4, 0.6, seed=21)
    second = random_dag_batch(48, 700, 8, 0.85, seed=22)
    for which, batch in (("first", first), ("second", second)):
        want = schedule_batch(batch, N.MASK_ALL).host()   # existing entry point, caller-allocated workspace
        got = _static_call(batch, N.MASK_ALL)
        assert_results_equal(got, want, f"static scratch, {which} call")   # bit-exact, like every other parity test
    # and a single-algorithm mask right after, still on the same array
    got = _static_call(first, 1 << N.ALGO_MRU)
    want = schedule_batch(first, 1 << N.ALGO_MRU).host()
    assert_results_equal(got, want, "static scratch, third call (MRU only)")


def test_static_scratch_thread_per_dag_kernels_on_stale_contents(oracle_mod):
    """Without the `cached` output a batch that fits takes the thread-per-DAG kernels on the built-in array as well
    (record streams, verdict bytes, routing bytes, param classes all live in it). They run here right after a call of
    another shape has left its data everywhere in the array, are checked against the CPU oracle bit for bit, and the
    launch counters show that it really was those kernels."""
    lib = N.load()
    before = random_dag_batch(160, 150, 4, 0.6, seed=31)          # leaves other offsets and sizes behind
    _static_call(before, N.MASK_ALL)
    batch = random_dag_batch(40, 600, 8, 0.85, seed=32)
    want = oracle_mod.schedule_batch(batch, N.MASK_ALL, nthreads=8)
    ms = (C.c_double * N.NUM_KERNELS)()
    cnt = (C.c_int64 * N.NUM_KERNELS)()
    lib.dls_profile_enable(1)
    lib.dls_profile_read(ms, cnt)
    for i in range(N.NUM_KERNELS):
        cnt[i] = 0
    for _ in range(2):   # twice: the second call reads an array full of the first one's own leftovers
        got = _static_call(batch, N.MASK_ALL, peak=True, cached=False)
        assert_results_equal(got, want, "static scratch, thread-per-DAG kernels")
    lib.dls_profile_read(ms, cnt)
    lib.dls_profile_enable(0)
    names = {lib.dls_kernel_name(k).decode(): int(cnt[k]) for k in range(N.NUM_KERNELS)}
    assert names["stream_build_kernel"] == 2 and names["sched_stream_mru_kernel"] == 2 and names["unpermute_kernel"] == 2, names


def test_static_scratch_refuses_batches_beyond_its_size():
    lib = N.load()
    big = random_dag_batch(2048, 1000, 8, 0.9, seed=23)   # needs ~50 MB of scratch; the built-in array is 32 MiB
    dev = torch.device("cuda", torch.cuda.current_device())
    d = DeviceBatch(big, dev)
    assert lib.dls_workspace_bytes(C.byref(d.struct), N.MASK_ALL) > N.STATIC_SCRATCH_BYTES
    out = BatchResult(big, N.MASK_ALL, dev, peak=False, cached=False)
    before = lib.dls_launch_count()
    rc = lib.dls_schedule_batch_static(C.byref(d.struct), C.byref(out.struct), N.MASK_ALL, None)
    assert rc == N.ERR_TOO_LARGE
    msg = lib.dls_last_error().decode()
    assert "built-in scratch" in msg and str(N.STATIC_SCRATCH_BYTES) in msg
    assert lib.dls_launch_count() == before   # refused before anything was enqueued
