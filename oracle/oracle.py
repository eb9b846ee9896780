"""ctypes wrapper of oracle/libdls_oracle.so — TEST INFRASTRUCTURE, not product code.

Allowed importers: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
Parity status: PINNED (tests/test_oracle_golden.py against fixtures produced by the reference itself;
tests/test_oracle_vs_reference.py live against /root/reference when present).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from dls_b200 import _native as N
from dls_b200.batch import ResultArrays
from dls_b200.lowering import DagBatch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libdls_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "dls_oracle.c")
    hdr = os.path.join(HERE, "..", "include", "dls.h")
    if (not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(src)
            and os.path.getmtime(LIB) >= os.path.getmtime(hdr)):
        return LIB
    proc = subprocess.run(["make", "-B", "-C", HERE], capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + proc.stdout + proc.stderr)
    return LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        lib = C.CDLL(LIB)
        lib.dls_oracle_schedule_batch.restype = C.c_int
        lib.dls_oracle_schedule_batch.argtypes = [C.POINTER(N.DlsBatch), C.POINTER(N.DlsResult), C.c_int, C.c_int]
        lib.dls_oracle_priorities.restype = C.c_int
        lib.dls_oracle_priorities.argtypes = [C.POINTER(N.DlsBatch), C.POINTER(N.DlsPrio), C.c_int]
        _lib = lib
    return _lib


def schedule_batch(batch: DagBatch, mask: int = N.MASK_ALL, nthreads: int = 1, timeline: bool = False) -> ResultArrays:
    """Oracle twin of dls_b200.schedule_batch on host arrays."""
    lib = load()
    res = ResultArrays.empty(batch, mask, timeline=timeline)
    bs = batch.host_struct()
    rs = res.host_struct()
    rc = lib.dls_oracle_schedule_batch(C.byref(bs), C.byref(rs), mask, nthreads)
    if rc != 0:
        raise RuntimeError(f"oracle rc={rc}")
    return res


def priorities(batch: DagBatch, nthreads: int = 1) -> dict:
    lib = load()
    out = {
        "depth": np.zeros(batch.nt_topo, np.uint16), "outdeg": np.zeros(batch.nt_topo, np.uint16),
        "up": np.zeros(batch.nt_inst, np.float64), "order_dg": np.zeros(batch.nt_topo, np.uint16),
        "order_mru": np.zeros(batch.nt_topo, np.uint16), "order_cp": np.zeros(batch.nt_inst, np.uint16),
    }
    pr = N.DlsPrio()
    for k, a in out.items():
        setattr(pr, k, a.ctypes.data)
    bs = batch.host_struct()
    rc = lib.dls_oracle_priorities(C.byref(bs), C.byref(pr), nthreads)
    if rc != 0:
        raise RuntimeError(f"oracle rc={rc}")
    return out
