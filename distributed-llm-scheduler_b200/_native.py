"""ctypes view of the C ABI in include/dls.h and the loader/builder of the CUDA library.

The product path has NO CPU fallback: `load()` raises if `libdls_b200.so` is missing, and every compute entry
point needs a CUDA device. (The CPU oracle lives under oracle/ and is never imported from here.)
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import threading

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC_DIR = os.path.join(PKG_DIR, "csrc")
# DLS_B200_LIB: load another build of the same sources instead (kernel A/B measurements, tools/build_variant.py)
LIB_PATH = os.environ.get("DLS_B200_LIB") or os.path.join(PKG_DIR, "libdls_b200.so")
INCLUDE_DIR = os.path.join(REPO_ROOT, "include")

ALGO_DFS, ALGO_GREEDY, ALGO_CRITICAL, ALGO_MRU = 0, 1, 2, 3
NUM_ALGOS = 4
MASK_ALL = 0xF
MAX_TASKS, MAX_PARAMS, MAX_NODES = 65535, 65535, 32
F_DUR_PER_INSTANCE, F_MEM_PER_INSTANCE = 1, 2
TASK_FAILED, TASK_PENDING, SEQ_NONE = -1, -2, 0xFFFF
ST_BAD_TOPOLOGY, ST_CARRY = 1, 2
OK, ERR_BAD_ARG, ERR_TOO_LARGE, ERR_WORKSPACE, ERR_CUDA = 0, -1, -2, -3, -4

ALGO_NAMES = {ALGO_DFS: "DFS", ALGO_GREEDY: "Greedy", ALGO_CRITICAL: "Critical", ALGO_MRU: "MRU_spec"}


class DlsBatch(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("T", C.c_int32), ("n_max", C.c_int32), ("K_max", C.c_int32),
        ("P_max", C.c_int32), ("E_max", C.c_int32), ("R_max", C.c_int32), ("flags", C.c_int32),
        ("nt_topo", C.c_int64), ("nt_inst", C.c_int64), ("e_total", C.c_int64), ("r_total", C.c_int64),
        ("nk_inst", C.c_int64), ("np_inst", C.c_int64),
        ("task_off", C.c_void_p), ("pred_off", C.c_void_p), ("par_off", C.c_void_p), ("pid_off", C.c_void_p),
        ("pred_ptr", C.c_void_p), ("pred_idx", C.c_void_p), ("par_ptr", C.c_void_p), ("par_idx", C.c_void_p),
        ("rank", C.c_void_p), ("mem", C.c_void_p), ("dur", C.c_void_p),
        ("node_off", C.c_void_p), ("cap", C.c_void_p), ("spd", C.c_void_p),
    ]


class DlsSummary(C.Structure):
    _fields_ = [
        ("makespan", C.c_double), ("peak", C.c_double), ("completed", C.c_int32), ("failed", C.c_int32), ("pending", C.c_int32),
        ("hits", C.c_int32), ("misses", C.c_int32), ("rounds", C.c_int32), ("status", C.c_int32),
        ("reserved", C.c_int32),
    ]


SUMMARY_BYTES = C.sizeof(DlsSummary)  # 48
SUMMARY_DTYPE = [
    ("makespan", "<f8"), ("peak", "<f8"), ("completed", "<i4"), ("failed", "<i4"), ("pending", "<i4"), ("hits", "<i4"),
    ("misses", "<i4"), ("rounds", "<i4"), ("status", "<i4"), ("reserved", "<i4"),
]


class DlsResult(C.Structure):
    _fields_ = [
        ("node_of", C.c_void_p), ("seq_pos", C.c_void_p), ("summary", C.c_void_p), ("fin", C.c_void_p),
        ("cnt", C.c_void_p), ("avail", C.c_void_p), ("peak", C.c_void_p), ("cached", C.c_void_p), ("tfin", C.c_void_p),
    ]


class DlsGen(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("n", C.c_int32), ("K", C.c_int32), ("max_deps", C.c_int32), ("regime", C.c_double),
        ("seed", C.c_uint64), ("rank", C.c_void_p), ("param_pos", C.c_void_p),
    ]


class DlsPrio(C.Structure):
    _fields_ = [
        ("depth", C.c_void_p), ("outdeg", C.c_void_p), ("up", C.c_void_p), ("order_dg", C.c_void_p),
        ("order_mru", C.c_void_p), ("order_cp", C.c_void_p),
    ]


# Every symbol include/dls.h declares; tests check the built library exports all of them.
EXPORTED_SYMBOLS = (
    "dls_version", "dls_last_error", "dls_workspace_bytes", "dls_schedule_batch", "dls_priorities",
    "dls_launch_count", "dls_profile_enable", "dls_profile_read", "dls_kernel_name", "dls_schedule_batch_static", "dls_set_stream_path", "dls_set_replay_path", "dls_generate_counts", "dls_generate_fill", "dls_aggregate", "dls_rows_from_degrees", "dls_repeat_rank_row",
)
STATIC_SCRATCH_BYTES = 32 << 20
AGG_FIELDS = 8
AGG_NAMES = ("count", "completion_rate", "makespan", "avg_node_utilization", "load_balance_score",
             "param_cache_hits", "param_cache_misses", "with_completed")
NUM_KERNELS = 14

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",  # the reference's fp64 op order must be replayed without FMA contraction
    "-Xcompiler", "-fPIC", "-shared",
]


def sources():
    return sorted(os.path.join(CSRC_DIR, f) for f in os.listdir(CSRC_DIR) if f.endswith(".cu"))


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = sources() + [os.path.join(CSRC_DIR, f) for f in os.listdir(CSRC_DIR) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(INCLUDE_DIR, "dls.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, defines=(), out: str = None) -> str:
    """Compile csrc/*.cu for sm_100a into libdls_b200.so, in-tree (so it travels to the GPU box). `defines` / `out`
    build a variant library next to it (extra -D macros) for A/B measurements."""
    out = out or LIB_PATH
    if not force and not defines and out == LIB_PATH and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc] + NVCC_FLAGS + [f"-D{d}" for d in defines] + ["-I", INCLUDE_DIR, "-I", CSRC_DIR, "-o", out] + sources()
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + proc.stdout + proc.stderr)
    if verbose:
        print(proc.stderr)
    return out


_lib = None
_lock = threading.Lock()


def load() -> C.CDLL:
    """Load libdls_b200.so; raise loudly if it is not there (there is no other implementation)."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA library is the only implementation of the scheduling path "
                "(no CPU fallback). Build it with `python -c 'import __graft_entry__ as g; g.build()'`.")
        lib = C.CDLL(LIB_PATH)
        lib.dls_version.restype = C.c_int
        lib.dls_last_error.restype = C.c_char_p
        lib.dls_workspace_bytes.restype = C.c_size_t
        lib.dls_workspace_bytes.argtypes = [C.POINTER(DlsBatch), C.c_int]
        lib.dls_schedule_batch.restype = C.c_int
        lib.dls_schedule_batch.argtypes = [C.POINTER(DlsBatch), C.POINTER(DlsResult), C.c_int, C.c_void_p,
                                           C.c_size_t, C.c_void_p]
        lib.dls_schedule_batch_static.restype = C.c_int
        lib.dls_schedule_batch_static.argtypes = [C.POINTER(DlsBatch), C.POINTER(DlsResult), C.c_int, C.c_void_p]
        lib.dls_generate_counts.restype = C.c_int
        lib.dls_generate_counts.argtypes = [C.POINTER(DlsGen), C.c_void_p, C.c_void_p, C.c_void_p]
        lib.dls_generate_fill.restype = C.c_int
        lib.dls_generate_fill.argtypes = [C.POINTER(DlsGen), C.POINTER(DlsBatch), C.c_void_p]
        lib.dls_rows_from_degrees.restype = C.c_int
        lib.dls_rows_from_degrees.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        lib.dls_repeat_rank_row.restype = C.c_int
        lib.dls_repeat_rank_row.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        lib.dls_aggregate.restype = C.c_int
        lib.dls_aggregate.argtypes = [C.POINTER(DlsBatch), C.POINTER(DlsResult), C.c_int, C.c_void_p, C.c_int32,
                                      C.c_void_p, C.c_void_p]
        lib.dls_set_stream_path.restype = C.c_int
        lib.dls_set_stream_path.argtypes = [C.c_int]
        lib.dls_set_replay_path.restype = C.c_int
        lib.dls_set_replay_path.argtypes = [C.c_int]
        lib.dls_priorities.restype = C.c_int
        lib.dls_priorities.argtypes = [C.POINTER(DlsBatch), C.POINTER(DlsPrio), C.c_void_p, C.c_size_t, C.c_void_p]
        lib.dls_launch_count.restype = C.c_int64
        lib.dls_profile_enable.restype = C.c_int
        lib.dls_profile_enable.argtypes = [C.c_int]
        lib.dls_profile_read.restype = C.c_int
        lib.dls_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        lib.dls_kernel_name.restype = C.c_char_p
        lib.dls_kernel_name.argtypes = [C.c_int]
        _lib = lib
        return lib


class DlsError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"dls error {code}: {text}")
        self.code = code


def check(rc: int):
    if rc != OK:
        raise DlsError(rc, load().dls_last_error().decode("utf-8", "replace"))
