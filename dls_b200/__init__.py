"""Importable name of the package that lives in `distributed-llm-scheduler_b200/` (a hyphen cannot be imported).

`import dls_b200` resolves submodules (`dls_b200.schedulers`, `dls_b200.batch`, ...) from that directory.
"""
import os as _os

__path__.insert(0, _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                                 "distributed-llm-scheduler_b200"))

from . import _native  # noqa: E402
from .lowering import DagBatch, LoweredDag, lower_nodes, lower_tasks, set_order  # noqa: E402,F401
from .batch import (BatchResult, DeviceBatch, PipelinedScheduler, ResultArrays, algos_of, pin_batch, priorities,  # noqa: E402,F401
                    schedule_batch, schedule_dict, sequence_of)
from .schedulers import (SCHEDULERS, BaseScheduler, CriticalPathScheduler, DFSScheduler,  # noqa: E402,F401
                         GreedyScheduler, MRUScheduler, Node, Task)

ALGO_DFS, ALGO_GREEDY, ALGO_CRITICAL, ALGO_MRU = (_native.ALGO_DFS, _native.ALGO_GREEDY, _native.ALGO_CRITICAL,
                                                  _native.ALGO_MRU)
MASK_ALL = _native.MASK_ALL
