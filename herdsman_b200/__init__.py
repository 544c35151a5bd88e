"""herdsman_b200 -- B200-native gate-evaluation backend for herd-fhe/herdsman's map-task hot path.

Only what the path needs: `csrc/` (sm_100a CUDA kernels + the C ABI of include/herdb.h, built
in-tree as libherdb.so), `api` (ctypes binding and the Python mirror of IWorkerGroup / MapTask)
and `circuit` (mirror of herd::common::Circuit and the benchmark circuit builders).
"""
from . import api, circuit  # noqa: F401
from .api import Backend, CudaWorkerGroup, MapTask, TaskHandle, HerdbError  # noqa: F401
from .circuit import Circuit, DataType, Operation  # noqa: F401
