"""ctypes binding of libherdb.so (include/herdb.h) and the Python mirror of herdsman's worker seam.

`CudaWorkerGroup.schedule_task(MapTask) -> TaskHandle` mirrors
IWorkerGroup::schedule_task / TaskHandle of the reference
(include/execution/worker/i_worker_group.hpp:12-49, src/execution/worker/i_worker_group.cpp:4-20):
same names, same status enum, completion callback installed after scheduling.

There is no CPU fallback: if the CUDA library is missing or no device is present every entry
point raises.
"""
import ctypes as C
import dataclasses
import enum
import os
import threading

import numpy as np

from .circuit import Circuit, herdb_circuit, herdb_circuit_info, herdb_schedule_info

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HERDB_LIB") or os.path.join(_HERE, "libherdb.so")      # HERDB_LIB: an A/B build of the same library
_lib = None

N_STD128 = 512
N_TEST = 32
RING_N = 1024
NOALLOC_MAX_BATCH = 2368        # HERDB_NOALLOC_MAX_BATCH (include/herdb_params.h)
GATE = {"OR": 0, "AND": 1, "NOR": 2, "NAND": 3, "XOR_FAST": 4, "XNOR_FAST": 5}
EXPORTS = [
    "herdb_create", "herdb_destroy", "herdb_strerror", "herdb_last_error", "herdb_device_count",
    "herdb_load_key", "herdb_load_key_mem", "herdb_load_key_dev", "herdb_unload_key", "herdb_key_lwe_n",
    "herdb_map_async", "herdb_reduce_async", "herdb_reduce_rows", "herdb_wait_idle", "herdb_partition_read", "herdb_partition_write", "herdb_partition_rows",
    "herdb_eval_gate_batch", "herdb_eval_circuit_rows", "herdb_eval_pipeline_rows", "herdb_eval_gate_batch_dev", "herdb_eval_circuit_rows_dev",
    "herdb_circuit_analyze", "herdb_circuit_schedule", "herdb_circuit_eval_plain", "herdb_get_stats", "herdb_get_stream", "herdb_reset_stats", "herdb_measure_int_peak",
    "herdb_task_threads", "herdb_key_blob_size", "herdb_loaded_keys", "herdb_partition_count", "herdb_eval_gate_batch_dev_shape",
    "herdb_measure_l2_bandwidth", "herdb_eval_gate_batch_dev_noalloc", "herdb_task_window",
]


class HerdbError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        name = lib().herdb_strerror(code).decode()
        super().__init__("herdb error %d (%s)%s" % (code, name, ": " + detail if detail else ""))


class herdb_stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("gates_bootstrapped", C.c_uint64),
                ("blind_rotate_ms", C.c_double), ("keyswitch_ms", C.c_double),
                ("blind_rotate_launches", C.c_uint64), ("keyswitch_launches", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("key_load_ms", C.c_double), ("key_bcast_ms", C.c_double)]


DONE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int)


class herdb_input_frame(C.Structure):       # include/herdb.h: InputDataFramePtr of a ReduceTask
    _fields_ = [("df_uuid", C.c_char_p), ("partition", C.c_uint32), ("row_count", C.c_uint64)]


def lib():
    """Load libherdb.so; raises if the CUDA extension has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("herdsman_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; "
                               "g.build()'` or `make -C herdsman_b200/csrc`; there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, cp, sz, ci, u32, u64 = C.c_void_p, C.c_char_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64
        L.herdb_create.argtypes = [C.POINTER(vp), C.POINTER(ci), ci, cp, cp]
        L.herdb_destroy.argtypes = [vp]; L.herdb_destroy.restype = None
        L.herdb_strerror.argtypes = [ci]; L.herdb_strerror.restype = cp
        L.herdb_last_error.argtypes = [vp]; L.herdb_last_error.restype = cp
        L.herdb_device_count.argtypes = [vp]
        L.herdb_load_key.argtypes = [vp, cp, ci]
        L.herdb_load_key_mem.argtypes = [vp, cp, vp, sz]
        L.herdb_load_key_dev.argtypes = [vp, cp, ci, vp, sz]
        L.herdb_unload_key.argtypes = [vp, cp]
        L.herdb_key_lwe_n.argtypes = [vp, cp, C.POINTER(u32)]
        L.herdb_map_async.argtypes = [vp, cp, cp, u32, u64, cp, u32, ci, C.POINTER(herdb_circuit), DONE_FN, vp]
        L.herdb_reduce_async.argtypes = [vp, cp, C.POINTER(herdb_input_frame), sz, cp, u32, ci, C.POINTER(herdb_circuit), DONE_FN, vp]
        L.herdb_reduce_rows.argtypes = [vp, cp, C.POINTER(herdb_circuit), vp, sz, vp]
        L.herdb_wait_idle.argtypes = [vp]
        L.herdb_partition_read.argtypes = [cp, u64, sz, vp]
        L.herdb_partition_write.argtypes = [cp, u64, sz, vp]
        L.herdb_partition_rows.argtypes = [u64, u32, u32]; L.herdb_partition_rows.restype = u64
        L.herdb_eval_gate_batch.argtypes = [vp, cp, ci, ci, vp, vp, vp, sz]
        L.herdb_eval_gate_batch_dev.argtypes = [vp, cp, ci, ci, vp, vp, vp, sz]
        L.herdb_eval_circuit_rows.argtypes = [vp, cp, C.POINTER(herdb_circuit), vp, vp, sz]
        L.herdb_eval_pipeline_rows.argtypes = [vp, cp, C.POINTER(C.POINTER(herdb_circuit)), sz, vp, vp, sz]
        L.herdb_eval_circuit_rows_dev.argtypes = [vp, cp, ci, C.POINTER(herdb_circuit), vp, vp, sz]
        L.herdb_circuit_analyze.argtypes = [C.POINTER(herdb_circuit), C.POINTER(herdb_circuit_info)]
        L.herdb_circuit_schedule.argtypes = [C.POINTER(herdb_circuit), u64, u32, C.POINTER(herdb_schedule_info), vp, sz]
        L.herdb_circuit_eval_plain.argtypes = [C.POINTER(herdb_circuit), u64, u32, vp, vp]
        L.herdb_get_stats.argtypes = [vp, ci, C.POINTER(herdb_stats)]
        L.herdb_get_stream.argtypes = [vp, ci, C.POINTER(vp)]
        L.herdb_reset_stats.argtypes = [vp]
        L.herdb_measure_int_peak.argtypes = [vp, ci] + [C.POINTER(C.c_double)] * 5
        L.herdb_measure_l2_bandwidth.argtypes = [vp, ci, C.POINTER(C.c_double)]
        L.herdb_eval_gate_batch_dev_noalloc.argtypes = [vp, cp, ci, ci, vp, vp, vp, sz]
        L.herdb_task_threads.argtypes = [vp]
        L.herdb_task_window.argtypes = [vp]
        L.herdb_key_blob_size.argtypes = [u32]; L.herdb_key_blob_size.restype = sz
        L.herdb_loaded_keys.argtypes = [vp]
        L.herdb_partition_count.argtypes = [cp, sz, C.POINTER(u64)]
        L.herdb_eval_gate_batch_dev_shape.argtypes = [vp, cp, ci, ci, vp, vp, vp, sz, ci]
        _lib = L
    return _lib


def _ptr(x):
    """numpy array / torch tensor / int -> raw address"""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return x.ctypes.data
    return x.data_ptr()          # torch tensor (host pinned or device)


def analyze(circuit: Circuit):
    info = herdb_circuit_info()
    cc = circuit.to_c()
    rc = lib().herdb_circuit_analyze(C.byref(cc), C.byref(info))
    if rc:
        raise HerdbError(rc)
    return {k: getattr(info, k) for k, _ in herdb_circuit_info._fields_}


def schedule(circuit, rows=0, wave_jobs=0):
    """herdb_circuit_schedule: level widths and grid rounds of the schedule a launch of `rows` rows gets"""
    cc = circuit.to_c()
    info = herdb_schedule_info()
    cap = max(1, analyze(circuit)["levels"])
    widths = np.zeros(cap, np.uint32)
    rc = lib().herdb_circuit_schedule(C.byref(cc), rows, wave_jobs, C.byref(info), _ptr(widths), cap)
    if rc:
        raise HerdbError(rc)
    d = {k: getattr(info, k) for k, _ in herdb_schedule_info._fields_}
    d["level_widths"] = widths[:info.levels].tolist()
    return d


def eval_plain(circuit, in_bits, rows=0, wave_jobs=0):
    """herdb_circuit_eval_plain: the compiled levels (scheduled for `rows`) run on one row of plaintext bits"""
    cc = circuit.to_c()
    in_bits = np.ascontiguousarray(in_bits, np.uint8)
    out = np.zeros(max(1, circuit.out_bits), np.uint8)
    rc = lib().herdb_circuit_eval_plain(C.byref(cc), rows, wave_jobs, _ptr(in_bits), _ptr(out))
    if rc:
        raise HerdbError(rc)
    return out[:circuit.out_bits]


# ------------------------------------------------------------------------------------------------ partition files
def partition_write(path, rows):
    """rows: uint16 array [n_rows, bits, n+1] (or [n_rows, words]); framing [u32 len][payload] per row"""
    rows = np.ascontiguousarray(rows, np.uint16)
    n_rows = rows.shape[0]
    row_bytes = (rows.size // n_rows) * 2 if n_rows else 0
    rc = lib().herdb_partition_write(str(path).encode(), n_rows, row_bytes, _ptr(rows) if rows.size else None)
    if rc:
        raise HerdbError(rc, str(path))


def partition_read(path, n_rows, bits, n):
    out = np.zeros((n_rows, bits, n + 1), np.uint16)
    rc = lib().herdb_partition_read(str(path).encode(), n_rows, bits * (n + 1) * 2, _ptr(out) if out.size else None)
    if rc:
        raise HerdbError(rc, str(path))
    return out


def partition_rows(rows, partitions, k):
    return lib().herdb_partition_rows(rows, partitions, k)


def partition_count(path, bits, n):
    """rows in a partition file of `bits` ciphertexts per row (reads to the end of the file)"""
    cnt = C.c_uint64()
    rc = lib().herdb_partition_count(str(path).encode(), bits * (n + 1) * 2, C.byref(cnt))
    if rc:
        raise HerdbError(rc, str(path))
    return cnt.value


# ------------------------------------------------------------------------------------------------ backend
class Backend:
    """herdb_ctx: owns the devices, streams and device-resident keys."""

    def __init__(self, devices=(0,), storage_dir="", key_dir=""):
        self.L = lib()
        self.h = C.c_void_p()
        ids = (C.c_int * len(devices))(*devices)
        rc = self.L.herdb_create(C.byref(self.h), ids, len(devices), storage_dir.encode(), key_dir.encode())
        if rc:
            raise HerdbError(rc, "herdb_create failed (no CUDA device? this backend has no CPU fallback)")
        self._callbacks = {}
        self._cb_lock = threading.Lock()

    def close(self):
        if self.h:
            self.L.herdb_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc:
            raise HerdbError(rc, self.L.herdb_last_error(self.h).decode())

    # keys
    def load_key(self, session, schema_type=1):
        self._check(self.L.herdb_load_key(self.h, session.encode(), schema_type))

    def load_key_mem(self, session, blob):
        blob = np.ascontiguousarray(blob, np.uint8)
        self._check(self.L.herdb_load_key_mem(self.h, session.encode(), _ptr(blob), blob.size))

    def load_key_dev(self, session, d_blob, nbytes, device=0):
        self._check(self.L.herdb_load_key_dev(self.h, session.encode(), device, _ptr(d_blob), nbytes))

    def unload_key(self, session):
        self._check(self.L.herdb_unload_key(self.h, session.encode()))

    def loaded_keys(self):
        return self.L.herdb_loaded_keys(self.h)

    def task_threads(self):
        return self.L.herdb_task_threads(self.h)

    def task_window(self):
        return self.L.herdb_task_window(self.h)

    def key_lwe_n(self, session):
        n = C.c_uint32()
        self._check(self.L.herdb_key_lwe_n(self.h, session.encode(), C.byref(n)))
        return n.value

    # file-less evaluation
    def eval_gate_batch(self, session, gate, a, b, out=None, device=0):
        """host buffers (numpy or pinned torch): the end-to-end path, H2D + evaluate + D2H"""
        g = GATE[gate] if isinstance(gate, str) else int(gate)
        if out is None:
            out = np.empty_like(a)
        batch = a.shape[0] if a.ndim > 1 else 1
        self._check(self.L.herdb_eval_gate_batch(self.h, session.encode(), device, g, _ptr(a), _ptr(b), _ptr(out), batch))
        return out

    def eval_gate_batch_dev(self, session, gate, d_a, d_b, d_out, batch, device=0, shape=None):
        """shape: ciphertexts per CTA forced to 4, 2 or 1 (None / 0 = the launch policy)"""
        g = GATE[gate] if isinstance(gate, str) else int(gate)
        if shape:
            self._check(self.L.herdb_eval_gate_batch_dev_shape(self.h, session.encode(), device, g, _ptr(d_a), _ptr(d_b), _ptr(d_out), batch, shape))
        else:
            self._check(self.L.herdb_eval_gate_batch_dev(self.h, session.encode(), device, g, _ptr(d_a), _ptr(d_b), _ptr(d_out), batch))

    def eval_gate_batch_dev_noalloc(self, session, gate, d_a, d_b, d_out, batch, device=0):
        """herdb_eval_gate_batch_dev_noalloc: no allocation on the call path, batch <= NOALLOC_MAX_BATCH"""
        g = GATE[gate] if isinstance(gate, str) else int(gate)
        self._check(self.L.herdb_eval_gate_batch_dev_noalloc(self.h, session.encode(), device, g, _ptr(d_a), _ptr(d_b), _ptr(d_out), batch))

    def eval_circuit_rows(self, session, circuit: Circuit, rows_in, out=None):
        n = self.key_lwe_n(session)
        n_rows = rows_in.shape[0]
        if isinstance(rows_in, np.ndarray):
            rows_in = np.ascontiguousarray(rows_in)      # column slices of a wider frame are views
        if out is None:
            out = np.empty((n_rows, circuit.out_bits, n + 1), np.uint16)
        cc = circuit.to_c()
        self._check(self.L.herdb_eval_circuit_rows(self.h, session.encode(), C.byref(cc), _ptr(rows_in), _ptr(out), n_rows))
        return out

    def eval_pipeline_rows(self, session, circuits, rows_in):
        """chained map stages, intermediates resident in HBM (herdb_eval_pipeline_rows)"""
        n = self.key_lwe_n(session)
        rows_in = np.ascontiguousarray(rows_in)
        out = np.empty((rows_in.shape[0], circuits[-1].out_bits, n + 1), np.uint16)
        ccs = [c.to_c() for c in circuits]
        arr = (C.POINTER(herdb_circuit) * len(ccs))(*[C.pointer(c) for c in ccs])
        self._check(self.L.herdb_eval_pipeline_rows(self.h, session.encode(), arr, len(ccs), _ptr(rows_in), _ptr(out), rows_in.shape[0]))
        return out

    def reduce_rows(self, session, circuit: Circuit, rows_in):
        """fold n rows (output-row layout) to one row with a binary reducer circuit (herdb_reduce_rows)"""
        n = self.key_lwe_n(session)
        rows_in = np.ascontiguousarray(rows_in)
        out = np.empty((1, circuit.out_bits, n + 1), np.uint16)
        cc = circuit.to_c()
        self._check(self.L.herdb_reduce_rows(self.h, session.encode(), C.byref(cc), _ptr(rows_in), rows_in.shape[0], _ptr(out)))
        return out

    def eval_circuit_rows_dev(self, session, circuit: Circuit, d_rows_in, d_rows_out, n_rows, device=0):
        cc = circuit.to_c()
        self._check(self.L.herdb_eval_circuit_rows_dev(self.h, session.encode(), device, C.byref(cc), _ptr(d_rows_in), _ptr(d_rows_out), n_rows))

    # measurement
    def stats(self, device=0):
        s = herdb_stats()
        self._check(self.L.herdb_get_stats(self.h, device, C.byref(s)))
        return {k: getattr(s, k) for k, _ in herdb_stats._fields_}

    def stream(self, device=0):
        """raw cudaStream_t of the device's launch stream (wrap with torch.cuda.ExternalStream)"""
        p = C.c_void_p()
        self._check(self.L.herdb_get_stream(self.h, device, C.byref(p)))
        return p.value

    def reset_stats(self):
        self._check(self.L.herdb_reset_stats(self.h))

    def measure_int_peak(self, device=0):
        v = [C.c_double() for _ in range(5)]
        self._check(self.L.herdb_measure_int_peak(self.h, device, *[C.byref(x) for x in v]))
        return dict(zip(["imad_tops", "iadd3_tops", "imadhi_tops", "imadwide_tops", "shoup_mix_mul_tops"], [x.value for x in v]))

    def measure_l2_bandwidth(self, device=0):
        v = C.c_double()
        self._check(self.L.herdb_measure_l2_bandwidth(self.h, device, C.byref(v)))
        return v.value

    def wait_idle(self):
        self._check(self.L.herdb_wait_idle(self.h))


# ------------------------------------------------------------------------------------------------ the worker seam
@dataclasses.dataclass
class DataFramePtr:            # common/include/herd/common/model/worker/data_frame_ptr.hpp:9-13
    uuid: str
    partition: int


@dataclasses.dataclass
class InputDataFramePtr:       # data_frame_ptr.hpp:15-19
    pointer: DataFramePtr
    row_count: int


@dataclasses.dataclass
class CryptoKeyPtr:            # crypto_key_ptr.hpp:10-13 (SchemaType::BINFHE == 1, schema_type.hpp:9-13)
    schema_type: int = 1


@dataclasses.dataclass
class MapTask:                 # common/include/herd/common/model/worker/task.hpp:12-21
    session_uuid: str
    input_data_frame_ptr: InputDataFramePtr
    output_data_frame_ptr: DataFramePtr
    crypto_key_ptr: CryptoKeyPtr
    circuit: Circuit


@dataclasses.dataclass
class ReduceTask:              # task.hpp:23-32
    session_uuid: str
    input_data_frame_ptrs: list
    output_data_frame_ptr: DataFramePtr
    crypto_key_ptr: CryptoKeyPtr
    circuit: Circuit


class TaskHandle:
    """IWorkerGroup::TaskHandle (i_worker_group.hpp:15-42)"""

    class Status(enum.IntEnum):
        COMPLETED = 0
        PENDING = 1
        TIME_OUT = 2
        ERROR = 3

    def __init__(self):
        self._status = TaskHandle.Status.PENDING
        self._callback = None
        self._mutex = threading.Lock()
        self._completed = False
        self._event = threading.Event()

    def completed(self):
        return self._completed

    def status(self):
        return self._status

    def set_completion_callback(self, callback):
        with self._mutex:
            self._callback = callback
            fire = self._completed          # the reference loses a completion that races the install
        if fire:                            # (SURVEY.md 3.3); re-firing here is the safe superset
            callback(self._status)

    def _mark_completed(self, status):
        with self._mutex:
            self._status = TaskHandle.Status(status)
            self._completed = True
            cb = self._callback
        self._event.set()
        if cb:
            cb(self._status)

    def wait(self, timeout=None):
        self._event.wait(timeout)
        return self._status


class CudaWorkerGroup:
    """IWorkerGroup (i_worker_group.hpp:44-48) backed by the in-process CUDA backend."""

    def __init__(self, storage_dir, key_dir, devices=(0,)):
        self.backend = Backend(devices, storage_dir, key_dir)
        self._concurrent = self.backend.task_window()       # tasks to keep outstanding: task threads + a queue to coalesce from
        self._handles = {}
        self._next = 1
        self._lock = threading.Lock()

        def _done(user, status):
            with self._lock:
                h = self._handles.pop(user, None)
            if h is not None:
                h._mark_completed(status)
        self._done = DONE_FN(_done)

    def concurrent_workers(self):
        return self._concurrent

    def schedule_task(self, task) -> TaskHandle:
        assert task is not None, "monostate task"           # grpc_worker_group.cpp:74
        h = TaskHandle()
        with self._lock:
            tag = self._next; self._next += 1
            self._handles[tag] = h
        cc = task.circuit.to_c()
        B = self.backend
        if isinstance(task, ReduceTask):
            frames = (herdb_input_frame * max(1, len(task.input_data_frame_ptrs)))()
            for k, f in enumerate(task.input_data_frame_ptrs):
                frames[k] = herdb_input_frame(f.pointer.uuid.encode(), f.pointer.partition, f.row_count)
            rc = B.L.herdb_reduce_async(B.h, task.session_uuid.encode(), frames, len(task.input_data_frame_ptrs),
                                        task.output_data_frame_ptr.uuid.encode(), task.output_data_frame_ptr.partition,
                                        int(task.crypto_key_ptr.schema_type), C.byref(cc), self._done, tag)
            if rc:
                with self._lock:
                    self._handles.pop(tag, None)
                h._mark_completed(TaskHandle.Status.ERROR)
            return h
        rc = B.L.herdb_map_async(B.h, task.session_uuid.encode(),
                                 task.input_data_frame_ptr.pointer.uuid.encode(), task.input_data_frame_ptr.pointer.partition,
                                 task.input_data_frame_ptr.row_count,
                                 task.output_data_frame_ptr.uuid.encode(), task.output_data_frame_ptr.partition,
                                 int(task.crypto_key_ptr.schema_type), C.byref(cc), self._done, tag)
        if rc:
            with self._lock:
                self._handles.pop(tag, None)
            h._mark_completed(TaskHandle.Status.ERROR)
        return h

    def close(self):
        self.backend.close()
