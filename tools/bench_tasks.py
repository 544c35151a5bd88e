#!/usr/bin/env python3
"""tools/bench_tasks.py -- rows/s through the task seam ("including file I/O and excluding it").

A data frame of --rows rows in --partitions partition files is evaluated as herdsman would drive it: one MapTask per
partition through CudaWorkerGroup.schedule_task (the IWorkerGroup mirror over herdb_map_async), the executor's window
of concurrent_workers() tasks in flight, every stage reading the previous stage's partition files and writing its own.
The same rows then go through the host-buffer entry point (no files).  One JSON line:

  seam.rows_per_s       wall clock from the first schedule_task to the last completion (file reads and writes inside)
  host.rows_per_s       herdb_eval_circuit_rows / herdb_eval_pipeline_rows on the same rows held in (pageable) host memory
  seam_over_host        the ratio the task seam is held to (>= 0.97 on one GPU)

  --config lt32 | pipeline     --devices N (one process driving N GPUs: rows of every task cut into per-device slices)
"""
import argparse
import hashlib
import json
import os
import shutil
import sys
import tempfile
import time
import uuid

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from herdsman_b200 import api, circuit as cm  # noqa: E402
import clientsim as cs  # noqa: E402

N = 512


_OUT = None


def emit(line):
    os.write(_OUT, (json.dumps(line) + "\n").encode())


def main():
    global _OUT
    sys.stdout.flush()
    _OUT = os.dup(1); os.dup2(2, 1)                     # NCCL prints its version banner on fd 1: keep stdout for the JSON lines
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="pipeline", choices=["lt32", "pipeline", "add8"])
    ap.add_argument("--rows", type=int, default=16384)
    ap.add_argument("--partitions", type=int, default=4)
    ap.add_argument("--devices", type=int, default=1)
    ap.add_argument("--dir", default=None, help="where storage/ and keys/ go (default: a temp dir; use /dev/shm to take the disk out)")
    ap.add_argument("--big", action="store_true",
                    help="full-size runs (config 3 at 1M rows): rows are generated and written partition by partition in pieces, never held "
                         "in memory as a whole; one timed pass through the seam, no warm-up pass, no host-buffer pass; outputs are checked by "
                         "decryption partition by partition")
    args = ap.parse_args()
    root = tempfile.mkdtemp(prefix="herdb_tasks_", dir=args.dir)
    try:
        run(args, root)
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_big(args, root):
    """lt32 (config 3) at full size through the task seam"""
    sk, skr, blob = cs.keygen(N)
    session = str(uuid.uuid4())
    key_dir, storage = os.path.join(root, "keys"), os.path.join(root, "storage")
    os.makedirs(os.path.join(key_dir, session)); os.makedirs(os.path.join(storage, session))
    blob.tofile(os.path.join(key_dir, session, "1.key"))
    R, P = args.rows, args.partitions
    c = cm.compare_circuit(32)
    in_bytes = R * (4 + 64 * (N + 1) * 2)
    free = shutil.disk_usage(root).free
    if free < in_bytes * 1.1:
        raise SystemExit("bench_tasks --big: %d GB of input rows do not fit the %d GB free under %s" % (in_bytes >> 30, free >> 30, root))
    rng = np.random.default_rng(3)
    x = rng.integers(0, 1 << 32, R, dtype=np.uint64); y = rng.integers(0, 1 << 32, R, dtype=np.uint64)
    in_df, out_df = str(uuid.uuid4()), str(uuid.uuid4())
    os.makedirs(os.path.join(storage, session, in_df)); os.makedirs(os.path.join(storage, session, out_df))
    t0 = time.perf_counter()
    start, piece = 0, 8192
    row_words = 64 * (N + 1)
    for k in range(P):
        cnt = api.partition_rows(R, P, k)
        with open(os.path.join(storage, session, in_df, str(k)), "wb") as f:
            for s0 in range(start, start + cnt, piece):
                s1 = min(s0 + piece, start + cnt)
                mat = np.concatenate([cm.pack_words(x[s0:s1], 32), cm.pack_words(y[s0:s1], 32)], axis=1)
                framed = np.empty((s1 - s0, 2 + row_words), np.uint16)
                framed[:, 0] = (row_words * 2) & 0xFFFF; framed[:, 1] = (row_words * 2) >> 16       # [u32 length] little endian
                cts = cs.encrypt_bits(N, sk, mat, seed=3, first=s0 * 64)
                framed[:, 2:] = cts.reshape(s1 - s0, row_words)
                framed.tofile(f)
        start += cnt
    t_gen = time.perf_counter() - t0
    wg = api.CudaWorkerGroup(storage, key_dir, devices=tuple(range(args.devices)))
    be = wg.backend
    t0 = time.perf_counter()
    be.load_key(session)
    t_key = time.perf_counter() - t0
    key_stats = be.stats(0)
    window = wg.concurrent_workers()
    be.reset_stats()
    t0 = time.perf_counter()
    pending, k = [], 0
    while k < P or pending:
        while k < P and len(pending) < window:
            t = api.MapTask(session, api.InputDataFramePtr(api.DataFramePtr(in_df, k), api.partition_rows(R, P, k)),
                            api.DataFramePtr(out_df, k), api.CryptoKeyPtr(1), c)
            pending.append(wg.schedule_task(t)); k += 1
        if pending.pop(0).wait(7200) != api.TaskHandle.Status.COMPLETED:
            raise SystemExit("task failed: " + be.L.herdb_last_error(be.h).decode())
    t_seam = time.perf_counter() - t0
    st = [be.stats(d) for d in range(args.devices)]
    ok, start, h = True, 0, hashlib.sha256()
    for k in range(P):
        cnt = api.partition_rows(R, P, k)
        path = os.path.join(storage, session, out_df, str(k))
        got = api.partition_read(path, cnt, 1, N)
        h.update(got.tobytes())
        dec = cs.decrypt_bits(N, sk, got).reshape(cnt)
        ok = ok and bool(np.array_equal(dec, (x[start:start + cnt] < y[start:start + cnt]).astype(dec.dtype)))
        start += cnt
    line = {"config": "configs[2]: 32-bit compare at full size through the task seam", "rows": R, "partitions": P, "devices": args.devices,
            "gates_per_row": 187, "task_threads": window, "dir": root, "input_bytes": in_bytes, "generate_and_write_s": t_gen,
            "key_load": {"seconds": t_key, "key_load_ms": key_stats["key_load_ms"], "key_bcast_ms": key_stats["key_bcast_ms"],
                         "note": "324 MiB per device: host -> first device, one ncclBroadcast to the others, conversion on every device at once"},
            "seam": {"seconds": t_seam, "rows_per_s": R / t_seam, "gates_per_s": R * 187 / t_seam, "gates_per_s_per_device": R * 187 / t_seam / args.devices,
                     "file_bytes_read": in_bytes, "file_bytes_written": R * (4 + (N + 1) * 2),
                     "gates_per_device": [s["gates_bootstrapped"] for s in st]},
            "all_rows_decrypt_to_plaintext": ok, "output_sha256": h.hexdigest()}
    emit(line)
    wg.close()
    if not ok:
        raise SystemExit("bench_tasks: wrong rows")


def run(args, root):
    if args.big:
        return run_big(args, root)
    sk, skr, blob = cs.keygen(N)
    session = str(uuid.uuid4())
    key_dir, storage = os.path.join(root, "keys"), os.path.join(root, "storage")
    os.makedirs(os.path.join(key_dir, session)); os.makedirs(os.path.join(storage, session))
    blob.tofile(os.path.join(key_dir, session, "1.key"))
    R, P = args.rows, args.partitions
    rng = np.random.default_rng(5)
    if args.config == "lt32":
        x = rng.integers(0, 1 << 32, R, dtype=np.uint64); y = rng.integers(0, 1 << 32, R, dtype=np.uint64)
        stages, cols, bits, plain = [cm.compare_circuit(32)], (x, y), (32, 32), (x < y).astype(np.uint64)
    elif args.config == "add8":
        x = rng.integers(0, 256, R, dtype=np.uint64); y = rng.integers(0, 256, R, dtype=np.uint64)
        stages, cols, bits, plain = [cm.adder_circuit(8)], (x, y), (8, 8), (x + y) % 256
    else:
        a = rng.integers(0, 256, R, dtype=np.uint64); b = rng.integers(0, 256, R, dtype=np.uint64); c = rng.integers(0, 256, R, dtype=np.uint64)
        s = (a + b) % 256
        stages, cols, bits, plain = cm.pipeline_stages(8), (a, b, c), (8, 8, 8), np.where(s < c, s, c)
    rows_in = cs.encrypt_bits(N, sk, np.concatenate([cm.pack_words(v, w) for v, w in zip(cols, bits)], axis=1), seed=5)
    gates_per_row = sum(api.analyze(c)["boot_gates"] for c in stages)

    def new_frame():
        df = str(uuid.uuid4()); os.makedirs(os.path.join(storage, session, df)); return df

    in_df = new_frame()
    start = 0
    for k in range(P):
        cnt = api.partition_rows(R, P, k)
        api.partition_write(os.path.join(storage, session, in_df, str(k)), rows_in[start:start + cnt]); start += cnt

    wg = api.CudaWorkerGroup(storage, key_dir, devices=tuple(range(args.devices)))
    be = wg.backend
    be.load_key(session)                                  # key load is per session, not per task: outside both timings
    window = wg.concurrent_workers()

    def drive(df):
        """the executor's loop: keep `window` tasks in flight, a stage starts when the previous one has finished"""
        for c in stages:
            out_df = new_frame()
            pending, k = [], 0
            while k < P or pending:
                while k < P and len(pending) < window:
                    t = api.MapTask(session, api.InputDataFramePtr(api.DataFramePtr(df, k), api.partition_rows(R, P, k)),
                                    api.DataFramePtr(out_df, k), api.CryptoKeyPtr(1), c)
                    pending.append(wg.schedule_task(t)); k += 1
                h = pending.pop(0)
                if h.wait(3600) != api.TaskHandle.Status.COMPLETED:
                    raise SystemExit("task failed: " + be.L.herdb_last_error(be.h).decode())
            df = out_df
        return df

    drive(in_df)                                          # warm-up: allocations, page cache
    be.reset_stats()
    t0 = time.perf_counter()
    out_df = drive(in_df)
    t_seam = time.perf_counter() - t0
    st = [be.stats(d) for d in range(args.devices)]
    got = np.concatenate([api.partition_read(os.path.join(storage, session, out_df, str(k)), api.partition_rows(R, P, k), stages[-1].out_bits, N) for k in range(P)])
    ok = bool(np.array_equal(cm.unpack_words(cs.decrypt_bits(N, sk, got)).reshape(R), plain))

    ev = (lambda: be.eval_circuit_rows(session, stages[0], rows_in)) if len(stages) == 1 else (lambda: be.eval_pipeline_rows(session, stages, rows_in))
    ev()
    t0 = time.perf_counter()
    host = ev()
    t_host = time.perf_counter() - t0
    same = bool(np.array_equal(host, got))
    in_bytes = sum(c.in_bits for c in stages) * (N + 1) * 2 * R; out_bytes = sum(c.out_bits for c in stages) * (N + 1) * 2 * R
    line = {"config": args.config, "rows": R, "partitions": P, "devices": args.devices, "stages": len(stages), "gates_per_row": gates_per_row,
            "task_threads": be.task_threads(), "task_window": window, "dir": root,
            "seam": {"seconds": t_seam, "rows_per_s": R / t_seam, "gates_per_s": R * gates_per_row / t_seam,
                     "file_bytes_read": in_bytes, "file_bytes_written": out_bytes,
                     "kernel_ms_per_device": [s["blind_rotate_ms"] + s["keyswitch_ms"] for s in st]},
            "host": {"seconds": t_host, "rows_per_s": R / t_host, "gates_per_s": R * gates_per_row / t_host,
                     "note": "same rows in pageable host memory, no files; chained stages keep their intermediate rows in HBM"},
            "seam_over_host": t_host / t_seam,
            "decrypt_matches_plaintext": ok, "seam_rows_equal_host_rows": same,
            "output_sha256": hashlib.sha256(got.tobytes()).hexdigest()}
    emit(line)
    wg.close()
    if not (ok and same):
        raise SystemExit("bench_tasks: wrong rows")


if __name__ == "__main__":
    main()
