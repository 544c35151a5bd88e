#!/usr/bin/env python3
"""tools/bench_circuits.py -- the circuit configurations of BASELINE.json on the GPU backend (host-buffer
entry point herdb_eval_circuit_rows, i.e. end to end with H2D/D2H), one JSON line per configuration:

  add8      configs[0]  8-bit ripple-carry add, 1,024 rows
  lt32      configs[2]  32-bit compare (x < y), --rows-lt32 rows (default 16,384; the config names 1M)
  mul32     configs[3]  32-bit multiply, --rows-mul32 rows (default 1,024; the config names 256K)
  pipe8     configs[4]  add -> compare -> select as three chained stages, --rows-pipe rows
  sweep     configs[1]  single NAND gate, batch sweep 1 .. --sweep-max

Every run decrypts ALL outputs and compares them with plaintext evaluation; `--oracle-rows R` also compares
the first R rows bit-exactly with the CPU oracle (slow: ~50 ms per gate per core).
With --devices N the in-process multi-device context shards the rows over N GPUs of the box.
"""
import argparse
import json
import os
import sys
import time

os.environ.setdefault("OMP_WAIT_POLICY", "passive")     # oracle threads must not spin beside the CUDA launch path

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from herdsman_b200 import api, circuit as cm  # noqa: E402
import clientsim as cs

N = 512


def enc_words(sk, cols, bits, seed):
    mats = [cm.pack_words(c, b) for c, b in zip(cols, bits)]
    return cs.encrypt_bits(N, sk, np.concatenate(mats, axis=1), seed=seed)


def run_circuit(be, sk, name, circ, cols, bits, plain, seed, oracle_rows, orc):
    rows_in = enc_words(sk, cols, bits, seed)
    info = api.analyze(circ)
    rows = rows_in.shape[0]
    be.eval_circuit_rows("s", circ, rows_in[:min(rows, 64)])                      # warm-up (allocations)
    dt, spent = None, 0.0
    while dt is None or (spent < 3.0 and dt < 1.5):          # short configurations: best of a few runs
        be.reset_stats()
        t0 = time.perf_counter()
        out = be.eval_circuit_rows("s", circ, rows_in)
        d = time.perf_counter() - t0
        spent += d
        if dt is None or d < dt:
            dt, st = d, be.stats()
    dec = cm.unpack_words(cs.decrypt_bits(N, sk, out)).reshape(rows)
    ok = bool(np.array_equal(dec, plain))
    sm = be.device_sm_count() if hasattr(be, "device_sm_count") else 148
    sched = api.schedule(circ, rows, sm * 4)
    info = dict(info, levels=sched["levels"], slots=sched["slots"], max_level_width=sched["max_level_width"],
                grid_rounds_asap=sched["rounds_asap"], grid_rounds=sched["rounds"])
    line = {"config": name, "rows": rows, "circuit": info, "seconds": dt, "rows_per_s": rows / dt,
            "gates_per_s": rows * info["boot_gates"] / dt, "decrypt_matches_plaintext": ok,
            "kernel_ms": {"blind_rotate": st["blind_rotate_ms"], "keyswitch": st["keyswitch_ms"]},
            "launches": st["kernel_launches"]}
    if oracle_rows and orc is not None:
        r = min(oracle_rows, rows)
        t0 = time.perf_counter()
        ref = orc.circuit_rows(circ.to_c(), rows_in[:r], circ.out_bits)
        cdt = time.perf_counter() - t0
        line["oracle"] = {"rows": r, "bit_exact": bool(np.array_equal(out[:r], ref)), "seconds": cdt,
                          "rows_per_s": r / cdt, "cores": os.cpu_count()}
    print(json.dumps(line), flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--devices", type=int, default=1)
    ap.add_argument("--configs", default="add8,lt32,mul32,pipe8,sweep")
    ap.add_argument("--rows-add8", type=int, default=1024)
    ap.add_argument("--rows-lt32", type=int, default=16384)
    ap.add_argument("--rows-mul32", type=int, default=1024)
    ap.add_argument("--rows-pipe", type=int, default=8192)
    ap.add_argument("--sweep-max", type=int, default=1 << 17)
    ap.add_argument("--oracle-rows", type=int, default=0)
    args = ap.parse_args()
    be = api.Backend(devices=tuple(range(args.devices)))
    sk, skr, blob = cs.keygen(N, 0x48455244)
    be.load_key_mem("s", blob)
    orc = None
    if args.oracle_rows:
        from oracle import pyoracle as po
        orc = po.Oracle(blob)
    ok = True
    todo = args.configs.split(",")
    if "add8" in todo:
        rng = np.random.default_rng(1); r = args.rows_add8
        x = rng.integers(0, 256, r, dtype=np.uint64); y = rng.integers(0, 256, r, dtype=np.uint64)
        ok &= run_circuit(be, sk, "configs[0] add8", cm.adder_circuit(8), (x, y), (8, 8), (x + y) % 256, 1, args.oracle_rows, orc)
    if "lt32" in todo:
        rng = np.random.default_rng(3); r = args.rows_lt32
        x = rng.integers(0, 1 << 32, r, dtype=np.uint64); y = rng.integers(0, 1 << 32, r, dtype=np.uint64)
        ok &= run_circuit(be, sk, "configs[2] lt32", cm.compare_circuit(32), (x, y), (32, 32), (x < y).astype(np.uint64), 3,
                          min(args.oracle_rows, 2), orc)
    if "mul32" in todo:
        rng = np.random.default_rng(4); r = args.rows_mul32
        x = rng.integers(0, 1 << 32, r, dtype=np.uint64); y = rng.integers(0, 1 << 32, r, dtype=np.uint64)
        ok &= run_circuit(be, sk, "configs[3] mul32", cm.multiply_circuit(32), (x, y), (32, 32), (x * y) & 0xFFFFFFFF, 4, 0, None)
    if "pipe8" in todo:
        rng = np.random.default_rng(5); r = args.rows_pipe
        a = rng.integers(0, 256, r, dtype=np.uint64); b = rng.integers(0, 256, r, dtype=np.uint64); d = rng.integers(0, 256, r, dtype=np.uint64)
        abd = enc_words(sk, (a, b, d), (8, 8, 8), 5)
        stages = [cm.adder_circuit(8), cm.compare_circuit(8), cm.select_circuit(8)]
        gates = sum(api.analyze(c)["boot_gates"] for c in stages)
        be.reset_stats()
        t0 = time.perf_counter()
        s1 = be.eval_circuit_rows("s", stages[0], abd[:, :16])
        s2 = be.eval_circuit_rows("s", stages[1], np.concatenate([s1, abd[:, 16:]], axis=1))
        s3 = be.eval_circuit_rows("s", stages[2], np.concatenate([s2, s1, abd[:, 16:]], axis=1))
        dt = time.perf_counter() - t0
        sm = (a + b) % 256
        good = bool(np.array_equal(cm.unpack_words(cs.decrypt_bits(N, sk, s3)).reshape(r), np.where(sm < d, sm, d)))
        ok &= good
        print(json.dumps({"config": "configs[4] add->compare->select (3 stages, host round trip between stages)", "rows": r,
                          "boot_gates_per_row": gates, "seconds": dt, "rows_per_s": r / dt, "gates_per_s": r * gates / dt,
                          "decrypt_matches_plaintext": good}), flush=True)
        chain = cm.pipeline_stages(8)                      # same work, intermediate rows stay in HBM
        be.eval_pipeline_rows("s", chain, abd[:64])
        t0 = time.perf_counter()
        out = be.eval_pipeline_rows("s", chain, abd)
        dt = time.perf_counter() - t0
        good = bool(np.array_equal(cm.unpack_words(cs.decrypt_bits(N, sk, out)).reshape(r), np.where(sm < d, sm, d)))
        ok &= good
        print(json.dumps({"config": "configs[4] add->compare->select (3 stages, intermediate rows resident on the device)", "rows": r,
                          "boot_gates_per_row": gates, "seconds": dt, "rows_per_s": r / dt, "gates_per_s": r * gates / dt,
                          "decrypt_matches_plaintext": good, "bit_identical_to_staged": bool(np.array_equal(out, s3))}), flush=True)
    if "sweep" in todo:
        rng = np.random.default_rng(2)
        B = 1
        pts = []
        while B <= args.sweep_max:
            a = rng.integers(0, 2, B).astype(np.uint8); b = rng.integers(0, 2, B).astype(np.uint8)
            ca = cs.encrypt_bits(N, sk, a, seed=2, first=0); cb = cs.encrypt_bits(N, sk, b, seed=2, first=1 << 24)
            be.eval_gate_batch("s", "NAND", ca, cb)
            be.reset_stats()
            t0 = time.perf_counter()
            out = be.eval_gate_batch("s", "NAND", ca, cb)
            dt = time.perf_counter() - t0
            st = be.stats()
            good = bool(np.array_equal(cs.decrypt_bits(N, sk, out), 1 - (a & b)))
            ok &= good
            pts.append({"batch": B, "gates_per_s_wall": B / dt, "gates_per_s_kernels": B / ((st["blind_rotate_ms"] + st["keyswitch_ms"]) * 1e-3),
                        "ms_wall": dt * 1e3, "ok": good})
            B *= 4 if B >= 1024 else 2
        print(json.dumps({"config": "configs[1] NAND batch sweep", "points": pts}), flush=True)
    be.close()
    if not ok:
        raise SystemExit("bench_circuits: a decrypt check failed")


if __name__ == "__main__":
    main()
