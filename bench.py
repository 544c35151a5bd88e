#!/usr/bin/env python3
"""bench.py -- bootstrapped binary gates per second; circuit rows per second vs host CPU (BASELINE.json metric).

Headline workload (config.workload): BASELINE.json configs[1], the single-gate bootstrap microbench:
one step = one batch of NAND gates on fresh STD128 (n=512, N=1024, Q=2^27-2^11+1) LWE
ciphertexts per GPU, evaluated by the CUDA backend through the C ABI (include/herdb.h).

  value     whole-job gates/s with the input ciphertexts resident in HBM (herdb_eval_gate_batch_dev)
  e2e       the same through the host-buffer entry point (herdb_eval_gate_batch): H2D of the inputs
            from pinned host memory + evaluation + D2H of the outputs inside the timed region
  roofline  blind_rotate_kernel against the IMAD pipe: `frac` counts the issue slots the kernel EXECUTES;
            what the reference algorithm would need for the same gates is `effective_vs_reference_algorithm`
  circuits  (N = 1) rows/s of configs[0] (8-bit add, 1,024 rows), configs[2] (32-bit compare) and configs[4]
            (add -> compare -> select) through herdb_eval_circuit_rows / herdb_eval_pipeline_rows, with the oracle's
            rows/s on one thread and on all host threads and the rows compared bit-exactly
  strong    configs[2] at a FIXED total row count cut over the N ranks (sharding.row_slice): rows/s of the whole
            job and the sha256 of the gathered output rows, which must be the same at every N
  cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement"

  python bench.py --gpus 1 --steps 10 --warmup 3            (default)
  torchrun --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, rows/gates sharded,
                                                              key blob broadcast once over NCCL)
  python bench.py --impl reference ...                        (CPU arm: the oracle restating OpenFHE
                                                              binfhe on all host threads)
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_LWE = 512
SEED = 0x48455244
# algorithmic work per bootstrapped gate (SURVEY.md 8d / BASELINE.md 3)
MODMUL_PER_GATE = 46_137_344
# IMAD-pipe issue slots per gate of the reference algorithm: per CMUX step 51,200 Shoup
# butterflies x 4 slots (IMAD.HI = 2, IMAD, IMAD) + 32,768 key MACs x 2 (IMAD.WIDE) + 4,096 monomial
# products x 2; reductions of the 64-bit sums are overhead and not counted
IMAD_SLOTS_REFERENCE_PER_GATE = N_LWE * (51_200 * 4 + 32_768 * 2 + 4_096 * 2)
# what the kernel executes after the top-digit elimination: 2 inverse + 6 forward transforms per
# step instead of 2 + 8; the key MACs and monomial products are unchanged
IMAD_SLOTS_EXECUTED_PER_GATE = N_LWE * (40_960 * 4 + 32_768 * 2 + 4_096 * 2)
BSK_BYTES = N_LWE * 2 * 8 * 2 * 1024 * 4
KS_BYTES_PER_GATE = 1024 * 2 * (N_LWE + 1) * 2
CT_BYTES = (N_LWE + 1) * 2
WORKLOAD = "configs[1]: single-gate NAND bootstrap, STD128 (n=512,N=1024,Q=134215681,baseG=128,qKS=2^14,baseKS=128)"

_JSON_FD = None


def guard_stdout():
    """Libraries (NCCL's version banner, torchrun warnings) write to fd 1; the driver wants ONE JSON line there.
    Everything written to stdout from here on goes to stderr, the JSON line goes to the original fd."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def clocks_sampler(stop, out, gpu_index):
    q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    while not stop.is_set():
        try:
            r = subprocess.run(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5)
            f = [x.strip() for x in r.stdout.strip().split(",")]
            if len(f) >= 7:
                out.append(f)
        except Exception:
            pass
        stop.wait(0.2)


def summarize_clocks(samples):
    if not samples:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    sm = sorted(float(s[0]) for s in samples)
    reasons = []
    for i, name in [(3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap")]:
        if any(s[i].lower().startswith("active") for s in samples):
            reasons.append(name)
    return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(samples[0][1]), "reasons": reasons, "samples": len(samples)}


# ------------------------------------------------------------------------------------------------ the CPU arm
def run_reference(args, rank, world):
    """CPU arm: the oracle (port of OpenFHE binfhe gate evaluation) on all host threads."""
    if rank != 0:
        return
    from oracle import pyoracle as po
    cores = os.cpu_count() or 1
    sk, skr, blob = po.keygen(N_LWE, SEED)
    orc = po.Oracle(blob)
    # bounded sample: up to 16 gates per core and step (a gate costs ~40 ms on one core), fewer when many steps are
    # asked for, so that the whole run stays near two minutes; small samples understate the CPU (thread start-up)
    per_step = max(cores, args.ref_gates if args.ref_gates else cores * max(1, min(16, 3000 // max(1, args.steps))))
    rng = np.random.default_rng(2)
    bits_a = rng.integers(0, 2, per_step).astype(np.uint8); bits_b = rng.integers(0, 2, per_step).astype(np.uint8)
    ca = po.encrypt_bits(N_LWE, sk, bits_a, seed=2, first=0); cb = po.encrypt_bits(N_LWE, sk, bits_b, seed=2, first=1 << 24)
    for _ in range(args.warmup):
        orc.gate("NAND", ca[:cores], cb[:cores], threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = orc.gate("NAND", ca, cb, threads=cores)
    dt = time.perf_counter() - t0
    assert np.array_equal(po.decrypt_bits(N_LWE, sk, out), 1 - (bits_a & bits_b))
    v = per_step * args.steps / dt
    t0 = time.perf_counter()
    orc.gate("NAND", ca[:8], cb[:8], threads=1)
    v1 = 8 / (time.perf_counter() - t0)
    line = {
        "impl": "reference", "metric": "bootstrapped binary gates/sec", "value": v, "unit": "gates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 (u64 products)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "gates_per_step": per_step,
                   "note": "reference CPU path = oracle/ restating OpenFHE binfhe v1.0.x; OpenFHE itself is not in the reference tree"},
        "cpu_baseline": {"value": v, "unit": "gates/s", "cores": cores, "kind": "port",
                         "sample": "%d NAND gates per step x %d steps, OpenMP over gates" % (per_step, args.steps),
                         "one_thread": {"value": v1, "unit": "gates/s", "sample": "8 NAND gates, 1 thread (the reference's single CPU worker)"}},
        "e2e": {"value": v, "unit": "gates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------ circuits (rows/s)
def _enc_cols(cs, cm, sk, cols, bits, seed, first_row=0):
    """rows of UINT columns -> ciphertext rows; the generator is counter based, so a slice [first_row, ...) of a
    frame is the same ciphertexts whichever rank encrypts it"""
    mat = np.concatenate([cm.pack_words(c, b) for c, b in zip(cols, bits)], axis=1)
    return cs.encrypt_bits(N_LWE, sk, mat, seed=seed, first=first_row * mat.shape[1])


def _timed_rows(fn, be, reps=1):
    best = None
    for _ in range(reps):
        be.reset_stats()
        t0 = time.perf_counter()
        out = fn()
        dt = time.perf_counter() - t0
        if best is None or dt < best[0]:
            best = (dt, out, be.stats())
    return best


def circuits_block(be, sk, args):
    """configs[0], [2], [4] on one GPU through the host-buffer entry points, next to the oracle on the host cores"""
    from herdsman_b200 import api, circuit as cm
    from oracle import pyoracle as po
    import clientsim as cs
    cores = os.cpu_count() or 1
    _, _, blob = cs.keygen(N_LWE, SEED)
    orc = po.Oracle(blob)
    out = {}

    def run(name, cfg, stages, cols, bits, plain, seed, rows, cmp_rows, cpu_1t_rows):
        import torch
        rows_np = _enc_cols(cs, cm, sk, cols, bits, seed)
        pin = torch.empty(rows_np.shape, dtype=torch.uint16, pin_memory=True)       # page-locked host rows, as the e2e headline uses
        pin.numpy()[:] = rows_np
        rows_in = pin.numpy()
        del rows_np
        gates_per_row = sum(api.analyze(c)["boot_gates"] for c in stages)
        ev = (lambda r=rows_in: be.eval_circuit_rows("bench", stages[0], r)) if len(stages) == 1 else (lambda r=rows_in: be.eval_pipeline_rows("bench", stages, r))
        ev(rows_in[:64])                                           # allocations
        dt, got, st = _timed_rows(ev, be, reps=2 if rows <= 2048 else 1)
        dec = cm.unpack_words(cs.decrypt_bits(N_LWE, sk, got)).reshape(rows)
        if not np.array_equal(dec, plain):
            raise SystemExit("bench: %s outputs do not decrypt to the plaintext result" % name)

        def oracle_rows(r, threads):
            cur = rows_in[:r]
            t0 = time.perf_counter()
            for c in stages:
                cur = orc.circuit_rows(c.to_c(), cur, c.out_bits, threads=threads)
            return cur, time.perf_counter() - t0
        ref, t_all = oracle_rows(cmp_rows, cores)
        exact = int((got[:cmp_rows].reshape(cmp_rows, -1) == ref.reshape(cmp_rows, -1)).all(axis=1).sum())
        if exact != cmp_rows:
            raise SystemExit("bench: %s GPU rows differ from the oracle's" % name)
        line = {"config": cfg, "rows": rows, "gates_per_row": gates_per_row, "seconds": dt, "rows_per_s": rows / dt, "gates_per_s": rows * gates_per_row / dt,
                "kernel_ms": st["blind_rotate_ms"] + st["keyswitch_ms"],
                "entry_point": "herdb_eval_circuit_rows" if len(stages) == 1 else "herdb_eval_pipeline_rows (intermediate rows stay in HBM)",
                "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"],
                "cpu": {"all_threads": {"rows_per_s": cmp_rows / t_all, "threads": cores, "rows": cmp_rows}},
                "parity": "GPU == oracle bit-exact on %d/%d rows; all %d rows decrypt to the plaintext result" % (exact, cmp_rows, rows)}
        if cpu_1t_rows:
            _, t1 = oracle_rows(cpu_1t_rows, 1)
            line["cpu"]["one_thread"] = {"rows_per_s": cpu_1t_rows / t1, "rows": cpu_1t_rows, "note": "the reference's single CPU worker"}
        line["speedup_vs_cpu_all_threads"] = line["rows_per_s"] / line["cpu"]["all_threads"]["rows_per_s"]
        out[name] = line

    rng = np.random.default_rng(1)
    R = 1024
    x = rng.integers(0, 256, R, dtype=np.uint64); y = rng.integers(0, 256, R, dtype=np.uint64)
    run("add8", "configs[0]: 8-bit add, 1,024 rows", [cm.adder_circuit(8)], (x, y), (8, 8), (x + y) % 256, 1, R, 2 * cores, 16 if args.circuits == "full" else 4)
    R = args.circuit_rows
    rng = np.random.default_rng(3)
    x = rng.integers(0, 1 << 32, R, dtype=np.uint64); y = rng.integers(0, 1 << 32, R, dtype=np.uint64)
    run("lt32", "configs[2]: 32-bit compare, %d rows (the config names 1M)" % R, [cm.compare_circuit(32)], (x, y), (32, 32), (x < y).astype(np.uint64), 3, R, cores, 0)
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, R, dtype=np.uint64); b = rng.integers(0, 256, R, dtype=np.uint64); c = rng.integers(0, 256, R, dtype=np.uint64)
    sm = (a + b) % 256
    run("pipeline", "configs[4]: add -> compare -> select, %d rows (the config names 4M)" % R, cm.pipeline_stages(8), (a, b, c), (8, 8, 8), np.where(sm < c, sm, c), 5, R, cores, 0)
    return out


def strong_block(be, sk, args, rank, world, dist, dev, barrier):
    """configs[2] at a fixed total row count: rank r evaluates rows sharding.row_slice(total, world, r); rows/s of the
    whole job (max over ranks), sha256 of the gathered output rows (identical at every N)"""
    import torch
    from herdsman_b200 import circuit as cm
    from herdsman_b200.sharding import row_slice
    import clientsim as cs
    total = args.strong_rows
    r0, r1 = row_slice(total, world, rank)
    rng = np.random.default_rng(3)
    x = rng.integers(0, 1 << 32, total, dtype=np.uint64); y = rng.integers(0, 1 << 32, total, dtype=np.uint64)
    c = cm.compare_circuit(32)
    w = N_LWE + 1
    h_in = torch.empty((r1 - r0, 64, w), dtype=torch.uint16, pin_memory=True)
    h_in.numpy()[:] = _enc_cols(cs, cm, sk, (x[r0:r1], y[r0:r1]), (32, 32), 3, first_row=r0)
    h_out = torch.empty((r1 - r0, 1, w), dtype=torch.uint16, pin_memory=True)
    be.eval_circuit_rows("bench", c, h_in[:64], out=h_out[:64])
    barrier()
    t0 = time.perf_counter()
    be.eval_circuit_rows("bench", c, h_in, out=h_out)
    dt = time.perf_counter() - t0
    barrier()
    dec = cs.decrypt_bits(N_LWE, sk, h_out.numpy()).reshape(r1 - r0)
    if not np.array_equal(dec, (x[r0:r1] < y[r0:r1]).astype(dec.dtype)):
        raise SystemExit("bench: strong-scaling rows do not decrypt to x < y")
    if dist is not None:
        t = torch.tensor([dt], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        if total % world == 0:
            mine = h_out.to(dev)
            gathered = torch.empty((total, 1, w), dtype=torch.uint16, device=dev)
            dist.all_gather_into_tensor(gathered.view(torch.uint8), mine.view(torch.uint8))     # NCCL has no 16-bit integer type
            all_rows = gathered.cpu().numpy()
        else:
            parts = [None] * world
            dist.all_gather_object(parts, h_out.numpy())
            all_rows = np.concatenate(parts)
    else:
        all_rows = h_out.numpy()
    return {"config": "configs[2]: 32-bit compare, %d rows in total cut over %d GPU(s) (the config names 1M)" % (total, world), "scaling": "strong",
            "rows": total, "gates_per_row": 187, "seconds": dt, "rows_per_s": total / dt, "gates_per_s": total * 187 / dt,
            "output_sha256": hashlib.sha256(np.ascontiguousarray(all_rows).tobytes()).hexdigest(),
            "note": "host-buffer entry point per rank (pinned rows, H2D + D2H inside); the hash covers all output rows in row order and must not depend on N"}


# ------------------------------------------------------------------------------------------------ the GPU arm
def run_ours(args, rank, world, local_rank):
    import torch
    from herdsman_b200 import api
    import clientsim as cs
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    be = api.Backend(devices=(local_rank,))          # raises without a CUDA device: no CPU fallback

    # ---- keys: generated once (client stand-in) on rank 0, broadcast over NVLink (NCCL), converted on each GPU
    blob_size = api.lib().herdb_key_blob_size(N_LWE)
    sk = np.zeros(N_LWE, np.int8)
    if rank == 0:
        sk, skr, blob = cs.keygen(N_LWE, SEED)
        d_blob = torch.from_numpy(blob).to(dev)
    else:
        d_blob = torch.empty(blob_size, dtype=torch.uint8, device=dev)
    if world > 1:
        dist.broadcast(d_blob, src=0)
        t_sk = torch.from_numpy(sk.copy()).to(dev); dist.broadcast(t_sk, src=0); sk = t_sk.cpu().numpy()
    torch.cuda.synchronize()
    be.load_key_dev("bench", d_blob, blob_size)
    del d_blob
    torch.cuda.empty_cache()

    # ---- synthetic encrypted inputs (per rank), pinned on the host and resident in HBM
    # whole waves of the persistent kernel: one CTA per SM, 4 ciphertexts per CTA
    wave = 4 * torch.cuda.get_device_properties(dev).multi_processor_count
    B = args.batch if args.batch else -(-32768 // wave) * wave
    w = N_LWE + 1
    rng = np.random.default_rng(100 + rank)
    bits_a = rng.integers(0, 2, B).astype(np.uint8); bits_b = rng.integers(0, 2, B).astype(np.uint8)
    h_a = torch.empty((B, w), dtype=torch.uint16, pin_memory=True); h_b = torch.empty_like(h_a).pin_memory()
    h_out = torch.empty((B, w), dtype=torch.uint16, pin_memory=True)
    cs.encrypt_bits(N_LWE, sk, bits_a, seed=2 + rank, first=0, out=h_a.numpy())
    cs.encrypt_bits(N_LWE, sk, bits_b, seed=2 + rank, first=1 << 24, out=h_b.numpy())
    d_a = h_a.to(dev); d_b = h_b.to(dev); d_out = torch.zeros((B, w), dtype=torch.uint16, device=dev)
    ext = torch.cuda.ExternalStream(be.stream(), device=dev)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(ext)
        for _ in range(steps):
            fn()
        e1.record(ext)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        barrier()
        if dist is not None:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    dev_step = lambda: be.eval_gate_batch_dev("bench", "NAND", d_a, d_b, d_out, B)      # noqa: E731
    e2e_step = lambda: be.eval_gate_batch("bench", "NAND", h_a, h_b, h_out)             # noqa: E731

    peaks = be.measure_int_peak() if rank == 0 else {}
    l2_gbs = be.measure_l2_bandwidth() if rank == 0 else None
    for _ in range(max(3, args.warmup)):
        dev_step()
    stop = threading.Event(); samples = []
    th = threading.Thread(target=clocks_sampler, args=(stop, samples, local_rank), daemon=True); th.start()
    be.reset_stats()
    ms = timed(dev_step, args.steps)
    st = be.stats()
    stop.set(); th.join()
    clocks = summarize_clocks(samples)
    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)

    # correctness of what was timed: every output decrypts to NAND of the inputs
    ok_dev = np.array_equal(cs.decrypt_bits(N_LWE, sk, d_out.cpu().numpy()), 1 - (bits_a & bits_b))
    ok_e2e = np.array_equal(cs.decrypt_bits(N_LWE, sk, h_out.numpy()), 1 - (bits_a & bits_b))
    if not (ok_dev and ok_e2e):
        raise SystemExit("bench: decrypted outputs do not match NAND of the inputs")
    del d_a, d_b, d_out, h_a, h_b, h_out
    torch.cuda.empty_cache()

    strong = None
    if args.strong_rows and not (world == 1 and args.circuits == "off"):
        try:                                            # never lose the headline line to a side block
            strong = strong_block(be, sk, args, rank, world, dist, dev, barrier)
        except Exception as e:                          # noqa: BLE001
            strong = {"error": "%s: %s" % (type(e).__name__, e)}
            if dist is not None:                        # the ranks may no longer be in step: skip the final barrier
                dist = None
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    gates = world * B * args.steps
    value = gates / (ms * 1e-3)
    e2e_value = gates / (ms_e2e * 1e-3)
    br_ms = st["blind_rotate_ms"] / max(1, st["blind_rotate_launches"])
    ks_ms = st["keyswitch_ms"] / max(1, st["keyswitch_launches"])
    jobs_per_launch = st["gates_bootstrapped"] / max(1, st["blind_rotate_launches"])
    imad_peak = peaks.get("imad_tops") or 18.4
    reference_slots = jobs_per_launch * IMAD_SLOTS_REFERENCE_PER_GATE / (br_ms * 1e-3) / 1e12
    executed_slots = jobs_per_launch * IMAD_SLOTS_EXECUTED_PER_GATE / (br_ms * 1e-3) / 1e12
    # DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/), scaled
    # to this launch's job count; null when no capture is committed
    traffic = None
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["blind_rotate_kernel"]
        traffic = cap["dram_bytes"] * jobs_per_launch / cap["jobs"]
    except Exception:
        pass
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]; peak_src = "measured"
    except Exception:
        hbm_peak = 6650.0; peak_src = "fallback"
    br_bytes = jobs_per_launch * (2 * CT_BYTES + 2 * 1025) + BSK_BYTES      # BSK streamed once per launch (R = jobs/launch)
    ks_bytes = jobs_per_launch * (KS_BYTES_PER_GATE + 2 * 1025 + CT_BYTES)
    line = {
        "metric": "bootstrapped binary gates/sec", "value": value, "unit": "gates/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 (u64 products)", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "gates_per_step_per_gpu": B, "parallelism": "gates sharded over %d GPU(s), no data-path collective; key blob NCCL-broadcast once" % world,
                   "l2": "working set per step (BSK 64 MiB + KSK 260 MiB + monomials 4 MiB + %.0f MiB ciphertexts) exceeds the 126 MB L2" % (B * (3 * CT_BYTES + 2064) / 2**20)},
        "e2e": {"value": e2e_value, "unit": "gates/s", "h2d_bytes_per_step": 2 * B * CT_BYTES * world, "d2h_bytes_per_step": B * CT_BYTES * world,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(st["kernel_launches"]) * world,
        "clocks": clocks,
        "roofline": {"bound": "int32-imad", "kernel": "blind_rotate_kernel", "achieved": executed_slots, "peak": imad_peak,
                     "unit": "T IMAD-slots/s", "frac": executed_slots / imad_peak, "peak_source": "measured in this run (herdb_measure_int_peak)",
                     "traffic": traffic,
                     "note": "`achieved` = issue slots the kernel really executes (2 inverse + 6 forward NTTs, 32 key MACs and the monomial products per "
                             "point and step) / kernel time; slots the top-digit elimination never issues are not counted",
                     "effective_vs_reference_algorithm": {"achieved": reference_slots, "frac": reference_slots / imad_peak,
                                                          "note": "the same gates costed as the reference algorithm does them (2 + 8 NTTs per step, SURVEY.md 8d); work, not pipe utilisation"},
                     "modmul_per_s": jobs_per_launch * MODMUL_PER_GATE / (br_ms * 1e-3),
                     "kernel_ms": br_ms, "jobs_per_launch": jobs_per_launch, "share_of_step": st["blind_rotate_ms"] / ms,
                     "int_peaks_tops": peaks,
                     "hbm": {"bound": "hbm", "achieved": br_bytes / (br_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": br_bytes / (br_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": peak_src},
                     "keyswitch_kernel": {"bound": "l2", "achieved": ks_bytes / (ks_ms * 1e-3) / 1e9 if ks_ms else None, "peak": l2_gbs, "unit": "GB/s",
                                          "frac": (ks_bytes / (ks_ms * 1e-3) / 1e9 / l2_gbs) if ks_ms and l2_gbs else None, "kernel_ms": ks_ms,
                                          "peak_source": "measured in this run (herdb_measure_l2_bandwidth: 32 MiB re-read by all SMs)",
                                          "note": "2.1 MB of key-switch rows per gate; jobs sweep i in lock-step, so rows are served by L2, not HBM"}},
        "kernels": {"blind_rotate_ms": br_ms, "keyswitch_ms": ks_ms},
    }
    if strong is not None:
        line["strong"] = strong
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(be, sk, args)
    if world == 1 and args.circuits != "off":
        try:
            line["circuits"] = circuits_block(be, sk, args)
        except Exception as e:                          # noqa: BLE001
            line["circuits"] = {"error": "%s: %s" % (type(e).__name__, e)}
    emit(line)


def cpu_baseline(be, sk, args):
    """oracle (port of OpenFHE binfhe EvalBinGate) on the box's host cores, bounded sample; also the parity check"""
    from oracle import pyoracle as po
    import clientsim as cs
    cores = os.cpu_count() or 1
    _, _, blob = cs.keygen(N_LWE, SEED)
    orc = po.Oracle(blob)
    cnt = cores * args.cpu_gates_per_core
    rng = np.random.default_rng(7)
    a = rng.integers(0, 2, cnt).astype(np.uint8); b = rng.integers(0, 2, cnt).astype(np.uint8)
    ca = cs.encrypt_bits(N_LWE, sk, a, seed=77, first=0); cb = cs.encrypt_bits(N_LWE, sk, b, seed=77, first=1 << 24)
    orc.gate("NAND", ca[:cores], cb[:cores], threads=cores)
    t0 = time.perf_counter()
    ref = orc.gate("NAND", ca, cb, threads=cores)
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    orc.gate("NAND", ca[:48], cb[:48], threads=1)
    dt1 = time.perf_counter() - t0
    got = be.eval_gate_batch("bench", "NAND", ca, cb)
    return {"value": cnt / dt, "unit": "gates/s", "cores": cores, "kind": "port",
            "sample": "%d NAND gates (same keys, STD128), OpenMP over gates, %.1f s" % (cnt, dt),
            "one_thread": {"value": 48 / dt1, "unit": "gates/s", "sample": "48 NAND gates on 1 thread (the reference's single CPU worker), %.1f s" % dt1},
            "parity": "GPU == oracle bit-exact on %d/%d sampled gates" % (int((got == ref).all(axis=1).sum()), cnt)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="gates per step per GPU (default: 32768 rounded up to whole waves of 4 x SM count)")
    ap.add_argument("--ref-gates", type=int, default=0, help="reference arm: gates per step (default up to 16 x cores)")
    ap.add_argument("--cpu-gates-per-core", type=int, default=128)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--circuits", default="on", choices=["on", "full", "off"],
                    help="circuit rows/s block at N = 1 (outside the timed headline loop); full: 16 one-thread oracle rows of the adder instead of 4")
    ap.add_argument("--circuit-rows", type=int, default=16384, help="rows of the compare and pipeline circuits in the circuits block")
    ap.add_argument("--strong-rows", type=int, default=16384, help="total rows of the fixed-size (strong scaling) compare run; 0 = skip")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    guard_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
