"""tools/probe_overhead.py -- wall clock of Backend.eval_circuit_rows against the CUDA-event time of its kernels,
repeated, for a small circuit (where per-call host overhead shows)."""
import os, sys, time, json
import numpy as np
os.environ.setdefault("OMP_WAIT_POLICY", "passive")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from herdsman_b200 import api, circuit as cm  # noqa: E402
import clientsim as cs

N = 512
be = api.Backend(devices=(0,))
sk, skr, blob = cs.keygen(N, 7)
be.load_key_mem("s", blob)
rng = np.random.default_rng(1)
for name, circ, rows in (("add8", cm.adder_circuit(8), 1024), ("add8", cm.adder_circuit(8), 1184), ("lt32", cm.compare_circuit(32), 1024)):
    bits = rng.integers(0, 2, (rows, circ.in_bits), dtype=np.uint8)
    rows_in = cs.encrypt_bits(N, sk, bits, seed=3)
    out = np.empty((rows, circ.out_bits, N + 1), np.uint16)
    be.eval_circuit_rows("s", circ, rows_in[:64])
    for rep in range(4):
        be.reset_stats()
        t0 = time.perf_counter()
        be.eval_circuit_rows("s", circ, rows_in, out=out)
        dt = time.perf_counter() - t0
        st = be.stats()
        print(json.dumps({"circuit": name, "rows": rows, "rep": rep, "wall_ms": round(dt * 1e3, 2),
                          "kernels_ms": round(st["blind_rotate_ms"] + st["keyswitch_ms"], 2), "launches": st["kernel_launches"]}), flush=True)
