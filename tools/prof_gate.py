"""Small single-gate workload for ncu: STD128 key, `batch` NAND gates, `reps` launches."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from herdsman_b200 import api
import clientsim as cs

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 592
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
n = int(sys.argv[3]) if len(sys.argv) > 3 else 512
b = api.Backend()
sk, skr, blob = cs.keygen(n)
b.load_key_mem("s", blob)
bits = np.random.default_rng(0).integers(0, 2, batch).astype(np.uint8)
ca = cs.encrypt_bits(n, sk, bits, first=0); cb = cs.encrypt_bits(n, sk, bits, first=1 << 20)
for _ in range(reps):
    b.reset_stats()
    out = b.eval_gate_batch("s", "NAND", ca, cb)
    st = b.stats()
    print("batch %d: blind_rotate %.3f ms, keyswitch %.3f ms -> %.0f gates/s (kernels)" % (
        batch, st["blind_rotate_ms"], st["keyswitch_ms"], batch / ((st["blind_rotate_ms"] + st["keyswitch_ms"]) * 1e-3)))
assert (cs.decrypt_bits(n, sk, out) == 1 - bits).all()
print("ok")
