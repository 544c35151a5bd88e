"""Small-job CTA shapes for ncu: STD128 key, one full grid round at C = 2 (296 gates) and at C = 1 (148 gates), twice each."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from herdsman_b200 import api  # noqa: E402
import clientsim as cs  # noqa: E402

N = 512
be = api.Backend()
sk, skr, blob = cs.keygen(N)
be.load_key_mem("s", blob)
for shape, batch in ((2, 296), (1, 148), (2, 296), (1, 148)):
    bits = np.random.default_rng(batch).integers(0, 2, batch).astype(np.uint8)
    ca = cs.encrypt_bits(N, sk, bits, first=0); cb = cs.encrypt_bits(N, sk, bits, first=1 << 20)
    d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda(); d_out = torch.zeros_like(d_a)
    be.reset_stats()
    be.eval_gate_batch_dev("s", "NAND", d_a, d_b, d_out, batch, shape=shape)
    st = be.stats()
    assert (cs.decrypt_bits(N, sk, d_out.cpu().numpy()) == 1 - bits).all()
    print("C=%d batch %d: blind_rotate %.3f ms, keyswitch %.3f ms" % (shape, batch, st["blind_rotate_ms"], st["keyswitch_ms"]))
print("ok")
