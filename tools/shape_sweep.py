#!/usr/bin/env python3
"""tools/shape_sweep.py -- blind-rotation CTA shapes on small launches (STD128, device-resident NAND batches):
kernel time (CUDA events inside the library) of a batch at the forced shapes C = 4, 2, 1 and under the launch policy.
One JSON line per batch."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from herdsman_b200 import api  # noqa: E402
import clientsim as cs  # noqa: E402

N = 512
batches = [int(v) for v in sys.argv[1:]] or [1, 2, 4, 16, 64, 128, 148, 256, 296, 400, 512, 592, 640, 740, 888, 1024, 1184, 2048, 4096]
be = api.Backend()
sk, skr, blob = cs.keygen(N)
be.load_key_mem("s", blob)
for B in batches:
    bits = np.random.default_rng(B).integers(0, 2, B).astype(np.uint8)
    ca = cs.encrypt_bits(N, sk, bits, first=0); cb = cs.encrypt_bits(N, sk, bits, first=1 << 20)
    d_a = torch.from_numpy(ca).cuda(); d_b = torch.from_numpy(cb).cuda(); d_out = torch.zeros_like(d_a)
    line = {"batch": B}
    ref = None
    for shape in (4, 2, 1, 0):
        if shape in (2, 1) and B > 1200:
            continue
        best = None
        for rep in range(3):
            be.reset_stats()
            be.eval_gate_batch_dev("s", "NAND", d_a, d_b, d_out, B, shape=shape)
            st = be.stats()
            ms = st["blind_rotate_ms"] + st["keyswitch_ms"]
            best = ms if best is None or ms < best else best
        out = d_out.cpu().numpy()
        if ref is None:
            ref = out
            assert (cs.decrypt_bits(N, sk, out) == 1 - bits).all()
        assert np.array_equal(out, ref), (B, shape)
        line["policy" if shape == 0 else "C%d" % shape] = {"ms": round(best, 4), "gates_per_s": round(B / (best * 1e-3), 1)}
    print(json.dumps(line), flush=True)
