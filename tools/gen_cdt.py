#!/usr/bin/env python3
"""Generate the 64-bit cumulative table for the discrete Gaussian (sigma = 3.19)
used by the synthetic key generator / encryptor (client stand-in) and by the
oracle.  Output: the HERDB_DGG_CDT initialiser pasted into include/herdb_params.h.

x in [-K, K], K = 40 (> 12 sigma); CDT[i] = floor(2^64 * P(X <= -K + i)) for
i < 2K, last entry saturates to 2^64-1.  Sampling: draw u uniform in [0, 2^64);
x = -K + (number of table entries <= u).
"""
from decimal import Decimal, getcontext
getcontext().prec = 60
SIGMA = Decimal("3.19")
K = 40
w = [(-(Decimal(x) ** 2) / (2 * SIGMA * SIGMA)).exp() for x in range(-K, K + 1)]
tot = sum(w)
acc = Decimal(0)
out = []
for i, wi in enumerate(w):
    acc += wi
    v = int((acc / tot) * (1 << 64))
    out.append(min(v, (1 << 64) - 1))
out[-1] = (1 << 64) - 1
print("#define HERDB_DGG_K %d" % K)
print("#define HERDB_DGG_CDT_INIT { \\")
for i in range(0, len(out), 3):
    print("  " + ", ".join("0x%016xULL" % v for v in out[i:i + 3]) + (", \\" if i + 3 < len(out) else " \\"))
print("}")
