#!/usr/bin/env python3
"""Regenerates tests/golden/golden_v1.npz.

There are NO golden vectors in the reference (test/CMakeLists.txt:20 has an empty source list)
and OpenFHE is not installed, so these vectors are produced by the CPU oracle itself
(oracle/oracle.c): they pin the oracle and the key/ciphertext generator against regressions and
give the GPU tests fixed inputs/outputs; they do NOT pin parity with OpenFHE (parity unpinned).
Keys are regenerated from the seed (21 MB for n=32), only their SHA-256 is stored.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po          # noqa: E402
from herdsman_b200 import circuit as cm    # noqa: E402

SEED = 0x48455244


def main():
    out = {}
    for n in (32, 512):
        sk, skr, blob = po.keygen(n, SEED)
        out["key_sha256_n%d" % n] = np.frombuffer(hashlib.sha256(blob.tobytes()).digest(), np.uint8)
        out["sk_n%d" % n] = sk
    n = 32
    sk, skr, blob = po.keygen(n, SEED)
    orc = po.Oracle(blob)
    a = np.array([0, 0, 1, 1, 1, 0], np.uint8); b = np.array([0, 1, 0, 1, 1, 0], np.uint8)
    ca = po.encrypt_bits(n, sk, a, seed=11, first=0); cb = po.encrypt_bits(n, sk, b, seed=11, first=64)
    out["bits_a"], out["bits_b"], out["ct_a"], out["ct_b"] = a, b, ca, cb
    for g in po.GATE:
        out["gate_" + g] = orc.gate(g, ca, cb)
    out["not_a"] = orc.not_(ca)
    out["trace_nand_step1"] = orc.blind_rotate_trace("NAND", ca[3], cb[3], 1)
    out["trace_nand_full"] = orc.blind_rotate_trace("NAND", ca[3], cb[3], n)
    out["extract_nand"] = orc.bootstrap_extract("NAND", ca[3], cb[3])
    c = cm.adder_circuit(8)
    x = np.array([200, 17]); y = np.array([100, 5])
    bits = np.concatenate([cm.pack_words(x, 8), cm.pack_words(y, 8)], axis=1)
    rows = po.encrypt_bits(n, sk, bits, seed=12).reshape(2, 16, n + 1)
    out["add8_rows_in"] = rows
    out["add8_rows_out"] = orc.circuit_rows(c.to_c(), rows, 8)
    out["add8_plain"] = (x + y) % 256
    # one STD128 gate (the production parameter set)
    n = 512
    sk, skr, blob = po.keygen(n, SEED)
    orc = po.Oracle(blob)
    ca = po.encrypt_bits(n, sk, [1, 0], seed=13, first=0); cb = po.encrypt_bits(n, sk, [1, 1], seed=13, first=8)
    out["std128_ct_a"], out["std128_ct_b"] = ca, cb
    out["std128_nand"] = orc.gate("NAND", ca, cb)
    out["std128_and"] = orc.gate("AND", ca, cb)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
