"""The blind-rotation CTA body (herdsman_b200/csrc/br_core.cuh) compiled by g++ and run by 256 host
threads (tests/emu), against the oracle: accumulator traces after k CMUX steps and the extracted
ciphertext must be bit-identical.  This is how the kernel's indexing, shared-memory swizzles, lazy
ranges and table layouts are checked in the GPU-less build container."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import pyoracle as po

EMU_DIR = os.path.join(os.path.dirname(__file__), "emu")


@pytest.fixture(scope="module")
def emu():
    subprocess.run(["make", "-C", EMU_DIR], check=True, capture_output=True)
    L = C.CDLL(os.path.join(EMU_DIR, "libemu.so"))
    L.emu_blind_rotate.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
    L.emu_blind_rotate_shape.argtypes = [C.c_int] + L.emu_blind_rotate.argtypes
    return L


def _run(emu, blob, gate, ca, cb, steps, shape=4):
    B = ca.shape[0]
    ext = np.zeros((B, 1025), np.uint16); trace = np.zeros((B, 2, 1024), np.uint32)
    rc = emu.emu_blind_rotate_shape(shape, blob.ctypes.data, blob.size, po.GATE[gate], ca.ctypes.data, cb.ctypes.data, B, steps,
                                    ext.ctypes.data, trace.ctypes.data)
    assert rc == 0
    return ext, trace


@pytest.mark.parametrize("gate,steps", [("NAND", 0), ("NAND", 1), ("NAND", 2), ("NAND", 32), ("AND", 32), ("OR", 32),
                                        ("NOR", 32), ("XOR_FAST", 32), ("XNOR_FAST", 5)])
def test_emulated_kernel_body_matches_oracle(emu, test_key, gate, steps):
    n, sk, blob = test_key["n"], test_key["sk"], test_key["blob"]
    o = po.Oracle(blob)
    rng = np.random.default_rng(5)
    B = 6                                              # two groups of 4, the second one ragged
    a = rng.integers(0, 2, B).astype(np.uint8); b = rng.integers(0, 2, B).astype(np.uint8)
    ca = po.encrypt_bits(n, sk, a, first=0); cb = po.encrypt_bits(n, sk, b, first=100)
    ext, trace = _run(emu, blob, gate, ca, cb, steps)
    for i in range(B):
        assert np.array_equal(o.blind_rotate_trace(gate, ca[i], cb[i], steps).reshape(2, 1024), trace[i]), (gate, steps, i)
        if steps == n:
            assert np.array_equal(o.bootstrap_extract(gate, ca[i], cb[i]), ext[i])


@pytest.mark.parametrize("shape,gate,steps", [(2, "NAND", 0), (2, "NAND", 3), (2, "OR", 32), (2, "XOR_FAST", 32),
                                              (1, "NAND", 1), (1, "AND", 32), (1, "XNOR_FAST", 32)])
def test_emulated_small_job_shapes_match_oracle(emu, test_key, shape, gate, steps):
    """the C = 2 and C = 1 CTA shapes (128 threads, whole key slice staged by bulk copy, two chunks of evaluation
    points per thread in phase B; at C = 1 warps 2 and 3 have no polynomial in phase A)"""
    n, sk, blob = test_key["n"], test_key["sk"], test_key["blob"]
    o = po.Oracle(blob)
    rng = np.random.default_rng(6)
    B = 3                                              # C = 2: a full group and a ragged one
    a = rng.integers(0, 2, B).astype(np.uint8); b = rng.integers(0, 2, B).astype(np.uint8)
    ca = po.encrypt_bits(n, sk, a, first=0); cb = po.encrypt_bits(n, sk, b, first=100)
    ext, trace = _run(emu, blob, gate, ca, cb, steps, shape)
    for i in range(B):
        assert np.array_equal(o.blind_rotate_trace(gate, ca[i], cb[i], steps).reshape(2, 1024), trace[i]), (shape, gate, steps, i)
        if steps == n:
            assert np.array_equal(o.bootstrap_extract(gate, ca[i], cb[i]), ext[i])


def test_emulated_kernel_body_extreme_inputs(emu, test_key):
    """a_i = 0 (monomial X^0 - 1 = 0: the step is a no-op), a_i = q-1, b at the window edges"""
    n, blob = test_key["n"], test_key["blob"]
    o = po.Oracle(blob)
    ca = np.zeros((4, n + 1), np.uint16); cb = np.zeros((4, n + 1), np.uint16)
    ca[1, :n] = 1023; cb[1, :n] = 1023
    ca[2, n] = 383; cb[2, n] = 1
    ca[3, :n] = 512; ca[3, n] = 895
    ext, trace = _run(emu, blob, "NAND", ca, cb, n)
    for i in range(4):
        assert np.array_equal(o.blind_rotate_trace("NAND", ca[i], cb[i], n).reshape(2, 1024), trace[i])
        assert np.array_equal(o.bootstrap_extract("NAND", ca[i], cb[i]), ext[i])


def test_fp64_fold_residue_and_range(emu):
    """fold62_f64 (phase B): for every x < 2^62 the result is congruent to x + addend mod Q and lies in
    addend + (0.2499 Q, 1.7501 Q) -- the bound the accumulator update and the monomial products rely on."""
    Q = po.Q if hasattr(po, "Q") else 134215681
    rng = np.random.default_rng(11)
    xs = [0, 1, Q - 1, Q, Q + 1, (1 << 62) - 1, (1 << 62) - Q, (1 << 61), (1 << 43) - 1, 1 << 43, (1 << 11) - 1, 1 << 11]
    # quotient boundaries: x/Q close to k + 1/2 (where the rounding of the DFMA decides) and to integers
    for k in list(rng.integers(0, (1 << 62) // Q, 2000)) + [0, 1, (1 << 62) // Q - 1]:
        k = int(k)
        for dx in (-2, -1, 0, 1, 2):
            xs += [k * Q + Q // 2 + dx, k * Q + dx]
    xs += [int(v) for v in rng.integers(0, 1 << 62, 200000, dtype=np.uint64)]
    x = np.array([v for v in xs if 0 <= v < (1 << 62)], np.uint64)
    add = rng.integers(0, 2 * Q, x.size).astype(np.uint32); add[::3] = 0
    out = np.zeros(x.size, np.uint32)
    emu.emu_fold62_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    emu.emu_fold62_f64(x.ctypes.data, add.ctypes.data, out.ctypes.data, x.size)
    r = out.astype(np.int64) - add.astype(np.int64)
    assert np.array_equal(r % Q, (x % np.uint64(Q)).astype(np.int64))
    assert r.min() > 0.2499 * Q and r.max() < 1.7501 * Q
