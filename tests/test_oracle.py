"""CPU tests of the oracle (oracle/oracle.c): algebraic self-checks, decrypt == plaintext gate, and
the committed golden vectors.  The oracle restates OpenFHE binfhe v1.0.x (not in the reference tree;
PARITY UNPINNED -- see oracle/oracle.h), so these are what stands behind it."""
import hashlib
import os

import numpy as np
import pytest

from oracle import pyoracle as po

GOLD = os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz")
PLAIN = {"OR": lambda a, b: a | b, "AND": lambda a, b: a & b, "NOR": lambda a, b: 1 - (a | b),
         "NAND": lambda a, b: 1 - (a & b), "XOR_FAST": lambda a, b: a ^ b, "XNOR_FAST": lambda a, b: 1 - (a ^ b)}


@pytest.fixture(scope="module")
def orc32(test_key):
    return po.Oracle(test_key["blob"])


def test_parameters():
    Q = po.Q
    assert Q == 2**27 - 2**11 + 1 and Q % 2048 == 1
    assert all(Q % p for p in range(2, 11600))          # prime (sqrt(Q) < 11586)
    o = po.Oracle()
    psi = po.lib().orc_psi(o.h)
    assert psi == 282116                                  # minimum primitive 2N-th root
    assert pow(psi, 1024, Q) == Q - 1
    # no smaller primitive 2N-th root exists: all primitive roots are the odd powers of any one of them
    g = next(w for w in (pow(c, (Q - 1) // 2048, Q) for c in range(2, 100)) if pow(w, 1024, Q) == Q - 1)
    g2, x, best = g * g % Q, g, g
    for _ in range(1024):
        best = min(best, x); x = x * g2 % Q
    assert best == psi


def test_ntt_roundtrip_and_negacyclic_product():
    o = po.Oracle()
    rng = np.random.default_rng(0)
    for _ in range(3):
        p = rng.integers(0, po.Q, 1024, dtype=np.uint32); q = rng.integers(0, po.Q, 1024, dtype=np.uint32)
        assert np.array_equal(o.intt(o.ntt(p)), p)
        ref = np.zeros(1024, np.uint32)
        po.lib().orc_negacyclic_mul_schoolbook(p.ctypes.data, q.ctypes.data, ref.ctypes.data)
        prod = (o.ntt(p).astype(np.uint64) * o.ntt(q).astype(np.uint64) % po.Q).astype(np.uint32)
        assert np.array_equal(o.intt(prod), ref)
    # X * X^(N-1) = -1
    x = np.zeros(1024, np.uint32); x[1] = 1
    y = np.zeros(1024, np.uint32); y[1023] = 1
    prod = (o.ntt(x).astype(np.uint64) * o.ntt(y).astype(np.uint64) % po.Q).astype(np.uint32)
    want = np.zeros(1024, np.uint32); want[0] = po.Q - 1
    assert np.array_equal(o.intt(prod), want)


def test_modswitch_integer_form_equals_openfhe_double_form_exhaustively():
    L = po.lib()
    assert L.orc_round_mismatches(po.QKS, po.Q, 0, po.Q) == 0          # Q -> qKS, every v in [0, Q)
    assert L.orc_round_mismatches(po.LWE_Q, po.QKS, 0, po.QKS) == 0    # qKS -> q, every v


def test_signed_digit_decompose():
    import ctypes as C
    rng = np.random.default_rng(1)
    vals = list(rng.integers(0, po.Q, 2000)) + [0, 1, 63, 64, 65, po.Q // 2 - 1, po.Q // 2, po.Q // 2 + 1, po.Q - 1, po.Q - 64, po.Q - 65]
    d = (C.c_int32 * 4)()
    for t in vals:
        po.lib().orc_signed_digit_decompose(int(t), d)
        digs = list(d)
        assert all(-64 <= x <= 63 for x in digs)
        centred = int(t) if int(t) < po.Q // 2 else int(t) - po.Q
        assert sum(x * 128**l for l, x in enumerate(digs)) == centred      # 4 digits of 7 bits cover |t| < 2^26


@pytest.mark.parametrize("gate", list(PLAIN))
def test_truth_tables_decrypt(orc32, test_key, gate):
    n, sk = test_key["n"], test_key["sk"]
    a = np.array([0, 0, 1, 1] * 4, np.uint8); b = np.array([0, 1, 0, 1] * 4, np.uint8)
    ca = po.encrypt_bits(n, sk, a, seed=3, first=0); cb = po.encrypt_bits(n, sk, b, seed=3, first=100)
    out = orc32.gate(gate, ca, cb)
    assert np.array_equal(po.decrypt_bits(n, sk, out), PLAIN[gate](a, b))
    # outputs are fresh enough to be chained
    out2 = orc32.gate("AND", out, ca)
    assert np.array_equal(po.decrypt_bits(n, sk, out2), PLAIN[gate](a, b) & a)


def test_not_and_constants(orc32, test_key):
    n, sk = test_key["n"], test_key["sk"]
    a = np.array([0, 1, 1, 0], np.uint8)
    ca = po.encrypt_bits(n, sk, a, seed=4)
    assert np.array_equal(po.decrypt_bits(n, sk, orc32.not_(ca)), 1 - a)
    assert np.array_equal(orc32.not_(orc32.not_(ca)), ca)
    for v in (0, 1):
        ct = np.zeros(n + 1, np.uint16)
        po.lib().orc_eval_const(orc32.h, v, ct.ctypes.data)
        assert po.decrypt_bits(n, sk, ct)[0] == v and not ct[:n].any()


def test_std128_gate_decrypts_and_noise_is_small(std128_key):
    n, sk = std128_key["n"], std128_key["sk"]
    o = po.Oracle(std128_key["blob"])
    a = np.array([0, 0, 1, 1, 1, 0, 1, 0], np.uint8); b = np.array([0, 1, 0, 1, 1, 1, 0, 0], np.uint8)
    ca = po.encrypt_bits(n, sk, a, seed=5, first=0); cb = po.encrypt_bits(n, sk, b, seed=5, first=100)
    out = o.gate("NAND", ca, cb)
    dec = po.decrypt_bits(n, sk, out)
    assert np.array_equal(dec, 1 - (a & b))
    noise = [abs(po.lib().orc_noise(n, sk.ctypes.data, out[i].ctypes.data, int(dec[i]))) for i in range(8)]
    assert max(noise) < 64          # decryption threshold is q/8 = 128


def test_golden_vectors(test_key):
    g = np.load(GOLD)
    sk, skr, blob = po.keygen(32, 0x48455244)
    assert hashlib.sha256(blob.tobytes()).digest() == g["key_sha256_n32"].tobytes()
    assert np.array_equal(sk, g["sk_n32"])
    o = po.Oracle(blob)
    ca, cb = g["ct_a"], g["ct_b"]
    assert np.array_equal(po.encrypt_bits(32, sk, g["bits_a"], seed=11, first=0), ca)
    for name in po.GATE:
        assert np.array_equal(o.gate(name, ca, cb), g["gate_" + name]), name
    assert np.array_equal(o.not_(ca), g["not_a"])
    assert np.array_equal(o.blind_rotate_trace("NAND", ca[3], cb[3], 1), g["trace_nand_step1"])
    assert np.array_equal(o.blind_rotate_trace("NAND", ca[3], cb[3], 32), g["trace_nand_full"])
    assert np.array_equal(o.bootstrap_extract("NAND", ca[3], cb[3]), g["extract_nand"])


def test_golden_std128(std128_key):
    g = np.load(GOLD)
    assert hashlib.sha256(std128_key["blob"].tobytes()).digest() == g["key_sha256_n512"].tobytes()
    o = po.Oracle(std128_key["blob"])
    assert np.array_equal(o.gate("NAND", g["std128_ct_a"], g["std128_ct_b"]), g["std128_nand"])
    assert np.array_equal(o.gate("AND", g["std128_ct_a"], g["std128_ct_b"]), g["std128_and"])


def test_bad_key_blob_rejected(test_key):
    o = po.Oracle()
    bad = test_key["blob"][:4096].copy()
    with pytest.raises(ValueError):
        o.load_key(bad)
    bad2 = test_key["blob"].copy(); bad2[0] ^= 1
    with pytest.raises(ValueError):
        o.load_key(bad2)
