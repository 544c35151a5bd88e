"""A second, structurally different restatement of the gate bootstrap, written from SURVEY.md 8a (rows a7-a13)
alone: numpy, coefficient domain, SCHOOLBOOK negacyclic products (no NTT anywhere), OpenFHE's double-precision
rounding formula for the mod switches.  It reads the same flat key blob (whose bootstrapping key is in the
coefficient domain, include/herdb_params.h) and must agree with oracle/oracle.c bit for bit.

The oracle's parity with OpenFHE is unpinned (nothing to pin it against in the reference tree); what this
test adds is that two independent implementations of the published algorithm -- one NTT-based in C, one
schoolbook in numpy -- produce identical ciphertexts, so a convention slip inside either (twiddle order,
digit order, monomial sign, rounding) would show."""
import numpy as np
import pytest

from herdsman_b200 import api
import clientsim as cs
from oracle import pyoracle as po

N, Q, QLWE, QKS, BG, DG, BKS, DKS = 1024, (1 << 27) - (1 << 11) + 1, 1024, 1 << 14, 128, 4, 128, 2
GATE_CONST = {"OR": 5 * QLWE // 8, "AND": 7 * QLWE // 8, "NOR": QLWE // 8, "NAND": 3 * QLWE // 8,
              "XOR_FAST": 5 * QLWE // 8, "XNOR_FAST": QLWE // 8}


def parse_blob(blob):
    hdr = np.frombuffer(blob[:64].tobytes(), np.uint32)
    n = int(hdr[3])
    bsk_words = n * 2 * 2 * DG * 2 * N
    bsk = np.frombuffer(blob[64:64 + 4 * bsk_words].tobytes(), np.uint32).reshape(n, 2, 2 * DG, 2, N).astype(np.int64)
    ksk = np.frombuffer(blob[64 + 4 * bsk_words:].tobytes(), np.uint16)[:N * BKS * DKS * (n + 1)].reshape(N, BKS, DKS, n + 1).astype(np.int64)
    return n, bsk, ksk


def negacyclic(small, big):
    """small: |coefficients| <= 64 (int64); big: residues in [0, Q).  Exact: 64 * 2^27 * 1024 < 2^44."""
    full = np.convolve(small, big)
    res = full[:N].copy()
    res[:N - 1] -= full[N:]
    return res % Q


def times_monomial_minus_one(p, e):
    """(X^e - 1) * p in Z_Q[X]/(X^N + 1), 0 <= e < 2N"""
    e %= 2 * N
    sign = 1
    if e >= N:
        e -= N; sign = -1
    rot = np.concatenate([-p[N - e:], p[:N - e]]) if e else p.copy()
    return (sign * rot - p) % Q


def signed_digits(poly):
    """SignedDigitDecompose (a11): centred value, four signed base-128 digits, each in [-64, 63]"""
    d = np.where(poly < Q // 2, poly, poly - Q).astype(np.int64)
    out = []
    for _ in range(DG):
        r = ((d + BG // 2) % BG) - BG // 2
        out.append(r)
        d = (d - r) // BG
    assert not d.any()
    return out


def round_q(v, to, frm):
    """ModSwitch / RoundqQ (a12) exactly as OpenFHE writes it: floor(0.5 + double(v) * double(to) / double(frm)) mod to"""
    return np.floor(0.5 + v.astype(np.float64) * float(to) / float(frm)).astype(np.int64) % to


def eval_bin_gate(gate, n, bsk, ksk, ct1, ct2):
    ct1 = ct1.astype(np.int64); ct2 = ct2.astype(np.int64)
    prep = (2 * (ct1 - ct2)) % QLWE if gate in ("XOR_FAST", "XNOR_FAST") else (ct1 + ct2) % QLWE       # a7 / a8
    a, b = prep[:n], int(prep[n])
    # a9: accumulator (0, m)
    q1 = GATE_CONST[gate]; q2 = (q1 + QLWE // 2) % QLWE
    m = np.zeros(N, np.int64)
    for j in range(QLWE // 2):
        t = (b - j) % QLWE
        inside = (q1 <= t < q2) if q1 < q2 else not (q2 <= t < q1)
        m[j * (2 * N // QLWE)] = (Q - (Q // 8 + 1)) if inside else (Q // 8 + 1)
    acc = [np.zeros(N, np.int64), m]
    # a10: n CMUX steps, two RGSW keys per step
    for i in range(n):
        idx = ((QLWE - int(a[i])) % QLWE) * (2 * N // QLWE)
        dct = [None] * (2 * DG)
        for j in range(2):
            for l, r in enumerate(signed_digits(acc[j])):
                dct[j + 2 * l] = r                                   # interleaved order (a11)
        upd = [np.zeros(N, np.int64), np.zeros(N, np.int64)]
        for s, e in ((0, idx), (1, (2 * N - idx) % (2 * N))):
            for j in range(2):
                prod = np.zeros(N, np.int64)
                for r in range(2 * DG):
                    prod = (prod + negacyclic(dct[r], bsk[i, s, r, j])) % Q
                upd[j] = (upd[j] + times_monomial_minus_one(prod, e)) % Q
        acc = [(acc[j] + upd[j]) % Q for j in range(2)]
    # a8: extraction, ModSwitch Q -> qKS
    a_ext = np.concatenate([acc[0][:1], (-acc[0][:0:-1]) % Q])
    b_ext = (int(acc[1][0]) + Q // 8 + 1) % Q
    ext = round_q(np.concatenate([a_ext, [b_ext]]), QKS, Q)
    # a13: key switch
    out = np.zeros(n + 1, np.int64); out[n] = ext[N]
    for i in range(N):
        t = int(ext[i])
        for j in range(DKS):
            out -= ksk[i, t % BKS, j]
            t //= BKS
    return round_q(out % QKS, QLWE, QKS).astype(np.uint16)


@pytest.mark.parametrize("gate", ["NAND", "OR", "XOR_FAST"])
def test_schoolbook_restatement_agrees_with_oracle(test_key, gate):
    n, bsk, ksk = parse_blob(test_key["blob"])
    assert n == test_key["n"]
    orc = po.Oracle(test_key["blob"])
    bits_a = np.array([0, 1, 1], np.uint8); bits_b = np.array([1, 1, 0], np.uint8)
    ca = cs.encrypt_bits(n, test_key["sk"], bits_a, seed=91, first=0)
    cb = cs.encrypt_bits(n, test_key["sk"], bits_b, seed=91, first=100)
    ref = orc.gate(gate, ca, cb)
    for k in range(len(bits_a)):
        got = eval_bin_gate(gate, n, bsk, ksk, ca[k], cb[k])
        assert np.array_equal(got, ref[k]), "schoolbook restatement and oracle.c differ"
    plain = {"NAND": 1 - (bits_a & bits_b), "OR": bits_a | bits_b, "XOR_FAST": bits_a ^ bits_b}[gate]
    assert np.array_equal(cs.decrypt_bits(n, test_key["sk"], ref), plain)
