"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY:
importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never from the herdsman_b200 package."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

N = 1024
Q = 134215681
LWE_Q = 1024
QKS = 16384
GATE = {"OR": 0, "AND": 1, "NOR": 2, "NAND": 3, "XOR_FAST": 4, "XNOR_FAST": 5}


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        vp, u64, u32, sz, ci = C.c_void_p, C.c_uint64, C.c_uint32, C.c_size_t, C.c_int
        L.orc_create.restype = vp
        L.orc_destroy.argtypes = [vp]
        L.orc_load_key.argtypes = [vp, vp, sz]
        L.orc_lwe_n.argtypes = [vp]; L.orc_lwe_n.restype = u32
        L.orc_psi.argtypes = [vp]; L.orc_psi.restype = u32
        L.orc_eval_not.argtypes = [vp, vp, vp]
        L.orc_eval_const.argtypes = [vp, ci, vp]
        L.orc_eval_bin_gate.argtypes = [vp, ci, vp, vp, vp]
        L.orc_eval_gate_batch.argtypes = [vp, ci, vp, vp, vp, sz, ci]
        L.orc_eval_circuit_rows.argtypes = [vp, vp, vp, vp, sz, ci]
        L.orc_eval_circuit_plain.argtypes = [vp, vp, vp, sz]
        L.orc_ntt_forward.argtypes = [vp, vp]
        L.orc_ntt_inverse.argtypes = [vp, vp]
        L.orc_negacyclic_mul_schoolbook.argtypes = [vp, vp, vp]
        L.orc_blind_rotate_trace.argtypes = [vp, ci, vp, vp, u32, vp]
        L.orc_bootstrap_extract.argtypes = [vp, ci, vp, vp, vp]
        L.orc_round_qQ_double.argtypes = [u32, u32, u32]; L.orc_round_qQ_double.restype = u32
        L.orc_round_qQ_int.argtypes = [u32, u32, u32]; L.orc_round_qQ_int.restype = u32
        L.orc_round_mismatches.argtypes = [u32, u32, u32, u32]; L.orc_round_mismatches.restype = u64
        L.orc_signed_digit_decompose.argtypes = [u32, vp]
        L.orc_key_blob_size.argtypes = [u32]; L.orc_key_blob_size.restype = sz
        L.orc_keygen.argtypes = [u32, u64, vp, vp, vp, sz]
        L.orc_encrypt_bits.argtypes = [u32, vp, u64, u64, vp, sz, vp]
        L.orc_decrypt_bits.argtypes = [u32, vp, vp, sz, vp]
        L.orc_noise.argtypes = [u32, vp, vp, ci]; L.orc_noise.restype = C.c_int32
        L.orc_rng.argtypes = [u64, u64, u64]; L.orc_rng.restype = u64
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def keygen(n, seed=0x48455244):
    L = lib()
    sk = np.zeros(n, np.int8); skr = np.zeros(N, np.int8)
    blob = np.zeros(L.orc_key_blob_size(n), np.uint8)
    rc = L.orc_keygen(n, seed, _p(sk), _p(skr), _p(blob), blob.size)
    assert rc == 0
    return sk, skr, blob


def encrypt_bits(n, sk, bits, seed=1, first=0):
    bits = np.ascontiguousarray(bits, np.uint8).ravel()
    cts = np.zeros((bits.size, n + 1), np.uint16)
    lib().orc_encrypt_bits(n, _p(sk), seed, first, _p(bits), bits.size, _p(cts))
    return cts


def decrypt_bits(n, sk, cts):
    cts = np.ascontiguousarray(cts, np.uint16).reshape(-1, n + 1)
    bits = np.zeros(cts.shape[0], np.uint8)
    lib().orc_decrypt_bits(n, _p(sk), _p(cts), cts.shape[0], _p(bits))
    return bits


class Oracle:
    """Server-side oracle: loads a key blob and evaluates gates / circuits like the herd worker would."""

    def __init__(self, blob=None):
        self.L = lib()
        self.h = self.L.orc_create()
        self.n = 0
        if blob is not None:
            self.load_key(blob)

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_key(self, blob):
        blob = np.ascontiguousarray(blob, np.uint8)
        rc = self.L.orc_load_key(self.h, _p(blob), blob.size)
        if rc != 0:
            raise ValueError("oracle: bad key blob (%d)" % rc)
        self.n = self.L.orc_lwe_n(self.h)

    def gate(self, gate, a, b, threads=0):
        a = np.ascontiguousarray(a, np.uint16).reshape(-1, self.n + 1)
        b = np.ascontiguousarray(b, np.uint16).reshape(-1, self.n + 1)
        out = np.zeros_like(a)
        g = GATE[gate] if isinstance(gate, str) else int(gate)
        self.L.orc_eval_gate_batch(self.h, g, _p(a), _p(b), _p(out), a.shape[0], threads)
        return out

    def not_(self, a):
        a = np.ascontiguousarray(a, np.uint16).reshape(-1, self.n + 1)
        out = np.zeros_like(a)
        for i in range(a.shape[0]):
            self.L.orc_eval_not(self.h, _p(a[i]), _p(out[i]))
        return out

    def circuit_rows(self, ccirc, rows_in, out_bits, threads=0):
        """ccirc: ctypes herdb_circuit (herdsman_b200.circuit.Circuit.to_c()); rows_in [rows, in_bits, n+1]"""
        rows_in = np.ascontiguousarray(rows_in, np.uint16)
        n_rows = rows_in.shape[0]
        out = np.zeros((n_rows, out_bits, self.n + 1), np.uint16)
        rc = self.L.orc_eval_circuit_rows(self.h, C.byref(ccirc), _p(rows_in), _p(out), n_rows, threads)
        if rc != 0:
            raise ValueError("oracle: malformed circuit")
        return out

    def circuit_plain(self, ccirc, bits_in, out_bits):
        bits_in = np.ascontiguousarray(bits_in, np.uint8)
        out = np.zeros((bits_in.shape[0], out_bits), np.uint8)
        rc = self.L.orc_eval_circuit_plain(C.byref(ccirc), _p(bits_in), _p(out), bits_in.shape[0])
        if rc != 0:
            raise ValueError("oracle: malformed circuit")
        return out

    def ntt(self, poly):
        p = np.ascontiguousarray(poly, np.uint32).copy(); self.L.orc_ntt_forward(self.h, _p(p)); return p

    def intt(self, poly):
        p = np.ascontiguousarray(poly, np.uint32).copy(); self.L.orc_ntt_inverse(self.h, _p(p)); return p

    def blind_rotate_trace(self, gate, a, b, steps):
        acc = np.zeros(2 * N, np.uint32)
        a = np.ascontiguousarray(a, np.uint16); b = np.ascontiguousarray(b, np.uint16)
        self.L.orc_blind_rotate_trace(self.h, GATE[gate], _p(a), _p(b), steps, _p(acc))
        return acc

    def bootstrap_extract(self, gate, a, b):
        ext = np.zeros(N + 1, np.uint16)
        a = np.ascontiguousarray(a, np.uint16); b = np.ascontiguousarray(b, np.uint16)
        self.L.orc_bootstrap_extract(self.h, GATE[gate], _p(a), _p(b), _p(ext))
        return ext
