"""Synthetic stand-in for the herd client (key generation, encryption, decryption in this repo's flat formats).
TEST / BENCH INFRASTRUCTURE ONLY -- see herd_clientsim.h.  Not imported by the herdsman_b200 package."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
RING_N = 1024


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libherdclientsim.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        L = C.CDLL(path)
        vp, sz, u32, u64 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64
        L.clientsim_key_blob_size.argtypes = [u32]; L.clientsim_key_blob_size.restype = sz
        L.clientsim_keygen.argtypes = [u32, u64, vp, vp, vp, sz]
        L.clientsim_encrypt_bits.argtypes = [u32, vp, u64, u64, vp, sz, vp]
        L.clientsim_decrypt_bits.argtypes = [u32, vp, vp, sz, vp]
        _LIB = L
    return _LIB


def _ptr(x):
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return x.ctypes.data
    return x.data_ptr()          # torch tensor (pinned host)


def key_blob_size(n):
    return lib().clientsim_key_blob_size(n)


def keygen(n, seed=0x48455244):
    """(sk_lwe, sk_ring, key blob)"""
    L = lib()
    sk = np.zeros(n, np.int8); skr = np.zeros(RING_N, np.int8)
    blob = np.zeros(L.clientsim_key_blob_size(n), np.uint8)
    if L.clientsim_keygen(n, seed, _ptr(sk), _ptr(skr), _ptr(blob), blob.size):
        raise ValueError("clientsim_keygen: bad argument")
    return sk, skr, blob


def encrypt_bits(n, sk, bits, seed=1, first=0, out=None):
    bits = np.ascontiguousarray(bits, np.uint8)
    cts = out if out is not None else np.empty(bits.shape + (n + 1,), np.uint16)
    if lib().clientsim_encrypt_bits(n, _ptr(sk), seed, first, _ptr(bits), bits.size, _ptr(cts)):
        raise ValueError("clientsim_encrypt_bits: bad argument")
    return cts


def decrypt_bits(n, sk, cts):
    cts = np.ascontiguousarray(cts, np.uint16)
    bits = np.zeros(cts.shape[:-1], np.uint8)
    if lib().clientsim_decrypt_bits(n, _ptr(sk), _ptr(cts), bits.size, _ptr(bits)):
        raise ValueError("clientsim_decrypt_bits: bad argument")
    return bits
