// client.cpp -- synthetic stand-in for the herd client (libherd, not in the reference tree):
// key generation, bit encryption and decryption in the flat formats of herdb_params.h.
// TEST / BENCH INFRASTRUCTURE ONLY, built into its own library (clientsim/libherdclientsim.so), not
// into libherdb.so: herdsman never sees secret keys, the generator is a counter-based mixer seeded
// with 64 bits (NOT a CSPRNG), and the flat key / row formats are this repo's own, not the
// OpenFHE-serialised objects a real herd client uploads ("Formats").
//
// Follows OpenFHE binfhe v1.0.x KeyGen / KeySwitchGen / KeyGenCGGI / Encrypt / Decrypt
// (lwe-pke.cpp, rgsw-acc-cggi.cpp) with the deterministic generator of herdb_params.h in
// place of OpenFHE's PRNG.  Written independently of oracle/oracle.c; tests check that both
// produce identical blobs from the same seed.
#include <cstring>
#include <vector>
#include "herd_clientsim.h"
#include "../herdsman_b200/csrc/modarith.cuh"

namespace {
using namespace herdb;

inline u64 mix64(u64 z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
inline u64 rng(u64 seed, u64 stream, u64 idx) { return mix64(mix64(seed ^ (stream * HERDB_RNG_C1)) + idx * HERDB_RNG_C2); }
inline u32 uniform(u64 r, u32 m) { return (u32)(((unsigned __int128)r * m) >> 64); }
inline int ternary(u64 r) { return (int)((r >> 33) % 3u) - 1; }
const u64 kCdt[2 * HERDB_DGG_K + 1] = HERDB_DGG_CDT_INIT;
inline int gauss(u64 r) {
    int c = 0;
    for (int i = 0; i < 2 * HERDB_DGG_K; ++i) c += (kCdt[i] <= r);
    return c - HERDB_DGG_K;
}

struct HostNtt {
    std::vector<u32> tw, itw;
    HostNtt() : tw(kN), itw(kN) {
        std::vector<u32> pw(2 * kN), ipw(2 * kN);
        pw[0] = ipw[0] = 1;
        for (u32 j = 1; j < 2 * kN; ++j) { pw[j] = cmulmod(pw[j - 1], kPsi); ipw[j] = cmulmod(ipw[j - 1], kPsiInv); }
        for (u32 k = 0; k < kN; ++k) { tw[k] = pw[cbitrev10(k)]; itw[k] = ipw[cbitrev10(k)]; }
    }
    void fwd(u32* a) const {
        u32 t = kN;
        for (u32 m = 1; m < kN; m <<= 1) {
            t >>= 1;
            for (u32 i = 0; i < m; ++i) {
                const u64 w = tw[m + i];
                for (u32 j = 2 * i * t; j < 2 * i * t + t; ++j) {
                    u32 u = a[j], v = (u32)(a[j + t] * w % kQ);
                    a[j] = u + v >= kQ ? u + v - kQ : u + v;
                    a[j + t] = u >= v ? u - v : u + kQ - v;
                }
            }
        }
    }
    void inv(u32* a) const {
        u32 t = 1;
        for (u32 m = kN >> 1; m >= 1; m >>= 1) {
            for (u32 i = 0; i < m; ++i) {
                const u64 w = itw[m + i];
                for (u32 j = 2 * i * t; j < 2 * i * t + t; ++j) {
                    u32 u = a[j], v = a[j + t];
                    a[j] = u + v >= kQ ? u + v - kQ : u + v;
                    a[j + t] = (u32)((u64)(u >= v ? u - v : u + kQ - v) * w % kQ);
                }
            }
            t <<= 1;
        }
        for (u32 j = 0; j < kN; ++j) a[j] = (u32)((u64)a[j] * kNInv % kQ);
    }
};
}  // namespace

extern "C" {

size_t clientsim_key_blob_size(uint32_t n) {
    return sizeof(herdb_key_header) + (size_t)n * 2 * kDG2 * 2 * kN * 4 + (size_t)kN * kBaseKS * kDKS * (n + 1) * 2;
}

int clientsim_keygen(uint32_t n, uint64_t seed, int8_t* sk_lwe, int8_t* sk_ring, void* blob, size_t blob_len) {
    if (!sk_lwe || !sk_ring || !blob || n == 0 || n > HERDB_N_STD128) return -1;
    if (blob_len < clientsim_key_blob_size(n)) return -1;
    const HostNtt ntt;
    herdb_key_header h;
    memset(&h, 0, sizeof h);
    h.magic = HERDB_KEY_MAGIC; h.version = HERDB_KEY_VERSION;
    h.n = n; h.ring_n = kN; h.q = kLweQ; h.Q = kQ; h.qKS = kQKS;
    h.baseG_log = HERDB_BASEG_LOG; h.digitsG = kDG; h.baseKS_log = HERDB_BASEKS_LOG; h.digitsKS = kDKS;
    memcpy(blob, &h, sizeof h);
    for (u32 i = 0; i < n; ++i) sk_lwe[i] = (int8_t)ternary(rng(seed, HERDB_STREAM_SK_LWE, i));
    for (u32 k = 0; k < kN; ++k) sk_ring[k] = (int8_t)ternary(rng(seed, HERDB_STREAM_SK_RING, k));
    std::vector<u32> z(kN);
    for (u32 k = 0; k < kN; ++k) z[k] = sk_ring[k] < 0 ? kQ - 1 : (u32)sk_ring[k];
    ntt.fwd(z.data());

    u32* bsk = reinterpret_cast<u32*>(static_cast<uint8_t*>(blob) + sizeof h);
    const long rgsw_count = (long)n * 2;
#pragma omp parallel for schedule(static)
    for (long t = 0; t < rgsw_count; ++t) {
        const u32 i = (u32)(t >> 1), s = (u32)(t & 1);
        const bool m = s == 0 ? sk_lwe[i] == 1 : sk_lwe[i] == -1;
        std::vector<u32> tmp(kN);
        for (u32 r = 0; r < kDG2; ++r) {
            const u64 base = (((u64)i * 2 + s) * kDG2 + r) * kN;
            u32* a = bsk + base * 2;
            u32* b = a + kN;
            for (u32 k = 0; k < kN; ++k) {
                a[k] = uniform(rng(seed, HERDB_STREAM_BSK_A, base + k), kQ);
                const int e = gauss(rng(seed, HERDB_STREAM_BSK_E, base + k));
                b[k] = e < 0 ? kQ + e : (u32)e;
                tmp[k] = a[k];
            }
            ntt.fwd(tmp.data());
            for (u32 k = 0; k < kN; ++k) tmp[k] = cmulmod(tmp[k], z[k]);
            ntt.inv(tmp.data());
            for (u32 k = 0; k < kN; ++k) { u32 v = b[k] + tmp[k]; b[k] = v >= kQ ? v - kQ : v; }
            if (m) {
                u32* col = (r & 1) ? b : a;
                u32 v = col[0] + (u32)((1ull << (HERDB_BASEG_LOG * (r >> 1))) % kQ);
                col[0] = v >= kQ ? v - kQ : v;
            }
        }
    }
    u16* ksk = reinterpret_cast<u16*>(bsk + (size_t)n * 2 * kDG2 * 2 * kN);
    const long ks_rows = (long)kN * kBaseKS * kDKS;
#pragma omp parallel for schedule(static)
    for (long row = 0; row < ks_rows; ++row) {
        const u32 j = (u32)(row % kDKS), v = (u32)((row / kDKS) % kBaseKS), i = (u32)(row / (kDKS * kBaseKS));
        u16* a = ksk + (size_t)row * (n + 1);
        int64_t b = gauss(rng(seed, HERDB_STREAM_KSK_E, (u64)row)) + (int64_t)sk_ring[i] * (int64_t)(v << (HERDB_BASEKS_LOG * j));
        for (u32 k = 0; k < n; ++k) {
            a[k] = (u16)uniform(rng(seed, HERDB_STREAM_KSK_A, (u64)row * n + k), kQKS);
            b += (int64_t)a[k] * sk_lwe[k];
        }
        a[n] = (u16)(((b % (int64_t)kQKS) + kQKS) % kQKS);
    }
    return 0;
}

int clientsim_encrypt_bits(uint32_t n, const int8_t* sk_lwe, uint64_t seed, uint64_t first_ct_index,
                              const uint8_t* bits, size_t count, uint16_t* cts) {
    if (!sk_lwe || !bits || !cts || n == 0 || n > HERDB_N_STD128) return -1;
#pragma omp parallel for schedule(static)
    for (long c = 0; c < (long)count; ++c) {
        u16* ct = cts + (size_t)c * (n + 1);
        const u64 ci = first_ct_index + (u64)c;
        int64_t b = (int64_t)(bits[c] & 1) * (kLweQ / HERDB_PT_MOD) + gauss(rng(seed, HERDB_STREAM_CT_E, ci));
        for (u32 k = 0; k < n; ++k) {
            ct[k] = (u16)uniform(rng(seed, HERDB_STREAM_CT_A, ci * n + k), kLweQ);
            b += (int64_t)ct[k] * sk_lwe[k];
        }
        ct[n] = (u16)(((b % (int64_t)kLweQ) + kLweQ) % kLweQ);
    }
    return 0;
}

int clientsim_decrypt_bits(uint32_t n, const int8_t* sk_lwe, const uint16_t* cts, size_t count, uint8_t* bits) {
    if (!sk_lwe || !bits || !cts || n == 0 || n > HERDB_N_STD128) return -1;
#pragma omp parallel for schedule(static)
    for (long c = 0; c < (long)count; ++c) {
        const u16* ct = cts + (size_t)c * (n + 1);
        int64_t r = ct[n];
        for (u32 k = 0; k < n; ++k) r -= (int64_t)ct[k] * sk_lwe[k];
        r += kLweQ / (2 * HERDB_PT_MOD);
        const u32 rr = (u32)(((r % (int64_t)kLweQ) + kLweQ) % kLweQ);
        bits[c] = (uint8_t)((HERDB_PT_MOD * rr) / kLweQ);
    }
    return 0;
}

}  // extern "C"
