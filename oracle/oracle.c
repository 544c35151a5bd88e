/*
 * oracle.c -- CPU oracle (test infrastructure; see oracle.h for the scope note and
 * the PARITY UNPINNED statement).  Plain C11 + OpenMP, 64-bit integer arithmetic
 * mirroring OpenFHE's NativeInteger (MATHBACKEND=2..., NATIVE_SIZE=64).
 *
 * Each function names the OpenFHE v1.0.x routine it restates; OpenFHE is pinned by
 * the reference at /root/reference/common/action/install-openfhe/action.yml:9-12
 * (v1.0.3) and /root/reference/common/nix/openfhe/default.nix:5-12 (v1.0.4) and is
 * linked at /root/reference/CMakeLists.txt:73,210-211 -- it is not in the tree.
 * In-tree facts followed: circuit model common/include/herd/common/model/circuit/
 * {circuit,nodes,operation}.hpp, operand order dag.hpp:208-230, bit widths
 * common/src/native_type_mapping.cpp:8-35.
 */
#include "oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define N    HERDB_RING_N
#define LOGN HERDB_RING_LOGN
#define QQ   ((uint64_t)HERDB_Q)
#define DG   HERDB_DIGITS_G
#define DG2  (2u * HERDB_DIGITS_G)
#define LQ   HERDB_LWE_Q
#define QKS  HERDB_QKS
#define BKS  (1u << HERDB_BASEKS_LOG)
#define DKS  HERDB_DIGITS_KS

/* ------------------------------------------------------------------ generator (herdb_params.h spec) */
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
uint64_t orc_rng(uint64_t seed, uint64_t stream, uint64_t idx) {
    return mix64(mix64(seed ^ (stream * HERDB_RNG_C1)) + idx * HERDB_RNG_C2);
}
static inline uint32_t rng_uniform(uint64_t r, uint32_t m) {
    return (uint32_t)(((unsigned __int128)r * m) >> 64);
}
static inline int rng_ternary(uint64_t r) { return (int)((r >> 33) % 3u) - 1; }
static const uint64_t DGG_CDT[2 * HERDB_DGG_K + 1] = HERDB_DGG_CDT_INIT;
static inline int rng_gauss(uint64_t r) {
    int c = 0;
    for (int i = 0; i < 2 * HERDB_DGG_K; ++i) c += (DGG_CDT[i] <= r);
    return c - HERDB_DGG_K;
}

/* ------------------------------------------------------------------ modular arithmetic (NativeInteger) */
static inline uint64_t mod_add(uint64_t a, uint64_t b, uint64_t m) { uint64_t s = a + b; return s >= m ? s - m : s; }
static inline uint64_t mod_sub(uint64_t a, uint64_t b, uint64_t m) { return a >= b ? a - b : a + m - b; }
static inline uint64_t mod_mul(uint64_t a, uint64_t b) { return (a * b) % QQ; }   /* a,b < 2^32: no overflow */
static uint64_t mod_pow(uint64_t b, uint64_t e) {
    uint64_t r = 1; b %= QQ;
    while (e) { if (e & 1) r = mod_mul(r, b); b = mod_mul(b, b); e >>= 1; }
    return r;
}
static inline uint32_t bitrev(uint32_t x, unsigned bits) {
    uint32_t r = 0;
    for (unsigned i = 0; i < bits; ++i) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

/* RootOfUnity(2N, Q): OpenFHE returns the minimum primitive 2N-th root so the choice is
 * deterministic (core/lib/math/nbtheory-impl.h).  Q-1 = 2^11 * 3*5*17*257. */
static uint32_t min_primitive_root_2N(void) {
    const uint64_t order = 2 * N;
    const uint64_t cof = (QQ - 1) / order;
    uint64_t g = 0;
    for (uint64_t c = 2; c < QQ; ++c) {          /* any element of exact order 2N */
        uint64_t w = mod_pow(c, cof);
        if (mod_pow(w, order / 2) == QQ - 1) { g = w; break; }
    }
    uint64_t best = g, x = g, g2 = mod_mul(g, g);
    for (uint64_t k = 1; k < order; k += 2) {     /* odd powers are exactly the primitive roots */
        if (x < best) best = x;
        x = mod_mul(x, g2);
    }
    return (uint32_t)best;
}

/* ------------------------------------------------------------------ context */
struct orc_ctx {
    uint32_t n;                     /* LWE dimension of the loaded key (0 = none) */
    uint32_t psi;
    uint32_t* tw;                   /* psi^bitrev(k), k < N                        */
    uint32_t* tw_shoup;             /* floor(tw * 2^32 / Q)  (ModMulFastConst)     */
    uint32_t* itw;                  /* psi^-bitrev(k)                              */
    uint32_t* itw_shoup;
    uint32_t ninv, ninv_shoup;
    uint32_t* monomial;             /* [2N][N] evaluation form of X^i - 1 (RingGSWCryptoParams::PreCompute) */
    uint32_t* bsk;                  /* [n][2][DG2][2][N] evaluation form            */
    uint16_t* ksk;                  /* [N][BKS][DKS][n+1]                           */
};

static inline uint32_t shoup_pre(uint32_t w) { return (uint32_t)(((uint64_t)w << 32) / QQ); }
/* ModMulFastConst: w*y mod Q with precomputed w' */
static inline uint32_t mul_shoup(uint32_t w, uint32_t wp, uint32_t y) {
    uint32_t q = (uint32_t)(((uint64_t)wp * y) >> 32);
    uint32_t r = (uint32_t)((uint64_t)w * y - (uint64_t)q * QQ);
    return r >= QQ ? r - (uint32_t)QQ : r;
}

/* ChineseRemainderTransformFTT::ForwardTransformToBitReverse (Cooley-Tukey, natural in,
 * bit-reversed out, merged psi powers).  Butterflies are Harvey's lazy form (values kept in
 * [0,4Q), one exact correction at the end) -- the residues are identical to per-butterfly
 * ModAdd/ModSub; this only keeps the CPU baseline honest (branch-free inner loop). */
static inline uint32_t mul_shoup_lazy(uint32_t w, uint32_t wp, uint32_t y) {     /* in [0,2Q) */
    uint32_t q = (uint32_t)(((uint64_t)wp * y) >> 32);
    return w * y - q * (uint32_t)QQ;
}
static inline uint32_t csub(uint32_t x, uint32_t m) { uint32_t d = x - m; return d < x ? d : x; } /* x>=m ? x-m : x */
void orc_ntt_forward(const orc_ctx* c, uint32_t* a) {
    const uint32_t Q2 = 2 * (uint32_t)QQ;
    uint32_t t = N;
    for (uint32_t m = 1; m < N; m <<= 1) {
        t >>= 1;
        for (uint32_t i = 0; i < m; ++i) {
            uint32_t w = c->tw[m + i], wp = c->tw_shoup[m + i];
            uint32_t* x = a + 2 * i * t; uint32_t* y = x + t;
            for (uint32_t j = 0; j < t; ++j) {
                uint32_t u = csub(x[j], Q2), v = mul_shoup_lazy(w, wp, y[j]);
                x[j] = u + v;
                y[j] = u - v + Q2;
            }
        }
    }
    for (uint32_t j = 0; j < N; ++j) a[j] = csub(csub(a[j], Q2), (uint32_t)QQ);
}
/* InverseTransformFromBitReverse (Gentleman-Sande, bit-reversed in, natural out, * N^-1) */
void orc_ntt_inverse(const orc_ctx* c, uint32_t* a) {
    const uint32_t Q2 = 2 * (uint32_t)QQ;
    uint32_t t = 1;
    for (uint32_t m = N >> 1; m >= 1; m >>= 1) {
        for (uint32_t i = 0; i < m; ++i) {
            uint32_t w = c->itw[m + i], wp = c->itw_shoup[m + i];
            uint32_t* x = a + 2 * i * t; uint32_t* y = x + t;
            for (uint32_t j = 0; j < t; ++j) {
                uint32_t u = x[j], v = y[j];                 /* both in [0,2Q) */
                x[j] = csub(u + v, Q2);
                y[j] = mul_shoup_lazy(w, wp, u - v + Q2);
            }
        }
        t <<= 1;
    }
    for (uint32_t j = 0; j < N; ++j) a[j] = csub(mul_shoup_lazy(c->ninv, c->ninv_shoup, a[j]), (uint32_t)QQ);
}

void orc_negacyclic_mul_schoolbook(const uint32_t* a, const uint32_t* b, uint32_t* out) {
    for (uint32_t k = 0; k < N; ++k) {
        uint64_t pos = 0, neg = 0;
        for (uint32_t i = 0; i < N; ++i) {
            uint64_t p = mod_mul(a[i], b[(k + N - i) % N]);
            if (i <= k) pos = mod_add(pos, p, QQ); else neg = mod_add(neg, p, QQ);
        }
        out[k] = (uint32_t)mod_sub(pos, neg, QQ);
    }
}

orc_ctx* orc_create(void) {
    orc_ctx* c = (orc_ctx*)calloc(1, sizeof(orc_ctx));
    if (!c) return NULL;
    c->psi = min_primitive_root_2N();
    c->tw = malloc(N * 4); c->tw_shoup = malloc(N * 4); c->itw = malloc(N * 4); c->itw_shoup = malloc(N * 4);
    uint64_t psi_inv = mod_pow(c->psi, 2 * N - 1);
    for (uint32_t k = 0; k < N; ++k) {
        uint32_t e = bitrev(k, LOGN);
        c->tw[k] = (uint32_t)mod_pow(c->psi, e);   c->tw_shoup[k] = shoup_pre(c->tw[k]);
        c->itw[k] = (uint32_t)mod_pow(psi_inv, e); c->itw_shoup[k] = shoup_pre(c->itw[k]);
    }
    c->ninv = (uint32_t)mod_pow(N, QQ - 2); c->ninv_shoup = shoup_pre(c->ninv);
    /* m_monomials: for i < N: X^i - 1; for i >= N: -X^(i-N) - 1; stored in EVALUATION form */
    c->monomial = malloc((size_t)2 * N * N * 4);
    for (uint32_t i = 0; i < 2 * N; ++i) {
        uint32_t* p = c->monomial + (size_t)i * N;
        memset(p, 0, N * 4);
        if (i < N) p[i] = 1; else p[i - N] = (uint32_t)QQ - 1;
        p[0] = (uint32_t)mod_sub(p[0], 1, QQ);
        orc_ntt_forward(c, p);
    }
    return c;
}
void orc_destroy(orc_ctx* c) {
    if (!c) return;
    free(c->tw); free(c->tw_shoup); free(c->itw); free(c->itw_shoup);
    free(c->monomial); free(c->bsk); free(c->ksk); free(c);
}
uint32_t orc_psi(const orc_ctx* c) { return c->psi; }
uint32_t orc_lwe_n(const orc_ctx* c) { return c->n; }

/* ------------------------------------------------------------------ client side: keys and ciphertexts */
size_t orc_key_blob_size(uint32_t n) {
    return sizeof(herdb_key_header) + (size_t)n * 2 * DG2 * 2 * N * 4 + (size_t)N * BKS * DKS * (n + 1) * 2;
}

/* RingGSWAccumulatorCGGI::KeyGenCGGI: RGSW(m) = 2*dG rows (a_r, a_r*z + e_r), plus
 * m * Gpow[r>>1] added to coefficient 0 of column (r & 1).  Stored here in COEFFICIENT form. */
static void rgsw_encrypt_coeff(const orc_ctx* c, const uint32_t* z_ntt, int m, uint64_t seed,
                               uint64_t base_idx, uint32_t* out /* [DG2][2][N] */) {
    uint32_t tmp[N];
    for (uint32_t r = 0; r < DG2; ++r) {
        uint32_t* a = out + (size_t)(r * 2 + 0) * N;
        uint32_t* b = out + (size_t)(r * 2 + 1) * N;
        for (uint32_t k = 0; k < N; ++k) {
            uint64_t idx = base_idx + (uint64_t)r * N + k;
            a[k] = rng_uniform(orc_rng(seed, HERDB_STREAM_BSK_A, idx), (uint32_t)QQ);
            int e = rng_gauss(orc_rng(seed, HERDB_STREAM_BSK_E, idx));
            b[k] = (uint32_t)(e >= 0 ? e : (int64_t)QQ + e);
        }
        memcpy(tmp, a, sizeof tmp);
        orc_ntt_forward(c, tmp);
        for (uint32_t k = 0; k < N; ++k) tmp[k] = (uint32_t)mod_mul(tmp[k], z_ntt[k]);
        orc_ntt_inverse(c, tmp);
        for (uint32_t k = 0; k < N; ++k) b[k] = (uint32_t)mod_add(b[k], tmp[k], QQ);
        if (m) {
            uint64_t g = 1ull << (HERDB_BASEG_LOG * (r >> 1));          /* Gpow[r>>1] = baseG^(r>>1) mod Q */
            uint32_t* col = (r & 1) ? b : a;
            col[0] = (uint32_t)mod_add(col[0], g % QQ, QQ);
        }
    }
}

int orc_keygen(uint32_t n, uint64_t seed, int8_t* sk_lwe, int8_t* sk_ring, void* blob, size_t blob_len) {
    if (blob_len < orc_key_blob_size(n)) return -1;
    orc_ctx* c = orc_create();
    herdb_key_header* h = (herdb_key_header*)blob;
    memset(h, 0, sizeof *h);
    h->magic = HERDB_KEY_MAGIC; h->version = HERDB_KEY_VERSION;
    h->n = n; h->ring_n = N; h->q = LQ; h->Q = (uint32_t)QQ; h->qKS = QKS;
    h->baseG_log = HERDB_BASEG_LOG; h->digitsG = DG; h->baseKS_log = HERDB_BASEKS_LOG; h->digitsKS = DKS;
    /* LWEEncryptionScheme::KeyGen: uniform ternary secrets (n for LWE, N for the ring key) */
    for (uint32_t i = 0; i < n; ++i) sk_lwe[i] = (int8_t)rng_ternary(orc_rng(seed, HERDB_STREAM_SK_LWE, i));
    for (uint32_t k = 0; k < N; ++k) sk_ring[k] = (int8_t)rng_ternary(orc_rng(seed, HERDB_STREAM_SK_RING, k));
    uint32_t z_ntt[N];
    for (uint32_t k = 0; k < N; ++k) z_ntt[k] = (uint32_t)(sk_ring[k] >= 0 ? sk_ring[k] : (int64_t)QQ + sk_ring[k]);
    orc_ntt_forward(c, z_ntt);
    /* RingGSWAccumulatorCGGI::KeyGenAcc: per LWE secret coefficient two RGSW keys,
     * [0] = E(s_i == +1), [1] = E(s_i == -1) */
    uint32_t* bsk = (uint32_t*)((uint8_t*)blob + sizeof *h);
    #pragma omp parallel for schedule(static)
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t s = 0; s < 2; ++s) {
            int m = (s == 0) ? (sk_lwe[i] == 1) : (sk_lwe[i] == -1);
            uint64_t base = ((uint64_t)i * 2 + s) * DG2 * N;
            rgsw_encrypt_coeff(c, z_ntt, m, seed, base, bsk + base * 2);
        }
    /* LWEEncryptionScheme::KeySwitchGen: K[i][v][j] = LWE_qKS( v * baseKS^j * z_i ), b = e + m + <a, s> */
    uint16_t* ksk = (uint16_t*)(bsk + (size_t)n * 2 * DG2 * 2 * N);
    #pragma omp parallel for schedule(static)
    for (uint32_t i = 0; i < N; ++i)
        for (uint32_t v = 0; v < BKS; ++v)
            for (uint32_t j = 0; j < DKS; ++j) {
                uint64_t row = ((uint64_t)i * BKS + v) * DKS + j;
                uint16_t* a = ksk + row * (n + 1);
                int e = rng_gauss(orc_rng(seed, HERDB_STREAM_KSK_E, row));
                int64_t msg = (int64_t)sk_ring[i] * (int64_t)(v << (HERDB_BASEKS_LOG * j));
                int64_t b = e + msg;
                for (uint32_t k = 0; k < n; ++k) {
                    a[k] = (uint16_t)rng_uniform(orc_rng(seed, HERDB_STREAM_KSK_A, row * n + k), QKS);
                    b += (int64_t)a[k] * sk_lwe[k];
                }
                a[n] = (uint16_t)(((b % (int64_t)QKS) + QKS) % QKS);
            }
    orc_destroy(c);
    return 0;
}

/* LWEEncryptionScheme::Encrypt: b = m*q/p + e + <a,s> mod q */
void orc_encrypt_bits(uint32_t n, const int8_t* sk, uint64_t seed, uint64_t first, const uint8_t* bits,
                      size_t count, uint16_t* cts) {
    #pragma omp parallel for schedule(static)
    for (size_t c = 0; c < count; ++c) {
        uint16_t* ct = cts + c * (n + 1);
        int64_t b = (int64_t)(bits[c] & 1) * (LQ / HERDB_PT_MOD) + rng_gauss(orc_rng(seed, HERDB_STREAM_CT_E, first + c));
        for (uint32_t k = 0; k < n; ++k) {
            ct[k] = (uint16_t)rng_uniform(orc_rng(seed, HERDB_STREAM_CT_A, (first + c) * n + k), LQ);
            b += (int64_t)ct[k] * sk[k];
        }
        ct[n] = (uint16_t)(((b % (int64_t)LQ) + LQ) % LQ);
    }
}
static inline uint32_t lwe_phase(uint32_t n, const int8_t* sk, const uint16_t* ct) {
    int64_t r = ct[n];
    for (uint32_t k = 0; k < n; ++k) r -= (int64_t)ct[k] * sk[k];
    return (uint32_t)(((r % (int64_t)LQ) + LQ) % LQ);
}
/* LWEEncryptionScheme::Decrypt: r = b - <a,s> + q/(2p); result = p*r/q */
void orc_decrypt_bits(uint32_t n, const int8_t* sk, const uint16_t* cts, size_t count, uint8_t* bits) {
    for (size_t c = 0; c < count; ++c) {
        uint32_t r = (lwe_phase(n, sk, cts + c * (n + 1)) + LQ / (2 * HERDB_PT_MOD)) % LQ;
        bits[c] = (uint8_t)((HERDB_PT_MOD * r) / LQ);
    }
}
int32_t orc_noise(uint32_t n, const int8_t* sk, const uint16_t* ct, int bit) {
    int32_t d = (int32_t)lwe_phase(n, sk, ct) - (int32_t)((bit & 1) * (LQ / HERDB_PT_MOD));
    d %= (int32_t)LQ; if (d < 0) d += LQ;
    return d >= (int32_t)LQ / 2 ? d - (int32_t)LQ : d;
}

/* ------------------------------------------------------------------ server side: key import */
int orc_load_key(orc_ctx* c, const void* blob, size_t len) {
    const herdb_key_header* h = (const herdb_key_header*)blob;
    if (len < sizeof *h || h->magic != HERDB_KEY_MAGIC || h->version != HERDB_KEY_VERSION) return -4;
    if (h->ring_n != N || h->Q != QQ || h->q != LQ || h->qKS != QKS || h->baseG_log != HERDB_BASEG_LOG ||
        h->digitsG != DG || h->baseKS_log != HERDB_BASEKS_LOG || h->digitsKS != DKS) return -4;
    if (len < orc_key_blob_size(h->n)) return -4;
    uint32_t n = h->n;
    free(c->bsk); free(c->ksk);
    size_t bsk_words = (size_t)n * 2 * DG2 * 2 * N;
    c->bsk = malloc(bsk_words * 4);
    memcpy(c->bsk, (const uint8_t*)blob + sizeof *h, bsk_words * 4);
    #pragma omp parallel for schedule(static)
    for (size_t p = 0; p < bsk_words / N; ++p) orc_ntt_forward(c, c->bsk + p * N);   /* SetFormat(EVALUATION) */
    size_t ksk_words = (size_t)N * BKS * DKS * (n + 1);
    c->ksk = malloc(ksk_words * 2);
    memcpy(c->ksk, (const uint8_t*)blob + sizeof *h + bsk_words * 4, ksk_words * 2);
    c->n = n;
    return 0;
}

/* ------------------------------------------------------------------ server side: gate evaluation */

/* LWEEncryptionScheme::RoundqQ: floor(0.5 + v*q/Q) mod q, in double as OpenFHE does */
uint32_t orc_round_qQ_double(uint32_t v, uint32_t q, uint32_t Q) {
    return (uint32_t)((uint64_t)floor(0.5 + (double)v * (double)q / (double)Q) % q);
}
/* the integer form the CUDA kernels use; proven equal for odd Q / power-of-two ratios
 *  and checked exhaustively in tests/test_oracle.py */
uint32_t orc_round_qQ_int(uint32_t v, uint32_t q, uint32_t Q) {
    return (uint32_t)(((2 * (uint64_t)v * q + Q) / (2 * (uint64_t)Q)) % q);
}

/* number of v in [start, start+count) where the double and the integer form differ (tests) */
uint64_t orc_round_mismatches(uint32_t q, uint32_t Q, uint32_t start, uint32_t count) {
    uint64_t bad = 0;
    #pragma omp parallel for reduction(+:bad)
    for (uint64_t v = start; v < (uint64_t)start + count; ++v)
        bad += orc_round_qQ_double((uint32_t)v, q, Q) != orc_round_qQ_int((uint32_t)v, q, Q);
    return bad;
}

/* RingGSWAccumulator::SignedDigitDecompose, VARIANT A, for one coefficient */
void orc_signed_digit_decompose(uint32_t t, int32_t digits[DG]) {
    const int64_t Q_int = (int64_t)QQ;
    const uint64_t QHalf = QQ >> 1;
    const int gBits = HERDB_BASEG_LOG;
    const int gBitsMaxBits = 64 - gBits;
    int64_t d = 0;
    if (t < QHalf) d += (int64_t)t; else d += (int64_t)t - Q_int;
    for (uint32_t l = 0; l < DG; ++l) {
        int64_t r = (int64_t)((uint64_t)d << gBitsMaxBits);   /* sign-extend the low gBits */
        r >>= gBitsMaxBits;
        d -= r;
        d >>= gBits;
        digits[l] = (int32_t)r;
    }
}

/* BootstrapGateCore: accumulator initialisation.  acc[0] = 0, acc[1] = NTT(m) */
static const uint32_t GATE_CONST_EIGHTHS[6] = { 5, 7, 1, 3, 5, 1 };   /* OR AND NOR NAND XOR_FAST XNOR_FAST */
static void acc_init(const orc_ctx* c, int gate, uint32_t b, uint32_t* acc /* [2][N] eval */) {
    const uint32_t q = LQ, qHalf = q >> 1;
    uint32_t q1 = GATE_CONST_EIGHTHS[gate] * (q >> 3);
    uint32_t q2 = (q1 + qHalf) % q;
    uint32_t Q2p = (uint32_t)(QQ / (2 * HERDB_PT_MOD)) + 1;     /* Q/8 + 1 */
    uint32_t Q2pNeg = (uint32_t)QQ - Q2p;
    uint32_t factor = 2 * N / q;
    uint32_t* m = acc + N;
    memset(acc, 0, 2 * N * 4);
    for (uint32_t j = 0; j < qHalf; ++j) {
        uint32_t temp = (b + q - j) % q;
        if (q1 < q2) m[j * factor] = (temp >= q1 && temp < q2) ? Q2pNeg : Q2p;
        else         m[j * factor] = (temp >= q2 && temp < q1) ? Q2p : Q2pNeg;
    }
    orc_ntt_forward(c, m);
}

/* RingGSWAccumulatorCGGI::AddToAccCGGI: one CMUX step,
 * acc += (dct * ek1) * monomial(a) + (dct * ek2) * monomial(2N - a) */
static void add_to_acc_cggi(const orc_ctx* c, const uint32_t* ek1, const uint32_t* ek2, uint32_t a,
                            uint32_t* acc /* [2][N] eval */) {
    uint32_t ct[2][N];
    static _Thread_local uint32_t dct[DG2][N];
    memcpy(ct, acc, sizeof ct);
    orc_ntt_inverse(c, ct[0]);                      /* SetFormat(COEFFICIENT): 2 inverse NTTs */
    orc_ntt_inverse(c, ct[1]);
    for (uint32_t j = 0; j < 2; ++j)                /* SignedDigitDecompose: output index j + 2*l */
        for (uint32_t k = 0; k < N; ++k) {
            int32_t dig[DG];
            orc_signed_digit_decompose(ct[j][k], dig);
            for (uint32_t l = 0; l < DG; ++l)
                dct[j + 2 * l][k] = (uint32_t)(dig[l] >= 0 ? dig[l] : (int64_t)QQ + dig[l]);
        }
    for (uint32_t r = 0; r < DG2; ++r) orc_ntt_forward(c, dct[r]);   /* 2*dG forward NTTs */
    uint32_t M = 2 * N;
    uint32_t indexPos = a, indexNeg = (M - a) % M;  /* index 2N wraps to 0 */
    if (indexPos == M) indexPos = 0;
    const uint32_t* monomial = c->monomial + (size_t)indexPos * N;
    const uint32_t* monomialNeg = c->monomial + (size_t)indexNeg * N;
    for (uint32_t j = 0; j < 2; ++j)
        for (uint32_t k = 0; k < N; ++k) {
            /* temp1 = sum_l dct[l] * ev1[l][j]; products < 2^54, eight of them < 2^57: the sum is
             * accumulated exactly in 64 bits and reduced once (same residue as per-term ModMul/ModAdd) */
            uint64_t t1 = 0, t2 = 0;
            for (uint32_t l = 0; l < DG2; ++l) {
                t1 += (uint64_t)dct[l][k] * ek1[(size_t)(l * 2 + j) * N + k];
                t2 += (uint64_t)dct[l][k] * ek2[(size_t)(l * 2 + j) * N + k];
            }
            t1 %= QQ; t2 %= QQ;
            uint64_t v = acc[j * N + k] + (t1 * monomial[k] + t2 * monomialNeg[k]) % QQ;
            acc[j * N + k] = (uint32_t)(v >= QQ ? v - QQ : v);
        }
}

/* RingGSWAccumulatorCGGI::EvalAcc, first `steps` LWE coefficients */
static void eval_acc(const orc_ctx* c, const uint32_t* a /* n, mod q */, uint32_t steps, uint32_t* acc) {
    const uint32_t q = LQ, M = 2 * N;
    for (uint32_t i = 0; i < steps; ++i) {
        uint32_t idx = ((q - a[i]) % q) * (M / q);          /* mod.ModSub(a[i], mod) * (M / modInt) */
        const uint32_t* ek1 = c->bsk + ((size_t)i * 2 + 0) * DG2 * 2 * N;
        const uint32_t* ek2 = c->bsk + ((size_t)i * 2 + 1) * DG2 * 2 * N;
        add_to_acc_cggi(c, ek1, ek2, idx, acc);
    }
}

/* EvalBinGate's ciphertext preparation: ct1 + ct2 (or 2*(ct1 - ct2) for XOR_FAST/XNOR_FAST) */
static void gate_prep(const orc_ctx* c, int gate, const uint16_t* ct1, const uint16_t* ct2, uint32_t* prep) {
    for (uint32_t k = 0; k <= c->n; ++k) {
        if (gate == HERDB_GATE_XOR_FAST || gate == HERDB_GATE_XNOR_FAST)
            prep[k] = (2u * ((uint32_t)ct1[k] + LQ - ct2[k])) % LQ;
        else
            prep[k] = ((uint32_t)ct1[k] + ct2[k]) % LQ;
    }
}

void orc_blind_rotate_trace(const orc_ctx* c, int gate, const uint16_t* ct1, const uint16_t* ct2,
                            uint32_t steps, uint32_t* acc_coeff) {
    uint32_t prep[HERDB_N_STD128 + 1];
    gate_prep(c, gate, ct1, ct2, prep);
    acc_init(c, gate, prep[c->n], acc_coeff);
    eval_acc(c, prep, steps, acc_coeff);
    orc_ntt_inverse(c, acc_coeff);
    orc_ntt_inverse(c, acc_coeff + N);
}

/* EvalBinGate up to (and including) ModSwitch(qKS): Transpose(acc[0]) -> LWE a, acc[1][0] + Q/8+1 -> b */
void orc_bootstrap_extract(const orc_ctx* c, int gate, const uint16_t* ct1, const uint16_t* ct2, uint16_t* ext) {
    uint32_t acc[2 * N];
    orc_blind_rotate_trace(c, gate, ct1, ct2, c->n, acc);
    /* NativePoly::Transpose == automorphism X -> X^(2N-1): a'(0) = a(0), a'(k) = -a(N-k) */
    uint32_t at[N];
    at[0] = acc[0];
    for (uint32_t k = 1; k < N; ++k) at[k] = (uint32_t)mod_sub(0, acc[N - k], QQ);
    uint32_t b = (uint32_t)mod_add((uint32_t)(QQ / 8) + 1, acc[N], QQ);
    for (uint32_t k = 0; k < N; ++k) ext[k] = (uint16_t)orc_round_qQ_double(at[k], QKS, (uint32_t)QQ);
    ext[N] = (uint16_t)orc_round_qQ_double(b, QKS, (uint32_t)QQ);
}

/* LWEEncryptionScheme::KeySwitch then ModSwitch(q) */
static void keyswitch_modswitch(const orc_ctx* c, const uint16_t* ext, uint16_t* out) {
    uint32_t n = c->n;
    uint32_t a[HERDB_N_STD128];
    memset(a, 0, sizeof a);
    uint32_t b = ext[N];
    for (uint32_t i = 0; i < N; ++i) {
        uint32_t atmp = ext[i];
        for (uint32_t j = 0; j < DKS; ++j) {
            uint32_t a0 = atmp % BKS; atmp /= BKS;
            const uint16_t* row = c->ksk + (((size_t)i * BKS + a0) * DKS + j) * (n + 1);
            for (uint32_t k = 0; k < n; ++k) a[k] = (uint32_t)mod_sub(a[k], row[k], QKS);
            b = (uint32_t)mod_sub(b, row[n], QKS);
        }
    }
    for (uint32_t k = 0; k < n; ++k) out[k] = (uint16_t)orc_round_qQ_double(a[k], LQ, QKS);
    out[n] = (uint16_t)orc_round_qQ_double(b, LQ, QKS);
}

/* BinFHEScheme::EvalBinGate (non-XOR branch) */
void orc_eval_bin_gate(const orc_ctx* c, int gate, const uint16_t* ct1, const uint16_t* ct2, uint16_t* out) {
    uint16_t ext[N + 1];
    orc_bootstrap_extract(c, gate, ct1, ct2, ext);
    keyswitch_modswitch(c, ext, out);
}
/* BinFHEScheme::EvalNOT: a' = -a, b' = q/4 - b */
void orc_eval_not(const orc_ctx* c, const uint16_t* ct, uint16_t* out) {
    for (uint32_t k = 0; k < c->n; ++k) out[k] = (uint16_t)(ct[k] == 0 ? 0 : LQ - ct[k]);
    out[c->n] = (uint16_t)mod_sub(LQ >> 2, ct[c->n], LQ);
}
/* EvalConstant: trivial (noiseless) ciphertext a = 0, b = value * q/4 */
void orc_eval_const(const orc_ctx* c, int value, uint16_t* out) {
    memset(out, 0, (c->n + 1) * 2);
    out[c->n] = (uint16_t)((value & 1) * (LQ / HERDB_PT_MOD));
}
void orc_eval_gate_batch(const orc_ctx* c, int gate, const uint16_t* ct1, const uint16_t* ct2, uint16_t* out,
                         size_t batch, int threads) {
    size_t w = c->n + 1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    #pragma omp parallel for schedule(dynamic, 1)
    for (size_t i = 0; i < batch; ++i) orc_eval_bin_gate(c, gate, ct1 + i * w, ct2 + i * w, out + i * w);
}

/* ------------------------------------------------------------------ per-row circuit interpreter
 * The herd worker (not in the reference tree) walks Circuit::circuit_graph per row
 * (SURVEY.md 3.5); node kinds from nodes.hpp:12-43, operand order = DAG::parents()
 * = edge insertion order (dag.hpp:208-230), bit widths native_type_mapping.cpp:8-35. */
static const uint32_t TYPE_BITS[9] = { 1, 8, 16, 32, 64, 8, 16, 32, 64 };

typedef struct {
    uint32_t n_nodes;
    uint32_t* order;        /* topological order */
    uint32_t* par_off;      /* n_nodes + 1 */
    uint32_t* par;          /* parents, in edge order */
    uint32_t* in_off;       /* bit offset of every input field */
    uint32_t* out_off;      /* bit offset of every output field */
    uint32_t in_bits, out_bits;
} circ_plan;

static void plan_free(circ_plan* p) { free(p->order); free(p->par_off); free(p->par); free(p->in_off); free(p->out_off); }

static int plan_build(const herdb_circuit* c, circ_plan* p) {
    memset(p, 0, sizeof *p);
    uint32_t nn = (uint32_t)c->n_nodes;
    p->n_nodes = nn;
    p->order = malloc((nn + 1) * 4); p->par_off = calloc(nn + 2, 4); p->par = malloc((c->n_edges + 1) * 4);
    size_t n_in_fields = c->in_tuple_offsets[c->n_tuples];
    p->in_off = malloc((n_in_fields + 1) * 4); p->out_off = malloc((c->n_out_fields + 1) * 4);
    uint32_t off = 0;
    for (size_t f = 0; f < n_in_fields; ++f) { if (c->in_field_types[f] > 8) return -1; p->in_off[f] = off; off += TYPE_BITS[c->in_field_types[f]]; }
    p->in_bits = off; off = 0;
    for (size_t f = 0; f < c->n_out_fields; ++f) { if (c->out_field_types[f] > 8) return -1; p->out_off[f] = off; off += TYPE_BITS[c->out_field_types[f]]; }
    p->out_bits = off;
    for (size_t e = 0; e < c->n_edges; ++e) { if (c->edges[e].from >= nn || c->edges[e].to >= nn) return -1; p->par_off[c->edges[e].to + 1]++; }
    for (uint32_t i = 0; i < nn; ++i) p->par_off[i + 1] += p->par_off[i];
    uint32_t* fill = calloc(nn + 1, 4);
    for (size_t e = 0; e < c->n_edges; ++e) { uint32_t t = c->edges[e].to; p->par[p->par_off[t] + fill[t]++] = c->edges[e].from; }
    /* Kahn, smallest node id first */
    uint32_t* indeg = calloc(nn + 1, 4); uint8_t* done = calloc(nn + 1, 1);
    for (uint32_t i = 0; i < nn; ++i) indeg[i] = p->par_off[i + 1] - p->par_off[i];
    uint32_t cnt = 0;
    while (cnt < nn) {
        int progress = 0;
        for (uint32_t i = 0; i < nn; ++i) {
            if (done[i] || indeg[i]) continue;
            done[i] = 1; p->order[cnt++] = i; progress = 1;
            for (size_t e = 0; e < c->n_edges; ++e) if (c->edges[e].from == i) indeg[c->edges[e].to]--;
        }
        if (!progress) { free(fill); free(indeg); free(done); return -1; }   /* cycle */
    }
    free(fill); free(indeg); free(done);
    /* arity checks */
    for (uint32_t i = 0; i < nn; ++i) {
        uint32_t np = p->par_off[i + 1] - p->par_off[i];
        const herdb_node* nd = &c->nodes[i];
        uint32_t want = 0;
        if (nd->kind == HERDB_NODE_OUTPUT) want = 1;
        else if (nd->kind == HERDB_NODE_OPERATION)
            want = nd->op == HERDB_OP_NOT ? 1 : nd->op == HERDB_OP_MUX ? 3 : 2;
        if (np != want) return -1;
        if (nd->kind == HERDB_NODE_INPUT) {
            if (nd->tuple >= c->n_tuples) return -1;
            uint32_t f = c->in_tuple_offsets[nd->tuple] + nd->field;
            if (f >= c->in_tuple_offsets[nd->tuple + 1] || nd->bit >= TYPE_BITS[c->in_field_types[f]]) return -1;
        }
        if (nd->kind == HERDB_NODE_OUTPUT && (nd->field >= c->n_out_fields || nd->bit >= TYPE_BITS[c->out_field_types[nd->field]])) return -1;
    }
    return 0;
}

static void eval_row(const orc_ctx* c, const herdb_circuit* circ, const circ_plan* p, const uint16_t* row_in,
                     uint16_t* row_out, uint16_t* vals /* n_nodes * w */, uint16_t* tmp /* 4 * w */) {
    size_t w = c->n + 1;
    for (uint32_t oi = 0; oi < p->n_nodes; ++oi) {
        uint32_t id = p->order[oi];
        const herdb_node* nd = &circ->nodes[id];
        const uint32_t* par = p->par + p->par_off[id];
        uint16_t* v = vals + (size_t)id * w;
        switch (nd->kind) {
        case HERDB_NODE_INPUT:
            memcpy(v, row_in + (size_t)(p->in_off[circ->in_tuple_offsets[nd->tuple] + nd->field] + nd->bit) * w, w * 2);
            break;
        case HERDB_NODE_CONSTANT: orc_eval_const(c, nd->const_val, v); break;
        case HERDB_NODE_OUTPUT:
            memcpy(row_out + (size_t)(p->out_off[nd->field] + nd->bit) * w, vals + (size_t)par[0] * w, w * 2);
            break;
        case HERDB_NODE_OPERATION: {
            const uint16_t* x = vals + (size_t)par[0] * w;
            const uint16_t* y = nd->op == HERDB_OP_NOT ? NULL : vals + (size_t)par[1] * w;
            switch (nd->op) {
            case HERDB_OP_AND:  orc_eval_bin_gate(c, HERDB_GATE_AND, x, y, v); break;
            case HERDB_OP_OR:   orc_eval_bin_gate(c, HERDB_GATE_OR, x, y, v); break;
            case HERDB_OP_NAND: orc_eval_bin_gate(c, HERDB_GATE_NAND, x, y, v); break;
            case HERDB_OP_NOT:  orc_eval_not(c, x, v); break;
            case HERDB_OP_XOR:  /* OpenFHE v1.0.x EvalBinGate(XOR): OR(AND(a, !b), AND(!a, b)) */
                if (nd->flags & HERDB_NODE_FAST) { orc_eval_bin_gate(c, HERDB_GATE_XOR_FAST, x, y, v); break; }   /* BINGATE XOR_FAST */
                orc_eval_not(c, x, tmp); orc_eval_not(c, y, tmp + w);
                orc_eval_bin_gate(c, HERDB_GATE_AND, x, tmp + w, tmp + 2 * w);
                orc_eval_bin_gate(c, HERDB_GATE_AND, tmp, y, tmp + 3 * w);
                orc_eval_bin_gate(c, HERDB_GATE_OR, tmp + 2 * w, tmp + 3 * w, v);
                break;
            case HERDB_OP_MUX: { /* (sel, a, b): OR(AND(sel, a), AND(!sel, b)) */
                const uint16_t* z = vals + (size_t)par[2] * w;
                orc_eval_not(c, x, tmp);
                orc_eval_bin_gate(c, HERDB_GATE_AND, x, y, tmp + w);
                orc_eval_bin_gate(c, HERDB_GATE_AND, tmp, z, tmp + 2 * w);
                orc_eval_bin_gate(c, HERDB_GATE_OR, tmp + w, tmp + 2 * w, v);
                break; }
            }
            break; }
        }
    }
}

int orc_eval_circuit_rows(const orc_ctx* c, const herdb_circuit* circ, const uint16_t* rows_in, uint16_t* rows_out,
                          size_t n_rows, int threads) {
    circ_plan p;
    if (plan_build(circ, &p)) { plan_free(&p); return -1; }
    size_t w = c->n + 1;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    #pragma omp parallel
    {
        uint16_t* vals = malloc((size_t)p.n_nodes * w * 2 + 2);
        uint16_t* tmp = malloc(4 * w * 2);
        #pragma omp for schedule(dynamic, 1)
        for (size_t r = 0; r < n_rows; ++r)
            eval_row(c, circ, &p, rows_in + r * p.in_bits * w, rows_out + r * p.out_bits * w, vals, tmp);
        free(vals); free(tmp);
    }
    plan_free(&p);
    return 0;
}

int orc_eval_circuit_plain(const herdb_circuit* circ, const uint8_t* bits_in, uint8_t* bits_out, size_t n_rows) {
    circ_plan p;
    if (plan_build(circ, &p)) { plan_free(&p); return -1; }
    uint8_t* v = malloc(p.n_nodes + 1);
    for (size_t r = 0; r < n_rows; ++r) {
        const uint8_t* in = bits_in + r * p.in_bits; uint8_t* out = bits_out + r * p.out_bits;
        for (uint32_t oi = 0; oi < p.n_nodes; ++oi) {
            uint32_t id = p.order[oi];
            const herdb_node* nd = &circ->nodes[id];
            const uint32_t* par = p.par + p.par_off[id];
            switch (nd->kind) {
            case HERDB_NODE_INPUT: v[id] = in[p.in_off[circ->in_tuple_offsets[nd->tuple] + nd->field] + nd->bit] & 1; break;
            case HERDB_NODE_CONSTANT: v[id] = nd->const_val & 1; break;
            case HERDB_NODE_OUTPUT: out[p.out_off[nd->field] + nd->bit] = v[par[0]]; break;
            case HERDB_NODE_OPERATION:
                switch (nd->op) {
                case HERDB_OP_AND: v[id] = v[par[0]] & v[par[1]]; break;
                case HERDB_OP_OR: v[id] = v[par[0]] | v[par[1]]; break;
                case HERDB_OP_NAND: v[id] = !(v[par[0]] & v[par[1]]); break;
                case HERDB_OP_NOT: v[id] = !v[par[0]]; break;
                case HERDB_OP_XOR: v[id] = v[par[0]] ^ v[par[1]]; break;
                case HERDB_OP_MUX: v[id] = v[par[0]] ? v[par[1]] : v[par[2]]; break;
                }
                break;
            }
        }
    }
    free(v); plan_free(&p);
    return 0;
}
