/*
 * oracle.h -- CPU oracle for the gate-evaluation hot path.  TEST INFRASTRUCTURE:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (libherdb.so) never does.
 *
 * PARITY UNPINNED: the arithmetic herdsman depends on lives in OpenFHE binfhe
 * v1.0.3/v1.0.4 (un-vendored; /root/reference/CMakeLists.txt:73,210-211,
 * common/action/install-openfhe/action.yml:9-12) and in the herd worker
 * program, neither of which is under /root/reference, and the reference has no
 * tests or golden vectors (test/CMakeLists.txt:20 has an empty source list).
 * This file restates OpenFHE v1.0.x's published GINX/CGGI algorithm
 * (binfhe/lib/{binfhe-base-scheme,lwe-pke,rgsw-acc,rgsw-acc-cggi}.cpp,
 * core math: ChineseRemainderTransformFTT) function by function; it is validated
 * semantically (decrypt == plaintext gate, algebraic identities) and against a
 * second, structurally different restatement (numpy, schoolbook products, no NTT,
 * OpenFHE's double-precision rounding: tests/test_oracle_independent.py), not
 * against OpenFHE outputs.
 */
#ifndef HERDB_ORACLE_H
#define HERDB_ORACLE_H

#include <stddef.h>
#include <stdint.h>
#include "../include/herdb.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_ctx orc_ctx;

/* server side (what the herd worker + OpenFHE do) */
orc_ctx* orc_create(void);
void     orc_destroy(orc_ctx*);
int      orc_load_key(orc_ctx*, const void* blob, size_t len);      /* 0 or negative */
uint32_t orc_lwe_n(const orc_ctx*);
void     orc_eval_not(const orc_ctx*, const uint16_t* ct, uint16_t* out);
void     orc_eval_const(const orc_ctx*, int value, uint16_t* out);
void     orc_eval_bin_gate(const orc_ctx*, int gate, const uint16_t* ct1, const uint16_t* ct2, uint16_t* out);
void     orc_eval_gate_batch(const orc_ctx*, int gate, const uint16_t* ct1, const uint16_t* ct2,
                             uint16_t* out, size_t batch, int threads);
/* per-row circuit interpreter (the inferred herd worker loop) */
int      orc_eval_circuit_rows(const orc_ctx*, const herdb_circuit*, const uint16_t* rows_in,
                               uint16_t* rows_out, size_t n_rows, int threads);
/* plaintext evaluation of the same circuit on packed bits (one byte per bit) */
int      orc_eval_circuit_plain(const herdb_circuit*, const uint8_t* bits_in, uint8_t* bits_out, size_t n_rows);

/* introspection for kernel debugging / unit tests */
uint32_t orc_psi(const orc_ctx*);
void     orc_ntt_forward(const orc_ctx*, uint32_t* poly /* N, in place, natural -> bit-reversed */);
void     orc_ntt_inverse(const orc_ctx*, uint32_t* poly /* N, in place, bit-reversed -> natural */);
void     orc_negacyclic_mul_schoolbook(const uint32_t* a, const uint32_t* b, uint32_t* out);
/* accumulator (coefficient domain, 2*N words: a then b) after `steps` CMUX steps of the
 * bootstrap of ct1+ct2 under `gate`; steps = n gives the final accumulator */
void     orc_blind_rotate_trace(const orc_ctx*, int gate, const uint16_t* ct1, const uint16_t* ct2,
                                uint32_t steps, uint32_t* acc_coeff);
/* extracted + mod-switched (Q -> qKS) ciphertext before key switching: N+1 u16 */
void     orc_bootstrap_extract(const orc_ctx*, int gate, const uint16_t* ct1, const uint16_t* ct2, uint16_t* ext);
uint32_t orc_round_qQ_double(uint32_t v, uint32_t q, uint32_t Q);    /* OpenFHE RoundqQ, double */
uint32_t orc_round_qQ_int(uint32_t v, uint32_t q, uint32_t Q);       /* integer form used on GPU */
uint64_t orc_round_mismatches(uint32_t q, uint32_t Q, uint32_t start, uint32_t count);
void     orc_signed_digit_decompose(uint32_t t, int32_t digits[HERDB_DIGITS_G]);

/* client side (what libherd does; synthetic) */
size_t   orc_key_blob_size(uint32_t n);
int      orc_keygen(uint32_t n, uint64_t seed, int8_t* sk_lwe, int8_t* sk_ring, void* blob, size_t blob_len);
void     orc_encrypt_bits(uint32_t n, const int8_t* sk_lwe, uint64_t seed, uint64_t first_ct_index,
                          const uint8_t* bits, size_t count, uint16_t* cts);
void     orc_decrypt_bits(uint32_t n, const int8_t* sk_lwe, const uint16_t* cts, size_t count, uint8_t* bits);
/* decryption noise of one ciphertext relative to message bit m, centred, in units of q */
int32_t  orc_noise(uint32_t n, const int8_t* sk_lwe, const uint16_t* ct, int bit);
uint64_t orc_rng(uint64_t seed, uint64_t stream, uint64_t idx);

#ifdef __cplusplus
}
#endif
#endif
