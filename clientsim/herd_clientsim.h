/*
 * herd_clientsim.h -- synthetic stand-in for the herd client (libherd, which is not in the reference tree).
 * TEST / BENCH INFRASTRUCTURE: herdsman never sees secret keys; a real client generates keys and encrypts rows
 * with OpenFHE and uploads OpenFHE-serialised blobs (key_service.cpp:38, storage_service.cpp:133-135).  These
 * helpers produce keys and ciphertexts in the flat formats of include/herdb_params.h from a 64-bit seed with a
 * counter-based mixer (NOT a CSPRNG) so that tests and benchmarks have inputs.  Not part of libherdb.so.
 * All functions return 0 or -1 (bad argument).
 */
#ifndef HERD_CLIENTSIM_H
#define HERD_CLIENTSIM_H
#include <stddef.h>
#include <stdint.h>
#include "../include/herdb_params.h"
#ifdef __cplusplus
extern "C" {
#endif
size_t clientsim_key_blob_size(uint32_t n);
int clientsim_keygen(uint32_t n, uint64_t seed, int8_t* sk_lwe /* n */, int8_t* sk_ring /* N */, void* blob, size_t blob_len);
int clientsim_encrypt_bits(uint32_t n, const int8_t* sk_lwe, uint64_t seed, uint64_t first_ct_index,
                           const uint8_t* bits, size_t count, uint16_t* cts);
int clientsim_decrypt_bits(uint32_t n, const int8_t* sk_lwe, const uint16_t* cts, size_t count, uint8_t* bits);
#ifdef __cplusplus
}
#endif
#endif
