/* Minimal stand-in for libuuid's <uuid/uuid.h> (absent from this image): just what herd::common::UUID uses
 * (common/src/uuid.cpp).  TEST INFRASTRUCTURE for tests/adapter. */
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
typedef unsigned char uuid_t[16];
#define UUID_STR_LEN 37
static inline void uuid_generate(uuid_t out) { for (int i = 0; i < 16; ++i) out[i] = (unsigned char)(rand() & 0xFF); out[6] = (out[6] & 0x0F) | 0x40; out[8] = (out[8] & 0x3F) | 0x80; }
static inline int uuid_parse(const char* in, uuid_t uu) {
    if (strlen(in) != 36) return -1;
    int k = 0;
    for (int i = 0; i < 36; ++i) {
        if (i == 8 || i == 13 || i == 18 || i == 23) { if (in[i] != '-') return -1; continue; }
        unsigned v;
        const char c = in[i];
        if (c >= '0' && c <= '9') v = (unsigned)(c - '0'); else if (c >= 'a' && c <= 'f') v = (unsigned)(c - 'a' + 10);
        else if (c >= 'A' && c <= 'F') v = (unsigned)(c - 'A' + 10); else return -1;
        if (k & 1) uu[k >> 1] = (unsigned char)(uu[k >> 1] | v); else uu[k >> 1] = (unsigned char)(v << 4);
        ++k;
    }
    return 0;
}
static inline void uuid_unparse(const uuid_t uu, char* out) {
    static const char* hex = "0123456789abcdef";
    int p = 0;
    for (int i = 0; i < 16; ++i) {
        if (i == 4 || i == 6 || i == 8 || i == 10) out[p++] = '-';
        out[p++] = hex[uu[i] >> 4]; out[p++] = hex[uu[i] & 15];
    }
}
static inline int uuid_compare(const uuid_t a, const uuid_t b) { return memcmp(a, b, 16); }
