// host_tables.h -- host-side construction of the constant tables the kernels read:
// lane twiddles (ntt_core.cuh) and the monomial table NTT(X^(2*row) - 1), row < 1024.
// OpenFHE precomputes the same monomials in RingGSWCryptoParams::PreCompute (m_monomials);
// here only the even exponents are needed because 2N/q = 2.
#pragma once
#include <vector>
#include "ntt_core.cuh"

namespace herdb {

// evaluation point of output position x of the forward NTT: psi^(2*bitrev(x)+1)
inline void build_monomial_table(std::vector<u32>& mono) {
    mono.resize((size_t)kN * kN);
    std::vector<u32> pw(2 * kN);
    pw[0] = 1;
    for (u32 j = 1; j < 2 * kN; ++j) pw[j] = cmulmod(pw[j - 1], kPsi);
    for (u32 row = 0; row < kN; ++row)
        for (u32 x = 0; x < kN; ++x) {
            u32 e = (2 * cbitrev10(x) + 1) * (2 * row) % (2 * kN);
            mono[(size_t)row * kN + x] = (pw[e] + kQ - 1) % kQ;
        }
}

// reference (sequential) forward NTT in the kernels' convention: natural in, bit-reversed out
inline void ntt_forward_host(u32* a) {
    u32 t = kN;
    for (u32 m = 1; m < kN; m <<= 1) {
        t >>= 1;
        for (u32 i = 0; i < m; ++i) {
            u32 w = ctw_fwd(m + i);
            for (u32 j = 2 * i * t; j < 2 * i * t + t; ++j) {
                u32 u = a[j], v = cmulmod(a[j + t], w);
                a[j] = (u + v) % kQ;
                a[j + t] = (u + kQ - v) % kQ;
            }
        }
    }
}

}  // namespace herdb
