// ntt_core.cuh -- register-resident negacyclic NTT of one 1024-coefficient polynomial by
// one warp: 32 coefficients per lane, two passes of five radix-2 stages each, one
// shared-memory transposition between the passes (swizzled, bank-conflict free).
//
// Replaces OpenFHE's ChineseRemainderTransformFTT<NativeVector>::ForwardTransformToBitReverse
// / InverseTransformFromBitReverse.  Forward = Cooley-Tukey, natural in,
// bit-reversed out; inverse = Gentleman-Sande, bit-reversed in, natural out.  The 1/N of
// the inverse is NOT applied here: it is folded into the bootstrapping key at load time.
//
// Lane layouts (T = lane id):
//   "strided"    x[k] <-> coefficient T + 32*k   (stages with stride >= 32: twiddle depends on k only)
//   "contiguous" x[e] <-> coefficient 32*T + e   (stages with stride <= 16: per-lane twiddles)
#pragma once
#include "modarith.cuh"

namespace herdb {

// per-lane twiddle tables, [16 chunks][32 lanes][4 words]: chunk g of lane T holds the
// (w, w') pairs 2g and 2g+1 of that lane's 31-pair list (pair 31 is padding)
constexpr u32 kLaneTwWords = 16 * 32 * 4;
// forward pass 2 list order: stage s=5 (1 pair), s=6 (2), s=7 (4), s=8 (8), s=9 (16)
// inverse pass 1 list order: t=1 (16 pairs), t=2 (8), t=4 (4), t=8 (2), t=16 (1)
inline void build_lane_twiddles(u32* fwd2, u32* inv1) {
    for (u32 T = 0; T < 32; ++T) {
        u32 fl[32][2] = {}, il[32][2] = {};
        u32 p = 0;
        for (u32 s = 5; s < 10; ++s) {
            u32 t = 16u >> (s - 5);
            for (u32 i = 0; i < 16 / t; ++i, ++p) {
                fl[p][0] = ctw_fwd((1u << s) + T * (16 / t) + i);
                fl[p][1] = cshoup(fl[p][0]);
            }
        }
        p = 0;
        for (u32 t = 1; t <= 16; t <<= 1)
            for (u32 i = 0; i < 16 / t; ++i, ++p) {
                il[p][0] = ctw_inv(512 / t + T * (16 / t) + i);
                il[p][1] = cshoup(il[p][0]);
            }
        for (u32 g = 0; g < 16; ++g)
            for (u32 w = 0; w < 4; ++w) {
                fwd2[(g * 32 + T) * 4 + w] = fl[2 * g + (w >> 1)][w & 1];
                inv1[(g * 32 + T) * 4 + w] = il[2 * g + (w >> 1)][w & 1];
            }
    }
}

// ---------------------------------------------------------------- forward (CT), lazy: no corrections.
// Input < B  ->  output < B + 2Q per stage; callers start below 2Q and end below 22Q < 2^32.
HERDB_HD void fwd_pass1(u32 (&x)[32]) {           // strided layout, stages 0..4 (strides 512..32)
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int h = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = (1 << s) + k / (2 * h);
            u32 v = mul_shoup_lazy(HERDB_TWF(idx, 0), HERDB_TWF(idx, 1), x[k + h]);
            u32 u = x[k];
            x[k] = u + v;
            x[k + h] = u - v + kQ2;
        }
    }
}
HERDB_HD void fwd_pass2(u32 (&x)[32], const u32 (&tw)[62]) {   // contiguous layout, stages 5..9
    int off = 0;
#pragma unroll
    for (int s = 5; s < 10; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            u32 v = mul_shoup_lazy(tw[2 * p], tw[2 * p + 1], x[e + t]);
            u32 u = x[e];
            x[e] = u + v;
            x[e + t] = u - v + kQ2;
        }
        off += 16 / t;
    }
}

// forward pass 2 with the lane twiddles streamed from the chunked table (no twiddle registers)
HERDB_HD void fwd_pass2_stream(u32 (&x)[32], const u32* lane_tw, u32 lane) {
    int off = 0;
#pragma unroll
    for (int s = 5; s < 10; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const u32* tp = lane_tw + ((p >> 1) * 32 + lane) * 4 + (p & 1) * 2;
            u32 v = mul_shoup_lazy(tp[0], tp[1], x[e + t]);
            u32 u = x[e];
            x[e] = u + v;
            x[e + t] = u - v + kQ2;
        }
        off += 16 / t;
    }
}

// M digit polynomials of the same accumulator component are transformed together by one warp: M independent
// butterflies per twiddle give the scheduler M times the instruction-level parallelism (a warp shares its
// scheduler with only one other warp) and every streamed twiddle pair is loaded once for the M of them.
// Digit l is given IN PLACE: x[l][k] = (r_l + 64) * 2^(7l), the 7-bit field of the decomposed value
// masked but not shifted down.  The transform is linear, so the warp computes 2^(7l) * NTT(r_l); the factor
// 2^(-7l) sits in the folded key rows (fold_top_digit).  Saves the shift of every digit extraction; the offset
// 64 * 2^(7l) goes into the add constants of stage 0 as before.  Inputs < 2^21, outputs < 3Q + 2^21 + 8Q.
// Centred forward butterfly, four instructions instead of five.  With the CEILING companion w'' = ceil(w * 2^32 / Q)
// = w' + 1 the quotient q = mulhi(w'', y) is floor(w*y/Q) or one more (w''*y/2^32 - w*y/Q lies in [0, 1) for every
// y < 2^32), so the Shoup remainder rho = w*y - q*Q lies in [-Q, Q) instead of [0, 2Q) and the difference needs
// no offset:  x' = (w*y + u) - q*Q = u + rho  (two IMADs, u as the addend of the first),  y' = u + u - x' = u - rho
// (ONE three-input add).  Every stage widens the range of the values by Q on either side; stage 0 of the digit
// transforms starts them at 11 Q (add constants), so after ten stages they lie in [Q, 22 Q + 2^21): non-negative,
// below 2^32, same upper bound as the [0, 2Q) form.  Ceiling companions: c_tw_fwd_up for the uniform stages, the
// lane table in shared memory gets its +1 when the CTA copies it (br_core.cuh).
HERDB_HD void bfly_fwd_centred(u32& a, u32& b, u32 w, u32 wpu) {
    const u32 q = mulhi32(wpu, b);
    const u32 xn = w * b + a - q * kQ;
    b = a + a - xn;
    a = xn;
}
constexpr u32 kFwdStart = 11u * kQ;      // offset of the values after stage 0 (a multiple of Q: invisible mod Q)
// ranges of a digit transform: after stage 0 in [kFwdStart - Q, kFwdStart + 2Q + 2^21), nine more stages of +-Q
constexpr u64 kFwdOutMin = (u64)kFwdStart - 10ull * kQ;
constexpr u64 kFwdOutMax = (u64)kFwdStart + 11ull * kQ + (1ull << 21);       // exclusive
static_assert(kFwdStart >= 10ull * kQ && kFwdOutMax < (1ull << 32), "digit transforms stay non-negative and below 2^32");
constexpr u32 dig_field_c(int l) { return kQ - (u32)((64ull << (7 * l)) % kQ); }
constexpr u32 dig_field_k1(int l) { return (dig_field_c(l) + cmulmod(ctw_fwd(1), dig_field_c(l))) % kQ; }
constexpr u32 dig_field_k2(int l) { return (dig_field_c(l) + kQ - cmulmod(ctw_fwd(1), dig_field_c(l))) % kQ + kQ2; }
constexpr u32 kDigF1_0 = dig_field_k1(0), kDigF1_1 = dig_field_k1(1), kDigF1_2 = dig_field_k1(2);
constexpr u32 kDigF2_0 = dig_field_k2(0), kDigF2_1 = dig_field_k2(1), kDigF2_2 = dig_field_k2(2);
template <int M>
HERDB_HD void fwd_pass1_fields_multi(u32 (&x)[M][32]) {
    static_assert(M <= 3, "fields of digits 0..2");
#pragma unroll
    for (int k = 0; k < 16; ++k)
#pragma unroll
        for (int m = 0; m < M; ++m) {
            const u32 k1 = m == 0 ? kDigF1_0 : m == 1 ? kDigF1_1 : kDigF1_2;
            const u32 k2 = m == 0 ? kDigF2_0 : m == 1 ? kDigF2_1 : kDigF2_2;
            u32 v = mul_shoup_lazy(HERDB_TWF(1, 0), HERDB_TWF(1, 1), x[m][k + 16]);
            u32 u = x[m][k];
            x[m][k] = u + v + (k1 + kFwdStart - kQ);          // v - Q in [-Q, Q): values in [10 Q, 13 Q + 2^21)
            x[m][k + 16] = u - v + (k2 - kQ2 + kFwdStart + kQ);
        }
#pragma unroll
    for (int s = 1; s < 5; ++s) {
        const int h = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = (1 << s) + k / (2 * h);
#pragma unroll
            for (int m = 0; m < M; ++m) {
                bfly_fwd_centred(x[m][k], x[m][k + h], HERDB_TWFU(idx, 0), HERDB_TWFU(idx, 1));
            }
        }
    }
}
// (in the centred form `lane_tw` must hold the CEILING companions)
template <int M>
HERDB_HD void fwd_pass2_stream_multi(u32 (&x)[M][32], const u32* lane_tw, u32 lane) {
    int off = 0;
#pragma unroll
    for (int s = 5; s < 10; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const u32* tp = lane_tw + ((p >> 1) * 32 + lane) * 4 + (p & 1) * 2;
            const u32 w = tp[0], wp = tp[1];
#pragma unroll
            for (int m = 0; m < M; ++m) {
                bfly_fwd_centred(x[m][e], x[m][e + t], w, wp);
            }
        }
        off += 16 / t;
    }
}

// ---------------------------------------------------------------- inverse (GS), statically lazy
// Every value carries a compile-time bound b (in units of 2Q: value < b * 2Q).  A butterfly (u, v) -> (u + v,
// w * (u - v + b_v * 2Q)) gives the sum the bound b_u + b_v and RESETS the product to 1 (Shoup takes any
// multiplicand below 2^32 and returns [0, 2Q)).  A conditional subtraction csub(x, m * 2Q), m = ceil(b / 2),
// halves a bound and is spent only where the next sum would pass 16 units (32 Q < 2^32) or where the caller's
// output bound demands it: an element that becomes a product before its bound matters never pays.  Everything
// is unrolled, so the bounds are constants and only the needed corrections are emitted: 16 + 36 per
// polynomial for the two passes (bounds 1 -> 2 -> 1) instead of 80 + 80 with the invariant [0, 2Q).
constexpr int kInvBoundMax = 16;
HERDB_HD void inv_reduce_to(u32& x, int& b, int lim) {        // at most four halvings: 16 -> 1
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (b > lim) { const int m = (b + 1) / 2; x = csub(x, (u32)m * kQ2); b = m; }
}
template <int B0, int BOUT>
HERDB_HD void inv_pass1(u32 (&x)[32], const u32* lane_tw /* smem or global, chunked [g][32][4] */, u32 lane) {
    int b[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) b[e] = B0;
    int off = 0;
#pragma unroll
    for (int t = 1; t <= 16; t <<= 1) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const u32* tp = lane_tw + ((p >> 1) * 32 + lane) * 4 + (p & 1) * 2;
            u32 u = x[e], v = x[e + t];
            x[e] = u + v;
            x[e + t] = mul_shoup_lazy(tp[0], tp[1], u - v + (u32)b[e + t] * kQ2);
            b[e] += b[e + t];
            b[e + t] = 1;
            inv_reduce_to(x[e], b[e], t == 16 ? BOUT : kInvBoundMax / 2);
        }
        off += 16 / t;
    }
}
template <int B0, int BOUT>
HERDB_HD void inv_pass2(u32 (&x)[32]) {           // strided layout, strides 32..512
    int b[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) b[k] = B0;
#pragma unroll
    for (int h = 1; h <= 16; h <<= 1) {
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = 16 / h + k / (2 * h);
            u32 u = x[k], v = x[k + h];
            x[k] = u + v;
            x[k + h] = mul_shoup_lazy(HERDB_TWI(idx, 0), HERDB_TWI(idx, 1), u - v + (u32)b[k + h] * kQ2);
            b[k] += b[k + h];
            b[k + h] = 1;
            inv_reduce_to(x[k], b[k], h == 16 ? BOUT : kInvBoundMax / 2);
        }
    }
}

// ---------------------------------------------------------------- shared-memory movement (one polynomial = 1024 words)
HERDB_HD void load_contig(const u32* poly, u32 lane, u32 (&x)[32]) {      // 8 x 128-bit, conflict free
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        V4 q = *reinterpret_cast<const V4*>(poly + 32 * lane + 4 * (c ^ (lane & 7)));
        x[4 * c + 0] = q.v[0]; x[4 * c + 1] = q.v[1]; x[4 * c + 2] = q.v[2]; x[4 * c + 3] = q.v[3];
    }
}
HERDB_HD void store_contig(u32* poly, u32 lane, const u32 (&x)[32]) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        V4 q;
        q.v[0] = x[4 * c + 0]; q.v[1] = x[4 * c + 1]; q.v[2] = x[4 * c + 2]; q.v[3] = x[4 * c + 3];
        *reinterpret_cast<V4*>(poly + 32 * lane + 4 * (c ^ (lane & 7))) = q;
    }
}
HERDB_HD void load_strided(const u32* poly, u32 lane, u32 (&x)[32]) {     // 32 x 32-bit, conflict free
#pragma unroll
    for (int k = 0; k < 32; ++k) x[k] = poly[32 * k + ((((lane >> 2) ^ (k & 7)) << 2) | (lane & 3))];
}
HERDB_HD void store_strided(u32* poly, u32 lane, const u32 (&x)[32]) {
#pragma unroll
    for (int k = 0; k < 32; ++k) poly[32 * k + ((((lane >> 2) ^ (k & 7)) << 2) | (lane & 3))] = x[k];
}

}  // namespace herdb
