// modarith.cuh -- 32-bit lazy modular arithmetic over Z_Q, Q = 2^27 - 2^11 + 1, and the
// compile-time twiddle tables of the negacyclic NTT (N = 1024).
//
// Replaces (on the integer pipes of sm_100a) OpenFHE's NativeInteger ModMulFastConst /
// ModAddFast / ModSubFast used by binfhe's AddToAccCGGI.
// Everything here is exact: values are kept as lazily reduced residues in u32 and every
// function documents the range it accepts and returns.
//
// Host/device: the same source is compiled by g++ for the CPU emulation test
// (tests/emu) and by nvcc for the kernels.
#pragma once
#include <stdint.h>
#include "../../include/herdb_params.h"

#if defined(__CUDACC__)
#define HERDB_HD __host__ __device__ __forceinline__
#define HERDB_D __device__ __forceinline__
#else
#define HERDB_HD inline
#define HERDB_D inline
#endif

namespace herdb {

typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;

constexpr u32 kQ = HERDB_Q;
constexpr u32 kQ2 = 2u * HERDB_Q;
constexpr u32 kN = HERDB_RING_N;
constexpr u32 kLogN = HERDB_RING_LOGN;
constexpr u32 kDG = HERDB_DIGITS_G;
constexpr u32 kDG2 = 2u * HERDB_DIGITS_G;
constexpr u32 kLweQ = HERDB_LWE_Q;
constexpr u32 kQKS = HERDB_QKS;
constexpr u32 kBaseKS = 1u << HERDB_BASEKS_LOG;
constexpr u32 kDKS = HERDB_DIGITS_KS;
constexpr u32 kPsi = 282116u;            // minimum primitive 2N-th root of unity mod Q (checked in tests)

static_assert(kQ == (1u << 27) - (1u << 11) + 1u, "kernels are specialised for Q = 2^27 - 2^11 + 1");
static_assert(kN == 1024 && kLweQ == 1024, "kernels are specialised for N = q = 1024");

// ---------------------------------------------------------------- constexpr helpers (host side / compile time)
constexpr u32 cmulmod(u64 a, u64 b) { return (u32)((a * b) % kQ); }
constexpr u32 cpowmod(u64 b, u64 e) {
    u64 r = 1; b %= kQ;
    while (e) { if (e & 1) r = (r * b) % kQ; b = (b * b) % kQ; e >>= 1; }
    return (u32)r;
}
constexpr u32 cbitrev10(u32 x) {
    u32 r = 0;
    for (int i = 0; i < 10; ++i) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}
constexpr u32 cshoup(u32 w) { return (u32)(((u64)w << 32) / kQ); }
constexpr u32 kPsiInv = cpowmod(kPsi, 2 * kN - 1);
constexpr u32 kNInv = cpowmod(kN, kQ - 2);
constexpr u32 kNInvShoup = cshoup(kNInv);
static_assert(cpowmod(kPsi, kN) == kQ - 1, "psi must be a primitive 2N-th root");
static_assert(cmulmod(kNInv, kN) == 1, "N^-1");
// forward twiddle k: psi^bitrev(k); inverse twiddle k: psi^-bitrev(k)
constexpr u32 ctw_fwd(u32 k) { return cpowmod(kPsi, cbitrev10(k)); }
constexpr u32 ctw_inv(u32 k) { return cpowmod(kPsiInv, cbitrev10(k)); }

// Top-digit elimination (br_core.cuh, "phase B"): the signed digits of a coefficient satisfy
// t = r0 + B*r1 + B^2*r2 + B^3*r3 (mod Q), B = 2^7, so by linearity of the NTT
//   NTT(r3) = B^-3 * (NTT(t) - NTT(r0) - B*NTT(r1) - B^2*NTT(r2))
// and NTT(t) is the accumulator itself.  Substituting into sum_l NTT(r_l) * K_l gives the same element
// of Z_Q from three digit transforms and the accumulator, against key rows folded once at load time:
//   K'_l = K_l - B^(l-3) * K_3  (l = 0,1,2),   K'_3 = B^-3 * K_3 (multiplies the accumulator).
constexpr u32 kBG = 1u << HERDB_BASEG_LOG;
constexpr u32 kBGInv1 = cpowmod(kBG, kQ - 2);
constexpr u32 kBGInv2 = cmulmod(kBGInv1, kBGInv1);
constexpr u32 kBGInv3 = cmulmod(kBGInv2, kBGInv1);
static_assert(cmulmod(kBGInv3, cmulmod(cmulmod(kBG, kBG), kBG)) == 1, "B^-3");
static_assert(HERDB_DIGITS_G == 4, "top-digit elimination is written for 4 gadget digits");
constexpr u32 kDGT = HERDB_DIGITS_G - 1;   // digit polynomials actually transformed per accumulator component

// Stages whose twiddle depends only on the register index (not on the lane) read their
// (w, w') pairs from constant memory: the operand folds into IMAD/IMAD.HI as c[bank][imm].
struct UniformTw { u32 v[32][2]; };
constexpr UniformTw make_uniform_fwd() {
    UniformTw t{};
    for (u32 k = 1; k < 32; ++k) { t.v[k][0] = ctw_fwd(k); t.v[k][1] = cshoup(t.v[k][0]); }
    return t;
}
constexpr UniformTw make_uniform_fwd_up() {      // ceiling companions ceil(w * 2^32 / Q) = w' + 1 (centred butterflies, ntt_core.cuh)
    UniformTw t{};
    for (u32 k = 1; k < 32; ++k) { t.v[k][0] = ctw_fwd(k); t.v[k][1] = cshoup(t.v[k][0]) + 1u; }
    return t;
}
constexpr UniformTw make_uniform_inv() {
    UniformTw t{};
    for (u32 k = 1; k < 32; ++k) { t.v[k][0] = ctw_inv(k); t.v[k][1] = cshoup(t.v[k][0]); }
    return t;
}
#if defined(__CUDACC__)
__constant__ UniformTw c_tw_fwd = make_uniform_fwd();
__constant__ UniformTw c_tw_inv = make_uniform_inv();
__constant__ UniformTw c_tw_fwd_up = make_uniform_fwd_up();
#endif
static constexpr UniformTw h_tw_fwd = make_uniform_fwd();
static constexpr UniformTw h_tw_inv = make_uniform_inv();
static constexpr UniformTw h_tw_fwd_up = make_uniform_fwd_up();

#if defined(__CUDA_ARCH__)
#define HERDB_TWF(k, i) c_tw_fwd.v[k][i]
#define HERDB_TWI(k, i) c_tw_inv.v[k][i]
#define HERDB_TWFU(k, i) c_tw_fwd_up.v[k][i]
#else
#define HERDB_TWFU(k, i) h_tw_fwd_up.v[k][i]
#define HERDB_TWF(k, i) h_tw_fwd.v[k][i]
#define HERDB_TWI(k, i) h_tw_inv.v[k][i]
#endif

// ---------------------------------------------------------------- run-time primitives
HERDB_HD u32 mulhi32(u32 a, u32 b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (u32)(((u64)a * b) >> 32);
#endif
}
HERDB_HD u32 umin32(u32 a, u32 b) { return a < b ? a : b; }

// w*y mod Q, lazily: any y < 2^32, w < Q, wp = floor(w*2^32/Q); result in [0, 2Q)
HERDB_HD u32 mul_shoup_lazy(u32 w, u32 wp, u32 y) {
    u32 q = mulhi32(wp, y);
    return w * y - q * kQ;
}
// x >= m ? x - m : x   (for x < 2m): unsigned wrap makes x-m huge when x < m
HERDB_HD u32 csub(u32 x, u32 m) { return umin32(x, x - m); }
// any x < 2^32 -> canonical residue in [0, Q): x - (x>>27)*Q lands in [0, 2Q) because
// x - (x>>27)*Q = (x mod 2^27) + 2047*(x>>27) < 2^27 + 2047*31 < 2Q
HERDB_HD u32 canon32(u32 x) { return csub(x - (x >> 27) * kQ, kQ); }
// Fold of a 62-bit sum with the quotient taken on the FP64 pipe, which is otherwise idle (DFMA issues beside IMAD on
// B200: profiles/round1/r01m_fp64_microbench.txt).  For x < 2^62 the register pair {lo = (x >> 11) mod 2^32,
// hi = 0x43300000 | (x >> 43)} read as a double is exactly 2^52 + y, y = floor(x / 2^11) < 2^51.  With the
// integer ci = round(2^63 / Q) (37 bits), M = ci / 2^52 and C = 1.5 * 2^52 - ci - 1 (both exact doubles)
//   fma(pair, M, C) = 1.5 * 2^52 + y * ci / 2^52 - 1, rounded once to an integer: its low word is q - 1
//   (mod 2^32) with |q - x/Q| <= 0.5 (rounding) + 0.25 (ci) + 2^-16 (the 11 dropped bits), so
//   r = lo32(x) - (q - 1) * Q (mod 2^32) = x - q * Q + Q exactly, in (0.2499 Q, 1.7501 Q).
// Two shifts, one DFMA and one IMAD replace shift + mask + IMAD.HI + 2 IMAD + add; the residue class is exact,
// only the representative differs from a Shoup fold's.  fold62_f64(x, a) = r + a for an addend a < 2^30.
HERDB_HD double f64_words(u32 hi, u32 lo) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double((int)hi, (int)lo);
#else
    u64 b = ((u64)hi << 32) | lo; double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}
HERDB_HD u32 f64_low(double d) {
#if defined(__CUDA_ARCH__)
    return (u32)__double2loint(d);
#else
    u64 b; __builtin_memcpy(&b, &d, 8); return (u32)b;
#endif
}
HERDB_HD double f64_fma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
constexpr double kTwo52d = 4503599627370496.0;
constexpr u64 kFoldCi = ((1ull << 63) + kQ / 2) / kQ;               // round(2^63 / Q)
static_assert(kFoldCi < (1ull << 37), "ci = round(2^63 / Q) has 37 bits");
constexpr double kFoldMd = (double)kFoldCi / kTwo52d;
constexpr double kFoldCd = 1.5 * kTwo52d - (double)kFoldCi - 1.0;
HERDB_HD u32 fold62_f64(u64 x, u32 addend = 0) {     // x < 2^62 -> addend + residue in (0.2499 Q, 1.7501 Q)
    const double d = f64_words(0x43300000u | (u32)(x >> 43), (u32)(x >> 11));
    const u32 q1 = f64_low(f64_fma(d, kFoldMd, kFoldCd));
    return (u32)x + addend - q1 * kQ;
}
// exact a*b mod Q for canonical a, b (key load only; 64-bit remainder)
HERDB_HD u32 mulmod_slow(u32 a, u32 b) { return (u32)(((u64)a * b) % kQ); }
// Folds the four gadget rows of one (LWE index, sign, accumulator component, column, evaluation point):
// in  k[l] = K_l / N   (canonical; what the forward transform of the key load produces)
// out k[l] = B^-l * (K_l - B^(l-3) K_3) / N for l < 3,  k[3] = B^-3 * K_3  (NOT divided by N: it multiplies the
//     stored accumulator, which already carries the 1/N).  The B^-l is there because phase A transforms the
//     digit fields in place, B^l * r_l (ntt_core.cuh, fwd_pass1_fields_multi): no shift in the digit extraction.
HERDB_HD void fold_top_digit(u32 (&k)[4]) {
    const u32 t = k[3];
    k[0] = csub(k[0] + kQ - mulmod_slow(t, kBGInv3), kQ);
    k[1] = mulmod_slow(csub(k[1] + kQ - mulmod_slow(t, kBGInv2), kQ), kBGInv1);
    k[2] = mulmod_slow(csub(k[2] + kQ - mulmod_slow(t, kBGInv1), kQ), kBGInv2);
    k[3] = mulmod_slow(mulmod_slow(t, kN % kQ), kBGInv3);
}
// Position of evaluation point x inside a bootstrapping-key row: the four points of an aligned group are stored
// rotated by one word (point 4g + p at word 4g + (p + 3) % 4).  A thread of phase B loads its four key words and
// its four operand words with one 128-bit load each, i.e. into two register quads, and multiplies point p of one
// with point p of the other: with both quads in natural order the two multiplicands of every IMAD.WIDE would have
// the same register index modulo 4 (same register bank, next to the 64-bit accumulator pair that spans both);
// the rotation puts them in different banks.
HERDB_HD u32 key_pos(u32 x) { return (x & ~3u) | ((x + 3u) & 3u); }
// bank-conflict-free layout of a 1024-word polynomial in shared memory: word x lives at
// swz(x); rows of 32 words, the 16-byte chunk index is XORed with the row index
HERDB_HD u32 swz(u32 x) { return x ^ (((x >> 5) & 7u) << 2); }

// LWE mod switch Q -> qKS = 2^14 (OpenFHE RoundqQ, integer form): floor((2*v*qKS + Q) / 2Q) mod qKS
HERDB_HD u32 modswitch_Q_to_qKS(u32 v) {
    return (u32)((((u64)v << 15) + kQ) / (2ull * kQ)) & (kQKS - 1);
}
// qKS -> q = 2^10: floor(v/16 + 0.5) mod q
HERDB_HD u32 modswitch_qKS_to_q(u32 v) { return ((v + 8u) >> 4) & (kLweQ - 1); }

struct alignas(16) V4 { u32 v[4]; };

}  // namespace herdb
