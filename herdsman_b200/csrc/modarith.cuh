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

// 2^31 mod Q and its Shoup companion: folds the high part of a 62-bit value
constexpr u32 kR31 = (u32)((1ull << 31) % kQ);
constexpr u32 kR31s = cshoup(kR31);

// Stages whose twiddle depends only on the register index (not on the lane) read their
// (w, w') pairs from constant memory: the operand folds into IMAD/IMAD.HI as c[bank][imm].
struct UniformTw { u32 v[32][2]; };
constexpr UniformTw make_uniform_fwd() {
    UniformTw t{};
    for (u32 k = 1; k < 32; ++k) { t.v[k][0] = ctw_fwd(k); t.v[k][1] = cshoup(t.v[k][0]); }
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
#endif
static constexpr UniformTw h_tw_fwd = make_uniform_fwd();
static constexpr UniformTw h_tw_inv = make_uniform_inv();

#if defined(__CUDA_ARCH__)
#define HERDB_TWF(k, i) c_tw_fwd.v[k][i]
#define HERDB_TWI(k, i) c_tw_inv.v[k][i]
#else
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
// any x < 2^62 -> residue < 2Q + 2^31 < 2^32 (not canonical)
HERDB_HD u32 fold62(u64 x) {
    u32 hi = (u32)(x >> 31), lo = (u32)x & 0x7FFFFFFFu;
    return mul_shoup_lazy(kR31, kR31s, hi) + lo;
}
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

// ================================================================================================
// FP64-assisted modular multiplication.  On B200 IMAD.HI (and IMAD.WIDE) cost two issue slots of the
// FMA pipe, so a Shoup product (IMAD.HI, IMAD, IMAD) occupies it for 4 slots.  The quotient can be
// computed on the otherwise idle FP64 pipe instead: for y < 2^32 held in the low word of a register
// pair whose high word is 0x43300000, the pair read as a double is exactly 2^52 + y, and
//     fma(2^52 + y, c, 2^52 - 2^52*c) = 2^52 + y*c   (one rounding, to an integer since ulp = 1)
// so the low word of the result is round(y*c).  With c = round(w/Q * 2^52) / 2^52 (2^52*c is an
// integer, hence the addend is exact) that is round(w*y/Q) up to 2^-20, and
//     r = w*y - q*Q  (two IMADs, mod 2^32)  is the SIGNED residue of w*y with |r| <= 0.51 Q.
// Everything stays exact: r == w*y (mod Q); only its representative differs from Shoup's.
namespace herdb {

constexpr double kTwo52 = 4503599627370496.0;
struct F64Tw { u32 w; u32 pad; double c; double k; };       // w, c = round(w/Q*2^52)/2^52, k = 2^52 - 2^52*c

inline F64Tw make_f64tw(u32 w) {
    F64Tw t;
    t.w = w; t.pad = 0;
    const unsigned __int128 num = ((unsigned __int128)w << 52);
    const u64 ci = (u64)((num + kQ / 2) / kQ);               // round(w * 2^52 / Q), < 2^52
    t.c = (double)ci / kTwo52;                                // exact: ci < 2^53
    t.k = kTwo52 - (double)ci;                                // exact integer
    return t;
}

HERDB_HD double pack52(u32 lo) {                              // the double 2^52 + lo
#if defined(__CUDA_ARCH__)
    return __hiloint2double(0x43300000, (int)lo);
#else
    u64 b = 0x4330000000000000ull | lo; double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}
HERDB_HD u32 lo52(double d) {
#if defined(__CUDA_ARCH__)
    return (u32)__double2loint(d);
#else
    u64 b; __builtin_memcpy(&b, &d, 8); return (u32)b;
#endif
}
HERDB_HD double fma64(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return __builtin_fma(a, b, c);
#endif
}
// uniform-twiddle tables in the FP64 form (index = twiddle index 1..31 as in UniformTw)
struct UniformTw64 { F64Tw v[32]; };
inline UniformTw64 make_uniform64(bool inverse) {
    UniformTw64 t;
    t.v[0] = make_f64tw(1);
    for (u32 k = 1; k < 32; ++k) t.v[k] = make_f64tw(inverse ? ctw_inv(k) : ctw_fwd(k));
    return t;
}
#if defined(__CUDACC__)
__constant__ UniformTw64 c_tw64_fwd;        // filled by upload_constants() (kernels.cu) once per device
__constant__ UniformTw64 c_tw64_inv;
#endif
inline const UniformTw64& host_tw64(bool inverse) {
    static const UniformTw64 f = make_uniform64(false), i = make_uniform64(true);
    return inverse ? i : f;
}
#if defined(__CUDA_ARCH__)
#define HERDB_TW64F(k) c_tw64_fwd.v[k]
#define HERDB_TW64I(k) c_tw64_inv.v[k]
#else
#define HERDB_TW64F(k) host_tw64(false).v[k]
#define HERDB_TW64I(k) host_tw64(true).v[k]
#endif

// signed residue of w*y (as u32 two's complement), |r| <= 0.51 Q, any y < 2^32
HERDB_HD u32 mul_f64(u32 w, double c, double k, u32 y) {
    const u32 q = lo52(fma64(pack52(y), c, k));
    return w * y - q * kQ;
}
// the same without the per-twiddle addend: (2^52 + y) - 2^52 is exact, then fma(y, c, 2^52)
HERDB_HD u32 mul_f64_2op(u32 w, double c, u32 y) {
    const u32 q = lo52(fma64(pack52(y) - kTwo52, c, kTwo52));
    return w * y - q * kQ;
}

}  // namespace herdb

// ================================================================================================
// FP64 quotient, modulus-2Q form (the one the blind-rotation kernel uses).  With c = ci / 2^52,
// ci = round(w * 2^52 / 2Q) < 2^51, and Y = 2^52 + y (y < 2^32 in the low word of a register pair
// whose high word is 0x43300000):
//     fma(Y, c, 2^52) = 2^52 + ci + round(y*c)          (one rounding, ulp = 1, sum < 2^53)
// so its low word is  low = lo32(ci) + q  (mod 2^32),  q = round(y*w/2Q) +- 2^-21, and
//     r0 = w*y - low*2Q   (two IMADs, mod 2^32),   r0 + E == w*y - q*2Q,   E = lo32(ci)*2Q mod 2^32
// is the SIGNED residue of w*y modulo 2Q (hence modulo Q), |r0 + E| <= Q + 128.  The addend of the
// DFMA is the same for every twiddle, E rides on the butterfly's adds (a constant-bank operand for
// lane-uniform twiddles).  Exactness never depends on the rounding; only the range bound does.
namespace herdb {
constexpr u32 kF64Slack = 128;                                // |r| <= Q + kF64Slack
struct F64U { u32 w, E, nE, pad; double c; };                 // nE = -E mod 2^32
inline F64U make_f64u(u32 w) {
    F64U t;
    const unsigned __int128 num = ((unsigned __int128)w << 52);
    const u64 ci = (u64)((num + kQ) / (2ull * kQ));           // round(w * 2^52 / 2Q)
    t.w = w; t.c = (double)ci / kTwo52;                       // exact (ci < 2^53)
    t.E = (u32)ci * kQ2; t.nE = 0u - t.E; t.pad = 0;
    return t;
}
HERDB_HD double with_lo(double d, u32 lo) {                   // replace the low word of a register pair
#if defined(__CUDA_ARCH__)
    return __hiloint2double(__double2hiint(d), (int)lo);
#else
    u64 b; __builtin_memcpy(&b, &d, 8); b = (b & 0xFFFFFFFF00000000ull) | lo; __builtin_memcpy(&d, &b, 8); return d;
#endif
}
HERDB_HD double mk_double(u32 hi, u32 lo) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double((int)hi, (int)lo);
#else
    u64 b = ((u64)hi << 32) | lo; double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}
// r0 of the text: y in the low word of pair yd
HERDB_HD u32 mul_q2(u32 w, double c, double yd) {
    const u32 low = lo52(fma64(yd, c, kTwo52));
    return w * lo52(yd) - low * kQ2;
}
}  // namespace herdb
