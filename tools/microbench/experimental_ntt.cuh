// tools/microbench/experimental_ntt.cuh -- NTT butterfly variants that were measured on B200 and NOT adopted by
// the product kernel (DESIGN.md "What was tried"): quotients on the FP64 pipe (two forms), a hybrid FP64/Shoup
// forward pass, two polynomials per warp in lock-step, and the earlier forms of the product code the centred
// butterflies and the FP64 folds replaced ([0, 2Q) forward pass 1 on raw digits, the Shoup fold of a 62-bit sum).
// Used only by the micro-benchmarks in this directory; kept so the measurements in profiles/round1/r01* can be reproduced.
#pragma once
#include "../../herdsman_b200/csrc/ntt_core.cuh"
namespace herdb {
// 2^31 mod Q and its Shoup companion: folds the high part of a 62-bit value
constexpr u32 kR31 = (u32)((1ull << 31) % kQ);
constexpr u32 kR31s = cshoup(kR31);

// any x < 2^62 -> residue < 2Q + 2^31 < 2^32 (not canonical)
HERDB_HD u32 fold62(u64 x) {
    u32 hi = (u32)(x >> 31), lo = (u32)x & 0x7FFFFFFFu;
    return mul_shoup_lazy(kR31, kR31s, hi) + lo;
}
// Forward pass 1 for a signed-digit polynomial given as RAW unsigned digits u = r + 64 in [0,127]
// (true coefficient r = u - 64, i.e. u + C with C = Q - 64 as a residue).  Stage 0 has a single
// twiddle W1, so the offset is absorbed into its two add constants:
//   X' = (u_x + C) + W1*(u_y + C) = u_x + W1*u_y + K1,   K1 = (C + W1*C) mod Q
//   Y' = (u_x + C) - W1*(u_y + C) = u_x - W1*u_y + K2,   K2 = ((C - W1*C) mod Q) + 2Q
// Outputs of stage 0 are < 3Q + 127; the remaining stages are the generic lazy ones.
constexpr u32 kDigC = kQ - 64;
constexpr u32 kDigCw = cmulmod(ctw_fwd(1), kDigC);
constexpr u32 kDigK1 = (kDigC + kDigCw) % kQ;
constexpr u32 kDigK2 = (kDigC + kQ - kDigCw) % kQ + kQ2;
HERDB_HD void fwd_pass1_digits(u32 (&x)[32]) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        u32 v = mul_shoup_lazy(HERDB_TWF(1, 0), HERDB_TWF(1, 1), x[k + 16]);
        u32 u = x[k];
        x[k] = u + v + kDigK1;
        x[k + 16] = u - v + kDigK2;
    }
#pragma unroll
    for (int s = 1; s < 5; ++s) {
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
// M digit polynomials of the same accumulator component transformed together by one warp: M independent
// butterflies per twiddle give the scheduler M times the instruction-level parallelism (a warp shares its
// scheduler with only one other warp) and every streamed twiddle pair is loaded once for the M of them.
template <int M>
HERDB_HD void fwd_pass1_digits_multi(u32 (&x)[M][32]) {
#pragma unroll
    for (int k = 0; k < 16; ++k)
#pragma unroll
        for (int m = 0; m < M; ++m) {
            u32 v = mul_shoup_lazy(HERDB_TWF(1, 0), HERDB_TWF(1, 1), x[m][k + 16]);
            u32 u = x[m][k];
            x[m][k] = u + v + kDigK1;
            x[m][k + 16] = u - v + kDigK2;
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
                u32 v = mul_shoup_lazy(HERDB_TWF(idx, 0), HERDB_TWF(idx, 1), x[m][k + h]);
                u32 u = x[m][k];
                x[m][k] = u + v;
                x[m][k + h] = u - v + kQ2;
            }
        }
    }
}
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


namespace herdb {

// The same pass with the quotient of every product taken on the FP64 pipe (modarith.cuh, mul_f64):
// each product is the signed residue r, |r| <= 0.51 Q, so a stage moves a value by at most 0.51 Q in
// either direction.  Stage 0 adds 8Q on top of the digit offset: outputs of stage 0 lie in
// [7.49Q, 9.51Q + 127], outputs of the pass in [5.4Q, 11.6Q]; the five lazy Shoup stages of pass 2
// then end below 21.6Q < 2^32, and 8 * 21.6Q * Q < 2^62 keeps the 64-bit MAC exact.
constexpr u32 kDigK1f = (kDigC + kDigCw) % kQ + 8 * kQ;
constexpr u32 kDigK2f = (kDigC + kQ - kDigCw) % kQ + 8 * kQ;
HERDB_HD void fwd_pass1_digits_f64(u32 (&x)[32]) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        u32 r = mul_f64(HERDB_TW64F(1).w, HERDB_TW64F(1).c, HERDB_TW64F(1).k, x[k + 16]);
        u32 u = x[k];
        x[k] = u + r + kDigK1f;
        x[k + 16] = u - r + kDigK2f;
    }
#pragma unroll
    for (int s = 1; s < 5; ++s) {
        const int h = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = (1 << s) + k / (2 * h);
            u32 r = mul_f64(HERDB_TW64F(idx).w, HERDB_TW64F(idx).c, HERDB_TW64F(idx).k, x[k + h]);
            u32 u = x[k];
            x[k] = u + r;
            x[k + h] = u - r;
        }
    }
}
// Two polynomials transformed in lock-step by one warp: every twiddle is fetched once and the two
// independent dependency chains double the instruction-level parallelism (2 warps per scheduler).
HERDB_HD void fwd_pass1_digits_x2(u32 (&x)[32], u32 (&y)[32]) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        u32 v = mul_shoup_lazy(HERDB_TWF(1, 0), HERDB_TWF(1, 1), x[k + 16]);
        u32 v2 = mul_shoup_lazy(HERDB_TWF(1, 0), HERDB_TWF(1, 1), y[k + 16]);
        u32 u = x[k], u2 = y[k];
        x[k] = u + v + kDigK1; x[k + 16] = u - v + kDigK2;
        y[k] = u2 + v2 + kDigK1; y[k + 16] = u2 - v2 + kDigK2;
    }
#pragma unroll
    for (int s = 1; s < 5; ++s) {
        const int h = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = (1 << s) + k / (2 * h);
            u32 v = mul_shoup_lazy(HERDB_TWF(idx, 0), HERDB_TWF(idx, 1), x[k + h]);
            u32 v2 = mul_shoup_lazy(HERDB_TWF(idx, 0), HERDB_TWF(idx, 1), y[k + h]);
            u32 u = x[k], u2 = y[k];
            x[k] = u + v; x[k + h] = u - v + kQ2;
            y[k] = u2 + v2; y[k + h] = u2 - v2 + kQ2;
        }
    }
}
HERDB_HD void fwd_pass2_stream_x2(u32 (&x)[32], u32 (&y)[32], const u32* lane_tw, u32 lane) {
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
            u32 v = mul_shoup_lazy(w, wp, x[e + t]);
            u32 v2 = mul_shoup_lazy(w, wp, y[e + t]);
            u32 u = x[e], u2 = y[e];
            x[e] = u + v; x[e + t] = u - v + kQ2;
            y[e] = u2 + v2; y[e + t] = u2 - v2 + kQ2;
        }
        off += 16 / t;
    }
}

// FP64-quotient form of the same pass: the product r is a signed residue, |r| <= 0.51 Q, and r + Q keeps
// the [0, 2Q) invariant (the +Q rides on the first IMAD's addend)
HERDB_HD void inv_pass2_f64(u32 (&x)[32]) {
#pragma unroll
    for (int h = 1; h <= 16; h <<= 1) {
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = 16 / h + k / (2 * h);
            u32 u = x[k], v = x[k + h];
            x[k] = csub(u + v, kQ2);
            x[k + h] = mul_f64(HERDB_TW64I(idx).w, HERDB_TW64I(idx).c, HERDB_TW64I(idx).k, u - v + kQ2) + kQ;
        }
    }
}

// ---------------------------------------------------------------- FP64-quotient passes, modulus-2Q form (modarith.cuh)
// Uniform twiddles come from constant memory as (w, E, -E, c); values live in register pairs whose
// high word stays 0x43300000 so that a value is a DFMA operand without any conversion.
struct UniformF64 { F64U v[32]; u32 dig_xp, dig_xm, pad0, pad1; };
constexpr u32 kF64Base = 9 * kQ;                              // centre of the lazy range of the forward NTT (a multiple of Q)
inline UniformF64 make_uniform_f64(bool inverse) {
    UniformF64 t;
    t.v[0] = make_f64u(1);
    for (u32 k = 1; k < 32; ++k) t.v[k] = make_f64u(inverse ? ctw_inv(k) : ctw_fwd(k));
    // stage 0 of a digit polynomial (raw digits u = r + 64): the -64 offsets and the range centre ride on E
    t.dig_xp = t.v[1].E + (kDigC + kDigCw) % kQ + kF64Base;
    t.dig_xm = t.v[1].nE + (kDigC + kQ - kDigCw) % kQ + kF64Base;
    t.pad0 = t.pad1 = 0;
    return t;
}
#if defined(__CUDACC__)
__constant__ UniformF64 c_f64u_fwd;         // filled by upload_constants() (kernels.cu) once per device
__constant__ UniformF64 c_f64u_inv;
#endif
inline const UniformF64& host_f64u(bool inverse) {
    static const UniformF64 f = make_uniform_f64(false), i = make_uniform_f64(true);
    return inverse ? i : f;
}
#if defined(__CUDA_ARCH__)
#define HERDB_F64UF c_f64u_fwd
#define HERDB_F64UI c_f64u_inv
#else
#define HERDB_F64UF host_f64u(false)
#define HERDB_F64UI host_f64u(true)
#endif

// one CT butterfly on register pairs: a' = a + w*b + ea, b' = a - w*b + eb   (ea/eb carry +-E)
HERDB_HD void bfly_q2(double& a, double& b, u32 w, double c, u32 ea, u32 eb) {
    const u32 r0 = mul_q2(w, c, b);
    const u32 u = lo52(a);
    a = with_lo(a, u + r0 + ea);
    b = with_lo(b, u - r0 + eb);
}
// Forward pass 1 of a digit polynomial given as raw digits in [0,127].  Every stage moves a value by
// at most Q + 128 either way: outputs lie in kF64Base + [-5Q - 640, 5Q + 640 + 127 + Q].
HERDB_HD void fwd_pass1_digits_q2(u32 (&x)[32]) {
    double xd[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) xd[k] = pack52(x[k]);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        bfly_q2(xd[k], xd[k + 16], HERDB_F64UF.v[1].w, HERDB_F64UF.v[1].c, HERDB_F64UF.dig_xp, HERDB_F64UF.dig_xm);
#pragma unroll
    for (int s = 1; s < 5; ++s) {
        const int h = 16 >> s;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = (1 << s) + k / (2 * h);
            bfly_q2(xd[k], xd[k + h], HERDB_F64UF.v[idx].w, HERDB_F64UF.v[idx].c, HERDB_F64UF.v[idx].E, HERDB_F64UF.v[idx].nE);
        }
    }
#pragma unroll
    for (int k = 0; k < 32; ++k) x[k] = lo52(xd[k]);
}
// Forward pass 2, hybrid: stages 5..7 (7 lane twiddles, FP64 form: w, E, c) and stages 8..9 (24 lane
// twiddles, Shoup form: w, w').  tf = 7 x {w, E, c_lo, c_hi}, ts = 24 x {w, w'}.
constexpr int kFwd2F64Tw = 7, kFwd2ShoupTw = 24;
constexpr u32 kLaneTwHybChunks = kFwd2F64Tw + kFwd2ShoupTw / 2;          // 19 x 16 bytes per lane
constexpr u32 kLaneTwHybWords = kLaneTwHybChunks * 32 * 4;
HERDB_HD void fwd_pass2_hyb(u32 (&x)[32], const u32 (&tf)[4 * kFwd2F64Tw], const u32 (&ts)[2 * kFwd2ShoupTw]) {
    double xd[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) xd[e] = pack52(x[e]);
    int off = 0;
#pragma unroll
    for (int s = 5; s < 8; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            bfly_q2(xd[e], xd[e + t], tf[4 * p], mk_double(tf[4 * p + 3], tf[4 * p + 2]), tf[4 * p + 1], 0u - tf[4 * p + 1]);
        }
        off += 16 / t;
    }
#pragma unroll
    for (int e = 0; e < 32; ++e) x[e] = lo52(xd[e]);
    off = 0;
#pragma unroll
    for (int s = 8; s < 10; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            u32 v = mul_shoup_lazy(ts[2 * p], ts[2 * p + 1], x[e + t]);
            u32 u = x[e];
            x[e] = u + v;
            x[e + t] = u - v + kQ2;
        }
        off += 16 / t;
    }
}
// the same pass with the lane twiddles streamed from shared memory (table of build_lane_twiddles_hyb):
// no twiddle registers, for the 128-register (16 warps per SM) configuration
HERDB_HD void fwd_pass2_hyb_stream(u32 (&x)[32], const u32* tw, u32 lane) {
    double xd[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) xd[e] = pack52(x[e]);
    int off = 0;
#pragma unroll
    for (int s = 5; s < 8; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const V4 q = *reinterpret_cast<const V4*>(tw + (p * 32 + lane) * 4);
            bfly_q2(xd[e], xd[e + t], q.v[0], mk_double(q.v[3], q.v[2]), q.v[1], 0u - q.v[1]);
        }
        off += 16 / t;
    }
#pragma unroll
    for (int e = 0; e < 32; ++e) x[e] = lo52(xd[e]);
    off = 0;
#pragma unroll
    for (int s = 8; s < 10; ++s) {
        const int t = 16 >> (s - 5);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const u32* tp = tw + ((kFwd2F64Tw + (p >> 1)) * 32 + lane) * 4 + (p & 1) * 2;
            u32 v = mul_shoup_lazy(tp[0], tp[1], x[e + t]);
            u32 u = x[e];
            x[e] = u + v;
            x[e + t] = u - v + kQ2;
        }
        off += 16 / t;
    }
}
// lane table of the hybrid pass: [19 chunks][32 lanes][4 words]; chunks 0..6 = {w, E, c_lo, c_hi} of the
// stage 5/6/7 twiddles, chunks 7..18 = two Shoup pairs each of the stage 8/9 twiddles
inline void build_lane_twiddles_hyb(u32* out) {
    for (u32 T = 0; T < 32; ++T) {
        u32 p = 0;
        for (u32 s = 5; s < 8; ++s) {
            u32 t = 16u >> (s - 5);
            for (u32 i = 0; i < 16 / t; ++i, ++p) {
                F64U f = make_f64u(ctw_fwd((1u << s) + T * (16 / t) + i));
                u64 cb; __builtin_memcpy(&cb, &f.c, 8);
                u32* o = out + (p * 32 + T) * 4;
                o[0] = f.w; o[1] = f.E; o[2] = (u32)cb; o[3] = (u32)(cb >> 32);
            }
        }
        u32 q = 0;
        for (u32 s = 8; s < 10; ++s) {
            u32 t = 16u >> (s - 5);
            for (u32 i = 0; i < 16 / t; ++i, ++q) {
                u32 w = ctw_fwd((1u << s) + T * (16 / t) + i);
                u32* o = out + ((kFwd2F64Tw + (q >> 1)) * 32 + T) * 4 + (q & 1) * 2;
                o[0] = w; o[1] = cshoup(w);
            }
        }
    }
}

}  // namespace herdb
