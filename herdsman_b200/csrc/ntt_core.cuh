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

// ---------------------------------------------------------------- inverse (GS): invariant all values in [0, 2Q)
HERDB_HD void inv_pass1(u32 (&x)[32], const u32* lane_tw /* smem or global, chunked [g][32][4] */, u32 lane) {
    int off = 0;
#pragma unroll
    for (int t = 1; t <= 16; t <<= 1) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            if (e & t) continue;
            const int p = off + e / (2 * t);
            const u32* tp = lane_tw + ((p >> 1) * 32 + lane) * 4 + (p & 1) * 2;
            u32 u = x[e], v = x[e + t];
            x[e] = csub(u + v, kQ2);
            x[e + t] = mul_shoup_lazy(tp[0], tp[1], u - v + kQ2);
        }
        off += 16 / t;
    }
}
HERDB_HD void inv_pass2(u32 (&x)[32]) {           // strided layout, strides 32..512
#pragma unroll
    for (int h = 1; h <= 16; h <<= 1) {
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            if (k & h) continue;
            const int idx = 16 / h + k / (2 * h);
            u32 u = x[k], v = x[k + h];
            x[k] = csub(u + v, kQ2);
            x[k + h] = mul_shoup_lazy(HERDB_TWI(idx, 0), HERDB_TWI(idx, 1), u - v + kQ2);
        }
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
