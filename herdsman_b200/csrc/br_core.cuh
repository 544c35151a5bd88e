// br_core.cuh -- the programmable-bootstrap body executed by one CTA (256 threads) for a
// group of kC = 4 ciphertexts that share every bootstrapping-key read.
//
// Replaces, for the hot path of herdsman's map tasks (MapTask -> per-row EvalBinGate), the
// OpenFHE binfhe v1.0.x routines restated in oracle/oracle.c:
//   EvalBinGate ciphertext preparation, BootstrapGateCore (accumulator init),
//   RingGSWAccumulatorCGGI::EvalAcc / AddToAccCGGI (the CMUX loop), SignedDigitDecompose,
//   Transpose + sample extraction and ModSwitch(Q -> qKS)            (SURVEY.md 8a a7-a12).
// Key switching + ModSwitch(qKS -> q) is keyswitch.cu.
//
// Data residency (per CTA, shared memory):
//   acc  [kC][2][1024] u32   accumulator, NTT domain, pre-scaled by 1/N, lazily reduced to [0,2Q)
//   dig  [kC][6][1024] u32   NTT of the signed-digit polynomials l = 0,1,2 of both components, lazy residues
//                            < 2^32; the top digit l = 3 is never transformed (top-digit elimination,
//                            modarith.cuh): its key rows are folded so that they multiply acc itself
//   itw  [16][32][4]   u32   per-lane twiddles of the inverse pass 1
//   lwe  [kC][n+1]     u16   prepared LWE ciphertext (ct1 + ct2 etc.)
// Per step: phase A = one warp per (ciphertext, component): INTT, decompose, 3 forward NTTs;
//           phase B = one thread per 4 evaluation points: 32 key words (staged in shared memory by
//           cp.async, shared by the 4 ciphertexts) x (6 digit transforms + the 2 accumulator components)
//           multiply-accumulated exactly in 64 bits, folded, multiplied by the two monomials and added
//           into acc.
//
// The body is templated on an execution environment so that tests/emu can run the very same
// code on the CPU with host threads and barriers (there is no GPU in the build container).
#pragma once
#include "ntt_core.cuh"

namespace herdb {

// Ciphertexts per CTA is a template parameter C.  C = 4 is the throughput shape: 256 threads = 8 warps, one warp
// per (ciphertext, component) polynomial in phase A, 4 evaluation points per thread in phase B, key rows 2..7
// staged by cp.async into slots freed inside the MAC loop.  C = 2 and C = 1 are the small-job shapes (launches
// and launch tails that would leave SMs empty at C = 4): 128 threads, 8 evaluation points per thread in phase
// B; with fewer ciphertexts there is room for the whole 128 KiB key slice of a step, so all eight rows arrive
// by TMA bulk copies at step start and phase B has no staging at all.  At C = 1 only warps 0 and 1 have a
// polynomial in phase A; all four warps share phase B.
constexpr int kCMax = 4;
template <int C> struct BrShape { static constexpr u32 kThreads = C == 4 ? 256u : 128u; };
constexpr u32 kExtStride = 1032;            // u16 words per extracted ciphertext (N + 1, padded to 16 B)
constexpr u32 kLweMax = HERDB_N_STD128;     // largest LWE dimension supported by the smem layout

enum OperandFlags : uint8_t { kOpNeg = 1, kOpConst = 2, kOpConstVal = 4 };

// one bootstrapped gate of a circuit level; a job = (gate, row)
struct GateDesc {
    uint8_t gate;                // herdb_gate
    uint8_t flag_a, flag_b;      // OperandFlags
    uint8_t pool_a, pool_b, pool_out;
    uint8_t pad[2];
    u32 idx_a, idx_b, idx_out;   // ciphertext index inside a row of the pool
};
// ciphertext (row r, idx) of a pool at base + r*row_words + idx*(n+1).  Dense rows: row_words = cts_per_row*(n+1).
// Rows in herdsman's partition framing ([u32 length][payload], storage_service.cpp:19-28) are used in place:
// row_words = 2 + cts_per_row*(n+1) and base points behind the first length word.
struct Pool { u16* base; u32 row_words; u32 pad; };

struct BrParams {
    const u32* bsk;              // [n][2 sign][8 rows][2 cols][1024]  NTT domain, canonical, rows folded by fold_top_digit, words at key_pos
    const u32* mono;             // [1024][1024]   NTT(X^(2*row) - 1), canonical
    const u32* ftw2;             // [16][32][4]    forward pass-2 lane twiddles
    const u32* itw1;             // [16][32][4]    inverse pass-1 lane twiddles
    const GateDesc* gates;       // gates of this level
    Pool pools[4];
    u16* ext;                    // [jobs][kExtStride]  extracted, mod-switched (qKS) ciphertexts
    u32* trace;                  // optional debug: [jobs][2][1024] accumulator (coefficient domain) after `steps`
    u32 n;                       // LWE dimension
    u32 steps;                   // CMUX steps to run (== n except in debug traces)
    u32 rows;                    // rows in this chunk
    u32 jobs;                    // gates_in_level * rows;  job -> gate = job / rows, row = job % rows
    u32 job0, job1;              // this launch covers jobs [job0, job1): a level may be cut into a body of whole grid
                                 // rounds at C = 4 and a tail at a smaller shape
};

template <int C>
struct BrSmem {
    u32 acc[C][2][kN];
    u32 dig[C][2 * kDGT][kN];
    u32 itw[kLaneTwWords];       // inverse pass-1 lane twiddles
    u32 ftw[kLaneTwWords];       // forward pass-2 lane twiddles (streamed by fwd_pass2_stream)
    u32 kst[C == 4 ? 2 : 8][4][kN];   // C = 4: bootstrapping-key staging ring (cp.async), two rows x (sign, column) x 1024 points;
                                 // C < 4: the whole key slice of a step, eight rows (bulk copies)
    u32 mono[C][2][kN];          // the two monomial rows of this step per ciphertext (cp.async, land during phase A)
    u16 lwe[C][kLweMax + 8];
    u32 gate_b[C];               // prepared b and gate code per ciphertext
    u32 gate_code[C];
    alignas(8) unsigned long long bar;   // mbarrier of the step-start bulk copies (key rows 0, 1 and the monomial rows)
};

constexpr u32 kAccInitPos = cmulmod(kQ / 8 + 1, kNInv);             // +(Q/8+1) / N
constexpr u32 kAccInitNeg = cmulmod(kQ - (kQ / 8 + 1), kNInv);      // -(Q/8+1) / N
constexpr u32 kGateConstEighths[6] = {5, 7, 1, 3, 5, 1};            // OR AND NOR NAND XOR_FAST XNOR_FAST

HERDB_HD u32 gate_const_eighths(u32 gate) {
    // RingGSWCryptoParams::m_gateConst = {5,7,1,3,5,1} * q/8
    return gate == 0 ? 5u : gate == 1 ? 7u : gate == 2 ? 1u : gate == 3 ? 3u : gate == 4 ? 5u : 1u;
}

// one LWE word of an operand after the free transforms (EvalNOT / trivial constant)
HERDB_HD u32 operand_word(const Pool* pools, u32 pool, u32 idx, u32 flags, u32 row, u32 n, u32 k) {
    if (flags & kOpConst) return k == n ? ((flags & kOpConstVal) ? kLweQ / HERDB_PT_MOD : 0u) : 0u;
    u32 v = pools[pool].base[(size_t)row * pools[pool].row_words + (size_t)idx * (n + 1) + k];
    if (flags & kOpNeg) v = (k == n) ? ((kLweQ >> 2) + kLweQ - v) & (kLweQ - 1) : (kLweQ - v) & (kLweQ - 1);
    return v;
}

// ------------------------------------------------------------------------------------------------
template <int kC, class Env>
// `bar_phase0`: parity of the number of bulk-copy phases S.bar has completed before this group (the caller
// initialises S.bar once, with one expected arrival, and runs groups back to back: steps of earlier groups)
HERDB_HD void blind_rotate_group(Env& env, const BrParams& P, BrSmem<kC>& S, u32 group, u32 bar_phase0) {
    constexpr u32 kThreads = BrShape<kC>::kThreads;
    constexpr int kChunks = (int)(kN / 4 / kThreads);              // chunks of 4 evaluation points per thread in phase B
    constexpr bool kRing = kC == 4;                                // key rows 2..7 through the cp.async ring (else: all rows by bulk copy)
    static_assert(kC == 4 || kC == 2 || kC == 1, "CTA shapes: 4 (throughput), 2 and 1 (small launches and tails)");
    static_assert(!kRing || kChunks == 1, "the ring recycles the four digit slots of four ciphertexts for the four (sign, column) pieces");
    const u32 tid = env.tid();
    const u32 lane = tid & 31, warp = tid >> 5;
    const bool has_poly = kC != 1 || warp < 2u;                    // phase A: warp = (ciphertext, component); C = 1 leaves warps 2, 3 without one
    const u32 n = P.n;

    // ---- ciphertext preparation: ct1 + ct2 (2*(ct1 - ct2) for XOR_FAST / XNOR_FAST), NOT folded in
    for (int c = 0; c < kC; ++c) {
        u32 job = P.job0 + group * kC + c; if (job >= P.job1) job = P.job1 - 1;
        const GateDesc g = P.gates[job / P.rows];
        const u32 row = job % P.rows;
        for (u32 k = tid; k <= n; k += kThreads) {
            u32 a = operand_word(P.pools, g.pool_a, g.idx_a, g.flag_a, row, n, k);
            u32 b = operand_word(P.pools, g.pool_b, g.idx_b, g.flag_b, row, n, k);
            u32 v = (g.gate >= HERDB_GATE_XOR_FAST) ? (2u * (a + kLweQ - b)) & (kLweQ - 1) : (a + b) & (kLweQ - 1);
            if (k == n) { S.gate_b[c] = v; S.gate_code[c] = g.gate; } else S.lwe[c][k] = (u16)v;
        }
    }
    // lane twiddles -> smem (once per group; 2 x 8 KB)
    // (the forward table is stored with CEILING Shoup companions, w' + 1: odd words; centred butterflies, ntt_core.cuh)
    for (u32 w = tid; w < kLaneTwWords; w += kThreads) { S.itw[w] = P.itw1[w]; S.ftw[w] = P.ftw2[w] + (w & 1u); }
    env.sync();
    // ---- BootstrapGateCore: acc = (0, NTT(m)), m[2j] = +-(Q/8+1) depending on (b - j) in [q1, q2)
    if (has_poly) {
        const int c = warp >> 1, j = warp & 1;
        u32 x[32];
        if (j == 0) {
#pragma unroll
            for (int k = 0; k < 32; ++k) x[k] = 0;
            store_contig(S.acc[c][0], lane, x);
        } else {
            const u32 b = S.gate_b[c];
            const u32 q1 = gate_const_eighths(S.gate_code[c]) * (kLweQ >> 3);
            const u32 q2 = (q1 + (kLweQ >> 1)) & (kLweQ - 1);
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                const u32 coef = lane + 32 * k;                        // coefficient index; factor 2N/q = 2
                const u32 temp = (b + kLweQ - (coef >> 1)) & (kLweQ - 1);
                const bool in = (q1 < q2) ? (temp >= q1 && temp < q2) : !(temp >= q2 && temp < q1);
                x[k] = (coef & 1) ? 0u : (in ? kAccInitNeg : kAccInitPos);
            }
            fwd_pass1(x);
            store_strided(S.dig[c][1], lane, x);
            env.syncwarp();
            load_contig(S.dig[c][1], lane, x);
            fwd_pass2_stream(x, P.ftw2, lane);                           // floor companions: straight from the global table
#pragma unroll
            for (int k = 0; k < 32; ++k) x[k] = canon32(x[k]);
            store_contig(S.acc[c][1], lane, x);
        }
    }
    env.sync();

    // ---- EvalAcc: n CMUX steps (AddToAccCGGI)
    for (u32 i = 0; i < P.steps; ++i) {
        // Key slice of this step.  C = 4: rows 0 and 1 and the eight monomial rows X^(+-a_i) - 1 (64 KiB, sixteen 4 KiB
        // pieces) are fetched by TMA bulk copies that ONE thread issues right after the barrier that ended the
        // previous step (every thread has finished reading the destinations); they land during phase A and are
        // awaited on an mbarrier at the start of phase B.  Rows 2..7 cannot be fetched that way: their
        // destinations are slots a thread frees in the middle of the MAC loop, so each thread copies exactly the
        // 16-byte pieces it will itself consume with cp.async (completion = cp.async.wait_group, no barrier):
        // into the ring stage and the digit slots it has already read.
        // C < 4: shared memory holds the whole slice, all eight rows come by bulk copy and nothing is recycled.
        const u32* Kstep = P.bsk + (size_t)i * (2 * kDG2 * 2 * kN);
        constexpr u32 kDigCt = 2 * kDGT * kN;           // words between the digit blocks of two ciphertexts
        if (tid == 0) {
            constexpr int kBulkRows = kRing ? 2 : (int)kDG2;
            env.bulk_begin(&S.bar, (kBulkRows * 4 + 2 * kC) * kN * 4);
#pragma unroll
            for (int r = 0; r < kBulkRows; ++r)
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int jo = 0; jo < 2; ++jo) env.bulk_copy(&S.kst[r][2 * s + jo][0], Kstep + ((s * kDG2 + r) * 2 + jo) * kN, kN * 4, &S.bar);
#pragma unroll
            for (int cc = 0; cc < kC; ++cc) {
                const u32 a = S.lwe[cc][i];
                const u32 row_pos = (kLweQ - a) & (kLweQ - 1);          // monomial X^(2*row_pos) - 1   (ek1, s = +1)
                const u32 row_neg = a;                                   // monomial X^(2N - 2*row_pos) - 1 (ek2, s = -1)
                env.bulk_copy(&S.mono[cc][0][0], P.mono + (size_t)row_pos * kN, kN * 4, &S.bar);
                env.bulk_copy(&S.mono[cc][1][0], P.mono + (size_t)row_neg * kN, kN * 4, &S.bar);
            }
        }
        // ===== phase A: warp = (ciphertext c, component j)
        if (has_poly) {
            const int c = warp >> 1, j = warp & 1;
            u32 x[32];
            load_contig(S.acc[c][j], lane, x);
            inv_pass1<1, 2>(x, S.itw, lane);
            store_contig(S.dig[c][j], lane, x);              // dig[c][j] doubles as transposition scratch
            env.syncwarp();
            load_strided(S.dig[c][j], lane, x);
            inv_pass2<2, 1>(x);
            // SignedDigitDecompose.  d = centred representative of the coefficient; OpenFHE peels signed
            // 7-bit digits r_l in [-64,63] iteratively.  Equivalently d + B = sum (r_l + 64) * 128^l with
            // B = 64*(1 + 128 + 128^2 + 128^3): the digits are the plain 7-bit fields of d + B (unique
            // representation; |d| < Q/2 < 2^26 keeps 0 < d + B < 2^28).
            constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
            u32 d[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                u32 t = csub(x[k], kQ);                       // canonical [0, Q)
                d[k] = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
            }
            env.syncwarp();                                   // all lanes done reading the scratch
            // digits l = 0,1,2 (the top digit is not needed: top-digit elimination), transformed together
            u32 y[kDGT][32];
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l)
#pragma unroll
                for (int k = 0; k < 32; ++k) y[l][k] = d[k] & (127u << (7 * l));   // the field in place: (r_l + 64) * 2^(7l)
            fwd_pass1_fields_multi<(int)kDGT>(y);                 // offset folded into stage 0, 2^(-7l) into the key rows
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l) store_strided(S.dig[c][j + 2 * l], lane, y[l]);   // output polynomial index j + 2*l
            env.syncwarp();
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l) load_contig(S.dig[c][j + 2 * l], lane, y[l]);
            fwd_pass2_stream_multi<(int)kDGT>(y, S.ftw, lane);   // lane twiddles streamed from smem, one load per 3 butterflies
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l) store_contig(S.dig[c][j + 2 * l], lane, y[l]);    // < 22Q < 2^32, not reduced
        }
        env.sync();
        // ===== phase B: thread = kChunks chunks of 4 evaluation points: 4*(tid + q*kThreads) ..+3
        auto phase_b = [&](const int q) {
            // exactness of the 64-bit sums: six digit transforms below kFwdOutMax and the two accumulator components
            // below 2Q, each times a key word below Q, stay below 2^62 (what fold62_f64 accepts)
            static_assert((6 * kFwdOutMax + 2 * 2ull * kQ) < ((1ull << 62) / kQ), "key MAC fits 62 bits");
            const u32 lin_b = 4 * (tid + (u32)q * kThreads);                 // staging ring and monomials are stored linearly
            const u32 phys_b = swz(lin_b);
            const u32* K = Kstep + lin_b;
            auto stage_key = [&](int r, u32* dst0, u32 dst_stride) {      // piece (s, jo) of row r -> dst0 + (2*s + jo)*dst_stride
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int jo = 0; jo < 2; ++jo) env.cp16(dst0 + (2 * s + jo) * dst_stride, K + ((s * kDG2 + r) * 2 + jo) * kN);
            };
            u64 mac[kC][2][2][4];
#pragma unroll
            for (int c = 0; c < kC; ++c)
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int jo = 0; jo < 2; ++jo)
#pragma unroll
                        for (int p = 0; p < 4; ++p) mac[c][s][jo][p] = 0;
            // Row r = j + 2l of the key multiplies the transform of digit l of component j for l < 3 and, for
            // l = 3 (r = 6, 7), the stored accumulator component j itself (folded rows, modarith.cuh).
            // Ring (C = 4), where row r's key pieces live: r = 0,2,6 ring 0; r = 1,4 ring 1; r = 3,5,7 the consumed digit slots 0,1,2
#pragma unroll
            for (int r = 0; r < (int)kDG2; ++r) {
                if (r == 0 && q == 0) env.mbar_wait(&S.bar, (bar_phase0 + i) & 1);            // the bulk copies of this step
                if (kRing && (r == 2 || r == 4)) env.cp_wait1();               // cp.async groups: {2,3} | {4,5} | {6,7}
                if (kRing && r == 6) env.cp_wait0();
                const u32* kbase = !kRing                          ? &S.kst[r % (kRing ? 2 : 8)][0][lin_b]
                                 : (r == 0 || r == 2 || r == 6)    ? &S.kst[0][0][lin_b]
                                 : (r == 1 || r == 4)              ? &S.kst[1][0][lin_b]
                                                                   : &S.dig[0][(r - 3) / 2][phys_b];
                const u32 kstride = (kRing && (r == 3 || r == 5 || r == 7)) ? kDigCt : kN;
                V4 kv[2][2];
#pragma unroll
                for (int s = 0; s < 2; ++s)
#pragma unroll
                    for (int jo = 0; jo < 2; ++jo) kv[s][jo] = *reinterpret_cast<const V4*>(kbase + (2 * s + jo) * kstride);
                u32 dep = 0;
#pragma unroll
                for (int c = 0; c < kC; ++c) {
                    V4 dv = *reinterpret_cast<const V4*>(r < 2 * (int)kDGT ? &S.dig[c][r][phys_b] : &S.acc[c][r - 2 * (int)kDGT][phys_b]);
                    dep |= dv.v[0];
#pragma unroll
                    for (int s = 0; s < 2; ++s)
#pragma unroll
                        for (int jo = 0; jo < 2; ++jo)
#pragma unroll
                            for (int p = 0; p < 4; ++p) mac[c][s][jo][p] += (u64)dv.v[p] * kv[s][jo].v[(p + 3) & 3];   // key words are stored rotated: key_pos
                }
                // digit r is consumed: its slots (and the ring stage it came from) take the next pieces.  The
                // asm operands make the copies wait for this iteration's shared-memory reads (dep, kv).
                if (kRing && r <= 2) {
                    env.cp_fence(dep ^ kv[1][1].v[0]);
                    stage_key(2 * r + 2, &S.kst[r & 1][0][lin_b], kN);
                    stage_key(2 * r + 3, &S.dig[0][r][phys_b], kDigCt);
                    env.cp_commit();
                }
            }
#pragma unroll
            for (int c = 0; c < kC; ++c) {
                const V4 mp = *reinterpret_cast<const V4*>(&S.mono[c][0][lin_b]);
                const V4 mn = *reinterpret_cast<const V4*>(&S.mono[c][1][lin_b]);
#pragma unroll
                for (int jo = 0; jo < 2; ++jo) {
                    V4 av = *reinterpret_cast<const V4*>(&S.acc[c][jo][phys_b]);
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        const u32 rp = fold62_f64(mac[c][0][jo][p]);       // < 1.76 Q: the sum below stays < 2^56
                        const u32 rn = fold62_f64(mac[c][1][jo][p]);
                        // acc + update in (0.2499 Q, 3.7501 Q) -> [0, 2Q): what the inverse transform and the MAC accept
                        av.v[p] = csub(fold62_f64((u64)mp.v[p] * rp + (u64)mn.v[p] * rn, av.v[p]), kQ2);
                    }
                    *reinterpret_cast<V4*>(&S.acc[c][jo][phys_b]) = av;
                }
            }
        };
        if constexpr (kChunks == 1) phase_b(0);
        else {
#pragma unroll 1
            for (int q = 0; q < kChunks; ++q) phase_b(q);
        }
        env.sync();
    }

    // ---- back to coefficients, Transpose(acc[0]) -> LWE a, acc[1][0] + Q/8 + 1 -> b, ModSwitch(qKS)
    if (has_poly) {
        const int c = warp >> 1, j = warp & 1;
        const u32 job = P.job0 + group * kC + c;
        u32 x[32];
        load_contig(S.acc[c][j], lane, x);
        inv_pass1<1, 2>(x, S.itw, lane);
        store_contig(S.dig[c][j], lane, x);
        env.syncwarp();
        load_strided(S.dig[c][j], lane, x);
        inv_pass2<2, 1>(x);
        if (job < P.job1) {
            u16* ext = P.ext + (size_t)job * kExtStride;
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                const u32 coef = lane + 32 * k;
                const u32 v = csub(x[k], kQ);
                if (P.trace) P.trace[((size_t)job * 2 + j) * kN + coef] = v;
                if (j == 0) {       // a'(0) = a(0), a'(N - k) = -a(k)
                    const u32 t = (coef == 0) ? v : (v == 0 ? 0u : kQ - v);
                    ext[(kN - coef) & (kN - 1)] = (u16)modswitch_Q_to_qKS(t);
                } else if (coef == 0) {
                    u32 bsum = v + (kQ / 8 + 1); if (bsum >= kQ) bsum -= kQ;
                    ext[kN] = (u16)modswitch_Q_to_qKS(bsum);
                }
            }
        }
    }
    env.sync();      // smem is reused by the next group
}

}  // namespace herdb
