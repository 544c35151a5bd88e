// tools/ntt_bench.cu -- phase A of the blind-rotation step in isolation (one warp per polynomial:
// inverse NTT, signed digit decomposition, four forward NTTs), classic Shoup passes (VAR 0) against
// the FP64-quotient hybrid passes (VAR 1).  Prints clocks per phase A per SM and checks that both
// variants produce congruent digit NTTs.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o ntt_bench ntt_bench.cu
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "experimental_ntt.cuh"
using namespace herdb;

struct Smem {
    u32 acc[4][2][kN];
    u32 dig[4][kDG2][kN];
    u32 itw[kLaneTwWords];
    u32 ftw[kLaneTwHybWords];
};

template <int VAR>
__global__ void __launch_bounds__(256, 1) k(const u32* ftw_g, const u32* itw_g, const u32* ftw_hyb_g, u32* out, int iters) {
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (u32 w = tid; w < kLaneTwWords; w += 256) S.itw[w] = itw_g[w];
    if (VAR == 0 || VAR >= 4) { for (u32 w = tid; w < kLaneTwWords; w += 256) S.ftw[w] = ftw_g[w]; }
    else { for (u32 w = tid; w < kLaneTwHybWords; w += 256) S.ftw[w] = ftw_hyb_g[w]; }
    for (u32 w = tid; w < 8 * kN; w += 256) (&S.acc[0][0][0])[w] = (w * 2654435761u + blockIdx.x * 97u) % kQ;
    __syncthreads();
    const int c = warp >> 1, j = warp & 1;
    for (int it = 0; it < iters; ++it) {
        u32 x[32];
        load_contig(S.acc[c][j], lane, x);
        inv_pass1<1, 2>(x, S.itw, lane);
        store_contig(S.dig[c][j], lane, x);
        __syncwarp();
        load_strided(S.dig[c][j], lane, x);
        inv_pass2<2, 1>(x);
        constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
        u32 d[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            u32 t = csub(x[k], kQ);
            d[k] = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
        }
        __syncwarp();
        if (VAR == 0) {
            u32 ftw[62];
#pragma unroll
            for (int g = 0; g < 16; ++g) {
                V4 q = *reinterpret_cast<const V4*>(&S.ftw[(g * 32 + lane) * 4]);
                ftw[4 * g] = q.v[0]; ftw[4 * g + 1] = q.v[1];
                if (g < 15) { ftw[4 * g + 2] = q.v[2]; ftw[4 * g + 3] = q.v[3]; }
            }
#pragma unroll
            for (int l = 0; l < 4; ++l) {
#pragma unroll
                for (int kk = 0; kk < 32; ++kk) x[kk] = (l == 3) ? (d[kk] >> (7 * l)) : ((d[kk] >> (7 * l)) & 127u);
                u32* o = S.dig[c][j + 2 * l];
                fwd_pass1_digits(x);
                store_strided(o, lane, x);
                __syncwarp();
                load_contig(o, lane, x);
                fwd_pass2(x, ftw);
                store_contig(o, lane, x);
            }
        } else if (VAR == 4) {
#pragma unroll
            for (int l = 0; l < 4; ++l) {
#pragma unroll
                for (int kk = 0; kk < 32; ++kk) x[kk] = (l == 3) ? (d[kk] >> (7 * l)) : ((d[kk] >> (7 * l)) & 127u);
                u32* o = S.dig[c][j + 2 * l];
                fwd_pass1_digits(x);
                store_strided(o, lane, x);
                __syncwarp();
                load_contig(o, lane, x);
                fwd_pass2_stream(x, S.ftw, lane);
                store_contig(o, lane, x);
            }
        } else if (VAR == 5) {
#pragma unroll
            for (int l = 0; l < 4; l += 2) {
                u32 y[32];
#pragma unroll
                for (int kk = 0; kk < 32; ++kk) {
                    x[kk] = (d[kk] >> (7 * l)) & 127u;
                    y[kk] = (l + 1 == 3) ? (d[kk] >> (7 * (l + 1))) : ((d[kk] >> (7 * (l + 1))) & 127u);
                }
                u32* o = S.dig[c][j + 2 * l];
                u32* o2 = S.dig[c][j + 2 * (l + 1)];
                fwd_pass1_digits_x2(x, y);
                store_strided(o, lane, x); store_strided(o2, lane, y);
                __syncwarp();
                load_contig(o, lane, x); load_contig(o2, lane, y);
                fwd_pass2_stream_x2(x, y, S.ftw, lane);
                store_contig(o, lane, x); store_contig(o2, lane, y);
            }
        } else {
            u32 tf[4 * kFwd2F64Tw], ts[2 * kFwd2ShoupTw];
#pragma unroll
            for (int g = 0; g < kFwd2F64Tw; ++g) {
                V4 q = *reinterpret_cast<const V4*>(&S.ftw[(g * 32 + lane) * 4]);
                tf[4 * g] = q.v[0]; tf[4 * g + 1] = q.v[1]; tf[4 * g + 2] = q.v[2]; tf[4 * g + 3] = q.v[3];
            }
#pragma unroll
            for (int g = 0; g < kFwd2ShoupTw / 2; ++g) {
                V4 q = *reinterpret_cast<const V4*>(&S.ftw[((kFwd2F64Tw + g) * 32 + lane) * 4]);
                ts[4 * g] = q.v[0]; ts[4 * g + 1] = q.v[1]; ts[4 * g + 2] = q.v[2]; ts[4 * g + 3] = q.v[3];
            }
#pragma unroll
            for (int l = 0; l < 4; ++l) {
#pragma unroll
                for (int kk = 0; kk < 32; ++kk) x[kk] = (l == 3) ? (d[kk] >> (7 * l)) : ((d[kk] >> (7 * l)) & 127u);
                u32* o = S.dig[c][j + 2 * l];
                fwd_pass1_digits_q2(x);
                store_strided(o, lane, x);
                __syncwarp();
                load_contig(o, lane, x);
                fwd_pass2_hyb(x, tf, ts);
                store_contig(o, lane, x);
            }
        }
        __syncthreads();
        // feed something back so that iterations differ
        if (it + 1 < iters) {
            u32 y[32];
            load_contig(S.dig[c][j + 2], lane, y);
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) y[kk] = canon32(y[kk]);
            store_contig(S.acc[c][j], lane, y);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0)
        for (u32 w = tid; w < 4 * kDG2 * kN; w += 256) out[w] = canon32((&S.dig[0][0][0])[w]);
}

// 16 warps per SM: warp = (ciphertext, component, half); both halves run the inverse NTT of their
// polynomial, each transforms two of its four digit polynomials; lane twiddles streamed from smem
template <int VAR>
__global__ void __launch_bounds__(512, 1) k16(const u32* ftw_g, const u32* itw_g, const u32* ftw_hyb_g, u32* out, int iters) {
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (u32 w = tid; w < kLaneTwWords; w += 512) S.itw[w] = itw_g[w];
    if (VAR == 3) { for (u32 w = tid; w < kLaneTwWords; w += 512) S.ftw[w] = ftw_g[w]; }
    else { for (u32 w = tid; w < kLaneTwHybWords; w += 512) S.ftw[w] = ftw_hyb_g[w]; }
    for (u32 w = tid; w < 8 * kN; w += 512) (&S.acc[0][0][0])[w] = (w * 2654435761u + blockIdx.x * 97u) % kQ;
    __syncthreads();
    const int c = warp >> 2, j = (warp >> 1) & 1, h = warp & 1;
    for (int it = 0; it < iters; ++it) {
        u32 x[32];
        u32* scratch = S.dig[c][j + 4 * h];
        load_contig(S.acc[c][j], lane, x);
        inv_pass1<1, 2>(x, S.itw, lane);
        store_contig(scratch, lane, x);
        __syncwarp();
        load_strided(scratch, lane, x);
        inv_pass2<2, 1>(x);
        constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
        u32 d[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            u32 t = csub(x[k], kQ);
            t = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
            d[k] = t >> (14 * h);                 // this warp's two 7-bit fields
        }
        __syncwarp();
#pragma unroll
        for (int l2 = 0; l2 < 2; ++l2) {
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) x[kk] = (l2 == 0) ? (d[kk] & 127u) : (h ? (d[kk] >> 7) : ((d[kk] >> 7) & 127u));
            u32* o = S.dig[c][j + 2 * (2 * h + l2)];
            if (VAR == 3) fwd_pass1_digits(x); else fwd_pass1_digits_q2(x);
            store_strided(o, lane, x);
            __syncwarp();
            load_contig(o, lane, x);
            if (VAR == 3) fwd_pass2_stream(x, S.ftw, lane); else fwd_pass2_hyb_stream(x, S.ftw, lane);
            store_contig(o, lane, x);
        }
        __syncthreads();
        if (it + 1 < iters && h == 0) {
            u32 y[32];
            load_contig(S.dig[c][j + 2], lane, y);
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) y[kk] = canon32(y[kk]);
            store_contig(S.acc[c][j], lane, y);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0)
        for (u32 w = tid; w < 4 * kDG2 * kN; w += 512) out[w] = canon32((&S.dig[0][0][0])[w]);
}

int main() {
    std::vector<u32> ftw(kLaneTwWords), itw(kLaneTwWords), hyb(kLaneTwHybWords);
    build_lane_twiddles(ftw.data(), itw.data());
    build_lane_twiddles_hyb(hyb.data());
    const UniformF64 uf = make_uniform_f64(false), ui = make_uniform_f64(true);
    cudaMemcpyToSymbol(c_f64u_fwd, &uf, sizeof uf); cudaMemcpyToSymbol(c_f64u_inv, &ui, sizeof ui);
    u32 *dftw, *ditw, *dhyb, *dout[6];
    cudaMalloc(&dftw, ftw.size() * 4); cudaMalloc(&ditw, itw.size() * 4); cudaMalloc(&dhyb, hyb.size() * 4);
    cudaMemcpy(dftw, ftw.data(), ftw.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(ditw, itw.data(), itw.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dhyb, hyb.data(), hyb.size() * 4, cudaMemcpyHostToDevice);
    const size_t outw = 4 * kDG2 * kN;
    for (int v = 0; v < 6; ++v) cudaMalloc(&dout[v], outw * 4);
    cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(k<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(k<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(k16<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(k16<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    int dev = 0, sms = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int iters = 200;
    for (int v = 0; v < 6; ++v) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
            cudaEventRecord(a);
            if (v == 0) k<0><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[0], iters);
            else if (v == 1) k<1><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[1], iters);
            else if (v == 2) k16<2><<<sms, 512, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[2], iters);
            else if (v == 3) k16<3><<<sms, 512, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[3], iters);
            else if (v == 4) k<4><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[4], iters);
            else k<5><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dhyb, dout[5], iters);
            cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            cudaError_t e = cudaGetLastError();
            if (rep) printf("VAR %d: %.3f ms, %.0f clk per phase A (8 warps x (1 INTT + 4 NTT)) at %d MHz  [%s]\n", v, ms,
                            ms * 1e-3 * khz * 1e3 / iters, khz / 1000, cudaGetErrorString(e));
        }
    }
    std::vector<u32> o0(outw), o1(outw);
    cudaMemcpy(o0.data(), dout[0], outw * 4, cudaMemcpyDeviceToHost);
    size_t allbad = 0;
    for (int v = 1; v < 6; ++v) {
        cudaMemcpy(o1.data(), dout[v], outw * 4, cudaMemcpyDeviceToHost);
        size_t bad = 0; for (size_t i = 0; i < outw; ++i) bad += o0[i] != o1[i];
        printf("mismatches VAR%d vs VAR0: %zu of %zu\n", v, bad, outw);
        allbad += bad;
    }
    return allbad != 0;
}
