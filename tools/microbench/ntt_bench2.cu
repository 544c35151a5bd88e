// tools/ntt_bench2.cu -- phase A of the blind-rotation step as the product kernel runs it today (inverse
// NTT, signed digit decomposition, three forward NTTs transformed together, lane twiddles streamed from
// shared memory), timed in isolation with 1, 2, 3 and 4 warps per scheduler.  Answers: how much of the
// IMAD pipe can this code use as a function of occupancy?  (Warps beyond 8 alias the shared-memory
// polynomials of warp % 8: the data is garbage, the instruction stream and its timing are the same.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o ntt_bench2 ntt_bench2.cu
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "experimental_ntt.cuh"
using namespace herdb;

struct Smem {
    u32 acc[4][2][kN];
    u32 dig[4][2 * kDGT][kN];
    u32 itw[kLaneTwWords];
    u32 ftw[kLaneTwWords];
};

template <int WARPS, int REGCAP>
__global__ void __launch_bounds__(WARPS * 32, 1) k(const u32* ftw_g, const u32* itw_g, u32* out, int iters, int skew = 0) {
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = (tid >> 5) & 7;
    for (u32 w = tid; w < kLaneTwWords; w += WARPS * 32) { S.itw[w] = itw_g[w]; S.ftw[w] = ftw_g[w]; }
    for (u32 w = tid; w < 8 * kN; w += WARPS * 32) (&S.acc[0][0][0])[w] = (w * 2654435761u + blockIdx.x * 97u) % kQ;
    __syncthreads();
    const int c = warp >> 1, j = warp & 1;
    for (int it = 0; it < iters; ++it) {
        if (skew && (warp & 4)) { const long long t0 = clock64(); while (clock64() - t0 < skew) { } }
        u32 x[32];
        load_contig(S.acc[c][j], lane, x);
        if (!(REGCAP & 2)) {
            inv_pass1<1, 2>(x, S.itw, lane);
            store_contig(S.dig[c][j], lane, x);
            __syncwarp();
            load_strided(S.dig[c][j], lane, x);
            inv_pass2<2, 1>(x);
        }
        constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
        u32 d[32];
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) {
            if (REGCAP & 4) { d[kk] = x[kk]; continue; }
            u32 t = csub(x[kk], kQ);
            d[kk] = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
        }
        __syncwarp();
        u32 y[kDGT][32];
#pragma unroll
        for (int l = 0; l < (int)kDGT; ++l)
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) y[l][kk] = (d[kk] >> (7 * l)) & 127u;
        fwd_pass1_digits_multi<(int)kDGT>(y);
        if (!(REGCAP & 1)) {
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l) store_strided(S.dig[c][j + 2 * l], lane, y[l]);
            __syncwarp();
#pragma unroll
            for (int l = 0; l < (int)kDGT; ++l) load_contig(S.dig[c][j + 2 * l], lane, y[l]);
        }
        fwd_pass2_stream_multi<(int)kDGT>(y, S.ftw, lane);
#pragma unroll
        for (int l = 0; l < (int)kDGT; ++l) store_contig(S.dig[c][j + 2 * l], lane, y[l]);
        if (REGCAP & 8) continue;
        __syncthreads();
        if (it + 1 < iters) {          // next accumulator: some canonical values
            u32 z[32];
            load_contig(S.dig[c][j + 2], lane, z);
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) z[kk] = canon32(z[kk]);
            store_contig(S.acc[c][j], lane, z);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && tid < 256)
        for (u32 w = tid; w < 4 * 2 * kDGT * kN; w += 256) out[w] = canon32((&S.dig[0][0][0])[w]);
}

// Helper-warp variant: 8 transform warps (inverse NTT, decomposition, forward NTTs of digits 0 and 1) plus 4
// helper warps, one per scheduler, that take the forward NTT of digit 2 of the two polynomials of their scheduler
// (digits handed over as packed bytes in shared memory, named barriers).  12 warps, 3 per scheduler.
struct Smem2 : Smem { u32 d2[8][kN / 4]; };
__global__ void __launch_bounds__(384, 1) k_helper(const u32* ftw_g, const u32* itw_g, u32* out, int iters) {
    extern __shared__ __align__(16) unsigned char raw[];
    Smem2& S = *reinterpret_cast<Smem2*>(raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (u32 w = tid; w < kLaneTwWords; w += 384) { S.itw[w] = itw_g[w]; S.ftw[w] = ftw_g[w]; }
    for (u32 w = tid; w < 8 * kN; w += 384) (&S.acc[0][0][0])[w] = (w * 2654435761u + blockIdx.x * 97u) % kQ;
    __syncthreads();
    for (int it = 0; it < iters; ++it) {
        if (warp < 8) {
            const int c = warp >> 1, j = warp & 1;
            u32 x[32];
            load_contig(S.acc[c][j], lane, x);
            inv_pass1<1, 2>(x, S.itw, lane);
            store_contig(S.dig[c][j], lane, x);
            __syncwarp();
            load_strided(S.dig[c][j], lane, x);
            inv_pass2<2, 1>(x);
            constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
            u32 y[2][32];
#pragma unroll
            for (int m = 0; m < 8; ++m) {
                u32 pk = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int kk = 4 * m + q;
                    u32 t = csub(x[kk], kQ);
                    const u32 d = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
                    y[0][kk] = d & 127u; y[1][kk] = (d >> 7) & 127u;
                    pk |= ((d >> 14) & 127u) << (8 * q);
                }
                S.d2[warp][m * 32 + lane] = pk;
            }
            asm volatile("bar.arrive %0, 64;" ::"r"(1 + (int)warp) : "memory");
            fwd_pass1_digits_multi<2>(y);
            store_strided(S.dig[c][j], lane, y[0]); store_strided(S.dig[c][j + 2], lane, y[1]);
            __syncwarp();
            load_contig(S.dig[c][j], lane, y[0]); load_contig(S.dig[c][j + 2], lane, y[1]);
            fwd_pass2_stream_multi<2>(y, S.ftw, lane);
            store_contig(S.dig[c][j], lane, y[0]); store_contig(S.dig[c][j + 2], lane, y[1]);
        } else {
            const int q = warp - 8;
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                const int pw = q + 4 * half, c = pw >> 1, j = pw & 1;
                asm volatile("bar.sync %0, 64;" ::"r"(1 + pw) : "memory");
                u32 y[1][32];
#pragma unroll
                for (int m = 0; m < 8; ++m) {
                    const u32 pk = S.d2[pw][m * 32 + lane];
#pragma unroll
                    for (int b = 0; b < 4; ++b) y[0][4 * m + b] = (pk >> (8 * b)) & 127u;
                }
                fwd_pass1_digits_multi<1>(y);
                store_strided(S.dig[c][j + 4], lane, y[0]);
                __syncwarp();
                load_contig(S.dig[c][j + 4], lane, y[0]);
                fwd_pass2_stream_multi<1>(y, S.ftw, lane);
                store_contig(S.dig[c][j + 4], lane, y[0]);
            }
        }
        __syncthreads();
        if (it + 1 < iters && warp < 8) {
            const int c = warp >> 1, j = warp & 1;
            u32 z[32];
            load_contig(S.dig[c][j + 2], lane, z);
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) z[kk] = canon32(z[kk]);
            store_contig(S.acc[c][j], lane, z);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && tid < 256)
        for (u32 w = tid; w < 4 * 2 * kDGT * kN; w += 256) out[w] = canon32((&S.dig[0][0][0])[w]);
}
void run_helper(const u32* dftw, const u32* ditw, u32* dout, int sms) {
    cudaFuncSetAttribute(k_helper, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem2));
    const int iters = 400;
    k_helper<<<sms, 384, sizeof(Smem2)>>>(dftw, ditw, dout, 4); cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k_helper<<<sms, 384, sizeof(Smem2)>>>(dftw, ditw, dout, iters);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k_helper);
    const double clk = ms * 1e-3 * 1.965e9 / iters, pipe = 2 * 4 * 160 * 8;
    printf("helper-warp variant, 12 warps/SM (%d regs, %zu B local): %8.3f ms  %7.0f clk per phase A, butterfly pipe need %6.0f -> %.1f %%  [%s]\n",
           fa.numRegs, (size_t)fa.localSizeBytes, ms, clk, pipe, 100.0 * pipe / clk, cudaGetErrorString(cudaGetLastError()));
}

template <int WARPS, int REGCAP>
void run(const u32* dftw, const u32* ditw, u32* dout, int sms) {
    cudaFuncSetAttribute(k<WARPS, REGCAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    const int iters = 400;
    k<WARPS, REGCAP><<<sms, WARPS * 32, sizeof(Smem)>>>(dftw, ditw, dout, 4); cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<WARPS, REGCAP><<<sms, WARPS * 32, sizeof(Smem)>>>(dftw, ditw, dout, iters);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k<WARPS, REGCAP>);
    const double clk = ms * 1e-3 * 1.965e9 / iters;
    const double pipe = (WARPS / 4.0) * ((REGCAP & 2) ? 3 : 4) * 160 * 8;            // warps per scheduler x 4 transforms x 160 warp-butterflies x 8 clocks
    printf("mask %2d  %2d warps/SM (%d regs, %zu B local): %8.3f ms  %7.0f clk per phase A, butterfly pipe need %6.0f -> %.1f %%  [%s]\n", REGCAP, WARPS,
           fa.numRegs, (size_t)fa.localSizeBytes, ms, clk, pipe, 100.0 * pipe / clk, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    std::vector<u32> ftw(kLaneTwWords), itw(kLaneTwWords);
    build_lane_twiddles(ftw.data(), itw.data());
    u32 *dftw, *ditw, *dout;
    cudaMalloc(&dftw, ftw.size() * 4); cudaMalloc(&ditw, itw.size() * 4); cudaMalloc(&dout, 4 * 2 * kDGT * kN * 4);
    cudaMemcpy(dftw, ftw.data(), ftw.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(ditw, itw.data(), itw.size() * 4, cudaMemcpyHostToDevice);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<4, 0>(dftw, ditw, dout, sms);
    run<8, 0>(dftw, ditw, dout, sms);
    run<12, 0>(dftw, ditw, dout, sms);
    run<16, 0>(dftw, ditw, dout, sms);
    // ablations (bit 0: no transposition between the forward passes, 1: no inverse transform, 2: no decomposition
    // arithmetic, 3: no barrier / refresh between iterations)
    run<8, 8>(dftw, ditw, dout, sms); run<8, 9>(dftw, ditw, dout, sms); run<8, 10>(dftw, ditw, dout, sms); run<8, 12>(dftw, ditw, dout, sms);
    run<8, 14>(dftw, ditw, dout, sms); run<8, 15>(dftw, ditw, dout, sms);
    run<12, 8>(dftw, ditw, dout, sms); run<12, 15>(dftw, ditw, dout, sms);
    run_helper(dftw, ditw, dout, sms);
    // second warp of every scheduler started `skew` clocks late in every iteration (8 warps, full phase A)
    for (int skew : {0, 200, 500, 1000, 2000, 4000}) {
        cudaFuncSetAttribute(k<8, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a);
        k<8, 0><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dout, 400, skew);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("skew %4d clk: %8.3f ms  %7.0f clk per phase A  [%s]\n", skew, ms, ms * 1e-3 * 1.965e9 / 400, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
