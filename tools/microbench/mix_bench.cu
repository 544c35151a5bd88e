// tools/mix_bench.cu -- does it pay to run the two phases of the blind-rotation step against each other?
// One CTA of 8 warps per SM, shared memory as in the product kernel.  Phase A is the product's own code
// (inverse NTT, decomposition, three forward NTTs; ntt_core.cuh); phase B is a stand-in with the product's
// instruction mix per thread and step (8 rows x 64 IMAD.WIDE from two 4x4 operand tiles in shared memory, then
// 32 outputs x (3 fold62 + 2 IMAD.WIDE + canon + accumulate)).
//   mode 0: all 8 warps run A, CTA barrier, all run B, CTA barrier                     (today's schedule)
//   mode 1: warps 0-3 run A while warps 4-7 run B, then swap; the halves only meet their own named barrier
//           (each scheduler always holds one warp of each phase)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mix_bench mix_bench.cu
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
    u32 kst[2][4][kN];
};

__device__ __forceinline__ void phase_a(Smem& S, int c, int j, u32 lane) {
    u32 x[32];
    load_contig(S.acc[c][j], lane, x);
    inv_pass1<1, 2>(x, S.itw, lane);
    store_contig(S.dig[c][j], lane, x);
    __syncwarp();
    load_strided(S.dig[c][j], lane, x);
    inv_pass2<2, 1>(x);
    constexpr u32 kB = 64u * (1u + 128u + 128u * 128u + 128u * 128u * 128u);
    u32 y[kDGT][32];
#pragma unroll
    for (int kk = 0; kk < 32; ++kk) {
        u32 t = csub(x[kk], kQ);
        const u32 d = (t < (kQ >> 1)) ? t + kB : t + (kB - kQ);
#pragma unroll
        for (int l = 0; l < (int)kDGT; ++l) y[l][kk] = (d >> (7 * l)) & 127u;
    }
    __syncwarp();
    fwd_pass1_digits_multi<(int)kDGT>(y);
#pragma unroll
    for (int l = 0; l < (int)kDGT; ++l) store_strided(S.dig[c][j + 2 * l], lane, y[l]);
    __syncwarp();
#pragma unroll
    for (int l = 0; l < (int)kDGT; ++l) load_contig(S.dig[c][j + 2 * l], lane, y[l]);
    fwd_pass2_stream_multi<(int)kDGT>(y, S.ftw, lane);
#pragma unroll
    for (int l = 0; l < (int)kDGT; ++l) store_contig(S.dig[c][j + 2 * l], lane, y[l]);
}

// tid2: 0..255 position of this thread in the point tiling (4 points x 4 ciphertexts per thread)
__device__ __forceinline__ void phase_b(Smem& S, u32 tid2, u32 salt) {
    unsigned long long mac[4][4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int s = 0; s < 4; ++s)
#pragma unroll
            for (int p = 0; p < 4; ++p) mac[c][s][p] = 0;
#pragma unroll 1
    for (int r = 0; r < 8; ++r) {
        V4 kv[4], dv[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) kv[s] = *reinterpret_cast<const V4*>(&S.kst[r & 1][s][4 * tid2]);
#pragma unroll
        for (int c = 0; c < 4; ++c) dv[c] = *reinterpret_cast<const V4*>(r < 6 ? &S.dig[c][r][4 * tid2] : &S.acc[c][r - 6][4 * tid2]);
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int p = 0; p < 4; ++p) mac[c][s][p] += (unsigned long long)dv[c].v[p] * kv[s].v[p];
    }
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            V4 av = *reinterpret_cast<const V4*>(&S.acc[c][k][4 * tid2]);
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const u32 rp = fold62(mac[c][k][p] >> 2), rn = fold62(mac[c][2 + k][p] >> 2);
                const u32 mp = (salt + p * 77u + c) & (kQ - 1), mn = (salt * 3u + p) & (kQ - 1);
                const u32 yv = fold62(((unsigned long long)mp * rp + (unsigned long long)mn * rn) >> 2);
                av.v[p] = canon32(av.v[p] + yv);
            }
            *reinterpret_cast<V4*>(&S.acc[c][k][4 * tid2]) = av;
        }
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(const u32* ftw_g, const u32* itw_g, u32* out, int iters) {
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (u32 w = tid; w < kLaneTwWords; w += 256) { S.itw[w] = itw_g[w]; S.ftw[w] = ftw_g[w]; }
    for (u32 w = tid; w < 8 * kN; w += 256) { (&S.acc[0][0][0])[w] = (w * 2654435761u + blockIdx.x * 97u) % kQ; (&S.kst[0][0][0])[w] = (w * 40503u + 11u) % kQ; }
    __syncthreads();
    if (MODE == 0) {
        for (int it = 0; it < iters; ++it) {
            phase_a(S, warp >> 1, warp & 1, lane);
            __syncthreads();
            phase_b(S, tid, it);
            __syncthreads();
        }
    } else {
        // half h works on ciphertexts {2h, 2h+1} in phase A (its 4 warps = 4 polynomials) and on half of the points in phase B
        const int h = warp >> 2, wl = warp & 3;
        const int bar = 1 + h;
        for (int it = 0; it < 2 * iters; ++it) {
            if (((it + h) & 1) == 0) {
                // this half's phase A: its 4 warps transform 4 of the 8 polynomials; twice per two sub-steps
#pragma unroll 1
                for (int rep = 0; rep < 2; ++rep) phase_a(S, 2 * h + (wl >> 1), wl & 1, lane);
            } else {
#pragma unroll 1
                for (int rep = 0; rep < 2; ++rep) phase_b(S, rep * 128 + (tid & 127), it);
            }
            asm volatile("bar.sync %0, 128;" ::"r"(bar) : "memory");
        }
    }
    if (blockIdx.x == 0)
        for (u32 w = tid; w < 8 * kN; w += 256) out[w] = (&S.acc[0][0][0])[w];
}

template <int MODE> void run(const u32* dftw, const u32* ditw, u32* dout, int sms) {
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    const int iters = 300;
    k<MODE><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dout, 4); cudaDeviceSynchronize();
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<sms, 256, sizeof(Smem)>>>(dftw, ditw, dout, iters);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k<MODE>);
    printf("mode %d (%d regs, %zu B local): %8.3f ms  %7.0f clk per step-equivalent (8 phase-A polynomials + 256 phase-B thread tiles)  [%s]\n",
           MODE, fa.numRegs, (size_t)fa.localSizeBytes, ms, ms * 1e-3 * 1.965e9 / iters / (MODE ? 2 : 1), cudaGetErrorString(cudaGetLastError()));
}

int main() {
    std::vector<u32> ftw(kLaneTwWords), itw(kLaneTwWords);
    build_lane_twiddles(ftw.data(), itw.data());
    u32 *dftw, *ditw, *dout;
    cudaMalloc(&dftw, ftw.size() * 4); cudaMalloc(&ditw, itw.size() * 4); cudaMalloc(&dout, 8 * kN * 4);
    cudaMemcpy(dftw, ftw.data(), ftw.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(ditw, itw.data(), itw.size() * 4, cudaMemcpyHostToDevice);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int rep = 0; rep < 2; ++rep) { run<0>(dftw, ditw, dout, sms); run<1>(dftw, ditw, dout, sms); }
    return 0;
}
