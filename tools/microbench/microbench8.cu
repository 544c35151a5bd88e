// tools/microbench8.cu -- forward pass 1 (5 stages x 16 butterflies on 32 registers, uniform twiddles):
//   M0  classic unsigned lazy Shoup            IMAD.HI, IMAD, IMAD, IADD, IADD3(+2Q)
//   M1  signed centred Shoup                   IMAD.HI.S32, IMAD, IMAD, IADD, IADD   (r in (-Q/4, 5Q/4), values signed)
//   M2  signed centred, q*Q by shifts          IMAD.HI.S32, IMAD, LEA, LEA, IADD3, IADD3
// Prints clocks per warp-butterfly per SMSP at 2 and 4 warps per scheduler and checks M1/M2 against M0.
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "experimental_ntt.cuh"
using namespace herdb;
struct STw { int w[32]; int wp[32]; };
__constant__ STw c_stw;

__device__ __forceinline__ void bfly_s(int& a, int& b, int w, int wp) {      // M1
    const int q = __mulhi(wp, b);
    const int r = w * b - q * (int)kQ;
    const int u = a;
    a = u + r; b = u - r;
}
__device__ __forceinline__ void bfly_sh(int& a, int& b, int w, int wp) {     // M2
    const int q = __mulhi(wp, b);
    const int t = w * b;
    const int v1 = q + (int)((u32)q << 27);
    const int t2 = t + (int)((u32)q << 11);
    const int u = a;
    a = u + t2 - v1; b = u - t2 + v1;
}
template <int MODE, int THREADS> __global__ void __launch_bounds__(THREADS, 1) k(u32* out, int iters) {
    u32 x[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = (threadIdx.x * 977u + i * 131u + blockIdx.x) % 127;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) fwd_pass1(x);
        else {
#pragma unroll
            for (int s = 0; s < 5; ++s) {
                const int h = 16 >> s;
#pragma unroll
                for (int kk = 0; kk < 32; ++kk) {
                    if (kk & h) continue;
                    const int idx = (1 << s) + kk / (2 * h);
                    int a = (int)x[kk], b = (int)x[kk + h];
                    if (MODE == 1) bfly_s(a, b, c_stw.w[idx], c_stw.wp[idx]); else bfly_sh(a, b, c_stw.w[idx], c_stw.wp[idx]);
                    x[kk] = (u32)a; x[kk + h] = (u32)b;
                }
            }
        }
        if (it + 1 < iters) {
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] = x[i] & 127u;       // cheap re-initialisation
        }
    }
    // canonical residues of the last pass
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        long long v = (MODE == 0) ? (long long)x[i] : (long long)(int)x[i];
        out[(blockIdx.x * THREADS + threadIdx.x) * 32 + i] = (u32)(((v % (long long)kQ) + kQ) % kQ);
    }
}
template <int MODE, int THREADS> float run(const char* name, u32* out, int sms) {
    const int iters = 2000;
    k<MODE, THREADS><<<sms, THREADS>>>(out, 8); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE, THREADS><<<sms, THREADS>>>(out, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double bf = (double)iters * 80 * THREADS * sms;
    printf("%-28s %3d thr  %.3f ms  %5.2f butterflies/clk/SM -> %.2f clk per warp-butterfly per SMSP [%s]\n", name, THREADS, ms,
           bf / (ms * 1e-3) / sms / 1.965e9, 128.0 / (bf / (ms * 1e-3) / sms / 1.965e9), cudaGetErrorString(cudaGetLastError()));
    return ms;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    STw h;
    for (u32 i = 0; i < 32; ++i) {
        long long w = i ? ctw_fwd(i) : 1;
        if (w > kQ / 2) w -= kQ;
        h.w[i] = (int)w;
        // round(w * 2^32 / Q)
        __int128 num = (__int128)w * ((__int128)1 << 32);
        __int128 qq = (num >= 0) ? (num + kQ / 2) / kQ : -((-num + kQ / 2) / kQ);
        h.wp[i] = (int)qq;
    }
    cudaMemcpyToSymbol(c_stw, &h, sizeof h);
    u32* out; cudaMalloc(&out, (size_t)sms * 512 * 32 * 4);
    std::vector<u32> r0((size_t)sms * 256 * 32), r1(r0.size());
    run<0, 256>("M0 classic", out, sms); cudaMemcpy(r0.data(), out, r0.size() * 4, cudaMemcpyDeviceToHost);
    run<1, 256>("M1 signed", out, sms); cudaMemcpy(r1.data(), out, r1.size() * 4, cudaMemcpyDeviceToHost);
    size_t bad = 0; for (size_t i = 0; i < r0.size(); ++i) bad += r0[i] != r1[i];
    printf("   M1 vs M0 mismatches: %zu\n", bad);
    run<2, 256>("M2 signed, shifts", out, sms); cudaMemcpy(r1.data(), out, r1.size() * 4, cudaMemcpyDeviceToHost);
    bad = 0; for (size_t i = 0; i < r0.size(); ++i) bad += r0[i] != r1[i];
    printf("   M2 vs M0 mismatches: %zu\n", bad);
    run<0, 512>("M0 classic", out, sms); run<1, 512>("M1 signed", out, sms); run<2, 512>("M2 signed, shifts", out, sms);
    return 0;
}
