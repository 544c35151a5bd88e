// tools/microbench5.cu -- forward pass 1 (5 stages x 16 butterflies on 32 registers, uniform twiddles)
// with the three quotient forms; butterflies per clock per SM at 2 warps per scheduler (as in the kernel)
#include <cstdio>
#include <cuda_runtime.h>
#include "experimental_ntt.cuh"
using namespace herdb;
struct UTw64 { F64Tw v[32]; };
__constant__ UTw64 c_tw64;
template <int MODE> __global__ void __launch_bounds__(256, 1) k(u32* out, int iters) {
    u32 x[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = 8 * kQ + (threadIdx.x * 977u + i * 131u + blockIdx.x) % 127;
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
                    u32 r = (MODE == 1) ? mul_f64(c_tw64.v[idx].w, c_tw64.v[idx].c, c_tw64.v[idx].k, x[kk + h])
                                        : mul_f64_2op(c_tw64.v[idx].w, c_tw64.v[idx].c, x[kk + h]);
                    u32 u = x[kk];
                    x[kk] = u + r; x[kk + h] = u - r;
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = (x[i] & 127u) + 8 * kQ;       // cheap re-initialisation (1 LOP + 1 ADD per 2.5 butterflies)
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) r ^= x[i];
    out[blockIdx.x * 256 + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, u32* out, int sms) {
    const int iters = 2000;
    k<MODE><<<sms, 256>>>(out, 8); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<sms, 256>>>(out, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double bf = (double)iters * 80 * 256 * sms;
    printf("%-30s %.3f ms  %5.2f butterflies/clk/SM -> %.2f clk per warp-butterfly per SMSP [%s]\n", name, ms,
           bf / (ms * 1e-3) / sms / 1.965e9, 128.0 / (bf / (ms * 1e-3) / sms / 1.965e9), cudaGetErrorString(cudaGetLastError()));
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    UTw64 h; for (u32 i = 0; i < 32; ++i) h.v[i] = make_f64tw(i ? ctw_fwd(i) : 1);
    cudaMemcpyToSymbol(c_tw64, &h, sizeof h);
    u32* out; cudaMalloc(&out, sms * 1024);
    run<0>("pass1 classic (IMAD.HI)", out, sms); run<1>("pass1 DFMA (1 FP64 op)", out, sms); run<2>("pass1 DADD+DFMA", out, sms);
    return 0;
}
