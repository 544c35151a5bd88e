// tools/microbench6.cu -- do DFMA (FP64 pipe) and IMAD (FMA-heavy pipe) issue concurrently on B200?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32;
template <int MODE, int WARPS> __global__ void __launch_bounds__(WARPS * 32) k(u32* out, double* dout, u32 seed, int iters) {
    u32 a[8], b = seed | 1u, c = seed * 2654435761u + threadIdx.x; double d[8], e = 1.0000001, f = 0.5;
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 97u + i * 13u + seed; d[i] = 1.0 + i + threadIdx.x; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0 || MODE == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 1 || MODE == 2) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(e), "d"(f));
                if (MODE == 3) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(c), "r"(b));
                                 asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(e), "d"(f)); }
                if (MODE == 4) { asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(e), "d"(f)); }
            }
    }
    u32 r = 0; double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { r ^= a[i]; s += d[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r; dout[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE, int WARPS> void run(const char* name) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    u32* out; double* dout; cudaMalloc(&out, sms * WARPS * 32 * 4); cudaMalloc(&dout, sms * WARPS * 32 * 8);
    int iters = 4096;
    k<MODE, WARPS><<<sms, WARPS * 32>>>(out, dout, 12345u, 16); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE, WARPS><<<sms, WARPS * 32>>>(out, dout, 12345u, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double groups = (double)iters * 64 * WARPS * 32 * sms;
    printf("%-26s warps/SM=%2d  %.3f ms  %6.2f groups/clk/SM -> %.2f clk per warp-group per SMSP\n", name, WARPS, ms,
           groups / (ms * 1e-3) / sms / 1.965e9, 128.0 / (groups / (ms * 1e-3) / sms / 1.965e9));
    cudaFree(out); cudaFree(dout);
}
int main() {
    run<0, 8>("IMAD"); run<1, 8>("DFMA"); run<2, 8>("IMAD + DFMA"); run<3, 8>("2 IMAD + DFMA"); run<4, 8>("IMAD.HI + DFMA");
    run<0, 32>("IMAD"); run<1, 32>("DFMA"); run<2, 32>("IMAD + DFMA"); run<3, 32>("2 IMAD + DFMA"); run<4, 32>("IMAD.HI + DFMA");
    return 0;
}
