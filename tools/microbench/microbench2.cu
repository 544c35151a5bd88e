// tools/microbench2.cu -- candidate NTT-butterfly / MAC instruction forms on sm_100a (B200):
// butterflies per clock per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench2 microbench2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
#define Q 134215681u
#define QINV_NEG 0x0u
#define ITERS 1024
#define CH 8
__device__ __forceinline__ u32 hi_wide(u32 a, u32 b) {   // hi word through IMAD.WIDE
    u64 p; asm("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"(a), "r"(b));
    u32 lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(p)); return hi;
}
template <int MODE> __global__ void __launch_bounds__(256) k(u32* out, u32 seed) {
    u32 y[CH], x[CH]; u32 w = (seed * 2654435761u) % Q, wp = (u32)(((u64)w << 32) / Q), qi = seed * 40503u | 1u;
    u64 acc[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { y[i] = threadIdx.x * 97u + i * 13u + seed; x[i] = y[i] ^ 0xabcdu; acc[i] = y[i]; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (MODE == 0) { u32 q = __umulhi(wp, y[i]); u32 v = w * y[i] - q * Q; u32 t = x[i]; x[i] = t + v; y[i] = t - v + 2 * Q; }
                if (MODE == 1) { u32 q = hi_wide(wp, y[i]); u32 v = w * y[i] - q * Q; u32 t = x[i]; x[i] = t + v; y[i] = t - v + 2 * Q; }
                if (MODE == 2) { // Montgomery via two wides: T = w*y; m = lo(T)*qi; T += m*Q; v = hi(T)
                    u64 T = (u64)w * y[i]; u32 m = (u32)T * qi; T += (u64)m * Q; u32 v = (u32)(T >> 32);
                    u32 t = x[i]; x[i] = t + v; y[i] = t - v + 2 * Q; }
                if (MODE == 3) { // adds absorbed: x' = (w*y + t) - q*Q ; y' = 2t + 2Q - x'
                    u32 q = hi_wide(wp, y[i]); u32 t = x[i]; u32 xn = w * y[i] + t; xn = xn - q * Q; x[i] = xn; y[i] = 2 * t + 2 * Q - xn; }
                if (MODE == 4) { // GS form with csub: s = csub(u+v), d = shoup(u - v + 2Q)
                    u32 a = x[i], b = y[i]; u32 s = a + b; s = min(s, s - 2 * Q); u32 d = a - b + 2 * Q;
                    u32 q = __umulhi(wp, d); x[i] = s; y[i] = w * d - q * Q; }
                if (MODE == 5) { // GS with wide-hi
                    u32 a = x[i], b = y[i]; u32 s = a + b; s = min(s, s - 2 * Q); u32 d = a - b + 2 * Q;
                    u32 q = hi_wide(wp, d); x[i] = s; y[i] = w * d - q * Q; }
                if (MODE == 6) { acc[i] += (u64)y[i] * w; y[i] += (u32)acc[i]; }     // MAC + 1 add
                if (MODE == 7) { acc[i] += (u64)y[i] * w; }                          // pure MAC
                if (MODE == 8) { // two independent MAC per add
                    acc[i] += (u64)y[i] * w; acc[i] += (u64)x[i] * wp; }
            }
        }
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) r ^= y[i] ^ x[i] ^ (u32)acc[i] ^ (u32)(acc[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, int bps) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    int blocks = sms * bps;
    u32* out; cudaMalloc(&out, blocks * 256 * 4);
    k<MODE><<<blocks, 256>>>(out, 12345u); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(out, 12345u); cudaEventRecord(e1); cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
    double groups = (double)ITERS * 4 * CH * 256 * blocks;
    printf("%-34s bps=%d  %.3f ms  %6.2f T/s  = %5.2f per clk per SM @%.2f GHz nominal\n", name, bps, best, groups / (best * 1e-3) / 1e12,
           groups / (best * 1e-3) / sms / (clk * 1e3), clk * 1e-6);
    cudaFree(out);
}
int main() {
    for (int bps : {2, 8}) {
        run<0>("CT shoup IMAD.HI", bps); run<1>("CT shoup WIDE-hi", bps); run<2>("CT montgomery 2xWIDE", bps); run<3>("CT WIDE-hi adds absorbed", bps);
        run<4>("GS shoup IMAD.HI", bps); run<5>("GS shoup WIDE-hi", bps); run<6>("MAC wide + add", bps); run<7>("MAC wide", bps); run<8>("2 MAC wide", bps);
    }
    return 0;
}
