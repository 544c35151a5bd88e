// tools/microbench7.cu -- what limits IMAD.WIDE (64-bit multiply-accumulate) on sm_100a?  Register-only
// loops at 2 warps per scheduler, as in phase B of blind_rotate:
//   P1  acc[i] += a*b            a, b loop invariant (operand reuse), 64 accumulators
//   P2  acc[c][j] += d[c]*k[j]   4 x 4 outer-product tile, 16 accumulators, operands rotate every iteration
//   P3  acc[c][j][p] += d[c][p]*k[j][p]  4 x 4 x 4 tile, 64 accumulators (the kernel's shape)
//   P4  P3 with the product split: lo += a*b (IMAD), hi += umulhi(a,b) (IMAD.HI)  -- not exact, rate only
//   P5  P3 with non-accumulating IMAD.WIDE and a separate 64-bit add (IADD3 + IADD3.X)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;

template <int MODE> __global__ void __launch_bounds__(256) k(u32* out, u32 seed, int iters) {
    u64 acc[4][4][4];
    u32 d[4][4], kk[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int p = 0; p < 4; ++p) { d[c][p] = seed * (c * 4 + p + 1) + threadIdx.x; kk[c][p] = seed * 77u * (c * 4 + p + 3) + threadIdx.x * 3u; }
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int p = 0; p < 4; ++p) acc[c][j][p] = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int p = 0; p < 4; ++p) {
                        if (MODE == 1) acc[c][j][p] += (u64)d[0][0] * kk[0][0];
                        else if (MODE == 2) acc[c][j][p & 1] += (u64)d[c][p] * kk[j][p];
                        else if (MODE == 3) acc[c][j][p] += (u64)d[c][p] * kk[j][p];
                        else if (MODE == 4) {
                            u32 lo = (u32)acc[c][j][p], hi = (u32)(acc[c][j][p] >> 32);
                            lo += d[c][p] * kk[j][p]; hi += __umulhi(d[c][p], kk[j][p]);
                            acc[c][j][p] = ((u64)hi << 32) | lo;
                        } else {
                            u64 pr = (u64)d[c][p] * kk[j][p];
                            asm volatile("" : "+l"(pr));
                            acc[c][j][p] += pr;
                        }
                    }
            // refresh all operands so that nothing is loop invariant (32 adds per 64 MACs)
            if (MODE != 1) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int p = 0; p < 4; ++p) { d[c][p] += 0x9E3779B9u + rep; kk[c][p] += 0x7F4A7C15u + rep; }
            }
        }
    }
    u64 r = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int p = 0; p < 4; ++p) r ^= acc[c][j][p];
    out[blockIdx.x * 256 + threadIdx.x] = (u32)r ^ (u32)(r >> 32);
}
template <int MODE> void run(const char* name, int bps = 1) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    sms *= bps;
    u32* out; cudaMalloc(&out, sms * 256 * 4);
    const int iters = 4000;
    k<MODE><<<sms, 256>>>(out, 12345u, 10); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<sms, 256>>>(out, 12345u, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double macs = (double)iters * 4 * 64 * 256 * sms;
    const double per_clk_sm = macs / (ms * 1e-3) / (sms / bps) / 1.965e9;
    printf("bps=%d %-44s %.3f ms  %.1f MAC/clk/SM -> %.2f clk per warp-MAC per SMSP\n", bps, name, ms, per_clk_sm, 128.0 / per_clk_sm);
    cudaFree(out);
}
int main() {
    run<1>("P1 invariant operands, 64 acc");
    run<2>("P2 4x4 tile, 16 acc");
    run<3>("P3 4x4x4 tile, 64 acc");
    run<4>("P4 split IMAD + IMAD.HI");
    run<5>("P5 IMAD.WIDE no addend + 64-bit add");
    run<3>("P3 4x4x4 tile, 64 acc", 2);
    run<4>("P4 split IMAD + IMAD.HI", 2);
    return 0;
}
