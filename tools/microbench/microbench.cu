// tools/microbench.cu -- integer-pipe issue-rate microbenchmarks for sm_100a (B200).
// Prints lane-ops per clock per SM for single instructions and for the butterfly mixes the
// NTT kernels use.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
#define Q 134215681u
#define ITERS 2048
#define CH 8

template <int MODE> __global__ void __launch_bounds__(256) k(u32* out, u32 seed, long long* cyc) {
    u32 a[CH], h[CH]; u32 b = seed | 1u, c = seed * 2654435761u + threadIdx.x, w = seed ^ 0x5bd1e995u;
    u64 acc[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { a[i] = threadIdx.x * 97u + i * 13u + seed; h[i] = a[i] ^ 0xabcdu; acc[i] = a[i]; }
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (MODE == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 1) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
                if (MODE == 2) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
                if (MODE == 3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
                if (MODE == 4) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
                if (MODE == 5) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(c));
                if (MODE == 6) asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));
                if (MODE == 7) {  // Shoup form A: HI + MUL + MAD + 2 ADD
                    u32 q = __umulhi(a[i], b); u32 v = a[i] * w - q * Q; u32 x = h[i]; h[i] = x + v; a[i] = x - v + 2 * Q;
                }
                if (MODE == 8) {  // Shoup form B: HI + MAD + shifts/adds
                    u32 q = __umulhi(a[i], b); u32 t = a[i] * w - q; u32 g = q << 11; u32 v = t + g - (g << 16);
                    u32 x = h[i]; h[i] = x + v; a[i] = x - v + 2 * Q;
                }
                if (MODE == 9) {  // IMAD + FFMA interleave: does FP32 issue beside IMAD?
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                    float f = __uint_as_float(h[i]); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(1.0001f), "f"(0.5f)); h[i] = __float_as_uint(f);
                }
                if (MODE == 10) { // IMAD + IADD interleave
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(c));
                }
                if (MODE == 11) { // IMAD + 2 IADD
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(c));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(b));
                }
                if (MODE == 12) { // IMAD.HI + IMAD.WIDE? check hi via wide
                    u64 p = (u64)a[i] * b; a[i] = (u32)(p >> 32) + (u32)p;
                }
                if (MODE == 13) { double d = __longlong_as_double(acc[i]); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d) : "d"(1.0000001), "d"(0.5)); acc[i] = __double_as_longlong(d); }
                if (MODE == 14) { // HI + 3 ADD
                    asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(c));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(b));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(w));
                }
                if (MODE == 15) { // mad.wide + 2 ADD
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(c));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(h[i]) : "r"(b));
                }
            }
        }
    }
    long long t1 = clock64();
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) r ^= a[i] ^ h[i] ^ (u32)acc[i] ^ (u32)(acc[i] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char* name, int stmts, int blocks_per_sm) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int blocks = sms * blocks_per_sm;
    u32* out; long long* cyc; cudaMalloc(&out, blocks * 256 * 4); cudaMalloc(&cyc, blocks * 8);
    k<MODE><<<blocks, 256>>>(out, 12345u, cyc); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(out, 12345u, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long hc[1]; cudaMemcpy(hc, cyc, 8, cudaMemcpyDeviceToHost);
    double groups = (double)ITERS * 4 * CH;                       // statement groups per thread
    double lane_groups_per_clk_sm = groups * 256 * blocks_per_sm / (double)hc[0];
    printf("%-28s %2d instr/group  groups/clk/SM %7.2f  instr/clk/SM %7.2f  (%.3f ms, %.2f T groups/s)\n", name, stmts,
           lane_groups_per_clk_sm, lane_groups_per_clk_sm * stmts, ms, groups * 256 * blocks / (ms * 1e-3) / 1e12);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int bps : {4, 8}) {
        printf("--- %d blocks/SM (256 thr)\n", bps);
        run<0>("IMAD", 1, bps); run<1>("IMAD.HI", 1, bps); run<2>("IMAD.WIDE acc64", 1, bps); run<3>("IADD", 1, bps);
        run<4>("LOP3", 1, bps); run<5>("SHF", 1, bps); run<6>("IMNMX", 1, bps);
        run<7>("shoup A (HI,MUL,MAD,2ADD)", 5, bps); run<8>("shoup B (HI,MAD,shift/add)", 7, bps);
        run<9>("IMAD+FFMA", 2, bps); run<10>("IMAD+IADD", 2, bps); run<11>("IMAD+2IADD", 3, bps);
        run<12>("mul64 hi+lo", 2, bps); run<13>("DFMA", 1, bps); run<14>("HI+3ADD", 4, bps); run<15>("WIDE+2ADD", 3, bps);
    }
    return 0;
}
