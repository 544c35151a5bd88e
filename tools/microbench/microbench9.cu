// tools/microbench9.cu -- what does a 64-bit multiply-accumulate (IMAD.WIDE.U32 Rd, Ra, Rb, Rd) cost in the
// operand patterns of phase B (br_core.cuh)?  One block per SM, 8 / 16 / 32 warps per SM.
//   T0  acc[i] += a[i] * b            (one shared multiplier, 16 accumulators)      -- the earlier "1 slot" result
//   T1  4x4 tile from shared memory, compiler order   acc[c][s][p] += d[c][p] * k[s][p]   (64 accumulators)
//   T2  same tile, 32-bit low accumulate + IMAD.HI high accumulate (3 slots expected)
//   T3  same tile, product without addend (IMAD.WIDE ..., RZ) then 64-bit add on the ALU pipe
//   T4  plain IMAD (32-bit) with the same operand pattern: the 1-slot reference
// Prints clocks per warp-instruction per SMSP (2.0 = IMAD rate).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
struct alignas(16) V4 { u32 v[4]; };

template <int MODE, int THREADS> __global__ void __launch_bounds__(THREADS, 1) k(u32* out, int iters) {
    extern __shared__ __align__(16) u32 sm[];
    const u32 tid = threadIdx.x;
    for (u32 w = tid; w < 8 * 4 * THREADS; w += THREADS) sm[w] = w * 2654435761u + blockIdx.x;
    __syncthreads();
    u32 r = 0;
    if (MODE == 0) {
        u64 acc[16]; u32 a[16]; u32 b = sm[tid];
#pragma unroll
        for (int i = 0; i < 16; ++i) { acc[i] = 0; a[i] = sm[tid + i * 7]; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int rep = 0; rep < 16; ++rep)
#pragma unroll
                for (int i = 0; i < 16; ++i) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) r ^= (u32)acc[i] ^ (u32)(acc[i] >> 32);
    } else {
        u64 acc[4][4][4]; u32 lo[4][4][4], hi[4][4][4];
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int p = 0; p < 4; ++p) { acc[c][s][p] = 0; lo[c][s][p] = 0; hi[c][s][p] = 0; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                V4 kv[4], dv[4];
#pragma unroll
                for (int s = 0; s < 4; ++s) kv[s] = *reinterpret_cast<const V4*>(&sm[(s * THREADS + ((tid + it + rr) % THREADS)) * 4]);
#pragma unroll
                for (int c = 0; c < 4; ++c) dv[c] = *reinterpret_cast<const V4*>(&sm[((4 + c) * THREADS + ((tid + it + rr) % THREADS)) * 4]);
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int s = 0; s < 4; ++s)
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            if (MODE == 1) acc[c][s][p] += (u64)dv[c].v[p] * kv[s].v[p];
                            if (MODE == 2) { lo[c][s][p] += dv[c].v[p] * kv[s].v[p]; hi[c][s][p] += __umulhi(dv[c].v[p], kv[s].v[p]); }
                            if (MODE == 3) {
                                asm volatile("{ .reg .u64 pr; .reg .u32 pl, ph; mul.wide.u32 pr, %2, %3; mov.b64 {pl, ph}, pr;\n"
                                             "add.cc.u32 %0, %0, pl; addc.u32 %1, %1, ph; }" : "+r"(lo[c][s][p]), "+r"(hi[c][s][p])
                                             : "r"(dv[c].v[p]), "r"(kv[s].v[p]));
                            }
                            if (MODE == 4) lo[c][s][p] = dv[c].v[p] * kv[s].v[p] + lo[c][s][p];
                        }
                if (MODE == 5) {        // multiplier (second operand) fixed for 4 consecutive instructions
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int s = 0; s < 4; ++s)
#pragma unroll
                            for (int c = 0; c < 4; ++c) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[c][s][p]) : "r"(dv[c].v[p]), "r"(kv[s].v[p]));
                }
                if (MODE == 6) {        // multiplicand (first operand) fixed for 4 consecutive instructions
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int c = 0; c < 4; ++c)
#pragma unroll
                            for (int s = 0; s < 4; ++s) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[c][s][p]) : "r"(dv[c].v[p]), "r"(kv[s].v[p]));
                }
                if (MODE == 7) {        // both operands change on every instruction (diagonal walk of the tile)
#pragma unroll
                    for (int p = 0; p < 4; ++p)
#pragma unroll
                        for (int d = 0; d < 4; ++d)
#pragma unroll
                            for (int c = 0; c < 4; ++c) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[c][(c + d) & 3][p]) : "r"(dv[c].v[p]), "r"(kv[(c + d) & 3].v[p]));
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int p = 0; p < 4; ++p) r ^= (u32)acc[c][s][p] ^ (u32)(acc[c][s][p] >> 32) ^ lo[c][s][p] ^ hi[c][s][p];
    }
    out[blockIdx.x * THREADS + tid] = r;
}
template <int MODE, int THREADS> void run(const char* name, u32* out, int sms, double instr_per_iter) {
    const int iters = 4000;
    const size_t smem = (size_t)8 * 4 * THREADS * 4 > 100 * 1024 ? (size_t)8 * 4 * THREADS * 4 : 120 * 1024;   // > half the SM: one block per SM
    cudaFuncSetAttribute(k<MODE, THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<MODE, THREADS><<<sms, THREADS, smem>>>(out, 8); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE, THREADS><<<sms, THREADS, smem>>>(out, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double clk = ms * 1e-3 * 1.965e9;
    const double warp_instr_per_smsp = instr_per_iter * iters * (THREADS / 32) / 4.0;
    printf("%-44s %4d thr  %8.3f ms  %.2f clk per warp-MAC per SMSP  [%s]\n", name, THREADS, ms, clk / warp_instr_per_smsp,
           cudaGetErrorString(cudaGetLastError()));
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    u32* out; cudaMalloc(&out, (size_t)sms * 1024 * 4);
    run<0, 256>("T0 acc[i] += a[i]*b (16 acc)", out, sms, 256); run<0, 512>("T0 acc[i] += a[i]*b (16 acc)", out, sms, 256);
    run<0, 1024>("T0 acc[i] += a[i]*b (16 acc)", out, sms, 256);
    run<1, 256>("T1 tile 4x4x4 IMAD.WIDE acc64", out, sms, 256); run<1, 512>("T1 tile 4x4x4 IMAD.WIDE acc64", out, sms, 256);
    run<2, 256>("T2 tile lo IMAD + hi IMAD.HI", out, sms, 256);
    run<3, 256>("T3 tile IMAD.WIDE(RZ) + 64-bit add on ALU", out, sms, 256); run<3, 512>("T3 tile IMAD.WIDE(RZ) + 64-bit add on ALU", out, sms, 256);
    run<4, 256>("T4 tile IMAD 32-bit", out, sms, 256); run<4, 512>("T4 tile IMAD 32-bit", out, sms, 256);
    run<5, 256>("T5 tile asm order: multiplier fixed x4", out, sms, 256); run<5, 512>("T5 tile asm order: multiplier fixed x4", out, sms, 256);
    run<6, 256>("T6 tile asm order: multiplicand fixed x4", out, sms, 256); run<6, 512>("T6 tile asm order: multiplicand fixed x4", out, sms, 256);
    run<7, 256>("T7 tile asm order: diagonal", out, sms, 256); run<7, 512>("T7 tile asm order: diagonal", out, sms, 256);
    return 0;
}
