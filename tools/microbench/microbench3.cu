// tools/microbench3.cu -- does register-bank parity of the IMAD.WIDE sources matter?  Imitates the
// phase-B MAC loop of blind_rotate: 64 u64 accumulators, digit quads from shared memory, key quads
// from global memory; variant 0 multiplies d[p]*k[p] (same register parity), variant 1 d[p]*k[p^1].
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
template <int VAR> __global__ void __launch_bounds__(256, 1) k(const uint4* __restrict__ keys, u32* out, int iters) {
    extern __shared__ uint4 dig_raw[]; uint4 (*dig)[8][256] = (uint4 (*)[8][256])dig_raw;
    for (int i = threadIdx.x; i < 4 * 8 * 256; i += 256) dig_raw[i] = make_uint4(i * 7 + 1, i * 11 + 3, i * 13 + 5, i * 17 + 7);
    __syncthreads();
    u64 acc[4][4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int p = 0; p < 4; ++p) acc[c][j][p] = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            uint4 kv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) kv[j] = __ldg(keys + ((it & 7) * 32 + r * 4 + j) * 256 + threadIdx.x);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4 d = dig[c][r][threadIdx.x];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (VAR == 0) {
                        acc[c][j][0] += (u64)d.x * kv[j].x; acc[c][j][1] += (u64)d.y * kv[j].y;
                        acc[c][j][2] += (u64)d.z * kv[j].z; acc[c][j][3] += (u64)d.w * kv[j].w;
                    } else {
                        acc[c][j][0] += (u64)d.x * kv[j].y; acc[c][j][1] += (u64)d.y * kv[j].x;
                        acc[c][j][2] += (u64)d.z * kv[j].w; acc[c][j][3] += (u64)d.w * kv[j].z;
                    }
                }
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
template <int VAR> void run(const char* name) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint4* keys; u32* out; cudaMalloc(&keys, 8 * 32 * 256 * 16); cudaMemset(keys, 0x5a, 8 * 32 * 256 * 16); cudaMalloc(&out, sms * 256 * 4);
    int iters = 2000;
    cudaFuncSetAttribute(k<VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072); k<VAR><<<sms, 256, 131072>>>(keys, out, 10); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<VAR><<<sms, 256, 131072>>>(keys, out, iters); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double wides = (double)iters * 8 * 64 * 256 * sms;
    printf("%-28s %.3f ms  %.2f T wide/s = %.1f per clk per SM (1.965 GHz)  -> %.2f clk per warp-WIDE per SMSP\n", name, ms,
           wides / (ms * 1e-3) / 1e12, wides / (ms * 1e-3) / sms / 1.965e9, 32.0 * 4 / (wides / (ms * 1e-3) / sms / 1.965e9));
}
int main() { run<0>("MAC same parity"); run<1>("MAC crossed parity"); return 0; }
