// tools/microbench4.cu -- Shoup butterfly with the quotient computed on the FP64 pipe.
//   F0  classic: q = mulhi(w', y)                               (IMAD.HI + 2 IMAD + 2 IADD)
//   F1  q = lo32(fma(2^52 + y, w/Q, 2^52 - 2^52*w/Q))           (DFMA + 2 IMAD + 2 IADD), data kept in
//       register pairs (lo = value, hi = 0x43300000) so that "2^52 + y" needs no instruction
//   F2  yd = (2^52 + y) - 2^52;  q = lo32(fma(yd, w/Q, 2^52))    (DADD + DFMA + 2 IMAD + 2 IADD)
// prints butterflies per clock per SM and checks that every result is congruent to the classic one
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <cuda_runtime.h>
typedef uint32_t u32; typedef uint64_t u64;
#define Q 134215681u
#define CH 16
#define ITERS 512
__device__ __forceinline__ double pack(u32 lo) { return __hiloint2double(0x43300000, (int)lo); }
__device__ __forceinline__ u32 lo32(double d) { return (u32)__double2loint(d); }
template <int MODE> __global__ void __launch_bounds__(256, 1) k(u32* out, const u32* tw, const double* twc, const double* twk, int iters) {
    u32 w[4], wp[4]; double c[4], kw[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { w[i] = tw[i]; wp[i] = (u32)(((u64)w[i] << 32) / Q); c[i] = twc[i]; kw[i] = twk[i]; }
    double xd[CH], yd[CH];            // register pairs: low word = value (offset by a multiple of Q), high word = 0x43300000
    u32 x[CH], y[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { x[i] = 8 * Q + (threadIdx.x * 977u + i * 131u) % Q; y[i] = 8 * Q + (threadIdx.x * 313u + i * 71u + blockIdx.x) % Q; xd[i] = pack(x[i]); yd[i] = pack(y[i]); }
    const double two52 = 4503599627370496.0;
    for (int it = 0; it < iters; ++it) {
        // 10 stages of 16 butterflies without corrections (as in the NTT passes), then one reduction of all values
#pragma unroll
        for (int s = 0; s < 10; ++s) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                const int j = (i + s) % CH;
                if (MODE == 0) {
                    u32 q = __umulhi(wp[s & 3], y[j]); u32 v = w[s & 3] * y[j] - q * Q; u32 t = x[i];
                    x[i] = t + v; y[j] = t - v + 2 * Q;
                } else {
                    u32 yv = lo32(yd[j]);
                    double t = (MODE == 1) ? fma(yd[j], c[s & 3], kw[s & 3]) : fma(yd[j] - two52, c[s & 3], two52);
                    u32 q = lo32(t);
                    u32 r = w[s & 3] * yv - q * Q;              // signed, |r| <= 0.75 Q
                    u32 xv = lo32(xd[i]);
                    xd[i] = __hiloint2double(__double2hiint(xd[i]), (int)(xv + r));
                    yd[j] = __hiloint2double(__double2hiint(yd[j]), (int)(xv - r));
                }
            }
        }
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (MODE == 0) {
                x[i] = x[i] - (x[i] >> 27) * Q; y[i] = y[i] - (y[i] >> 27) * Q;
            } else {
                u32 xv = lo32(xd[i]), yv = lo32(yd[i]);
                xv = xv - (xv >> 27) * Q + 8 * Q; yv = yv - (yv >> 27) * Q + 8 * Q;
                xd[i] = __hiloint2double(__double2hiint(xd[i]), (int)xv); yd[i] = __hiloint2double(__double2hiint(yd[i]), (int)yv);
            }
        }
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) r ^= (MODE == 0) ? (x[i] % Q) * 3 + (y[i] % Q) : (lo32(xd[i]) % Q) * 3 + (lo32(yd[i]) % Q);
    out[blockIdx.x * 256 + threadIdx.x] = r;
}
template <int MODE> float run(const char* name, u32* out, const u32* tw, const double* twc, const double* twk, int sms) {
    k<MODE><<<sms, 256>>>(out, tw, twc, twk, 8); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); k<MODE><<<sms, 256>>>(out, tw, twc, twk, ITERS); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double bf = (double)ITERS * 10 * CH * 256 * sms;
    printf("%-34s %.3f ms  %.2f T butterflies/s = %5.2f per clk per SM -> %.2f clk per warp-butterfly per SMSP   [%s]\n", name, ms,
           bf / (ms * 1e-3) / 1e12, bf / (ms * 1e-3) / sms / 1.965e9, 128.0 / (bf / (ms * 1e-3) / sms / 1.965e9), cudaGetErrorString(cudaGetLastError()));
    return ms;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    u32 htw[4] = {282116u, 77777777u, 1234567u, 134215680u}; double hc[4], hk[4];
    for (int i = 0; i < 4; ++i) { hc[i] = (double)htw[i] / (double)Q; hk[i] = 4503599627370496.0 - 4503599627370496.0 * hc[i]; }
    u32 *tw, *o0, *o1, *o2; double *twc, *twk;
    cudaMalloc(&tw, 16); cudaMalloc(&twc, 32); cudaMalloc(&twk, 32); cudaMalloc(&o0, sms * 1024); cudaMalloc(&o1, sms * 1024); cudaMalloc(&o2, sms * 1024);
    cudaMemcpy(tw, htw, 16, cudaMemcpyHostToDevice); cudaMemcpy(twc, hc, 32, cudaMemcpyHostToDevice); cudaMemcpy(twk, hk, 32, cudaMemcpyHostToDevice);
    run<0>("F0 IMAD.HI shoup", o0, tw, twc, twk, sms);
    run<1>("F1 DFMA quotient (1 FP64 op)", o1, tw, twc, twk, sms);
    run<2>("F2 DADD+DFMA quotient (2 FP64 ops)", o2, tw, twc, twk, sms);
    // congruence check on a short run
    k<0><<<sms, 256>>>(o0, tw, twc, twk, 3); k<1><<<sms, 256>>>(o1, tw, twc, twk, 3); k<2><<<sms, 256>>>(o2, tw, twc, twk, 3); cudaDeviceSynchronize();
    u32* h0 = new u32[sms * 256]; u32* h1 = new u32[sms * 256]; u32* h2 = new u32[sms * 256];
    cudaMemcpy(h0, o0, sms * 1024, cudaMemcpyDeviceToHost); cudaMemcpy(h1, o1, sms * 1024, cudaMemcpyDeviceToHost); cudaMemcpy(h2, o2, sms * 1024, cudaMemcpyDeviceToHost);
    int bad1 = 0, bad2 = 0; for (int i = 0; i < sms * 256; ++i) { bad1 += h0[i] != h1[i]; bad2 += h0[i] != h2[i]; }
    printf("residue mismatches vs classic: F1 %d, F2 %d (of %d)\n", bad1, bad2, sms * 256);
    return 0;
}
