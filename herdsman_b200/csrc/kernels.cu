// kernels.cu -- hand-written sm_100a kernels of the gate-evaluation hot path.
//
//   blind_rotate_kernel   K1+K2+K3 of SURVEY.md 2.1: ciphertext prep, accumulator init, the n
//                         CMUX steps, sample extraction and ModSwitch(Q->qKS)   (br_core.cuh)
//   keyswitch_kernel      K4: LWE key switch (OpenFHE LWEEncryptionScheme::KeySwitch) + ModSwitch(qKS->q)
//   gather_outputs_kernel OutputNode handling: copies / NOTs / constants into the output rows
//   bsk_convert_kernel    key load: coefficient-domain RGSW rows -> NTT domain, * 1/N
//   bsk_fold_kernel       key load: folds the gadget rows for the top-digit elimination
//   ksk_relayout_kernel   key load: pads key-switch rows to a 16-byte multiple
//   int_peak_kernel       measurement: raw IMAD / IADD3 / IMAD.HI / IMAD.WIDE issue rates
//
// Integer pipes only (IMAD / IMAD.HI / IMAD.WIDE / IADD3 / LOP3): the work is exact modular
// arithmetic, there is no tensor-core formulation (BASELINE.json north_star).
#include <cuda_runtime.h>
#include <stdint.h>
#include "br_core.cuh"
#include "kernels.h"

namespace herdb {

struct DevEnv {
    __device__ __forceinline__ u32 tid() const { return threadIdx.x; }
    __device__ __forceinline__ void sync() { __syncthreads(); }
    __device__ __forceinline__ void syncwarp() { __syncwarp(); }
    // 16-byte asynchronous global -> shared copy (LDGSTS, L2 only), grouped by commit / wait_group
    __device__ __forceinline__ void cp16(u32* smem_dst, const u32* gsrc) const {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((u32)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
    }
    __device__ __forceinline__ void cp_commit() const { asm volatile("cp.async.commit_group;" ::: "memory"); }
    __device__ __forceinline__ void cp_wait1() const { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
    __device__ __forceinline__ void cp_wait0() const { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
    // TMA bulk copies (cp.async.bulk, global -> shared) completing on an mbarrier
    __device__ __forceinline__ void mbar_init(unsigned long long* bar, u32 count) const {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((u32)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __device__ __forceinline__ void bulk_begin(unsigned long long* bar, u32 bytes) const {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((u32)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ void bulk_copy(void* smem_dst, const void* gsrc, u32 bytes, unsigned long long* bar) const {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"((u32)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes), "r"((u32)__cvta_generic_to_shared(bar)) : "memory");
    }
    __device__ __forceinline__ void mbar_wait(unsigned long long* bar, u32 parity) const {
        asm volatile("{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}"
                     ::"r"((u32)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
    }
    // orders the copies that follow after the shared-memory reads whose results feed `dep`
    __device__ __forceinline__ void cp_fence(u32 dep) const { asm volatile("" ::"r"(dep) : "memory"); }
};

// persistent over groups of C jobs: one CTA per SM (C = 4: 256 threads, 212 KB of shared memory; C = 2, 1: 128 threads)
template <int C>
__global__ void __launch_bounds__(BrShape<C>::kThreads, 1) blind_rotate_kernel(const __grid_constant__ BrParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BrSmem<C>& S = *reinterpret_cast<BrSmem<C>*>(smem_raw);
    DevEnv env;
    const u32 groups = (P.job1 - P.job0 + C - 1) / C;
    if (threadIdx.x == 0) env.mbar_init(&S.bar, 1);
    __syncthreads();
    u32 done = 0;                                                    // bulk-copy phases completed so far = steps of earlier groups
    for (u32 g = blockIdx.x; g < groups; g += gridDim.x, done += P.steps) blind_rotate_group<C>(env, P, S, g, done & 1);
}

// ------------------------------------------------------------------------------------------------
// Key switch: a_out = -sum_{i,j} KSK_A[i][digit_j(a_i)][j], b_out = b - sum KSK_B[...]  (mod qKS),
// then ModSwitch to q.  One CTA per job.  The 2N row numbers of the job are put in shared memory once;
// thread w then owns the u16 quad (4w .. 4w+3) of the padded rows and sums them with 64-bit loads.  The
// values are below 2^14, so a 32-bit word carries two independent 16-bit lanes: two rows are added per
// step and the lanes masked back to 14 bits (only the sum mod qKS = 2^14 is needed), no unpacking.
constexpr u32 kKsThreads = 160;
__global__ void __launch_bounds__(kKsThreads) keyswitch_kernel(const __grid_constant__ KsParams P) {
    __shared__ __align__(16) u32 row_s[2 * kN];
    __shared__ u32 b_s;
    static_assert(kDKS == 2 && kQKS == (1u << 14), "lane arithmetic below assumes two digits and qKS = 2^14");
    const u32 quads = P.ks_stride >> 2;
    const u32 w = threadIdx.x;
    constexpr u32 kLaneMask = (kQKS - 1) | ((kQKS - 1) << 16);
    for (u32 job = blockIdx.x; job < P.jobs; job += gridDim.x) {
        __syncthreads();
        const u16* ext = P.ext + (size_t)job * kExtStride;
        for (u32 i = threadIdx.x; i < kN; i += blockDim.x) {
            const u32 t = ext[i];
            row_s[2 * i] = (i * kBaseKS + (t & (kBaseKS - 1))) * kDKS + 0;
            row_s[2 * i + 1] = (i * kBaseKS + (t >> HERDB_BASEKS_LOG)) * kDKS + 1;
        }
        if (threadIdx.x == 0) b_s = ext[kN];
        __syncthreads();
        if (w < quads) {
            const uint2* base = reinterpret_cast<const uint2*>(P.ksk) + w;
            u32 acc0 = 0, acc1 = 0;
#pragma unroll 4
            for (u32 k = 0; k < 2 * kN; k += 4) {
                const uint4 r = *reinterpret_cast<const uint4*>(&row_s[k]);
                const uint2 x0 = __ldg(base + (size_t)r.x * quads), x1 = __ldg(base + (size_t)r.y * quads);
                const uint2 x2 = __ldg(base + (size_t)r.z * quads), x3 = __ldg(base + (size_t)r.w * quads);
                acc0 = (acc0 + x0.x + x1.x) & kLaneMask; acc1 = (acc1 + x0.y + x1.y) & kLaneMask;
                acc0 = (acc0 + x2.x + x3.x) & kLaneMask; acc1 = (acc1 + x2.y + x3.y) & kLaneMask;
            }
            const GateDesc g = P.gates[job / P.rows];
            const u32 row = job % P.rows;
            u16* out = P.pools[g.pool_out].base + (size_t)row * P.pools[g.pool_out].row_words + (size_t)g.idx_out * (P.n + 1);
            const u32 sums[4] = {acc0 & 0xFFFFu, acc0 >> 16, acc1 & 0xFFFFu, acc1 >> 16};
#pragma unroll
            for (u32 e = 0; e < 4; ++e) {
                const u32 k = 4 * w + e;
                if (k <= P.n) {
                    const u32 v = (k == P.n ? b_s - sums[e] : 0u - sums[e]) & (kQKS - 1);
                    out[k] = (u16)modswitch_qKS_to_q(v);
                }
            }
        }
    }
}

// OutputNode: output row[o] = operand (with NOT / constant folding applied).  The output pool may be in the
// partition framing: frame_len != 0 also writes the u32 length word in front of every row, so that the buffer
// is the byte image of the partition file (storage_service.cpp:19-28) and the host writes it out unchanged.
__global__ void gather_outputs_kernel(const __grid_constant__ GatherParams P) {
    const u32 w = P.n + 1;
    const Pool out = P.pools[kGatherOutPool];
    const size_t total = (size_t)P.rows * P.out_bits * w;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const u32 k = (u32)(t % w);
        const size_t ct = t / w;
        const u32 o = (u32)(ct % P.out_bits), row = (u32)(ct / P.out_bits);
        const OutDesc d = P.outs[o];
        u16* dst = out.base + (size_t)row * out.row_words;
        dst[(size_t)o * w + k] = (u16)operand_word(P.pools, d.pool, d.idx, d.flags, row, P.n, k);
        if (P.framed && o == 0 && k < 2) dst[(int)k - 2] = (u16)(P.frame_len >> (16 * k));     // little-endian host (x86-64 / aarch64)
    }
    if (P.framed && !P.out_bits)                                  // rows without payload still carry their length word
        for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < P.rows; r += (size_t)gridDim.x * blockDim.x) {
            u16* dst = out.base + r * out.row_words; dst[-2] = 0; dst[-1] = 0;
        }
}

// ------------------------------------------------------------------------------------------------
// key load
__global__ void __launch_bounds__(256) bsk_convert_kernel(const u32* __restrict__ in, u32* __restrict__ out,
                                                          const u32* __restrict__ ftw2, u32 polys) {
    __shared__ __align__(16) u32 scratch[8][kN];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 ftw[62];
#pragma unroll
    for (int p = 0; p < 31; ++p) {
        ftw[2 * p] = ftw2[((p >> 1) * 32 + lane) * 4 + (p & 1) * 2];
        ftw[2 * p + 1] = ftw2[((p >> 1) * 32 + lane) * 4 + (p & 1) * 2 + 1];
    }
    for (u32 poly = blockIdx.x * 8 + warp; poly < polys; poly += gridDim.x * 8) {
        u32 x[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) x[k] = in[(size_t)poly * kN + lane + 32 * k];
        fwd_pass1(x);
        __syncwarp();
        store_strided(scratch[warp], lane, x);
        __syncwarp();
        load_contig(scratch[warp], lane, x);
        fwd_pass2(x, ftw);
#pragma unroll
        for (int e = 0; e < 32; ++e) x[e] = canon32(mul_shoup_lazy(kNInv, kNInvShoup, x[e]));
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            uint4 q = make_uint4(x[4 * c], x[4 * c + 1], x[4 * c + 2], x[4 * c + 3]);
            *reinterpret_cast<uint4*>(out + (size_t)poly * kN + 32 * lane + 4 * c) = q;
        }
    }
}

// key load, second pass (in place): folds the four gadget rows of every (i, sign, component, column, point)
// for the top-digit elimination (modarith.cuh, fold_top_digit)
__global__ void __launch_bounds__(256) bsk_fold_kernel(u32* __restrict__ bsk, u32 n) {
    // one thread per aligned group of four points: the folded words are written back rotated (key_pos, modarith.cuh)
    const size_t total = (size_t)n * 2 * 2 * 2 * (kN / 4);          // (i, sign, component j, column, group)
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const u32 g = (u32)(t % (kN / 4));
        const u32 col = (u32)(t / (kN / 4)) & 1u, j = (u32)(t / (2 * (kN / 4))) & 1u;
        const size_t is = t / (4 * (kN / 4));                          // i * 2 + sign
        u32 k[4][4];
#pragma unroll
        for (u32 p = 0; p < 4; ++p) {
#pragma unroll
            for (u32 l = 0; l < 4; ++l) k[p][l] = bsk[((is * kDG2 + j + 2 * l) * 2 + col) * kN + 4 * g + p];
            fold_top_digit(k[p]);
        }
#pragma unroll
        for (u32 p = 0; p < 4; ++p)
#pragma unroll
            for (u32 l = 0; l < 4; ++l) bsk[((is * kDG2 + j + 2 * l) * 2 + col) * kN + key_pos(4 * g + p)] = k[p][l];
    }
}

__global__ void ksk_relayout_kernel(const u16* __restrict__ in, u16* __restrict__ out, u32 n, u32 stride, size_t rows) {
    const size_t total = rows * stride;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t r = t / stride; const u32 k = (u32)(t % stride);
        out[t] = k <= n ? in[r * (n + 1) + k] : (u16)0;
    }
}

// ------------------------------------------------------------------------------------------------
// integer-pipe peak: 8 independent dependency chains per thread, 4096 iterations
template <int MODE>
__global__ void __launch_bounds__(256) int_peak_kernel(u32* out, u32 seed, int iters) {
    u32 a[8], b = seed | 1u, c = seed * 2654435761u + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 97u + i * 13u + seed;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (MODE == 0) a[i] = a[i] * b + c;                                     // IMAD
                else if (MODE == 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c));   // IADD3
                else if (MODE == 2) a[i] = __umulhi(a[i], b) + c;                       // IMAD.HI
                else if (MODE == 3) {                                                   // IMAD.WIDE
                    u64 t = (u64)a[i] * b + (((u64)c << 32) | a[i]);
                    a[i] = (u32)t ^ (u32)(t >> 32);
                } else {                                                                 // Shoup butterfly mix
                    u32 q = __umulhi(a[i], b); u32 v = a[i] * c - q * kQ; a[i] = a[(i + 1) & 7] + v;
                }
            }
        }
    }
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// L2 read bandwidth: every CTA streams the same `words`-word buffer (sized to stay resident in L2) `reps` times with
// 128-bit loads.  The roof the key switch is held against: its 2.1 MB of key rows per job are served by L2.
__global__ void __launch_bounds__(256) l2_read_kernel(const uint4* __restrict__ buf, size_t vecs, int reps, u32* out) {
    u32 acc = 0;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < vecs; i += (size_t)gridDim.x * blockDim.x) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345678u) out[0] = acc;
}

// ------------------------------------------------------------------------------------------------
// Scratch for the entry point that allocates nothing (herdb_eval_gate_batch_dev_noalloc, include/herdb.h): the
// extraction buffer between blind rotation and key switch -- [jobs][kExtStride] u16, cudaMalloc'ed and grown on demand
// on every other path -- lives here in a file-scope array of fixed size, HERDB_NOALLOC_MAX_BATCH jobs (4.66 MiB; every
// GPU that loads the library holds it until exit), next to the one gate descriptor such a launch needs.
// In HOST code these names are host-side stand-ins, not device addresses: noalloc_setup() asks the runtime for the
// device addresses (cudaGetSymbolAddress, once per device, with that device current) and the kernels are given those.
// Nothing in the buffer has to start at zero: blind rotation writes words 0..N of every job before the key switch
// reads them and the padding words are never read, so the previous call's contents are harmless.  All calls on a
// device share the array: calls that overlap on the GPU (other streams, other contexts) race.
static __device__ __align__(256) u16 g_noalloc_ext[(size_t)HERDB_NOALLOC_MAX_BATCH * kExtStride];
static __device__ __align__(16) GateDesc g_noalloc_gate;

__global__ void noalloc_set_gate_kernel(GateDesc g) { g_noalloc_gate = g; }      // device code names the array directly

// ------------------------------------------------------------------------------------------------
// launchers (plain C++ signatures so backend.cpp does not need nvcc)
template <int C>
static cudaError_t launch_br(BrParams P, u32 job0, u32 job1, int sm_count, cudaStream_t stream) {
    if (job1 <= job0) return cudaSuccess;
    // per device attribute; cheap enough to set on every launch (launches last milliseconds)
    cudaError_t e = cudaFuncSetAttribute(blind_rotate_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BrSmem<C>));
    if (e != cudaSuccess) return e;
    P.job0 = job0; P.job1 = job1;
    const u32 groups = (job1 - job0 + C - 1) / C;
    const u32 resident = (u32)sm_count;
    const u32 grid = groups < resident ? groups : resident;
    blind_rotate_kernel<C><<<grid, BrShape<C>::kThreads, sizeof(BrSmem<C>), stream>>>(P);
    return cudaGetLastError();
}
// A level of P.jobs jobs = a body of whole grid rounds at C = 4 (one CTA per SM, four ciphertexts each) and a tail.
// A tail that would fill at most half (a quarter) of the SMs at C = 4 runs as its own launch at C = 2 (C = 1): twice
// (four times) the SMs, each with a shorter step.  `shape` forces one CTA shape (0 = this policy); launches_out
// (may be null) receives the number of kernels launched.
cudaError_t launch_blind_rotate(const BrParams& P, int sm_count, cudaStream_t stream, int shape, int* launches_out) {
    const u32 sm = (u32)sm_count, jobs = P.jobs;
    int launches = 0;
    cudaError_t e = cudaSuccess;
    if (shape == 4 || shape == 2 || shape == 1) {
        e = shape == 4 ? launch_br<4>(P, 0, jobs, sm_count, stream) : shape == 2 ? launch_br<2>(P, 0, jobs, sm_count, stream) : launch_br<1>(P, 0, jobs, sm_count, stream);
        launches = jobs ? 1 : 0;
    } else {
        const u32 wave = 4 * sm;
        u32 body = jobs / wave * wave;
        const u32 tail = jobs - body;
        if (tail > 2 * sm) body = jobs;                                  // the tail is worth a C = 4 round of its own
        if (body) { e = launch_br<4>(P, 0, body, sm_count, stream); ++launches; }
        if (e == cudaSuccess && body < jobs) {
            e = tail > sm ? launch_br<2>(P, body, jobs, sm_count, stream) : launch_br<1>(P, body, jobs, sm_count, stream);
            ++launches;
        }
    }
    if (launches_out) *launches_out = launches;
    return e;
}
cudaError_t launch_keyswitch(const KsParams& P, int sm_count, cudaStream_t stream) {
    if (!P.jobs) return cudaSuccess;
    if ((P.ks_stride & 3) || (P.ks_stride >> 2) > kKsThreads) return cudaErrorInvalidValue;
    const u32 max_grid = (u32)sm_count * 12;                     // 12 resident CTAs of 160 threads per SM
    const u32 grid = P.jobs < max_grid ? P.jobs : max_grid;
    keyswitch_kernel<<<grid, kKsThreads, 0, stream>>>(P);
    return cudaGetLastError();
}
cudaError_t launch_gather_outputs(const GatherParams& P, int sm_count, cudaStream_t stream) {
    if (!P.rows || (!P.out_bits && !P.framed)) return cudaSuccess;
    gather_outputs_kernel<<<sm_count * 8, 256, 0, stream>>>(P);
    return cudaGetLastError();
}
cudaError_t launch_bsk_convert(const u32* in, u32* out, const u32* ftw2, u32 polys, int sm_count, cudaStream_t stream) {
    bsk_convert_kernel<<<sm_count * 2, 256, 0, stream>>>(in, out, ftw2, polys);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    bsk_fold_kernel<<<sm_count * 8, 256, 0, stream>>>(out, polys / (2 * kDG2 * 2));
    return cudaGetLastError();
}
cudaError_t launch_ksk_relayout(const u16* in, u16* out, u32 n, u32 stride, size_t rows, int sm_count, cudaStream_t stream) {
    ksk_relayout_kernel<<<sm_count * 8, 256, 0, stream>>>(in, out, n, stride, rows);
    return cudaGetLastError();
}
cudaError_t noalloc_setup(u16** ext_out, GateDesc** gate_out) {               // the device in question is current
    cudaError_t e = cudaGetSymbolAddress((void**)ext_out, g_noalloc_ext);
    if (e != cudaSuccess) return e;
    return cudaGetSymbolAddress((void**)gate_out, g_noalloc_gate);
}
cudaError_t launch_noalloc_set_gate(const GateDesc& g, cudaStream_t stream) {
    noalloc_set_gate_kernel<<<1, 1, 0, stream>>>(g);
    return cudaGetLastError();
}
cudaError_t launch_l2_read(const void* buf, size_t bytes, int reps, u32* out, int sm_count, cudaStream_t stream) {
    l2_read_kernel<<<sm_count * 8, 256, 0, stream>>>(static_cast<const uint4*>(buf), bytes / 16, reps, out);
    return cudaGetLastError();
}
cudaError_t launch_int_peak(int mode, u32* out, int blocks, int iters, cudaStream_t stream) {
    switch (mode) {
    case 0: int_peak_kernel<0><<<blocks, 256, 0, stream>>>(out, 12345u, iters); break;
    case 1: int_peak_kernel<1><<<blocks, 256, 0, stream>>>(out, 12345u, iters); break;
    case 2: int_peak_kernel<2><<<blocks, 256, 0, stream>>>(out, 12345u, iters); break;
    case 3: int_peak_kernel<3><<<blocks, 256, 0, stream>>>(out, 12345u, iters); break;
    default: int_peak_kernel<4><<<blocks, 256, 0, stream>>>(out, 12345u, iters); break;
    }
    return cudaGetLastError();
}

}  // namespace herdb
