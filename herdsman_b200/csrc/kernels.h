// kernels.h -- launcher interface between the host backend (backend.cpp) and kernels.cu
#pragma once
#include <cuda_runtime.h>
#include "br_core.cuh"

namespace herdb {

struct KsParams {
    const u16* ksk;              // [N][baseKS][dKS][ks_stride] u16, rows padded to a 16-byte multiple
    const u16* ext;              // [jobs][kExtStride]
    const GateDesc* gates;
    Pool pools[4];
    u32 ks_stride;               // u16 words per row (even, >= n + 1)
    u32 n, rows, jobs;
};
struct OutDesc { u32 idx; uint8_t pool, flags, pad[2]; };
constexpr u32 kGatherOutPool = 3;   // == kPoolOut (circuit.h)
struct GatherParams {
    const OutDesc* outs;         // [out_bits]
    Pool pools[4];               // pools[kGatherOutPool] = the output rows
    u32 n, rows, out_bits;
    u32 framed;                  // != 0: output rows are in the partition framing: write the u32 length word before every row
    u32 frame_len;               // that length (payload bytes of an output row)
};

cudaError_t launch_blind_rotate(const BrParams& P, int sm_count, cudaStream_t stream, int shape = 0, int* launches_out = nullptr);
cudaError_t launch_keyswitch(const KsParams& P, int sm_count, cudaStream_t stream);
cudaError_t launch_gather_outputs(const GatherParams& P, int sm_count, cudaStream_t stream);
cudaError_t launch_bsk_convert(const u32* in, u32* out, const u32* ftw2, u32 polys, int sm_count, cudaStream_t stream);
cudaError_t launch_ksk_relayout(const u16* in, u16* out, u32 n, u32 stride, size_t rows, int sm_count, cudaStream_t stream);
// the no-allocation entry point (kernels.cu, "Scratch for the entry point that allocates nothing")
cudaError_t noalloc_setup(u16** ext_out, GateDesc** gate_out);
cudaError_t launch_noalloc_set_gate(const GateDesc& g, cudaStream_t stream);
cudaError_t launch_l2_read(const void* buf, size_t bytes, int reps, u32* out, int sm_count, cudaStream_t stream);
cudaError_t launch_int_peak(int mode, u32* out, int blocks, int iters, cudaStream_t stream);

}  // namespace herdb
