// FP32 -> FP16 pre-pass of the plugin's kFLOAT branch (one launch for Q, K and V).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ptx_sm100.cuh"

namespace lgattn {

using namespace ptx;

// FP32 -> FP16 (round-to-nearest-even) of Q, K, V in one launch; the kFLOAT plugin branch.
// Replaces QKV_convert_from_fp32_to_fp16 (…fp32out.cu:706-768) minus its padding (not needed).
__global__ void __launch_bounds__(256)
lg_attn_cvt_f32_f16(const float4* __restrict__ q, const float4* __restrict__ k,
                    const float4* __restrict__ v, uint4* __restrict__ qh, uint4* __restrict__ kh,
                    uint4* __restrict__ vh, size_t nq8, size_t nkv8) {
    // Programmatic dependent launch on both sides: the attention kernel that follows may run its
    // prologue now (it waits for this grid to complete before loading), and this grid waits for
    // whatever precedes it in the stream before touching memory.
    pdl_launch_dependents();
    pdl_wait_prior_grids();
    // one thread = 8 elements: two float4 loads, one 16-byte store
    const size_t total = nq8 + 2 * nkv8;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        const float4* src;
        uint4* dst;
        size_t off;
        if (i < nq8) {
            src = q; dst = qh; off = i;
        } else if (i < nq8 + nkv8) {
            src = k; dst = kh; off = i - nq8;
        } else {
            src = v; dst = vh; off = i - nq8 - nkv8;
        }
        const float4 a = __ldg(src + 2 * off);
        const float4 b = __ldg(src + 2 * off + 1);
        uint4 r;
        r.x = pack_half2(a.x, a.y);
        r.y = pack_half2(a.z, a.w);
        r.z = pack_half2(b.x, b.y);
        r.w = pack_half2(b.z, b.w);
        dst[off] = r;
    }
}

}  // namespace lgattn
