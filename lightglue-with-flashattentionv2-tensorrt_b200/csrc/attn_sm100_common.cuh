// Pieces shared by the three attention kernels (attn_sm100_kernel_mq.cuh, _w4.cuh, attn_sm100_kernel.cuh): the two tcgen05.mma
// sequences, the ragged-KV tail mask, the lazy O rescale and the per-thread output row store.  Everything here is called from
// inside a kernel's own warp-role code; the callers own the barriers.
#pragma once
#include "ptx_sm100.cuh"

#include "attn_sm100.h"

namespace lgattn {

using namespace ptx;

// (elected lane of a converged MMA warp)  S[128 x kN] = Q[128 x 64] K[kN x 64]^T.  Both operands are K-major SWIZZLE_128B tiles
// (128-byte rows, 8-row groups 1024 B apart); advancing 16 halves along K = +32 B inside the swizzle row = +2 in the descriptor.
template <int kN>
__device__ __forceinline__ void umma_qk(uint32_t t_s, uint32_t smem_q, uint32_t smem_k) {
    constexpr uint32_t idesc = make_idesc_f16_f32acc(kBlockM, kN, false);
    const uint64_t desc_q = make_smem_desc_sw128(smem_q, 16, 1024);
    const uint64_t desc_k = make_smem_desc_sw128(smem_k, 16, 1024);
#pragma unroll
    for (int kk = 0; kk < kHeadDim / 16; ++kk) mma_f16_ss(t_s, desc_q + 2u * kk, desc_k + 2u * kk, idesc, kk > 0 ? 1u : 0u);
}

// (elected lane)  O[128 x 64] (+)= P[128 x kKV] V[kKV x 64]: A = P read from TMEM (16 halves = 8 columns per k-step), B = V as an
// MN-major SWIZZLE_128B tile (V is [kv, d] row-major, so no transpose pass; 16 kv rows = 2048 B = +128 in the descriptor per k-step).
template <int kKV>
__device__ __forceinline__ void umma_pv(uint32_t t_o, uint32_t t_p, uint32_t smem_v, uint32_t v_tile_bytes, bool accumulate) {
    constexpr uint32_t idesc = make_idesc_f16_f32acc(kBlockM, kHeadDim, true);
    const uint64_t desc_v = make_smem_desc_sw128(smem_v, v_tile_bytes, 1024);
#pragma unroll
    for (int kk = 0; kk < kKV / 16; ++kk) mma_f16_ts(t_o, t_p + 8u * kk, desc_v + 128u * kk, idesc, (accumulate || kk > 0) ? 1u : 0u);
}

// Ragged KV tail: columns >= valid of this thread's S row slice become -inf (TMA zero-filled the K rows past n_kv, so S = 0 there).
template <int kCols>
__device__ __forceinline__ void mask_kv_tail(uint32_t (&sr)[kCols], int valid) {
    if (valid < kCols) {
#pragma unroll
        for (int c = 0; c < kCols; ++c)
            if (c >= valid) sr[c] = 0xff800000u;
    }
}

// Lazy O rescale: this thread's kCols FP32 accumulator columns in TMEM times alpha, 32 columns at a time.
template <int kCols>
__device__ __forceinline__ void rescale_o_tmem(uint32_t t_o, float alpha) {
#pragma unroll
    for (int h = 0; h < kCols / 32; ++h) {
        uint32_t orow[32];
        tmem_ld_x32(t_o + 32u * h, orow);
        tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 32; ++c) orow[c] = __float_as_uint(__uint_as_float(orow[c]) * alpha);
        tmem_st_x32(t_o + 32u * h, orow);
    }
}

// One thread's output row (kCols FP32 accumulator values, the whole row by default) times inv_l to global memory, 16-byte stores.
template <bool kOutF32, int kCols = kHeadDim>
__device__ __forceinline__ void store_o_row(void* out_row, const uint32_t (&orow)[kCols], float inv_l) {
    if constexpr (!kOutF32) {
        __half* o = reinterpret_cast<__half*>(out_row);
#pragma unroll
        for (int c = 0; c < kCols / 8; ++c) {
            uint4 v;
            v.x = pack_half2(__uint_as_float(orow[8 * c + 0]) * inv_l, __uint_as_float(orow[8 * c + 1]) * inv_l);
            v.y = pack_half2(__uint_as_float(orow[8 * c + 2]) * inv_l, __uint_as_float(orow[8 * c + 3]) * inv_l);
            v.z = pack_half2(__uint_as_float(orow[8 * c + 4]) * inv_l, __uint_as_float(orow[8 * c + 5]) * inv_l);
            v.w = pack_half2(__uint_as_float(orow[8 * c + 6]) * inv_l, __uint_as_float(orow[8 * c + 7]) * inv_l);
            st_global_v4(o + c * 8, v);
        }
    } else {
        float* o = reinterpret_cast<float*>(out_row);
#pragma unroll
        for (int c = 0; c < kCols / 4; ++c) {
            uint4 v;
            v.x = __float_as_uint(__uint_as_float(orow[4 * c + 0]) * inv_l);
            v.y = __float_as_uint(__uint_as_float(orow[4 * c + 1]) * inv_l);
            v.z = __float_as_uint(__uint_as_float(orow[4 * c + 2]) * inv_l);
            v.w = __float_as_uint(__uint_as_float(orow[4 * c + 3]) * inv_l);
            st_global_v4(o + c * 4, v);
        }
    }
}

}  // namespace lgattn
