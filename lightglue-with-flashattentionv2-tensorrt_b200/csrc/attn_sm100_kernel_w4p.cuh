// B200-native (sm_100a) fused multi-head attention for LightGlue, head_dim 64 — PERSISTENT throughput kernel.
//
//   O = softmax(Q K^T / 8) V      per (batch, head); Q [B,H,N,64], K/V [B,H,M,64], O [B,H,N,64]
//
// Same math, tile shapes, TMEM map and warp roles as the one-tile-per-CTA kernel in attn_sm100_kernel_w4.cuh
// (which replaces /root/reference/lightglue_attention_plugin/attention_headdim_64_fp16in_{fp16,fp32}out.cu);
// what changes is the CTA's lifetime.  The phase trace of that kernel (tools/trace_w4.py, CTA-level stamps)
// shows ~4 200 clk per 128-row Q tile outside the steady KV loop — 580 set-up (barrier init, TMEM alloc),
// 2 470 until the first S arrives (TMA Q + K from L2/HBM, first QK), 1 200 epilogue — plus 2 100 clk between a CTA
// retiring and its successor starting on that SM: 6 300 of 45 300 clk per tile.  Here 2 CTAs per SM stay
// resident and walk the Q tiles `blockIdx.x + i * gridDim.x`:
//   * barriers and TMEM are set up once; every barrier's phase is carried by a running KV-tile counter
//   * the K/V ring streams across Q-tile boundaries; the next Q tile is loaded into the (single) Q buffer as soon
//     as the last QK of the current tile has completed (q_empty), two softmax tiles before it is needed
//   * S(0) of the next Q tile is computed under the last softmax tile + epilogue of the current one; the first PV of
//     the next tile waits for the epilogue to have pulled O out of TMEM (o_empty)
//   * the epilogue stores straight from registers (one 128/256-byte output row per thread, 16-byte stores): no
//     shared-memory staging, because the K/V ring it used to alias is live
// so a Q-tile boundary costs the softmax warps one O read-back + store and nothing else.
//
// No CUTLASS; see ptx_sm100.cuh for the hand-written encodings.

#pragma once
#include "ptx_sm100.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "attn_sm100.h"
#include "softmax_math.cuh"

#ifndef LG_ATTN_SPIN_NS
#define LG_ATTN_SPIN_NS 64
#endif

namespace lgattn {
namespace w4p {

using namespace ptx;

constexpr int kThreads = 256;   // 4 softmax warps + TMA + MMA + TMEM-alloc + spare
constexpr uint32_t kTileBytes = kBlockN * kHeadDim * 2;  // 16 KiB: 128 rows x 128 B
constexpr uint32_t kSmemBytes = (1 + 2 * kStages) * kTileBytes + 1024 + 1024;
constexpr uint32_t kSmemQ = 0;
constexpr uint32_t kSmemK = kSmemQ + kTileBytes;               // kStages tiles
constexpr uint32_t kSmemV = kSmemK + kStages * kTileBytes;     // kStages tiles
constexpr uint32_t kSmemBar = kSmemV + kStages * kTileBytes;   // mbarriers + tmem slot
enum : uint32_t {
    kBarQFull = 0,
    kBarQEmpty,
    kBarKFull,                        // [kStages]
    kBarKEmpty = kBarKFull + kStages, // [kStages]
    kBarVFull = kBarKEmpty + kStages,
    kBarVEmpty = kBarVFull + kStages,
    kBarSFull = kBarVEmpty + kStages,
    kBarSEmpty,
    kBarPFull,
    kBarPVDone,
    kBarOEmpty,
    kNumBars
};
constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
constexpr uint32_t kSmemUsed = kSmemTmemSlot + 16;
static_assert(kSmemUsed + 1024 <= kSmemBytes, "shared memory budget");

constexpr uint32_t kTmemCols = 256;
constexpr uint32_t kColS = 0;
constexpr uint32_t kColO = 128;
constexpr uint32_t kColP = 192;

constexpr int kRegsSoftmax = 216;
constexpr int kRegsOther = 40;

#ifdef LG_ATTN_W4_TRACE
// same layout as w4's trace (tools/trace_w4.py): rows [0,296) = first Q tile of each CTA, [296,592) = second
constexpr int kTraceCtas = 592, kTraceTiles = 16, kTraceEvents = 8;
__device__ long long g_w4p_trace[kTraceCtas][kTraceTiles][kTraceEvents];
__device__ unsigned g_w4p_smid[kTraceCtas];
#define W4P_TRACE(j, ev)                                                                      \
    do {                                                                                      \
        if (warp == 0 && lane == 0 && trace_row >= 0 && (j) < kTraceTiles)                    \
            g_w4p_trace[trace_row][(j)][(ev)] = clock64();                                    \
    } while (0)
#else
#define W4P_TRACE(j, ev)
#endif

struct TileCoord {
    int tile_m, hi, bi, bk;
};

__device__ __forceinline__ TileCoord tile_coord(int w, int tiles_m, const OutLayout& ol) {
    TileCoord c;
    c.tile_m = w % tiles_m;
    const int bh = w / tiles_m;
    c.hi = bh % ol.heads;
    c.bi = bh / ol.heads;
    c.bk = c.bi + ol.kv_shift >= ol.batch ? c.bi + ol.kv_shift - ol.batch : c.bi + ol.kv_shift;
    return c;
}

template <bool kOutF32>
__global__ void __launch_bounds__(kThreads, 2)
lg_attn_hd64_fwd(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, void* __restrict__ out, int n_q, int n_kv,
                 int tiles_m, int total_tiles, const OutLayout ol) {
    extern __shared__ uint8_t smem_raw[];
    // the aligned base is pinned in a register: otherwise ptxas re-derives it (S2UR SR_CgaCtaId + ULEA + ...) in front of every
    // mbarrier operation of the KV loop (+1.6 % on the throughput kernel, same-box A/B)
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));
    const uint32_t bars = smem + kSmemBar;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int num_kv_tiles = (n_kv + kBlockN - 1) / kBlockN;
    const int first = blockIdx.x, stride = gridDim.x;

    if (warp == 4 && lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
        mbar_init(bar(kBarQFull), 1);
        mbar_init(bar(kBarQEmpty), 1);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(kBarKFull + s), 1);
            mbar_init(bar(kBarKEmpty + s), 1);
            mbar_init(bar(kBarVFull + s), 1);
            mbar_init(bar(kBarVEmpty + s), 1);
        }
        mbar_init(bar(kBarSFull), 1);
        mbar_init(bar(kBarSEmpty), 4);  // one arrival per softmax warp
        mbar_init(bar(kBarPFull), 4);
        mbar_init(bar(kBarPVDone), 1);
        mbar_init(bar(kBarOEmpty), 4);
        fence_mbar_init();
    }
    if (warp == 6) {
        tmem_alloc<kTmemCols>(smem + kSmemTmemSlot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + kSmemTmemSlot));

    if (warp >= 4) {
        setmaxnreg_dec<kRegsOther>();
        if (warp == 4 && lane == 0) {
            // ===================================================================== TMA producer
            int g = 0;  // running KV-tile counter: ring slot and barrier phases continue across Q tiles
            int it = 0;
            for (int w = first; w < total_tiles; w += stride, ++it) {
                const TileCoord tc = tile_coord(w, tiles_m, ol);
                if (it > 0) mbar_wait_backoff(bar(kBarQEmpty), (it - 1) & 1, LG_ATTN_SPIN_NS);  // last QK of the previous Q tile done
                mbar_arrive_expect_tx(bar(kBarQFull), kTileBytes);
                tma_load_4d(smem + kSmemQ, &tm_q, bar(kBarQFull), 0, tc.tile_m * kBlockM, tc.hi, tc.bi);
                for (int j = 0; j < num_kv_tiles; ++j, ++g) {
                    const int s = g % kStages;
                    const uint32_t par = ((g / kStages) - 1) & 1;
                    if (g >= kStages) mbar_wait_backoff(bar(kBarKEmpty + s), par, LG_ATTN_SPIN_NS);
                    mbar_arrive_expect_tx(bar(kBarKFull + s), kTileBytes);
                    tma_load_4d(smem + kSmemK + s * kTileBytes, &tm_k, bar(kBarKFull + s), 0, j * kBlockN, tc.hi, tc.bk);
                    if (g >= kStages) mbar_wait_backoff(bar(kBarVEmpty + s), par, LG_ATTN_SPIN_NS);
                    mbar_arrive_expect_tx(bar(kBarVFull + s), kTileBytes);
                    tma_load_4d(smem + kSmemV + s * kTileBytes, &tm_v, bar(kBarVFull + s), 0, j * kBlockN, tc.hi, tc.bk);
                }
            }
        } else if (warp == 5 && lane == 0) {
            // ===================================================================== MMA issuer
            constexpr uint32_t idesc_qk = make_idesc_f16_f32acc(kBlockM, kBlockN, false);
            constexpr uint32_t idesc_pv = make_idesc_f16_f32acc(kBlockM, kHeadDim, true);
            const uint32_t t_s = tmem_base + kColS;
            const uint32_t t_o = tmem_base + kColO;
            const uint32_t t_p = tmem_base + kColP;
            const uint64_t desc_q = make_smem_desc_sw128(smem + kSmemQ, 16, 1024);
            const int n_work = first < total_tiles ? (total_tiles - first + stride - 1) / stride : 0;

            // S = Q K^T of running KV tile g; `last_of_q`: no later MMA reads this Q tile
            auto issue_qk = [&](int g, bool last_of_q) {
                const int s = g % kStages;
                mbar_wait_backoff(bar(kBarKFull + s), (g / kStages) & 1, LG_ATTN_SPIN_NS);
                tc_fence_after();
                const uint64_t desc_k = make_smem_desc_sw128(smem + kSmemK + s * kTileBytes, 16, 1024);
#pragma unroll
                for (int kk = 0; kk < kHeadDim / 16; ++kk)
                    mma_f16_ss(t_s, desc_q + 2u * kk, desc_k + 2u * kk, idesc_qk, kk > 0 ? 1u : 0u);
                mma_commit(bar(kBarKEmpty + s));
                if (last_of_q) mma_commit(bar(kBarQEmpty));
                mma_commit(bar(kBarSFull));
            };
            auto issue_pv = [&](int g, int j, int it) {
                const int s = g % kStages;
                mbar_wait_backoff(bar(kBarVFull + s), (g / kStages) & 1, LG_ATTN_SPIN_NS);
                mbar_wait_backoff(bar(kBarPFull), g & 1, LG_ATTN_SPIN_NS);
                if (j == 0 && it > 0) mbar_wait_backoff(bar(kBarOEmpty), (it - 1) & 1, LG_ATTN_SPIN_NS);  // epilogue has read O
                tc_fence_after();
                const uint64_t desc_v = make_smem_desc_sw128(smem + kSmemV + s * kTileBytes, kTileBytes, 1024);
#pragma unroll
                for (int kk = 0; kk < kBlockN / 16; ++kk)
                    mma_f16_ts(t_o, t_p + 8u * kk, desc_v + 128u * kk, idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
                mma_commit(bar(kBarVEmpty + s));
                mma_commit(bar(kBarPVDone));
            };

            if (n_work > 0) {
                mbar_wait(bar(kBarQFull), 0);
                issue_qk(0, num_kv_tiles == 1);
            }
            int g = 0;
            for (int it = 0; it < n_work; ++it) {
                for (int j = 0; j < num_kv_tiles; ++j, ++g) {
                    const bool next_in_q = j + 1 < num_kv_tiles;
                    const bool next_q = !next_in_q && it + 1 < n_work;
                    bool qk_issued = false;
                    if (next_in_q || next_q) {
                        mbar_wait_backoff(bar(kBarSEmpty), g & 1, LG_ATTN_SPIN_NS);  // S(g) now lives in softmax registers
                        if (next_in_q) {
                            issue_qk(g + 1, j + 2 == num_kv_tiles);
                            qk_issued = true;
                        } else if (mbar_try_wait(bar(kBarQFull), (it + 1) & 1)) {
                            issue_qk(g + 1, num_kv_tiles == 1);  // first S of the next Q tile, under this tile's tail
                            qk_issued = true;
                        }
                    }
                    issue_pv(g, j, it);
                    if (next_q && !qk_issued) {
                        mbar_wait_backoff(bar(kBarQFull), (it + 1) & 1, LG_ATTN_SPIN_NS);
                        issue_qk(g + 1, num_kv_tiles == 1);
                    }
                }
            }
        }
    } else {
        // ========================================================================= softmax warpgroup
        setmaxnreg_inc<kRegsSoftmax>();
        const uint32_t lane_sel = static_cast<uint32_t>(warp * 32) << 16;
        const uint32_t t_s = tmem_base + lane_sel + kColS;
        const uint32_t t_o = tmem_base + lane_sel + kColO;
        const uint32_t t_p = tmem_base + lane_sel + kColP;

        int g = 0, it = 0;
        for (int w = first; w < total_tiles; w += stride, ++it) {
            const TileCoord tc = tile_coord(w, tiles_m, ol);
#ifdef LG_ATTN_W4_TRACE
            const int trace_row = (it < 2 && blockIdx.x < 296) ? static_cast<int>(blockIdx.x) + 296 * it : -1;
            if (warp == 0 && lane == 0 && trace_row >= 0) {
                unsigned smid;
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                g_w4p_smid[trace_row] = smid;
            }
#endif
            float m_used = -INFINITY;  // row max the exponent is referenced to (raw score units)
            float m_thr = INFINITY, neg_m = 0.f;  // rescale threshold m_used + tau/c, and -m_used*c
            float l_sum = 0.f;
            int kv_left = n_kv;

            for (int j = 0; j < num_kv_tiles; ++j, ++g) {
                W4P_TRACE(j, 0);
                if (j == 0) {  // later tiles of this Q tile: s_full(g) was already awaited under tile g-1 (see below)
                    mbar_wait(bar(kBarSFull), g & 1);
                    tc_fence_after();
                }
                W4P_TRACE(j, 1);
                uint32_t sr[kBlockN];
#pragma unroll
                for (int c = 0; c < kBlockN / 32; ++c) tmem_ld_x32(t_s + 32u * c, sr + 32 * c);
                tmem_wait_ld();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(kBarSEmpty));
                W4P_TRACE(j, 2);

                // ragged KV tail: columns past n_kv get -inf (TMA zero-filled K rows gave S = 0 there)
                const int valid = kv_left;  // running count instead of n_kv - j*128: no constant-bank reload on the critical path
                kv_left -= kBlockN;
                if (valid < kBlockN) {
#pragma unroll
                    for (int c = 0; c < kBlockN; ++c)
                        if (c >= valid) sr[c] = 0xff800000u;
                }

                const float mx = row_max<kBlockN>(sr);

                // lazy rescale decision (warp-uniform so the tcgen05.ld/st below stay aligned).  The threshold and the scaled
                // reference are carried across tiles, so the common no-rescale step costs one compare + vote.
                float alpha = 1.f;
                bool rescale = false;
                if (j == 0) {
                    m_used = mx;
                    m_thr = mx + kTau / kScaleLog2;
                    neg_m = -mx * kScaleLog2;
                } else {
                    rescale = __any_sync(0xffffffffu, mx > m_thr);
                    if (rescale) {
                        const float m_new = fmaxf(m_used, mx);
                        alpha = ex2_approx((m_used - m_new) * kScaleLog2);
                        m_used = m_new;
                        m_thr = m_new + kTau / kScaleLog2;
                        neg_m = -m_new * kScaleLog2;
                    }
                }
                W4P_TRACE(j, 3);

                // one monolithic pass: splitting it into column chunks with the P stores / barrier probes in between
                // (measured: 722-738 vs 797 TFLOP/s) costs ptxas the freedom to spread the MUFU issue over the tile
                uint32_t pk[kBlockN / 2];
                const float tile_sum = softmax_exp_pass<kBlockN>(sr, pk, neg_m);

                W4P_TRACE(j, 4);
                // one barrier round trip per tile: QK(g+1) is issued after PV(g-1) by the same thread and tcgen05.commit covers
                // all of that thread's earlier MMAs, so s_full(g+1) also means the P buffer is free and O(g-1) final
                if (j + 1 < num_kv_tiles) {
                    mbar_wait(bar(kBarSFull), (g + 1) & 1);
                    tc_fence_after();
                } else if (j > 0) {
                    mbar_wait(bar(kBarPVDone), (g - 1) & 1);
                    tc_fence_after();
                }
                W4P_TRACE(j, 5);
                tmem_st_x32(t_p, pk);
                tmem_st_x32(t_p + 32u, pk + 32);
                if (rescale) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t orow[32];
                        tmem_ld_x32(t_o + 32u * h, orow);
                        tmem_wait_ld();
#pragma unroll
                        for (int c = 0; c < 32; ++c)
                            orow[c] = __float_as_uint(__uint_as_float(orow[c]) * alpha);
                        tmem_st_x32(t_o + 32u * h, orow);
                    }
                    l_sum *= alpha;
                }
                l_sum += tile_sum;
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(kBarPFull));
                W4P_TRACE(j, 6);
            }

            // --------------------------------------------------------------------------- epilogue
            mbar_wait(bar(kBarPVDone), (g - 1) & 1);
            tc_fence_after();
            uint32_t orow[kHeadDim];
            tmem_ld_x32(t_o, orow);
            tmem_ld_x32(t_o + 32u, orow + 32);
            tmem_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarOEmpty));  // the next Q tile's first PV may overwrite O now
            const float inv_l = 1.0f / l_sum;
            const int row = tc.tile_m * kBlockM + warp * 32 + lane;
            if (row < n_q) {
                if constexpr (!kOutF32) {
                    __half* o = reinterpret_cast<__half*>(out) + (tc.bi * ol.sb + tc.hi * ol.sh + row * ol.sr);
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        uint4 v;
                        v.x = pack_half2(__uint_as_float(orow[8 * c + 0]) * inv_l, __uint_as_float(orow[8 * c + 1]) * inv_l);
                        v.y = pack_half2(__uint_as_float(orow[8 * c + 2]) * inv_l, __uint_as_float(orow[8 * c + 3]) * inv_l);
                        v.z = pack_half2(__uint_as_float(orow[8 * c + 4]) * inv_l, __uint_as_float(orow[8 * c + 5]) * inv_l);
                        v.w = pack_half2(__uint_as_float(orow[8 * c + 6]) * inv_l, __uint_as_float(orow[8 * c + 7]) * inv_l);
                        st_global_v4(o + c * 8, v);
                    }
                } else {
                    float* o = reinterpret_cast<float*>(out) + (tc.bi * ol.sb + tc.hi * ol.sh + row * ol.sr);
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        uint4 v;
                        v.x = __float_as_uint(__uint_as_float(orow[4 * c + 0]) * inv_l);
                        v.y = __float_as_uint(__uint_as_float(orow[4 * c + 1]) * inv_l);
                        v.z = __float_as_uint(__uint_as_float(orow[4 * c + 2]) * inv_l);
                        v.w = __float_as_uint(__uint_as_float(orow[4 * c + 3]) * inv_l);
                        st_global_v4(o + c * 4, v);
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 6) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

}  // namespace w4p
}  // namespace lgattn
