// B200-native (sm_100a) fused multi-head attention for LightGlue, head_dim 64.
//
//   O = softmax(Q K^T / 8) V      per (batch, head); Q [B,H,N,64], K/V [B,H,M,64], O [B,H,N,64]
//
// Replaces the reference hot path (all under /root/reference/lightglue_attention_plugin/):
//   attention_headdim_64_fp16in_fp16out.cu:253-733   FA2-style kernel, FP16 out
//   attention_headdim_64_fp16in_fp32out.cu:253-703   FA2-style kernel, FP32 out
//   …fp16out.cu:736-857 / …fp32out.cu:706-830         pad / convert / unpad copy kernels
// with ONE kernel: ragged N, M are handled by TMA out-of-bounds zero fill (loads), an in-kernel
// -inf mask on the last KV tile, and row-predicated 128-bit stores — no pad / unpad passes.
//
// Structure (one CTA = one 128-row Q tile of one (batch, head); 2 CTAs co-reside per SM):
//   warps 0-3  softmax warpgroup: one S row per thread (tcgen05.ld 32x32b), FP32 online softmax,
//              P written back to TMEM as packed FP16, lazy O rescale, epilogue
//   warp  4    TMA producer (one elected lane): Q once, K/V tiles through a 2-stage mbarrier ring
//   warp  5    tcgen05.mma issuer (one elected lane): S = Q K^T (SS), O += P V (TS, P from TMEM)
//   warp  6    TMEM allocator;  warp 7 idle (register donor)
// TMEM columns: S [0,128) fp32 | O [128,192) fp32 | P [192,256) packed fp16 pairs.
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
#include "attn_sm100_common.cuh"
#include "softmax_math.cuh"

// poll interval of the single-thread service warps (TMA producer, MMA issuer)
#ifndef LG_ATTN_SPIN_NS
#define LG_ATTN_SPIN_NS 64
#endif

namespace lgattn {
namespace w4 {  // the 4-softmax-warp kernel: whole rows per thread, best sustained throughput on full grids

using namespace ptx;

constexpr int kThreads = 256;   // 4 softmax warps + TMA + MMA + TMEM-alloc + spare
// Q tile + 2 x (K, V) tiles (16 KiB each) + barriers, + 1 KiB slack for manual 1024-B alignment
#ifndef LG_ATTN_W4_EXTRA_SMEM
#define LG_ATTN_W4_EXTRA_SMEM 0   // tuning only: > 31 KiB forces one CTA per SM
#endif
constexpr uint32_t kSmemBytes = (1 + 2 * kStages) * 16384 + 1024 + 1024 + LG_ATTN_W4_EXTRA_SMEM;

// ------------------------------------------------------------------------------------------------
// shared-memory map (offsets from a 1024-byte aligned base; SWIZZLE_128B tiles need that alignment)
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kTileBytes = kBlockN * kHeadDim * 2;  // 16 KiB: 128 rows x 128 B
constexpr uint32_t kSmemQ = 0;
constexpr uint32_t kSmemK = kSmemQ + kTileBytes;               // kStages tiles
constexpr uint32_t kSmemV = kSmemK + kStages * kTileBytes;     // kStages tiles
constexpr uint32_t kSmemBar = kSmemV + kStages * kTileBytes;   // mbarriers + tmem slot
constexpr uint32_t kSmemStage = kSmemK;                        // epilogue staging aliases K/V ring
// barrier indices (8 bytes each)
enum : uint32_t {
    kBarQFull = 0,
    kBarKFull = 1,                    // [kStages]
    kBarKEmpty = kBarKFull + kStages, // [kStages]
    kBarVFull = kBarKEmpty + kStages,
    kBarVEmpty = kBarVFull + kStages,
    kBarSFull = kBarVEmpty + kStages,
    kBarSEmpty,
    kBarPFull,
    kBarPVDone,
    kNumBars
};
constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
constexpr uint32_t kSmemUsed = kSmemTmemSlot + 16;
static_assert(kSmemUsed + 1024 <= kSmemBytes, "shared memory budget");

// TMEM column map
constexpr uint32_t kTmemCols = 256;
constexpr uint32_t kColS = 0;
constexpr uint32_t kColO = 128;
constexpr uint32_t kColP = 192;

constexpr int kRegsSoftmax = 216;

// Tuning builds (profiles/r1_throughput_kernel_analysis.txt): LG_ATTN_W4_MERGE_WAIT=1 one barrier wait per tile;
// LG_ATTN_WHATIF=1|2|3 drop the row max | the exponentials | all arithmetic (WRONG results, timing decomposition only).
// A variant that scanned the next tile's row max out of TMEM under chunked exponentials measured 738 vs 805 TFLOP/s and
// was removed (git history: "max-ahead").
// Measured and dropped (same-box A/B against 832-835 TFLOP/s at B64 N2048): S read back in two halves with half the row max
// under the second TMEM load 827; arriving lane from elect.sync instead of lane == 0 835 (no difference); S(j+1) read back
// right behind the P(j) store (loop-carried S registers) spills 72 B at any softmax / service register split.
#ifndef LG_ATTN_W4_PIN_SMEM
#define LG_ATTN_W4_PIN_SMEM 1
#endif
#ifndef LG_ATTN_W4_MERGE_WAIT
#define LG_ATTN_W4_MERGE_WAIT 1
#endif
// (storing the epilogue straight from registers, one output row per thread, instead of through the swizzled staging
// buffer measured 792 vs 804 TFLOP/s on the same box and was dropped.)

constexpr int kRegsOther = 40;

// Optional per-tile phase tracing (tuning builds only, -DLG_ATTN_W4_TRACE): clock64 stamps of softmax
// warp 0 / lane 0 of the first CTAs, read back through lg_attn_debug_read_w4_trace().
#ifdef LG_ATTN_W4_TRACE
constexpr int kW4TraceCtas = 592, kW4TraceTiles = 16, kW4TraceEvents = 8;
__device__ long long g_w4_trace[kW4TraceCtas][kW4TraceTiles][kW4TraceEvents];
__device__ unsigned g_w4_smid[kW4TraceCtas];
#define W4_TRACE(j, ev)                                                                       \
    do {                                                                                      \
        if (warp == 0 && lane == 0 && blockIdx.x < kW4TraceCtas && (j) < kW4TraceTiles)       \
            g_w4_trace[blockIdx.x][(j)][(ev)] = clock64();                                    \
    } while (0)
// CTA-level stamps live in the unused event slot 7 of tile rows 0..5: entry, set-up done, loop done, stores done
// (clock64) and %globaltimer at entry / exit.
#define W4_TRACE_CTA(slot, val)                                                               \
    do {                                                                                      \
        if (warp == 0 && lane == 0 && blockIdx.x < kW4TraceCtas) g_w4_trace[blockIdx.x][(slot)][7] = (val); \
    } while (0)
__device__ __forceinline__ long long w4_globaltimer() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#else
#define W4_TRACE(j, ev)
#define W4_TRACE_CTA(slot, val)
#endif

template <bool kOutF32>
__global__ void __launch_bounds__(kThreads, 2)
lg_attn_hd64_fwd(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, void* __restrict__ out, int n_q, int n_kv,
                 int tiles_m, const OutLayout ol) {
    extern __shared__ uint8_t smem_raw[];
#if LG_ATTN_W4_PIN_SMEM
    // keep the aligned base in a register: otherwise ptxas re-derives it (S2UR SR_CgaCtaId + ULEA + ...) in front of every
    // mbarrier probe of the KV loop, ~1.5 % of the softmax warps' time in the ncu source view
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));
#else
    const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
#endif
    const uint32_t bars = smem + kSmemBar;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int tile_m = blockIdx.x % tiles_m;
    const int bh = blockIdx.x / tiles_m;
    const int hi = bh % ol.heads, bi = bh / ol.heads;
    const int bk = bi + ol.kv_shift >= ol.batch ? bi + ol.kv_shift - ol.batch : bi + ol.kv_shift;  // batch entry of K / V
    const int num_kv_tiles = (n_kv + kBlockN - 1) / kBlockN;
    W4_TRACE_CTA(0, clock64());
    W4_TRACE_CTA(4, w4_globaltimer());

    // Service warps run converged and issue through one elected lane (elect.sync): inside a `lane == 0` branch ptxas cannot prove
    // that the operands of UTMALDG / UTCHMMA / UTCBAR are warp-uniform and wraps every one of them in an R2UR + BRA.U.ANY waterfall
    // loop (~80 clk each; see attn_sm100_kernel_mq.cuh).  smem_u is the un-pinned (provably uniform) copy of the aligned base.
    const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars_u = smem_u + kSmemBar;
    auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
    if (warp == 4 && elect_one()) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
        mbar_init(ubar(kBarQFull), 1);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(ubar(kBarKFull + s), 1);
            mbar_init(ubar(kBarKEmpty + s), 1);
            mbar_init(ubar(kBarVFull + s), 1);
            mbar_init(ubar(kBarVEmpty + s), 1);
        }
        mbar_init(ubar(kBarSFull), 1);
        mbar_init(ubar(kBarSEmpty), 4);  // one arrival per softmax warp
        mbar_init(ubar(kBarPFull), 4);
        mbar_init(ubar(kBarPVDone), 1);
        fence_mbar_init();
        // first loads go out before the TMEM allocation / CTA barrier below: the 2.5 k clk until the first S is mostly
        // their latency (CTA-level trace), and only this thread ever arms these barriers
        mbar_arrive_expect_tx(ubar(kBarQFull), kTileBytes);
        tma_load_4d(smem_u + kSmemQ, &tm_q, ubar(kBarQFull), 0, tile_m * kBlockM, hi, bi);
        mbar_arrive_expect_tx(ubar(kBarKFull), kTileBytes);
        tma_load_4d(smem_u + kSmemK, &tm_k, ubar(kBarKFull), 0, 0, hi, bk);
        mbar_arrive_expect_tx(ubar(kBarVFull), kTileBytes);
        tma_load_4d(smem_u + kSmemV, &tm_v, ubar(kBarVFull), 0, 0, hi, bk);
    }
    __syncwarp();
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
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        if (warp == 4) {
            // ===================================================================== TMA producer
            for (int j = 1; j < num_kv_tiles; ++j) {   // tile 0 (and Q) were issued in the prologue
                const int s = j % kStages;
                const uint32_t par = ((j / kStages) - 1) & 1;
                if (j >= kStages) mbar_wait_backoff(ubar(kBarKEmpty + s), par, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(kBarKFull + s), kTileBytes);
                    tma_load_4d(smem_u + kSmemK + s * kTileBytes, &tm_k, ubar(kBarKFull + s), 0, j * kBlockN, hi, bk);
                }
                __syncwarp();
                if (j >= kStages) mbar_wait_backoff(ubar(kBarVEmpty + s), par, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(kBarVFull + s), kTileBytes);
                    tma_load_4d(smem_u + kSmemV + s * kTileBytes, &tm_v, ubar(kBarVFull + s), 0, j * kBlockN, hi, bk);
                }
                __syncwarp();
            }
        } else if (warp == 5) {
            // ===================================================================== MMA issuer
            const uint32_t t_s = tmem_u + kColS;
            const uint32_t t_o = tmem_u + kColO;
            const uint32_t t_p = tmem_u + kColP;

            auto issue_qk = [&](int j) {
                const int s = j % kStages;
                mbar_wait_backoff(ubar(kBarKFull + s), (j / kStages) & 1, LG_ATTN_SPIN_NS);
                tc_fence_after();
                if (elect_one()) {
                    umma_qk<kBlockN>(t_s, smem_u + kSmemQ, smem_u + kSmemK + s * kTileBytes);
                    mma_commit(ubar(kBarKEmpty + s));
                    mma_commit(ubar(kBarSFull));
                }
                __syncwarp();
            };

            mbar_wait(ubar(kBarQFull), 0);
            issue_qk(0);
            for (int j = 0; j < num_kv_tiles; ++j) {
                if (j + 1 < num_kv_tiles) {
                    mbar_wait_backoff(ubar(kBarSEmpty), j & 1, LG_ATTN_SPIN_NS);  // S(j) now lives in softmax registers
                    issue_qk(j + 1);
                }
                const int s = j % kStages;
                mbar_wait_backoff(ubar(kBarVFull + s), (j / kStages) & 1, LG_ATTN_SPIN_NS);
                mbar_wait_backoff(ubar(kBarPFull), j & 1, LG_ATTN_SPIN_NS);
                tc_fence_after();
                if (elect_one()) {
                    umma_pv<kBlockN>(t_o, t_p, smem_u + kSmemV + s * kTileBytes, kTileBytes, j > 0);
                    mma_commit(ubar(kBarVEmpty + s));
                    mma_commit(ubar(kBarPVDone));
                }
                __syncwarp();
            }
        }
    } else {
        // ========================================================================= softmax warpgroup
        setmaxnreg_inc<kRegsSoftmax>();
        const uint32_t lane_sel = static_cast<uint32_t>(warp * 32) << 16;
        const uint32_t t_s = tmem_base + lane_sel + kColS;
        const uint32_t t_o = tmem_base + lane_sel + kColO;
        const uint32_t t_p = tmem_base + lane_sel + kColP;

        W4_TRACE_CTA(1, clock64());
        float m_used = -INFINITY;  // row max the exponent is referenced to (raw score units)
        float m_thr = INFINITY, neg_m = 0.f;  // rescale threshold m_used + tau/c, and -m_used*c
        float l_sum = 0.f;
        int kv_left = n_kv;

#ifdef LG_ATTN_W4_TRACE
        if (warp == 0 && lane == 0 && blockIdx.x < kW4TraceCtas) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            g_w4_smid[blockIdx.x] = smid;
        }
#endif
        for (int j = 0; j < num_kv_tiles; ++j) {
            W4_TRACE(j, 0);
#if LG_ATTN_W4_MERGE_WAIT
            if (j == 0)   // later tiles: s_full(j) was already awaited under tile j-1 (see below)
#endif
            {
                mbar_wait(bar(kBarSFull), j & 1);
                tc_fence_after();
            }
            W4_TRACE(j, 1);
            uint32_t sr[kBlockN];
#pragma unroll
            for (int c = 0; c < kBlockN / 32; ++c) tmem_ld_x32(t_s + 32u * c, sr + 32 * c);
            tmem_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarSEmpty));
            W4_TRACE(j, 2);

            // ragged KV tail: columns past n_kv get -inf (TMA zero-filled K rows gave S = 0 there)
            const int valid = kv_left;  // running count instead of n_kv - j*128: no constant-bank reload on the critical path
            kv_left -= kBlockN;
            mask_kv_tail<kBlockN>(sr, valid);

#if defined(LG_ATTN_WHATIF) && (LG_ATTN_WHATIF == 1 || LG_ATTN_WHATIF == 3)
            const float mx = 0.f;   // what-if build (WRONG results in general): no row-max pass
#else
            const float mx = row_max<kBlockN>(sr);
#endif

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
            W4_TRACE(j, 3);

            // P = exp2(s*c - m*c) as packed FP16 pairs + the FP32 row sum (softmax_math.cuh)
            uint32_t pk[kBlockN / 2];
#if defined(LG_ATTN_WHATIF) && LG_ATTN_WHATIF >= 2
            // what-if builds (WRONG results): 2 = no exponentials (one FFMA2 + pack per pair), 3 = no arithmetic at all
            float tile_sum = 0.f;
#pragma unroll
            for (int c = 0; c < kBlockN; c += 2) {
#if LG_ATTN_WHATIF == 2
                float x0, x1;
                unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(sr[c]), __uint_as_float(sr[c + 1])),
                                       pack_f32x2(kScaleLog2, kScaleLog2), pack_f32x2(neg_m, neg_m)), x0, x1);
                pk[c / 2] = pack_half2(x0, x1);
#else
                pk[c / 2] = sr[c] ^ sr[c + 1];
#endif
            }
            tile_sum = __uint_as_float(pk[0]) + neg_m;
#else
            const float tile_sum = softmax_exp_pass<kBlockN>(sr, pk, neg_m);
#endif

            W4_TRACE(j, 4);
#if LG_ATTN_W4_MERGE_WAIT
            // ONE barrier round trip per tile instead of two (each costs 50-180 clk of MIO latency while the co-resident
            // CTA keeps the MUFU queue full): the MMA thread issues QK(j+1) after PV(j-1), and a tcgen05.commit covers every
            // earlier MMA of that thread, so s_full(j+1) — needed at the top of the next tile anyway — also says that
            // PV(j-1) has finished reading P and writing O.  Only the last tile still waits on pv_done itself.
            if (j + 1 < num_kv_tiles) {
                mbar_wait(bar(kBarSFull), (j + 1) & 1);
                tc_fence_after();
            } else if (j > 0) {
                mbar_wait(bar(kBarPVDone), (j - 1) & 1);
                tc_fence_after();
            }
#else
            if (j > 0) {
                mbar_wait(bar(kBarPVDone), (j - 1) & 1);  // P buffer free, O(j-1) final
                tc_fence_after();
            }
#endif
            W4_TRACE(j, 5);
            tmem_st_x32(t_p, pk);
            tmem_st_x32(t_p + 32u, pk + 32);
            if (rescale) {
                rescale_o_tmem<kHeadDim>(t_o, alpha);
                l_sum *= alpha;
            }
            l_sum += tile_sum;
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarPFull));
            W4_TRACE(j, 6);
        }

        // ------------------------------------------------------------------------------- epilogue
        mbar_wait(bar(kBarPVDone), (num_kv_tiles - 1) & 1);
        tc_fence_after();
        W4_TRACE_CTA(2, clock64());
        uint32_t orow[kHeadDim];
        tmem_ld_x32(t_o, orow);
        tmem_ld_x32(t_o + 32u, orow + 32);
        tmem_wait_ld();
        const float inv_l = 1.0f / l_sum;
        const int row0 = tile_m * kBlockM + warp * 32;  // first row of this warp's 32-row slab

        if constexpr (!kOutF32) {
            // stage 32 rows x 128 B per warp (16-B chunks XOR-swizzled by row), then 512-B coalesced stores
            const uint32_t stage = smem + kSmemStage + warp * (32 * 128);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    w[e] = pack_half2(__uint_as_float(orow[8 * c + 2 * e]) * inv_l,
                                      __uint_as_float(orow[8 * c + 2 * e + 1]) * inv_l);
                st_shared_v4(stage + lane * 128 + ((c ^ (lane & 7)) << 4), w[0], w[1], w[2], w[3]);
            }
            __syncwarp();
            __half* o = reinterpret_cast<__half*>(out) + (bi * ol.sb + hi * ol.sh);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + (lane >> 3);
                const int c = lane & 7;
                const uint4 v = ld_shared_v4(stage + r * 128 + ((c ^ (r & 7)) << 4));
                if (row0 + r < n_q) st_global_v4(o + (row0 + r) * ol.sr + c * 8, v);
            }
        } else {
            // 32 rows x 256 B per warp
            const uint32_t stage = smem + kSmemStage + warp * (32 * 256);
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                st_shared_v4(stage + lane * 256 + ((c ^ (lane & 15)) << 4),
                             __float_as_uint(__uint_as_float(orow[4 * c + 0]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 1]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 2]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 3]) * inv_l));
            }
            __syncwarp();
            float* o = reinterpret_cast<float*>(out) + (bi * ol.sb + hi * ol.sh);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int r = i * 2 + (lane >> 4);
                const int c = lane & 15;
                const uint4 v = ld_shared_v4(stage + r * 256 + ((c ^ (r & 15)) << 4));
                if (row0 + r < n_q) st_global_v4(o + (row0 + r) * ol.sr + c * 4, v);
            }
        }
    }

    W4_TRACE_CTA(3, clock64());
    W4_TRACE_CTA(5, w4_globaltimer());
    tc_fence_before();
    __syncthreads();
    if (warp == 6) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

}  // namespace w4
}  // namespace lgattn
