// B200-native (sm_100a) first half of LightGlue's ffn stack as ONE kernel:
//
//   hid = GELU(LayerNorm_512(cat(x, context) @ W1'^T + b1') * gamma + beta)          FP16 in, FP32 accumulate / statistics, FP16 out
//
// (lightglue_pytorch_with_plugin/lightglue.py:131-136 and :181-186: ffn[0] Linear(512, 512), ffn[1] LayerNorm, ffn[2] GELU; the
// attention out-projection is folded into W1' on the host, matcher.py).  Round 2 ran this as a cuBLAS GEMM followed by
// layernorm_gelu_kernel in place: the second pass re-reads and re-writes [rows, 512] FP16 and is bound by FP32 issue, not by HBM
// (358 warp instructions per row; 11 % of a forward), and nothing else runs beside it.  Here the LayerNorm + GELU arithmetic is
// the epilogue of a hand-written tcgen05 GEMM and runs UNDER the next tile's MMAs:
//
//   * one persistent CTA per SM, tile = 128 rows x all 512 hidden columns: a row's LayerNorm statistics never leave the CTA.
//     The FP32 accumulator is 128 lanes x 512 columns = all of TMEM, so it cannot be double-buffered; instead the 16 epilogue
//     warps DRAIN it in one short pass (tcgen05.ld, + bias, shifted sum / sum of squares, pack to FP16: 128 values = 64
//     registers per thread), hand TMEM back to the MMA warp, and only then do the expensive part (normalise, affine, erf-GELU,
//     store) from registers while the tensor pipe works on the next tile.
//   * warp w of the 16: TMEM lane quarter w & 3 (row 32 (w & 3) + lane of the tile), column group w >> 2 (128 columns).  The four
//     threads that share a row combine their (mean, M2) partials with Chan's formula through shared memory + one 128-thread
//     named barrier.
//   * operands by TMA into a 2-stage ring of K = 64 slabs (A 128 x 64, B 512 x 64 as two 256-row boxes, SWIZZLE_128B, K-major:
//     the same descriptors as the attention kernels' Q / K tiles); per slab 4 x 2 tcgen05.mma of 128 x 256 x 16.
//   * output: each warp stages 32 rows x 32 columns (2 KiB, SWIZZLE_64B pattern: conflict-free 16-byte stores) and one elected
//     lane issues a TMA store — rows past `rows` are clipped by the tensor map.
//
// Roofline and what bounds it (measured): 2 * 128 * 512 * 512 FLOP per tile = 8 192 clk of tensor pipe, 640 KiB
// of operands per tile out of L2 (W1' is re-read by every tile).  Neither is the limit: the phase trace (-DLG_FFN_TRACE) shows the
// epilogue warps busy back to back — drain ~3 000 clk, apply / store ~12 000 per tile — with the next tile's MMAs issued within
// 8 400 - 11 700 clk of the hand-back: the kernel is bound by the instruction issue of its LayerNorm + GELU arithmetic (~20
// instructions per element at IPC 0.7), the GEMM is what hides.  ncu: tensor pipe 47 - 51 % active, issue slots 66 - 69 %.
//
// No CUTLASS; encodings in ptx_sm100.cuh.
#pragma once
#include "ptx_sm100.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "glue_math.cuh"

// Timing decompositions (WRONG results by construction; tuning builds only, like the attention kernels' LG_ATTN_WHATIF):
//   LG_FFN_WHATIF 1: no GELU (normalise + affine only)   2: no tcgen05.mma issued (commits stay)
//                 3: W1' tiles loaded for the first tile of a CTA only (what the re-streaming of W1' through L2 costs)
#if defined(LG_FFN_WHATIF) && !defined(LG_ATTN_TUNING_BUILD)
#error "LG_FFN_WHATIF gives wrong results: tuning builds only (-DLG_ATTN_TUNING_BUILD)"
#endif
#ifndef LG_FFN_WHATIF
#define LG_FFN_WHATIF 0
#endif

// -DLG_FFN_TRACE (tuning builds): CTA 0 prints clock64 stamps of its first tiles — epilogue warp 0: accumulator full seen, drained
// (TMEM handed back), statistics agreed, tile stored; MMA warp: accumulator free seen, last MMA of the tile issued.
#ifdef LG_FFN_TRACE
#include <cstdio>
#define FFN_TRACE(arr, idx, slot) do { if (blockIdx.x == 0 && lane == 0 && (idx) < 8) arr[(idx) * 4 + (slot)] = clock64(); } while (0)
#else
#define FFN_TRACE(arr, idx, slot)
#endif

namespace lgffn {

using namespace lgattn::ptx;

constexpr int kBM = 128;          // rows per tile (= tcgen05.mma M, one TMEM lane per row)
constexpr int kN = 512;           // hidden width = LayerNorm width
constexpr int kK = 512;           // cat(x, context)
constexpr int kBK = 64;           // K slab: one SWIZZLE_128B atom of halves
constexpr int kKSteps = kK / kBK;
constexpr int kStages = 2;
constexpr int kEpiWarps = 16;
constexpr int kServiceWarps = 4;  // TMA producer, MMA issuer (+ TMEM allocator), two register donors (setmaxnreg wants a whole warpgroup)
constexpr int kThreads = 32 * (kEpiWarps + kServiceWarps);
constexpr int kColsPerWarp = kN / (kEpiWarps / 4);   // 128
constexpr int kRound = 32;        // columns per staging round

constexpr uint32_t kABytes = kBM * kBK * 2;            // 16 KiB
constexpr uint32_t kBHalfBytes = 256 * kBK * 2;        // 32 KiB (TMA boxes are at most 256 rows)
constexpr uint32_t kStageBytes = kABytes + 2 * kBHalfBytes;
constexpr uint32_t kSmemRing = 0;
constexpr uint32_t kSmemStage = kSmemRing + kStages * kStageBytes;          // 16 x 2 KiB output staging
constexpr uint32_t kStageTile = kRound * 32 * 2;                            // 32 rows x 32 halves
// the CTA-pair variant's ring is 16 KiB shorter: with 16 KiB more here its staging is double-buffered (2 x 2 KiB per warp starting
// 16 KiB below kSmemStage), so that a round only waits for the TMA store before last
constexpr uint32_t kSmemParams = kSmemStage + kEpiWarps * kStageTile + 16384;   // bias | gamma | beta, 512 floats each
constexpr uint32_t kSmemStats = kSmemParams + 3 * kN * 4;                   // [tile parity][column group][row] (mean, M2)
constexpr uint32_t kSmemBar = kSmemStats + 2 * 4 * kBM * 8;
constexpr int kMaxStages = 3;      // the CTA-pair variant has stages of 48 KiB: three fit where the one-CTA variant has two of 80 KiB
enum : uint32_t { kBarFull = 0, kBarEmpty = kMaxStages, kBarAccFull = 2 * kMaxStages, kBarAccEmpty, kNumBars };
constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
constexpr uint32_t kSmemUsed = kSmemTmemSlot + 16;
constexpr uint32_t kSmemBytes = kSmemUsed + 1024;   // + slack for the 1024-byte alignment of the swizzled tiles
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");

constexpr int kRegsLaunch = (65536 / kThreads) / 8 * 8;   // 96
constexpr int kRegsService = 32;
constexpr int kRegsEpi = ((kThreads * kRegsLaunch - kServiceWarps * 32 * kRegsService) / (kEpiWarps * 32)) / 8 * 8;   // 112

__device__ __forceinline__ float2 unpack_half2(uint32_t v) {
    return __half22float2(*reinterpret_cast<const __half2*>(&v));
}

// kPair: CTA pairs (clusters of 2, cta_group::2).  A pair works on 256 rows at a time: rank r holds rows 128 r .. of A, its own
// accumulator lanes and HALF of every W1' tile (rows 256 h + 128 r .. of column half h), the leader (rank 0) issues
// tcgen05.mma of M = 256.  Shared-memory traffic per SM drops from 96 B/clk of operand reads + 80 of TMA writes (above the
// 128 B/clk the memory delivers: the one-CTA main loop runs at ~1 000 TFLOP/s) to 64 + 47.  `tiles` counts 128-row tiles; a pair
// walks super-tiles blockIdx.x / 2 + i * gridDim.x / 2, rank r takes tile 2 * super-tile + r (rows past the end: TMA zero fill in,
// clipped out).
template <bool kPair>
__global__ void __launch_bounds__(kThreads, 1)
ffn1_ln_gelu_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                    const __grid_constant__ CUtensorMap tm_h, const __half* __restrict__ bias, const float* __restrict__ gamma,
                    const float* __restrict__ beta, float eps, int tiles) {
    // pair layout of a ring stage: A (16 KiB) | W1' rows of column half 0 (128 x 64: 16 KiB) | of column half 1 (16 KiB)
    constexpr uint32_t kBQuarterBytes = 128 * kBK * 2;
    constexpr uint32_t kStageStride = kPair ? kABytes + 2 * kBQuarterBytes : kStageBytes;   // 48 | 80 KiB
    constexpr int kNumStages = kPair ? 3 : kStages;   // ring depth (both fit the 160 KiB ring region)
    static_assert(kNumStages <= kMaxStages && kNumStages * kStageStride <= kStages * kStageBytes, "ring region");
    constexpr uint32_t kStageTx = kPair ? 2 * (kABytes + 2 * kBQuarterBytes) : kStageBytes;   // bytes landing per stage (pair: both CTAs')
    const uint32_t rank = kPair ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    extern __shared__ uint8_t smem_raw[];
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));   // pin (see attn_sm100_kernel_mq.cuh)
    const uint32_t bars = smem + kSmemBar;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    constexpr int kWarpTMA = kEpiWarps, kWarpMMA = kEpiWarps + 1;

    if (warp == kWarpTMA && lane == 0) {
        prefetch_tmap(&tm_x);
        prefetch_tmap(&tm_w);
        prefetch_tmap(&tm_h);
#pragma unroll
        for (int s = 0; s < kNumStages; ++s) {
            mbar_init(bar(kBarFull + s), 1);
            mbar_init(bar(kBarEmpty + s), 1);
        }
        mbar_init(bar(kBarAccFull), 1);
        mbar_init(bar(kBarAccEmpty), kPair ? 2 * kEpiWarps : kEpiWarps);   // pair: the leader's barrier collects both CTAs' epilogue warps
        fence_mbar_init();
    }
    if (warp == kWarpMMA) {
        if constexpr (kPair) {
            tmem_alloc_pair<512>(smem + kSmemTmemSlot);
            tmem_relinquish_pair();
        } else {
            tmem_alloc<512>(smem + kSmemTmemSlot);
            tmem_relinquish();
        }
    }
    // bias (FP16 -> FP32), gamma, beta into shared memory: read as broadcast 16-byte loads by every epilogue thread
    if (threadIdx.x < kN) {
        st_shared_f32(smem + kSmemParams + 4u * threadIdx.x, __half2float(bias[threadIdx.x]));
        // gamma / 2, beta / 2: the affine step then yields y / 2, which is what gelu_erf_of_half takes
        st_shared_f32(smem + kSmemParams + 4u * (kN + threadIdx.x), 0.5f * gamma[threadIdx.x]);
        st_shared_f32(smem + kSmemParams + 4u * (2 * kN + threadIdx.x), 0.5f * beta[threadIdx.x]);
    }
    tc_fence_before();
    __syncthreads();
    if constexpr (kPair) {   // the peer's barriers must exist before anything (TMA bytes, commits, arrivals) is aimed at them
        cluster_arrive();
        cluster_wait();
    }
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + kSmemTmemSlot));

    // one-CTA variant: tile index; pair variant: super-tile index (256 rows), this CTA's tile is 2 * index + rank
    const int first = kPair ? blockIdx.x / 2 : blockIdx.x, stride = kPair ? gridDim.x / 2 : gridDim.x;
    const int work = kPair ? (tiles + 1) / 2 : tiles;
    auto tile_of = [&](int w) { return kPair ? 2 * w + static_cast<int>(rank) : w; };

    if (warp >= kEpiWarps) {
        setmaxnreg_dec<kRegsService>();
        // service warps run converged and issue through elect.sync (uniform datapath, see attn_sm100_kernel_mq.cuh)
        const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t bars_u = smem_u + kSmemBar;
        auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
        if (warp == kWarpTMA) {
            // ===================================================================== operand producer
            int g = 0;
            for (int w = first; w < work; w += stride) {
                const int tile = tile_of(w);
#pragma unroll 1
                for (int k = 0; k < kKSteps; ++k, ++g) {
                    const int s = g % kNumStages;
                    if (g >= kNumStages) mbar_wait_backoff(ubar(kBarEmpty + s), ((g / kNumStages) - 1) & 1, 64);
                    if (elect_one()) {
                        const uint32_t dst = smem_u + kSmemRing + s * kStageStride;
                        if constexpr (kPair) {
                            // both CTAs' bytes complete on the leader's barrier; only the leader announces them
                            if (leader) mbar_arrive_expect_tx(ubar(kBarFull + s), kStageTx);
                            tma_load_4d_pair(dst, &tm_x, ubar(kBarFull + s), k * kBK, tile * kBM, 0, 0);
                            tma_load_4d_pair(dst + kABytes, &tm_w, ubar(kBarFull + s), k * kBK, 128 * static_cast<int>(rank), 0, 0);
                            tma_load_4d_pair(dst + kABytes + kBQuarterBytes, &tm_w, ubar(kBarFull + s), k * kBK, 256 + 128 * static_cast<int>(rank), 0, 0);
                        } else {
                            const bool load_b = LG_FFN_WHATIF != 3 || g < kKSteps;
                            mbar_arrive_expect_tx(ubar(kBarFull + s), load_b ? kStageBytes : kABytes);
                            tma_load_4d(dst, &tm_x, ubar(kBarFull + s), k * kBK, tile * kBM, 0, 0);
                            if (load_b) {
                                tma_load_4d(dst + kABytes, &tm_w, ubar(kBarFull + s), k * kBK, 0, 0, 0);
                                tma_load_4d(dst + kABytes + kBHalfBytes, &tm_w, ubar(kBarFull + s), k * kBK, 256, 0, 0);
                            }
                        }
                    }
                    __syncwarp();
                }
            }
        } else if (warp == kWarpMMA && leader) {
            // ===================================================================== MMA issuer (pair variant: the leader CTA only)
            constexpr uint32_t idesc = make_idesc_f16_f32acc(kPair ? 2 * kBM : kBM, 256, false);
            int g = 0, it = 0;
#ifdef LG_FFN_TRACE
            long long tm[32] = {};
#endif
            for (int w = first; w < work; w += stride, ++it) {
                if (it > 0) {   // the epilogue warps (of both CTAs) have drained the accumulator
                    if constexpr (kPair) mbar_wait_acquire_cluster(ubar(kBarAccEmpty), (it - 1) & 1);
                    else mbar_wait(ubar(kBarAccEmpty), (it - 1) & 1);
                }
                FFN_TRACE(tm, it, 0);
                tc_fence_after();
#pragma unroll 1
                for (int k = 0; k < kKSteps; ++k, ++g) {
                    const int s = g % kNumStages;
                    mbar_wait(ubar(kBarFull + s), (g / kNumStages) & 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a = smem_u + kSmemRing + s * kStageStride;
                        const uint64_t desc_a = make_smem_desc_sw128(a, 16, 1024);
                        const uint64_t desc_b0 = make_smem_desc_sw128(a + kABytes, 16, 1024);
                        const uint64_t desc_b1 = make_smem_desc_sw128(a + kABytes + (kPair ? kBQuarterBytes : kBHalfBytes), 16, 1024);
#pragma unroll
                        for (int kk = 0; kk < kBK / 16; ++kk) {
                            const uint32_t acc = (k > 0 || kk > 0) ? 1u : 0u;
                            if constexpr (kPair) {
                                mma_f16_ss_pair(tmem_u, desc_a + 2u * kk, desc_b0 + 2u * kk, idesc, acc);
                                mma_f16_ss_pair(tmem_u + 256u, desc_a + 2u * kk, desc_b1 + 2u * kk, idesc, acc);
                            } else if (LG_FFN_WHATIF != 2) {
                                mma_f16_ss(tmem_u, desc_a + 2u * kk, desc_b0 + 2u * kk, idesc, acc);
                                mma_f16_ss(tmem_u + 256u, desc_a + 2u * kk, desc_b1 + 2u * kk, idesc, acc);
                            }
                        }
                        if constexpr (kPair) {
                            mma_commit_pair(ubar(kBarEmpty + s), 3);   // both producers may refill this stage
                            if (k == kKSteps - 1) mma_commit_pair(ubar(kBarAccFull), 3);
                        } else {
                            mma_commit(ubar(kBarEmpty + s));
                            if (k == kKSteps - 1) mma_commit(ubar(kBarAccFull));
                        }
                    }
                    __syncwarp();
                }
                FFN_TRACE(tm, it, 1);
            }
#ifdef LG_FFN_TRACE
            if (blockIdx.x == 0 && lane == 0)
                for (int i = 1; i < 6; ++i) printf("mma tile %d: acc free seen +%lld, issued +%lld\n", i, tm[4 * i] - tm[4], tm[4 * i + 1] - tm[4]);
#endif
        }
    } else {
        // ========================================================================= epilogue warps
        setmaxnreg_inc<kRegsEpi>();
        const int quarter = warp & 3, cg = warp >> 2;
        const int row = quarter * 32 + lane;   // row inside the tile
        const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(cg * kColsPerWarp);
        const uint32_t p_bias = smem + kSmemParams + 4u * (cg * kColsPerWarp);
        const uint32_t p_gamma = p_bias + 4u * kN, p_beta = p_bias + 8u * kN;
        // one-CTA variant: one 2 KiB staging tile per warp; pair variant: two (the ring's unused 16 KiB + the extra 16 KiB)
        constexpr uint32_t kStgBufs = kPair ? 2 : 1;
        const uint32_t stg0 = smem + kSmemStage - (kPair ? 16384u : 0u) + static_cast<uint32_t>(warp) * kStageTile * kStgBufs;
        const uint32_t swz = static_cast<uint32_t>((lane >> 1) & 3);
        uint32_t round = 0;

        int it = 0;
#ifdef LG_FFN_TRACE
        long long te[32] = {};
#endif
        for (int w = first; w < work; w += stride, ++it) {
            const int tile = tile_of(w);
            mbar_wait(bar(kBarAccFull), it & 1);
            tc_fence_after();
            FFN_TRACE(te, it, 0);
            // ---- drain: accumulator + bias -> FP16 pairs in registers, shifted statistics in FP32
            uint32_t pk[kColsPerWarp / 2];
            float shift = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int c = 0; c < kColsPerWarp / 32; ++c) {
                uint32_t v[32];
                tmem_ld_x32(t_acc + 32u * c, v);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 b = ld_shared_v4f(p_bias + 4u * (32 * c + i));
                    const float a0 = __uint_as_float(v[i]) + b.x, a1 = __uint_as_float(v[i + 1]) + b.y;
                    const float a2 = __uint_as_float(v[i + 2]) + b.z, a3 = __uint_as_float(v[i + 3]) + b.w;
                    if (c == 0 && i == 0) shift = a0;
                    const float d0 = a0 - shift, d1 = a1 - shift, d2 = a2 - shift, d3 = a3 - shift;
                    s1 += (d0 + d1) + (d2 + d3);
                    s2 = fmaf(d0, d0, s2);
                    s2 = fmaf(d1, d1, s2);
                    s2 = fmaf(d2, d2, s2);
                    s2 = fmaf(d3, d3, s2);
                    pk[16 * c + i / 2] = pack_half2(a0, a1);
                    pk[16 * c + i / 2 + 1] = pack_half2(a2, a3);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {   // the next tile's MMAs may overwrite the accumulator now
                if (kPair && !leader) mbar_arrive_remote_release(dsmem_addr(bar(kBarAccEmpty), 0));
                else mbar_arrive(bar(kBarAccEmpty));
            }

            FFN_TRACE(te, it, 1);
            // ---- row statistics: this thread's 128 columns -> (mean, M2), combined with the row's three other column groups
            constexpr float kInvCols = 1.0f / kColsPerWarp;
            const float dm = s1 * kInvCols;
            const float mean_i = shift + dm;
            const float m2_i = fmaxf(fmaf(-dm, s1, s2), 0.f);
            const uint32_t st_base = smem + kSmemStats + static_cast<uint32_t>((it & 1) * 4 * kBM + row) * 8u;
            st_shared_v2f(st_base + static_cast<uint32_t>(cg * kBM) * 8u, mean_i, m2_i);
            named_bar_sync(1u + static_cast<uint32_t>(quarter), 128u);
            const float2 p0 = ld_shared_v2f(st_base), p1 = ld_shared_v2f(st_base + kBM * 8u);
            const float2 p2 = ld_shared_v2f(st_base + 2u * kBM * 8u), p3 = ld_shared_v2f(st_base + 3u * kBM * 8u);
            const float mean = 0.25f * ((p0.x + p1.x) + (p2.x + p3.x));
            const float e0 = p0.x - mean, e1 = p1.x - mean, e2 = p2.x - mean, e3 = p3.x - mean;
            const float m2 = ((p0.y + p1.y) + (p2.y + p3.y)) + kColsPerWarp * ((e0 * e0 + e1 * e1) + (e2 * e2 + e3 * e3));
            const float rstd = rsqrtf(m2 * (1.0f / kN) + eps);
            const float nmr = -mean * rstd;
            FFN_TRACE(te, it, 2);

            // ---- apply: normalise, affine, GELU, FP16, 32 columns per staging round, one TMA store per round
#pragma unroll
            for (int c = 0; c < kColsPerWarp / kRound; ++c) {
                uint32_t o[kRound / 2];
#pragma unroll
                for (int i = 0; i < kRound; i += 4) {
                    const float4 g4 = ld_shared_v4f(p_gamma + 4u * (kRound * c + i));
                    const float4 b4 = ld_shared_v4f(p_beta + 4u * (kRound * c + i));
                    const float2 x01 = unpack_half2(pk[(kRound * c + i) / 2]), x23 = unpack_half2(pk[(kRound * c + i) / 2 + 1]);
                    auto act = [](float v) { return LG_FFN_WHATIF == 1 ? v : lgmatch::gelu_erf_of_half(v); };
                    const float y0 = act(fmaf(fmaf(x01.x, rstd, nmr), g4.x, b4.x));
                    const float y1 = act(fmaf(fmaf(x01.y, rstd, nmr), g4.y, b4.y));
                    const float y2 = act(fmaf(fmaf(x23.x, rstd, nmr), g4.z, b4.z));
                    const float y3 = act(fmaf(fmaf(x23.y, rstd, nmr), g4.w, b4.w));
                    o[i / 2] = pack_half2(y0, y1);
                    o[i / 2 + 1] = pack_half2(y2, y3);
                }
                // the previous round's TMA store must have read the staging tile.  (Coalesced global stores read back from the
                // staging tile instead — no asynchronous read to wait for — were measured slower: 176 vs 161 us at 262 144 rows.)
                if (elect_one()) bulk_wait_group_read<kStgBufs - 1>();
                __syncwarp();
                const uint32_t stg = stg0 + (kPair ? (round & 1u) * kStageTile : 0u);
                const uint32_t stg_row = stg + static_cast<uint32_t>(lane) * 64u;
                ++round;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    st_shared_v4(stg_row + ((static_cast<uint32_t>(j) ^ swz) << 4), o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (elect_one()) {
                    tma_store_4d(&tm_h, stg, cg * kColsPerWarp + kRound * c, tile * kBM + quarter * 32, 0, 0);
                    bulk_commit_group();
                }
                __syncwarp();
            }
            FFN_TRACE(te, it, 3);
        }
#ifdef LG_FFN_TRACE
        if (blockIdx.x == 0 && warp == 0 && lane == 0)
            for (int i = 1; i < 6; ++i)
                printf("epi tile %d: acc full seen +%lld, drained +%lld, stats +%lld, stored +%lld\n", i, te[4 * i] - te[4], te[4 * i + 1] - te[4],
                       te[4 * i + 2] - te[4], te[4 * i + 3] - te[4]);
#endif
        // the staging tiles must outlive the stores that read them
        if (elect_one()) bulk_wait_group_read0();
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    if constexpr (kPair) {   // neither CTA may leave (or free TMEM) while the other can still signal its barriers / use its operands
        cluster_arrive();
        cluster_wait();
    }
    if (warp == kWarpMMA) {
        tc_fence_after();
        if constexpr (kPair) tmem_dealloc_pair<512>(tmem_base);
        else tmem_dealloc<512>(tmem_base);
    }
}

// ================================================================================================================================
// Few rows (a single image pair: 16 or 32 tiles): the kernel above would leave most SMs idle, one whole tile of arithmetic per CTA on
// the critical path.  Here a CLUSTER of 4 CTAs shares a 128-row tile and each takes 128 of the 512 hidden columns (a quarter of
// the MMAs, of W1' and of the epilogue); the LayerNorm statistics cross the cluster once per tile: every CTA pushes its per-row
// (mean, M2) over 128 columns into all four CTAs' shared memory (st.shared::cluster), one release / acquire cluster barrier
// publishes them.  One tile per cluster (no persistence), programmatic dependent launch: barrier init, TMEM allocation and the
// parameter loads (weights, not produced by the previous kernel) run before griddepcontrol.wait.
namespace small {

constexpr int kSplit = 4;                 // cluster size = column quarters
constexpr int kNC = kN / kSplit;          // 128 columns per CTA
constexpr int kStagesS = 4;
constexpr int kColsPerWarpS = kNC / (kEpiWarps / 4);   // 32
constexpr uint32_t kBBytesS = kNC * kBK * 2;           // 16 KiB
constexpr uint32_t kStageBytesS = kABytes + kBBytesS;  // 32 KiB
constexpr uint32_t kSmemRingS = 0;
constexpr uint32_t kSmemStageS = kSmemRingS + kStagesS * kStageBytesS;       // 16 x 2 KiB output staging
constexpr uint32_t kSmemParamsS = kSmemStageS + kEpiWarps * kStageTile;      // bias | gamma | beta, this CTA's 128 columns each
constexpr uint32_t kSmemStatsS = kSmemParamsS + 3 * kNC * 4;                 // [column group][row] (mean, M2) inside the CTA
constexpr uint32_t kSmemXStatsS = kSmemStatsS + 4 * kBM * 8;                 // [source CTA][row] (mean, M2) across the cluster
constexpr uint32_t kSmemBarS = kSmemXStatsS + kSplit * kBM * 8;
enum : uint32_t { kBarFullS = 0, kBarEmptyS = kStagesS, kBarAccFullS = 2 * kStagesS, kNumBarsS };
constexpr uint32_t kSmemTmemSlotS = kSmemBarS + 8 * kNumBarsS;
constexpr uint32_t kSmemBytesS = kSmemTmemSlotS + 16 + 1024;
static_assert(kSmemBytesS <= 227 * 1024, "shared memory budget");

// Chan's parallel-variance combine of four partials over n values each: (mean, M2) of the 4 n values
__device__ __forceinline__ float2 combine4(float2 p0, float2 p1, float2 p2, float2 p3, float n) {
    const float mean = 0.25f * ((p0.x + p1.x) + (p2.x + p3.x));
    const float e0 = p0.x - mean, e1 = p1.x - mean, e2 = p2.x - mean, e3 = p3.x - mean;
    return make_float2(mean, ((p0.y + p1.y) + (p2.y + p3.y)) + n * ((e0 * e0 + e1 * e1) + (e2 * e2 + e3 * e3)));
}

__global__ void __launch_bounds__(kThreads, 1)
ffn1_ln_gelu_small_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                          const __grid_constant__ CUtensorMap tm_h, const __half* __restrict__ bias, const float* __restrict__ gamma,
                          const float* __restrict__ beta, float eps) {
    extern __shared__ uint8_t smem_raw[];
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));
    const uint32_t bars = smem + kSmemBarS;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    constexpr int kWarpTMA = kEpiWarps, kWarpMMA = kEpiWarps + 1;
    const int rank = static_cast<int>(cluster_ctarank());
    const int tile = blockIdx.x / kSplit;
    const int col0 = rank * kNC;   // this CTA's first hidden column

    cluster_arrive();   // barrier #1 (waited for in front of the statistics push): every CTA of the cluster is running
    if (warp == kWarpTMA && lane == 0) {
        prefetch_tmap(&tm_x);
        prefetch_tmap(&tm_w);
        prefetch_tmap(&tm_h);
#pragma unroll
        for (int s = 0; s < kStagesS; ++s) {
            mbar_init(bar(kBarFullS + s), 1);
            mbar_init(bar(kBarEmptyS + s), 1);
        }
        mbar_init(bar(kBarAccFullS), 1);
        fence_mbar_init();
    }
    if (warp == kWarpMMA) {
        tmem_alloc<kNC>(smem + kSmemTmemSlotS);
        tmem_relinquish();
    }
    if (threadIdx.x < kNC) {
        st_shared_f32(smem + kSmemParamsS + 4u * threadIdx.x, __half2float(bias[col0 + threadIdx.x]));
        st_shared_f32(smem + kSmemParamsS + 4u * (kNC + threadIdx.x), 0.5f * gamma[col0 + threadIdx.x]);   // halved: see above
        st_shared_f32(smem + kSmemParamsS + 4u * (2 * kNC + threadIdx.x), 0.5f * beta[col0 + threadIdx.x]);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + kSmemTmemSlotS));

    // everything above touched only weights and this CTA's own state; the activations belong to the previous kernel
    pdl_wait_prior_grids();
    if (threadIdx.x == 0) pdl_launch_dependents();

    if (warp >= kEpiWarps) {
        const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t bars_u = smem_u + kSmemBarS;
        auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
        if (warp == kWarpTMA) {
#pragma unroll 1
            for (int k = 0; k < kKSteps; ++k) {
                const int s = k % kStagesS;
                if (k >= kStagesS) mbar_wait_backoff(ubar(kBarEmptyS + s), ((k / kStagesS) - 1) & 1, 64);
                if (elect_one()) {
                    const uint32_t dst = smem_u + kSmemRingS + s * kStageBytesS;
                    mbar_arrive_expect_tx(ubar(kBarFullS + s), kStageBytesS);
                    tma_load_4d(dst, &tm_x, ubar(kBarFullS + s), k * kBK, tile * kBM, 0, 0);
                    tma_load_4d(dst + kABytes, &tm_w, ubar(kBarFullS + s), k * kBK, col0, 0, 0);
                }
                __syncwarp();
            }
        } else if (warp == kWarpMMA) {
            constexpr uint32_t idesc = make_idesc_f16_f32acc(kBM, kNC, false);
#pragma unroll 1
            for (int k = 0; k < kKSteps; ++k) {
                const int s = k % kStagesS;
                mbar_wait(ubar(kBarFullS + s), (k / kStagesS) & 1);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t a = smem_u + kSmemRingS + s * kStageBytesS;
                    const uint64_t desc_a = make_smem_desc_sw128(a, 16, 1024);
                    const uint64_t desc_b = make_smem_desc_sw128(a + kABytes, 16, 1024);
#pragma unroll
                    for (int kk = 0; kk < kBK / 16; ++kk)
                        mma_f16_ss(tmem_u, desc_a + 2u * kk, desc_b + 2u * kk, idesc, (k > 0 || kk > 0) ? 1u : 0u);
                    mma_commit(ubar(kBarEmptyS + s));
                    if (k == kKSteps - 1) mma_commit(ubar(kBarAccFullS));
                }
                __syncwarp();
            }
        }
        cluster_wait();     // #1
        cluster_arrive();   // #2: the service warps have nothing to publish
        cluster_wait();
    } else {
        const int quarter = warp & 3, cg = warp >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(cg * kColsPerWarpS);
        const uint32_t p_bias = smem + kSmemParamsS + 4u * (cg * kColsPerWarpS);
        const uint32_t p_gamma = p_bias + 4u * kNC, p_beta = p_bias + 8u * kNC;
        const uint32_t stg = smem + kSmemStageS + static_cast<uint32_t>(warp) * kStageTile;
        const uint32_t stg_row = stg + static_cast<uint32_t>(lane) * 64u;
        const uint32_t swz = static_cast<uint32_t>((lane >> 1) & 3);

        mbar_wait(bar(kBarAccFullS), 0);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld_x32(t_acc, v);
        tmem_wait_ld();
        float a[kColsPerWarpS];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < kColsPerWarpS; i += 4) {
            const float4 b = ld_shared_v4f(p_bias + 4u * i);
            a[i] = __uint_as_float(v[i]) + b.x;
            a[i + 1] = __uint_as_float(v[i + 1]) + b.y;
            a[i + 2] = __uint_as_float(v[i + 2]) + b.z;
            a[i + 3] = __uint_as_float(v[i + 3]) + b.w;
        }
        const float shift = a[0];
#pragma unroll
        for (int i = 0; i < kColsPerWarpS; ++i) {
            const float d = a[i] - shift;
            s1 += d;
            s2 = fmaf(d, d, s2);
            // hold the pre-activation at FP16 precision, like the large-shape kernel and the library-GEMM path (identical numerics
            // whichever path a shape takes)
            a[i] = __half2float(__float2half_rn(a[i]));
        }
        constexpr float kInv = 1.0f / kColsPerWarpS;
        const float dm = s1 * kInv;
        st_shared_v2f(smem + kSmemStatsS + static_cast<uint32_t>(cg * kBM + row) * 8u, shift + dm, fmaxf(fmaf(-dm, s1, s2), 0.f));
        named_bar_sync(1u + static_cast<uint32_t>(quarter), 128u);
        const uint32_t sb = smem + kSmemStatsS + static_cast<uint32_t>(row) * 8u;
        const float2 cta = combine4(ld_shared_v2f(sb), ld_shared_v2f(sb + kBM * 8u), ld_shared_v2f(sb + 2u * kBM * 8u),
                                    ld_shared_v2f(sb + 3u * kBM * 8u), static_cast<float>(kColsPerWarpS));
        cluster_wait();     // #1: the peers are running, their shared memory may be written
        if (cg == 0) {
            const uint32_t mine = smem + kSmemXStatsS + static_cast<uint32_t>(rank * kBM + row) * 8u;
#pragma unroll
            for (int r = 0; r < kSplit; ++r) st_dsmem_v2(dsmem_addr(mine, static_cast<uint32_t>(r)), cta.x, cta.y);
        }
        cluster_arrive();   // #2 (release): the four CTAs' partials of every row are published ...
        cluster_wait();     //    (acquire) ... and visible; nobody writes a peer's shared memory after this
        const uint32_t xb = smem + kSmemXStatsS + static_cast<uint32_t>(row) * 8u;
        const float2 full = combine4(ld_shared_v2f(xb), ld_shared_v2f(xb + kBM * 8u), ld_shared_v2f(xb + 2u * kBM * 8u),
                                     ld_shared_v2f(xb + 3u * kBM * 8u), static_cast<float>(kNC));
        const float rstd = rsqrtf(full.y * (1.0f / kN) + eps);
        const float nmr = -full.x * rstd;
        uint32_t o[kColsPerWarpS / 2];
#pragma unroll
        for (int i = 0; i < kColsPerWarpS; i += 4) {
            const float4 g4 = ld_shared_v4f(p_gamma + 4u * i);
            const float4 b4 = ld_shared_v4f(p_beta + 4u * i);
            const float y0 = lgmatch::gelu_erf_of_half(fmaf(fmaf(a[i], rstd, nmr), g4.x, b4.x));
            const float y1 = lgmatch::gelu_erf_of_half(fmaf(fmaf(a[i + 1], rstd, nmr), g4.y, b4.y));
            const float y2 = lgmatch::gelu_erf_of_half(fmaf(fmaf(a[i + 2], rstd, nmr), g4.z, b4.z));
            const float y3 = lgmatch::gelu_erf_of_half(fmaf(fmaf(a[i + 3], rstd, nmr), g4.w, b4.w));
            o[i / 2] = pack_half2(y0, y1);
            o[i / 2 + 1] = pack_half2(y2, y3);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            st_shared_v4(stg_row + ((static_cast<uint32_t>(j) ^ swz) << 4), o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (elect_one()) {
            tma_store_4d(&tm_h, stg, col0 + cg * kColsPerWarpS, tile * kBM + quarter * 32, 0, 0);
            bulk_commit_group();
            bulk_wait_group_read0();
        }
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kWarpMMA) {
        tc_fence_after();
        tmem_dealloc<kNC>(tmem_base);
    }
}

}  // namespace small

}  // namespace lgffn
