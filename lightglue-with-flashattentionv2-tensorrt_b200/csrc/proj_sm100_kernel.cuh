// B200-native (sm_100a) input projection of a LightGlue transformer block with the rotary embedding in its epilogue:
//
//   out[:, 256 t .. 256 t + 255] = x @ W[256 t .. 256 t + 255, :]^T + b,   t < n_blocks  (q | k | v for the self block, qk | v for the cross block)
//   blocks t < n_rot are then rotated in place:  (a, b) -> (a cos_i - b sin_i, b cos_i + a sin_i)  for every pair (2i, 2i+1) of a head
//   ((cos_i, sin_i) of row r at cs[r][i]: the table of lg_match_posenc.  An epilogue thread owns a row, so reading its own 256 bytes
//   of the table directly costs 32 cache-line tags per load instruction — measured 201 us at 262 144 rows where the bytes take 130;
//   a pair-major copy of the table, read contiguously, thrashes the L1 sets instead (248 us).  The warp therefore loads the 32
//   rows' angles of a 32-column chunk coalesced and redistributes them through its staging tile.)
//
// (lightglue_pytorch_with_plugin/lightglue.py:139-147: Wqkv, unflatten, apply_cached_rotary_emb on q and k; :190-193: to_qk / to_v.)
// Round 2 ran this as a cuBLAS GEMM + rotary_kernel in place: K = 256 is short, the library GEMM reaches about half of the HBM
// roofline on it (18.5 us for 32768 rows x 768 columns where the bytes take 10), and the rotary pass re-reads and re-writes two
// thirds of what the GEMM has just written.  Here:
//
//   * work item = (128-row tile, 256-column block); the FP32 accumulator of an item is 256 TMEM columns, so two of them ping-pong and
//     the epilogue of item i (bias, rotary, FP16, store) runs under the MMAs of item i + 1;
//   * persistent CTAs walk items blockIdx.x + i * gridDim.x: at any moment the grid works on ~148 consecutive items, so the three
//     column blocks of a row tile read the same A tile out of L2;
//   * operands by TMA through a 3-stage ring of K = 64 slabs (A 128 x 64 + B 256 x 64, SWIZZLE_128B, K-major), 4 tcgen05.mma of
//     128 x 256 x 16 per slab;
//   * epilogue warp w of 16: TMEM lane quarter w & 3 (row 32 (w & 3) + lane), head w >> 2 of the block (64 columns = one head, so
//     a thread owns whole rotary pairs and all 32 angles of its row); output staged 32 rows x 32 columns at a time in the SWIZZLE_64B
//     pattern and written by TMA (rows past `rows` are clipped by the tensor map).
//
// Roofline: HBM.  Per row: 512 B of x + 256 B of (cos, sin) in, 2 n_blocks * 256 B out; the main loop needs 3 072 clk per item on the
// shared-memory side (192 KiB written by TMA + 192 KiB read by the MMAs at 128 B/clk) against 2 048 on the tensor pipe.
#pragma once
#include "ptx_sm100.cuh"

#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace lgproj {

using namespace lgattn::ptx;

constexpr int kBM = 128;          // rows per item
constexpr int kK = 256;           // embedding width
constexpr int kBK = 64;
constexpr int kKSteps = kK / kBK;
constexpr int kMaxBlocks = 3;
constexpr int kEpiWarps = 16;
constexpr int kServiceWarps = 4;
constexpr int kThreads = 32 * (kEpiWarps + kServiceWarps);
constexpr uint32_t kABytes = kBM * kBK * 2;            // 16 KiB
constexpr uint32_t kStageTile = 32 * 128;              // per warp: 32 rows x 16 (cos, sin) pairs of a chunk, then 32 rows x 32 halves of output

// kBN = columns per work item: 256 (4 heads; the throughput shape) or 128 (few rows — a single pair is 16 row tiles: 96 items of
// 128 columns cover the GPU where 48 of 256 do not, and all four K slabs of an item are in flight at once).
template <int kBN>
struct Cfg {
    static_assert(kBN == 256 || kBN == 128, "columns per item");
    static constexpr int kStages = kBN == 256 ? 3 : 4;
    static constexpr int kColsPerWarp = kBN / (kEpiWarps / 4);   // 64 = one head | 32 = half a head
    static constexpr uint32_t kBBytes = kBN * kBK * 2;            // 32 | 16 KiB
    static constexpr uint32_t kStageBytes = kABytes + kBBytes;
    static constexpr uint32_t kSmemRing = 0;
    static constexpr uint32_t kSmemStage = kSmemRing + kStages * kStageBytes;            // 16 x 4 KiB staging
    static constexpr uint32_t kSmemBias = kSmemStage + kEpiWarps * kStageTile;           // float[768]
    static constexpr uint32_t kSmemBar = kSmemBias + kMaxBlocks * 256 * 4;
    static constexpr uint32_t kBarFull = 0, kBarEmpty = kStages, kBarAccFull = 2 * kStages, kBarAccEmpty = 2 * kStages + 2,
                              kNumBars = 2 * kStages + 4;
    static constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
    static constexpr uint32_t kSmemBytes = kSmemTmemSlot + 16 + 1024;
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
    static constexpr int kTmemCols = 2 * kBN;   // two accumulators
};

constexpr int kRegsLaunch = (65536 / kThreads) / 8 * 8;   // 96
constexpr int kRegsService = 32;
constexpr int kRegsEpi = ((kThreads * kRegsLaunch - kServiceWarps * 32 * kRegsService) / (kEpiWarps * 32)) / 8 * 8;   // 112

template <int kBN>
__global__ void __launch_bounds__(kThreads, 1)
proj_rotary_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                   const __grid_constant__ CUtensorMap tm_o, const __half* __restrict__ bias, const float4* __restrict__ cs,
                   long long rows, int n_blocks, int n_rot, int items, int early_trigger) {
    using C = Cfg<kBN>;
    constexpr int kStages = C::kStages, kColsPerWarp = C::kColsPerWarp;
    constexpr uint32_t kBBytes = C::kBBytes, kStageBytes = C::kStageBytes, kSmemRing = C::kSmemRing, kSmemStage = C::kSmemStage,
                       kSmemBias = C::kSmemBias, kSmemBar = C::kSmemBar, kBarFull = C::kBarFull, kBarEmpty = C::kBarEmpty,
                       kBarAccFull = C::kBarAccFull, kBarAccEmpty = C::kBarAccEmpty, kSmemTmemSlot = C::kSmemTmemSlot;
    (void)kBBytes;
    const int per_row_tile = n_blocks * 256 / kBN;   // items per 128-row tile
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
        prefetch_tmap(&tm_o);
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar(kBarFull + s), 1);
            mbar_init(bar(kBarEmpty + s), 1);
        }
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar(kBarAccFull + a), 1);
            mbar_init(bar(kBarAccEmpty + a), kEpiWarps);
        }
        fence_mbar_init();
    }
    if (warp == kWarpMMA) {
        tmem_alloc<C::kTmemCols>(smem + kSmemTmemSlot);
        tmem_relinquish();
    }
    for (int i = threadIdx.x; i < n_blocks * 256; i += kThreads) st_shared_f32(smem + kSmemBias + 4u * i, __half2float(bias[i]));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + kSmemTmemSlot));

    // weights and own state only above; the activations belong to the previous kernel (programmatic dependent launch)
    pdl_wait_prior_grids();
    if (threadIdx.x == 0 && early_trigger) pdl_launch_dependents();

    const int first = blockIdx.x, stride = gridDim.x;

    if (warp >= kEpiWarps) {
        setmaxnreg_dec<kRegsService>();
        const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t bars_u = smem_u + kSmemBar;
        auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
        if (warp == kWarpTMA) {
            // ===================================================================== operand producer
            int g = 0;
            for (int item = first; item < items; item += stride) {
                const int rt = item / per_row_tile, t = item - rt * per_row_tile;
#pragma unroll 1
                for (int k = 0; k < kKSteps; ++k, ++g) {
                    const int s = g % kStages;
                    if (g >= kStages) mbar_wait_backoff(ubar(kBarEmpty + s), ((g / kStages) - 1) & 1, 64);
                    if (elect_one()) {
                        const uint32_t dst = smem_u + kSmemRing + s * kStageBytes;
                        mbar_arrive_expect_tx(ubar(kBarFull + s), kStageBytes);
                        tma_load_4d(dst, &tm_x, ubar(kBarFull + s), k * kBK, rt * kBM, 0, 0);
                        tma_load_4d(dst + kABytes, &tm_w, ubar(kBarFull + s), k * kBK, t * kBN, 0, 0);
                    }
                    __syncwarp();
                }
            }
        } else if (warp == kWarpMMA) {
            // ===================================================================== MMA issuer
            constexpr uint32_t idesc = make_idesc_f16_f32acc(kBM, kBN, false);
            int g = 0, it = 0;
            for (int item = first; item < items; item += stride, ++it) {
                const int a = it & 1;   // accumulator buffer
                if (it >= 2) mbar_wait(ubar(kBarAccEmpty + a), ((it >> 1) - 1) & 1);   // the epilogue has drained this buffer
                tc_fence_after();
#pragma unroll 1
                for (int k = 0; k < kKSteps; ++k, ++g) {
                    const int s = g % kStages;
                    mbar_wait(ubar(kBarFull + s), (g / kStages) & 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t sa = smem_u + kSmemRing + s * kStageBytes;
                        const uint64_t desc_a = make_smem_desc_sw128(sa, 16, 1024);
                        const uint64_t desc_b = make_smem_desc_sw128(sa + kABytes, 16, 1024);
#pragma unroll
                        for (int kk = 0; kk < kBK / 16; ++kk)
                            mma_f16_ss(tmem_u + static_cast<uint32_t>(a * kBN), desc_a + 2u * kk, desc_b + 2u * kk, idesc,
                                       (k > 0 || kk > 0) ? 1u : 0u);
                        mma_commit(ubar(kBarEmpty + s));
                        if (k == kKSteps - 1) mma_commit(ubar(kBarAccFull + a));
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ========================================================================= epilogue warps
        setmaxnreg_inc<kRegsEpi>();
        const int quarter = warp & 3, head = warp >> 2;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(head * kColsPerWarp);
        const uint32_t stg = smem + kSmemStage + static_cast<uint32_t>(warp) * kStageTile;
        const uint32_t stg_row = stg + static_cast<uint32_t>(lane) * 64u;
        const uint32_t swz = static_cast<uint32_t>((lane >> 1) & 3);

        int it = 0;
        for (int item = first; item < items; item += stride, ++it) {
            const int rt = item / per_row_tile, t = item - rt * per_row_tile;
            const int a = it & 1;
            const long long row0 = static_cast<long long>(rt) * kBM + quarter * 32;   // first row of this warp's slab
            const uint32_t p_bias = smem + kSmemBias + 4u * static_cast<uint32_t>(t * kBN + head * kColsPerWarp);
            // The warp's 32 rows x 128 B of the angle table for a 32-column chunk are loaded coalesced (4 rows per instruction) into
            // registers one chunk AHEAD — chunk 0 before the wait for the accumulator, chunk 1 under the arithmetic of chunk 0 — so
            // that the global-load latency is never on the epilogue's critical path (it was: ~900 clk twice per item, which made
            // the epilogue longer than the main loop it is meant to hide under).
            const int col_base = t * kBN + head * kColsPerWarp;   // first column of this warp's slice in the output row
            const bool rot = col_base < n_rot * 256;
            float4 pre[8];
            auto fetch = [&](int c) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int r_in = (lane >> 3) + 4 * j, seg = lane & 7;
                    pre[j] = make_float4(1.f, 0.f, 1.f, 0.f);
                    if (row0 + r_in < rows) pre[j] = __ldg(cs + (row0 + r_in) * 16 + 8 * (((col_base >> 5) + c) & 1) + seg);   // chunk = half of a head
                }
            };
            if (rot) fetch(0);
            mbar_wait(bar(kBarAccFull + a), (it >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < kColsPerWarp / 32; ++c) {
                uint32_t v[32];
                tmem_ld_x32(t_lane + static_cast<uint32_t>(a * kBN + 32 * c), v);
                tmem_wait_ld();
                if (c == kColsPerWarp / 32 - 1) {
                    // both halves are in registers: the MMAs of the item after next may overwrite this accumulator
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(kBarAccEmpty + a));
                }
                // the previous round's TMA store must have read the staging tile before anything is written to it
                if (elect_one()) bulk_wait_group_read0();
                __syncwarp();
                if (rot) {
                    // into the staging tile with the 128-byte XOR swizzle; then every lane reads its own row conflict-free
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int r_in = (lane >> 3) + 4 * j, seg = lane & 7;
                        st_shared_v4(stg + static_cast<uint32_t>(r_in) * 128u + (static_cast<uint32_t>(seg ^ (r_in & 7)) << 4),
                                     __float_as_uint(pre[j].x), __float_as_uint(pre[j].y), __float_as_uint(pre[j].z), __float_as_uint(pre[j].w));
                    }
                    __syncwarp();
                    if (c + 1 < kColsPerWarp / 32) fetch(c + 1);
                }
                uint32_t o[16];
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const float4 b = ld_shared_v4f(p_bias + 4u * (32 * c + i));
                    float x0 = __uint_as_float(v[i]) + b.x, x1 = __uint_as_float(v[i + 1]) + b.y;
                    float x2 = __uint_as_float(v[i + 2]) + b.z, x3 = __uint_as_float(v[i + 3]) + b.w;
                    if (rot) {
                        // (cos, sin, cos, sin) of this row's rotary pairs (16 per chunk) i / 2 and the next
                        const uint4 u = ld_shared_v4(stg + static_cast<uint32_t>(lane) * 128u + (static_cast<uint32_t>((i / 4) ^ (lane & 7)) << 4));
                        const float4 r = make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w));
                        const float y0 = fmaf(x0, r.x, -x1 * r.y), y1 = fmaf(x1, r.x, x0 * r.y);
                        const float y2 = fmaf(x2, r.z, -x3 * r.w), y3 = fmaf(x3, r.z, x2 * r.w);
                        x0 = y0; x1 = y1; x2 = y2; x3 = y3;
                    }
                    o[i / 2] = pack_half2(x0, x1);
                    o[i / 2 + 1] = pack_half2(x2, x3);
                }
                __syncwarp();   // every lane has read its angles: the tile is rewritten with the output below
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    st_shared_v4(stg_row + ((static_cast<uint32_t>(j) ^ swz) << 4), o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (elect_one()) {
                    tma_store_4d(&tm_o, stg, col_base + 32 * c, rt * kBM + quarter * 32, 0, 0);
                    bulk_commit_group();
                }
                __syncwarp();
            }
        }
        if (elect_one()) bulk_wait_group_read0();
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kWarpMMA) {
        tc_fence_after();
        tmem_dealloc<C::kTmemCols>(tmem_base);
    }
}

}  // namespace lgproj
