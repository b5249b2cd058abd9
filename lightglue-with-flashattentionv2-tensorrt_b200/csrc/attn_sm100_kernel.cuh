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
// Structure (one CTA = one 128-row Q tile of one (batch, head) x one KV range; 2 CTAs per SM):
//   warps 0-7  softmax: warp w owns TMEM lanes 32*(w%4).. (one S row per thread, tcgen05.ld
//              32x32b) and the 64-column half (w/4) of every 128-wide S tile; the two halves of a
//              row exchange their partial row max through shared memory (one 64-thread named
//              barrier per tile).  FP32 online softmax, exp2 split between MUFU.EX2 and an
//              FMA-pipe polynomial, P written back to TMEM as packed FP16, lazy O rescale, epilogue.
//   warp  8    TMA producer (one elected lane): Q once, K/V tiles through a 2-stage mbarrier ring;
//              the first loads are issued before the CTA-wide setup barrier
//   warp  9    tcgen05.mma issuer (one elected lane): S = Q K^T (SS), O += P V (TS, P from TMEM)
//   warp  10   TMEM allocator;  warp 11 idle (register donor)
// TMEM columns: S [0,128) fp32 | O [128,192) fp32 | P [192,256) packed fp16 pairs.
//
// Split-KV (latency shapes, e.g. B=1 N=M=1024: 32 Q tiles for 148 SMs): `num_splits` CTAs of one
// thread-block cluster share a Q tile and take consecutive KV ranges.  Every thread stores its slice
// of the unnormalised partial (O, m, l) straight into the receive area of the cluster CTA that owns
// the row (st.shared::cluster), a release/acquire cluster barrier publishes them, and every CTA then
// rescales / sums / normalises / stores the 128/num_splits rows it owns.  One launch, no workspace,
// no second kernel.  (Measured timeline, tools/trace_timeline.py: the 24 KiB a CTA sends cross at the
// ~20 B/clk DSMEM rate, ~1.2 us — the largest single item after the two KV tiles themselves.)
//
// Programmatic dependent launch: the prologue (barrier init, TMEM alloc, descriptor prefetch) runs
// before griddepcontrol.wait, so in a back-to-back chain it overlaps the previous kernel's tail.
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
namespace w8 {  // the 8-softmax-warp kernel: latency shapes, split-KV clusters

using namespace ptx;

constexpr int kSoftmaxWarps = 8;                     // two column-half warpgroups over the same 128 rows
constexpr int kThreads = 32 * (kSoftmaxWarps + 4);   // + TMA, MMA, TMEM-alloc, spare
// FP32-input instantiation (the plugin's kFLOAT branch, one launch): warps 10..23 convert the FP32 tiles that TMA brings in
// into the swizzled FP16 operand tiles (replaces QKV_convert_from_fp32_to_fp16, …fp32out.cu:706-804).  Q and K(0) are on the
// critical path of the first S, hence the many converter warps: 128 row pairs / 14 warps = 9-10 short iterations each.
#ifndef LG_ATTN_CVT_WARPS
#define LG_ATTN_CVT_WARPS 14
#endif
constexpr int kCvtWarps = LG_ATTN_CVT_WARPS;
constexpr int kThreadsF32 = 32 * (kSoftmaxWarps + 2 + kCvtWarps);
static_assert((2 + kCvtWarps) % 4 == 0, "setmaxnreg wants whole warpgroups of service warps");
// Q tile + 2 x (K, V) tiles (16 KiB each) + barriers / row statistics, + 1 KiB slack for 1024-B alignment
constexpr uint32_t kSmemBytes = (1 + 2 * kStages) * 16384 + 5120 + 1024;

// ------------------------------------------------------------------------------------------------
// shared-memory map (offsets from a 1024-byte aligned base; SWIZZLE_128B tiles need that alignment)
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kTileBytes = kBlockN * kHeadDim * 2;  // 16 KiB: 128 rows x 128 B
constexpr uint32_t kSmemQ = 0;
constexpr uint32_t kSmemRing = kSmemQ + kTileBytes;            // stage s: K at +s*32K, V at +s*32K+16K
constexpr uint32_t kStageBytes = 2 * kTileBytes;
constexpr uint32_t kSmemBar = kSmemRing + kStages * kStageBytes;  // mbarriers, tmem slot, row stats
constexpr uint32_t kSmemStage = kSmemRing;                     // epilogue staging aliases the ring
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
    kBarComb,                         // split-KV: bytes pushed into this CTA's receive area
    kBarRawQ,                         // FP32 input: the raw FP32 tile has landed in its buffer (TMA transaction barriers)
    kBarRawK,                         // [kStages]
    kBarRawV = kBarRawK + kStages,    // [kStages]
    kNumBars = kBarRawV + kStages
};
constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
// row-max / row-sum exchange between the two column halves: [parity][half][row] floats
constexpr uint32_t kSmemExch = (kSmemTmemSlot + 16 + 15u) & ~15u;
// split-KV: (unused since the direct-store combine; kept so the receive areas keep their offsets)
constexpr uint32_t kSmemStat = kSmemExch + 2 * 2 * kBlockM * 4;
// split-KV receive area, (m, l) part: [source split][owned row] float2  (1 KiB for any split count)
constexpr uint32_t kSmemCombStat = kSmemStat + kBlockM * 8;
constexpr uint32_t kSmemUsed = kSmemCombStat + kBlockM * 8;
static_assert(kSmemUsed + 1024 <= kSmemBytes, "shared memory budget");
// split-KV receive area, partial-O part: [source split][owned row][64] fp32 = 32 KiB, carried only by
// split launches (kSmemSplitExtra more dynamic shared memory; never aliased with live tiles because a
// peer may finish first).  Cluster launches of a TMEM kernel run one CTA per SM anyway.
constexpr uint32_t kSmemCombO = (kSmemUsed + 127u) & ~127u;
constexpr uint32_t kSmemSplitExtra = kBlockM * kHeadDim * 4 + 128;
// (Sizing the receive area for FP16 partials only — 16 KiB, which would let two CTAs share an SM so that the next launch of a
// chain is resident before this one retires — was measured and dropped: 10.0 instead of 6.7 us per call; with two slots per SM the
// block scheduler packs the next kernel's 128 CTAs two to an SM instead of spreading them.)
constexpr uint32_t kCombBytes = kBlockM * kHeadDim * 4 + kBlockM * 8;  // what each CTA receives in total
constexpr int kMaxSplits = 8;  // portable cluster size
// FP32 input: every operand tile gets a 32 KiB buffer.  TMA fills it with the FP32 tile as two boxes of 128 rows x 32 floats
// (128-byte rows, no swizzle: [box][row][32 floats]); the converter warps then rewrite it IN PLACE — a warp owns whole rows, reads
// both halves of a row pair into registers and writes the 128-byte-swizzled FP16 rows over the first 16 KiB (= box 0's rows of the
// same row pair) — so there is no staging ring to wait for and all first loads are in flight at once, like the FP16 path.
// Q at 0, stage s: K at 32K + s*64K, V at 64K + s*64K; the barrier / statistics / split-KV blocks sit kSmemShiftF32 further up.
constexpr uint32_t kRawBytes = kBlockN * kHeadDim * 4;   // 32 KiB
constexpr uint32_t kSmemShiftF32 = (1 + 2 * kStages) * (kRawBytes - kTileBytes);
constexpr uint32_t kSmemBytesF32 = kSmemBytes + kSmemSplitExtra + kSmemShiftF32;
static_assert(kSmemBytesF32 <= 227 * 1024, "shared memory budget (FP32 input)");

// TMEM column map
constexpr uint32_t kTmemCols = 256;
constexpr uint32_t kColS = 0;
constexpr uint32_t kColO = 128;
constexpr uint32_t kColP = 192;

constexpr int kHalfN = kBlockN / 2;  // S columns per softmax warp

// register budget: 384 threads launch at 80; the 4 service warps drop to 32, the 8 softmax warps take 104
constexpr int kRegsSoftmax = 104;
constexpr int kRegsOther = 32;

// LG_ATTN_COMB_MBAR=1: the split-KV combine is published per receiving CTA — every sender thread arrives (release.cluster)
// on the OWNER's kBarComb after its DSMEM stores and the owner waits (acquire.cluster) for the 256 arrivals of its own rows —
// instead of a second cluster-wide barrier that also waits for every other CTA's unrelated stores.
// Measured at B1 N=M=1024 (same box, 4 splits): cluster barrier 6.90 us per launch; 1 = one remote arrive per thread 7.17;
// 2 = fence.acq_rel.cluster + one remote arrive per warp 7.51.  The hardware cluster barrier stays.
#ifndef LG_ATTN_COMB_MBAR
#define LG_ATTN_COMB_MBAR 0
#endif

// launch flags
constexpr int kFlagEarlyTrigger = 2; // single-wave grid: release the dependent launch right after setup

// Optional timeline tracing (tuning builds only, -DLG_ATTN_TRACE): %globaltimer stamps of the first
// CTAs of each launch, read back through lg_attn_debug_read_trace().
#ifdef LG_ATTN_TRACE
constexpr int kTraceLaunches = 64, kTraceCtas = 4, kTraceEvents = 14;
__device__ unsigned long long g_trace[kTraceLaunches][kTraceCtas][kTraceEvents];
__device__ __forceinline__ void trace_event(int seq, int ev) {
    if (blockIdx.x < kTraceCtas) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_trace[seq % kTraceLaunches][blockIdx.x][ev] = t;
    }
}
#define LG_TRACE(ev) trace_event(trace_seq, ev)
#define LG_TRACE_ARG , int trace_seq
#else
#define LG_TRACE(ev)
#define LG_TRACE_ARG
#endif

// kBigSmem only distinguishes the instantiation launched with the extra split-KV receive area, so that
// its larger MaxDynamicSharedMemorySize attribute does not cap the occupancy of the regular launches.
template <bool kOutF32, bool kBigSmem = false, bool kInF32 = false>
__global__ void __launch_bounds__(kInF32 ? kThreadsF32 : kThreads, kInF32 ? 1 : 2)
lg_attn_hd64_fwd(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, void* __restrict__ out, int n_q, int n_kv,
                 int tiles_m, int num_splits, int flags, const OutLayout ol LG_TRACE_ARG) {
    if (threadIdx.x == 0) LG_TRACE(0);
    extern __shared__ uint8_t smem_raw[];
    // the aligned base is pinned in a register: otherwise ptxas re-derives it (S2UR SR_CgaCtaId + ULEA + ...) in front of every
    // mbarrier operation of the KV loop (+1.6 % on the throughput kernel, same-box A/B)
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));
    constexpr uint32_t kOff = kInF32 ? kSmemShiftF32 : 0u;   // FP32 input: everything behind the (larger) tile buffers moves up
    const uint32_t bars = smem + kSmemBar + kOff;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    // Split-KV: the num_splits CTAs of a cluster share one Q tile and take consecutive KV ranges.
    const int split = num_splits > 1 ? static_cast<int>(cluster_ctarank()) : 0;
    const int work = blockIdx.x / num_splits;
    const int tile_m = work % tiles_m;
    const int bh = work / tiles_m;
    const int hi = bh % ol.heads, bi = bh / ol.heads;
    const int bk = bi + ol.kv_shift >= ol.batch ? bi + ol.kv_shift - ol.batch : bi + ol.kv_shift;  // batch entry of K / V
    const int total_kv_tiles = (n_kv + kBlockN - 1) / kBlockN;
    const int kv_base = total_kv_tiles / num_splits, kv_rem = total_kv_tiles % num_splits;
    const int num_kv_tiles = kv_base + (split < kv_rem ? 1 : 0);            // >= 1 (host: splits <= tiles)
    const int kv_tile0 = split * kv_base + (split < kv_rem ? split : kv_rem);
    const int prologue_tiles = num_kv_tiles < kStages ? num_kv_tiles : kStages;

    // Service warps run converged and issue through one elected lane (elect.sync): inside a `lane == 0` branch ptxas cannot prove
    // that the operands of UTMALDG / UTCHMMA / UTCBAR are warp-uniform and wraps every one of them in an R2UR + BRA.U.ANY waterfall
    // loop (~80 clk each; see attn_sm100_kernel_mq.cuh).  smem_u is the un-pinned (provably uniform) copy of the aligned base.
    const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars_u = smem_u + kSmemBar + kOff;
    auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
    auto usmem_k = [&](int s) {
        return kInF32 ? smem_u + kRawBytes + static_cast<uint32_t>(s) * 2u * kRawBytes : smem_u + kSmemRing + static_cast<uint32_t>(s) * kStageBytes;
    };
    auto usmem_v = [&](int s) { return usmem_k(s) + (kInF32 ? kRawBytes : kTileBytes); };
    // FP32 input: a raw tile = two 128-row x 32-float boxes into the tile's 32 KiB buffer (elected lane of the producer warp)
    auto load_raw = [&](uint32_t dst, const CUtensorMap* tm, uint32_t full, int row, int b) {
        mbar_arrive_expect_tx(full, kRawBytes);
        tma_load_4d(dst, tm, full, 0, row, hi, b);
        tma_load_4d(dst + kRawBytes / 2, tm, full, 32, row, hi, b);
    };
    if (warp == kSoftmaxWarps && elect_one()) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
        // FP16 input: the operand tiles are filled by TMA (one arrive.expect_tx each); FP32 input: by the converter warps
        constexpr uint32_t kFillers = kInF32 ? kCvtWarps : 1;
        mbar_init(ubar(kBarQFull), kFillers);
        for (int s = 0; s < kStages; ++s) {
            mbar_init(ubar(kBarKFull + s), kFillers);
            mbar_init(ubar(kBarKEmpty + s), 1);
            mbar_init(ubar(kBarVFull + s), kFillers);
            mbar_init(ubar(kBarVEmpty + s), 1);
        }
        if constexpr (kInF32) {
            mbar_init(ubar(kBarRawQ), 1);
            for (int s = 0; s < kStages; ++s) {
                mbar_init(ubar(kBarRawK + s), 1);
                mbar_init(ubar(kBarRawV + s), 1);
            }
        }
        mbar_init(ubar(kBarSFull), 1);
        mbar_init(ubar(kBarSEmpty), kSoftmaxWarps);  // one arrival per softmax warp
        mbar_init(ubar(kBarPFull), kSoftmaxWarps);
        mbar_init(ubar(kBarPVDone), 1);
#if LG_ATTN_COMB_MBAR == 2
        {   // one arrival per (sender warp or half-warp) whose rows this CTA owns
            const int rp = num_splits > 1 ? kBlockM / num_splits : kBlockM;
            mbar_init(ubar(kBarComb), static_cast<uint32_t>(2 * rp / (rp < 32 ? rp : 32) * num_splits));
        }
#else
        mbar_init(ubar(kBarComb), LG_ATTN_COMB_MBAR ? kSoftmaxWarps * 32 : 1);
#endif
        fence_mbar_init();
        // Everything above is independent of earlier kernels in the stream; the loads are not.
        pdl_wait_prior_grids();
        LG_TRACE(12);
        // First loads go out before the CTA-wide setup barrier (they only need this thread's inits).
        if constexpr (kInF32) {
            load_raw(smem_u + kSmemQ, &tm_q, ubar(kBarRawQ), tile_m * kBlockM, bi);
            for (int j = 0; j < prologue_tiles; ++j) {
                load_raw(usmem_k(j), &tm_k, ubar(kBarRawK + j), (kv_tile0 + j) * kBlockN, bk);
                load_raw(usmem_v(j), &tm_v, ubar(kBarRawV + j), (kv_tile0 + j) * kBlockN, bk);
            }
        } else {
            mbar_arrive_expect_tx(ubar(kBarQFull), kTileBytes);
            tma_load_4d(smem_u + kSmemQ, &tm_q, ubar(kBarQFull), 0, tile_m * kBlockM, hi, bi);
            for (int j = 0; j < prologue_tiles; ++j) {
                mbar_arrive_expect_tx(ubar(kBarKFull + j), kTileBytes);
                tma_load_4d(usmem_k(j), &tm_k, ubar(kBarKFull + j), 0, (kv_tile0 + j) * kBlockN, hi, bk);
                mbar_arrive_expect_tx(ubar(kBarVFull + j), kTileBytes);
                tma_load_4d(usmem_v(j), &tm_v, ubar(kBarVFull + j), 0, (kv_tile0 + j) * kBlockN, hi, bk);
            }
        }
        LG_TRACE(13);
    }
    __syncwarp();
    if (warp == kSoftmaxWarps + 2) {
        tmem_alloc<kTmemCols>(smem + kOff + kSmemTmemSlot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + kOff + kSmemTmemSlot));
    if (threadIdx.x == 0) {
        // Small (single-wave) grids let the next kernel in the stream start its prologue right away;
        // large grids trigger at their epilogue so the dependents do not take slots from their own tail.
        if (flags & kFlagEarlyTrigger) pdl_launch_dependents();
        LG_TRACE(1);
    }
    // split-KV cluster barrier #1 ("my kBarComb is armed"): arrive now, wait only right before the
    // first push into a peer, by which time every peer has long arrived.
    if (num_splits > 1) cluster_arrive();

    if (warp >= kSoftmaxWarps) {
        setmaxnreg_dec<kRegsOther>();
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        if (kInF32 && warp == kSoftmaxWarps) {
            // ===================================================================== TMA producer (raw FP32 tiles)
            for (int j = prologue_tiles; j < num_kv_tiles; ++j) {
                const int s = j % kStages;
                const uint32_t par = ((j / kStages) - 1) & 1;
                mbar_wait_backoff(ubar(kBarKEmpty + s), par, LG_ATTN_SPIN_NS);   // the MMAs that read this buffer's FP16 tile are done
                if (elect_one()) load_raw(usmem_k(s), &tm_k, ubar(kBarRawK + s), (kv_tile0 + j) * kBlockN, bk);
                __syncwarp();
                mbar_wait_backoff(ubar(kBarVEmpty + s), par, LG_ATTN_SPIN_NS);
                if (elect_one()) load_raw(usmem_v(s), &tm_v, ubar(kBarRawV + s), (kv_tile0 + j) * kBlockN, bk);
                __syncwarp();
            }
        } else if (kInF32 && warp >= kSoftmaxWarps + 2) {
            // ===================================================================== FP32 -> FP16 converters (warps 10..), in place
            // One warp iteration = one row pair (2p, 2p+1): lane l reads the float4 q4 = l % 8 of box (l / 8) % 2 of row 2p + l / 16
            // (consecutive lanes read consecutive 16 B), all 32 loads complete (__syncwarp), then every lane writes its four halves to
            // columns 32 box + 4 q4 of the 128-byte-swizzled FP16 row (16-byte chunk index XOR row % 8, as TMA SWIZZLE_128B would
            // have written it) — bytes that belong to box 0 of the SAME row pair, which this warp has just read.  Round to nearest
            // even, like QKV_convert_from_fp32_to_fp16 (…fp32out.cu:706-768).
            const int cw = warp - (kSoftmaxWarps + 2);
            const int total_loads = 1 + 2 * num_kv_tiles;   // 0 = Q, 1 + 2j = K(j), 2 + 2j = V(j)
            for (int i = 0; i < total_loads; ++i) {
                const int j = (i - 1) / 2, s = j % kStages;
                const uint32_t buf = i == 0 ? smem_u + kSmemQ : ((i & 1) ? usmem_k(s) : usmem_v(s));
                const uint32_t raw = i == 0 ? ubar(kBarRawQ) : ((i & 1) ? ubar(kBarRawK + s) : ubar(kBarRawV + s));
                const uint32_t full = i == 0 ? ubar(kBarQFull) : ((i & 1) ? ubar(kBarKFull + s) : ubar(kBarVFull + s));
                mbar_wait(raw, i == 0 ? 0u : static_cast<uint32_t>(j / kStages) & 1u);
                const int row_in_pair = lane >> 4, box = (lane >> 3) & 1, q4 = lane & 7;
#pragma unroll 2
                for (int p = cw; p < kBlockN / 2; p += kCvtWarps) {
                    const int r = 2 * p + row_in_pair;
                    const uint4 u = ld_shared_v4(buf + static_cast<uint32_t>(box) * (kRawBytes / 2) + static_cast<uint32_t>(r) * 128u +
                                                 static_cast<uint32_t>(q4) * 16u);
                    const uint32_t h0 = pack_half2(__uint_as_float(u.x), __uint_as_float(u.y));
                    const uint32_t h1 = pack_half2(__uint_as_float(u.z), __uint_as_float(u.w));
                    __syncwarp();   // every lane holds its part of the row pair before anything of it is overwritten
                    const uint32_t chunk = static_cast<uint32_t>(box * 4 + (q4 >> 1)) ^ static_cast<uint32_t>(r & 7);
                    st_shared_v2u(buf + static_cast<uint32_t>(r) * 128u + (chunk << 4) + static_cast<uint32_t>(q4 & 1) * 8u, h0, h1);
                }
                fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's async-proxy reads
                __syncwarp();
                if (lane == 0) mbar_arrive(full);
            }
        } else if (!kInF32 && warp == kSoftmaxWarps) {
            // ===================================================================== TMA producer
            for (int j = prologue_tiles; j < num_kv_tiles; ++j) {
                const int s = j % kStages;
                const uint32_t par = ((j / kStages) - 1) & 1;
                mbar_wait_backoff(ubar(kBarKEmpty + s), par, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(kBarKFull + s), kTileBytes);
                    tma_load_4d(usmem_k(s), &tm_k, ubar(kBarKFull + s), 0, (kv_tile0 + j) * kBlockN, hi, bk);
                }
                __syncwarp();
                mbar_wait_backoff(ubar(kBarVEmpty + s), par, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(kBarVFull + s), kTileBytes);
                    tma_load_4d(usmem_v(s), &tm_v, ubar(kBarVFull + s), 0, (kv_tile0 + j) * kBlockN, hi, bk);
                }
                __syncwarp();
            }
        } else if (warp == kSoftmaxWarps + 1) {
            // ===================================================================== MMA issuer
            const uint32_t t_s = tmem_u + kColS;
            const uint32_t t_o = tmem_u + kColO;
            const uint32_t t_p = tmem_u + kColP;

            auto issue_qk = [&](int j) {
                const int s = j % kStages;
                mbar_wait(ubar(kBarKFull + s), (j / kStages) & 1);
                tc_fence_after();
                if (elect_one()) {
                    umma_qk<kBlockN>(t_s, smem_u + kSmemQ, usmem_k(s));
                    mma_commit(ubar(kBarKEmpty + s));
                    mma_commit(ubar(kBarSFull));
                }
                __syncwarp();
            };

            mbar_wait(ubar(kBarQFull), 0);
            if (lane == 0) LG_TRACE(2);
            issue_qk(0);
            for (int j = 0; j < num_kv_tiles; ++j) {
                if (j + 1 < num_kv_tiles) {
                    mbar_wait(ubar(kBarSEmpty), j & 1);  // S(j) now lives in softmax registers
                    issue_qk(j + 1);
                }
                const int s = j % kStages;
                mbar_wait(ubar(kBarVFull + s), (j / kStages) & 1);
                mbar_wait(ubar(kBarPFull), j & 1);
                tc_fence_after();
                if (elect_one()) {
                    umma_pv<kBlockN>(t_o, t_p, usmem_v(s), kTileBytes, j > 0);
                    mma_commit(ubar(kBarVEmpty + s));
                    mma_commit(ubar(kBarPVDone));
                }
                __syncwarp();
            }
        }
        if (num_splits > 1) {  // service warps just take part in the cluster barrier(s)
            cluster_wait();
#if !LG_ATTN_COMB_MBAR
            cluster_arrive();
            cluster_wait();
#endif
        }
    } else {
        // ========================================================================= softmax warps
        setmaxnreg_inc<kRegsSoftmax>();
        const int quarter = warp & 3;          // TMEM lane quarter: rows 32*quarter + lane
        const int half = warp >> 2;            // S columns [64*half, 64*half + 64)
        const int row = quarter * 32 + lane;   // row inside the Q tile
        const uint32_t lane_sel = static_cast<uint32_t>(quarter * 32) << 16;
        const uint32_t t_s = tmem_base + lane_sel + kColS + static_cast<uint32_t>(half * kHalfN);
        const uint32_t t_o = tmem_base + lane_sel + kColO + static_cast<uint32_t>(half * 32);
        const uint32_t t_p = tmem_base + lane_sel + kColP + static_cast<uint32_t>(half * 32);
        const uint32_t exch = smem + kOff + kSmemExch;  // [parity][half][row] floats
        // named barrier shared with the other column half of the same rows: ids 1..4 (immediates, so
        // the CTA reserves 6 of the SM's 16 hardware barriers and two CTAs still fit on one SM)
        auto pair_sync = [&]() {
            switch (quarter) {
                case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
                case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
                case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
                default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
            }
        };
        auto exch_addr = [&](int parity, int h) {
            return exch + static_cast<uint32_t>(((parity * 2 + h) * kBlockM + row) * 4);
        };

        float m_used = -INFINITY;  // row max the exponent is referenced to (raw score units)
        float l_sum = 0.f;         // this half's share of the row sum

        for (int j = 0; j < num_kv_tiles; ++j) {
            mbar_wait(bar(kBarSFull), j & 1);
            if (j == 0 && threadIdx.x == 0) LG_TRACE(3);
            tc_fence_after();
            uint32_t sr[kHalfN];
            tmem_ld_x32(t_s, sr);
            tmem_ld_x32(t_s + 32u, sr + 32);
            tmem_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarSEmpty));

            // ragged KV tail: columns past n_kv get -inf (TMA zero-filled K rows gave S = 0 there)
            const int valid = n_kv - (kv_tile0 + j) * kBlockN - half * kHalfN;
            mask_kv_tail<kHalfN>(sr, valid);

            float mx = row_max<kHalfN>(sr);
            // combine with the other column half of the same rows
            st_shared_f32(exch_addr(j & 1, half), mx);
            pair_sync();
            mx = fmaxf(mx, ld_shared_f32(exch_addr(j & 1, half ^ 1)));

            // lazy rescale decision (warp-uniform so the tcgen05.ld/st below stay aligned; both
            // halves see identical row maxima, so they take the same branch)
            float alpha = 1.f;
            bool rescale = false;
            if (j == 0) {
                m_used = mx;
            } else {
                const bool need = (mx - m_used) * kScaleLog2 > kTau;
                rescale = __any_sync(0xffffffffu, need);
                if (rescale) {
                    const float m_new = fmaxf(m_used, mx);
                    alpha = ex2_approx((m_used - m_new) * kScaleLog2);
                    m_used = m_new;
                }
            }
            const float neg_m = -m_used * kScaleLog2;

            // P = exp2(s*c - m*c) as packed FP16 pairs + this half's FP32 row sum (softmax_math.cuh)
            uint32_t pk[kHalfN / 2];
            const float tile_sum = softmax_exp_pass<kHalfN>(sr, pk, neg_m);

            if (j > 0) {
                mbar_wait(bar(kBarPVDone), (j - 1) & 1);  // P buffer free, O(j-1) final
                tc_fence_after();
            }
            tmem_st_x32(t_p, pk);
            if (rescale) {
                rescale_o_tmem<32>(t_o, alpha);   // this warp's 32 of the 64 O columns
                l_sum *= alpha;
            }
            l_sum += tile_sum;
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(kBarPFull));
        }

        // ------------------------------------------------------------------------------- epilogue
        if (threadIdx.x == 0) {
            if (!(flags & kFlagEarlyTrigger)) pdl_launch_dependents();
            LG_TRACE(4);
        }
        mbar_wait(bar(kBarPVDone), (num_kv_tiles - 1) & 1);
        tc_fence_after();
        if (threadIdx.x == 0) LG_TRACE(5);
        // full row sum = this half's + the other half's (the parity slot last used by tile n-2 is free)
        st_shared_f32(exch_addr(num_kv_tiles & 1, half), l_sum);
        pair_sync();
        const float l_row = l_sum + ld_shared_f32(exch_addr(num_kv_tiles & 1, half ^ 1));
        uint32_t orow[32];  // this warp's 32 of the 64 output columns
        tmem_ld_x32(t_o, orow);
        tmem_wait_ld();
        const int row0 = tile_m * kBlockM + quarter * 32;  // first row of this warp's 32-row slab
        // All MMAs and TMA loads of this CTA are complete (PVDone of the last tile), so its K/V ring is
        // dead: stage through it (16-B chunks XOR-swizzled by row).

        if (num_splits > 1) {
            // ================================================================= split-KV combine
            // 1. every thread stores its UNNORMALISED partial O slice (fp32, 32 of the row's 64 columns) and
            //    the row's (m, l) straight into the receive area of the cluster CTA that owns the row
            //    (st.shared::cluster, 16-B chunks XOR-swizzled by owned row).  Direct DSMEM stores replace a
            //    local staging pass + bulk copies + a transaction barrier (7.8 -> 7.55 us per B=1 launch).
            const int rows_per = kBlockM / num_splits;
            const uint32_t recv_o = smem + kOff + kSmemCombO;
            const int owner = row / rows_per, lr_mine = row - owner * rows_per;
            const int slot_mine = split * rows_per + lr_mine;
            cluster_wait();                                // barrier #1: every peer CTA is running
            if (threadIdx.x == 0) LG_TRACE(11);
            if constexpr (kOutF32) {
                const uint32_t dst_o = dsmem_addr(recv_o + slot_mine * 256, owner);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    st_dsmem_v4(dst_o + (((half * 8 + c) ^ (lr_mine & 15)) << 4), orow[4 * c + 0], orow[4 * c + 1],
                                orow[4 * c + 2], orow[4 * c + 3]);
            } else {
                // FP16 output: the partial travels as FP16 of O_s / l_s (same magnitude as the result, so the
                // extra rounding is of the order of the output rounding itself) — half the DSMEM bytes, which
                // cross at only ~20 B/clk per SM
                const float inv_own = 1.0f / l_row;
                const uint32_t dst_o = dsmem_addr(recv_o + slot_mine * 128, owner);
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    st_dsmem_v4(dst_o + (((half * 4 + c) ^ (lr_mine & 7)) << 4),
                                pack_half2(__uint_as_float(orow[8 * c + 0]) * inv_own, __uint_as_float(orow[8 * c + 1]) * inv_own),
                                pack_half2(__uint_as_float(orow[8 * c + 2]) * inv_own, __uint_as_float(orow[8 * c + 3]) * inv_own),
                                pack_half2(__uint_as_float(orow[8 * c + 4]) * inv_own, __uint_as_float(orow[8 * c + 5]) * inv_own),
                                pack_half2(__uint_as_float(orow[8 * c + 6]) * inv_own, __uint_as_float(orow[8 * c + 7]) * inv_own));
            }
            if (half == 0) st_dsmem_v2(dsmem_addr(smem + kOff + kSmemCombStat + slot_mine * 8, owner), m_used, l_row);
            if (threadIdx.x == 0) LG_TRACE(6);
            // 2. cluster barrier #2 (release / acquire): every split's slices for the rows this CTA owns have
            //    landed.  After it nobody touches a peer's shared memory, so CTAs may exit independently.
#if LG_ATTN_COMB_MBAR == 2
            // one arriving lane per group of lanes that share an owner; the other lanes' stores are ordered before it by
            // their own cluster-scope fence + the warp barrier
            asm volatile("fence.acq_rel.cluster;" ::: "memory");
            __syncwarp();
            if ((lane & ((rows_per < 32 ? rows_per : 32) - 1)) == 0) mbar_arrive_remote_release(dsmem_addr(bar(kBarComb), owner));
            mbar_wait_acquire_cluster(bar(kBarComb), 0);
#elif LG_ATTN_COMB_MBAR
            mbar_arrive_remote_release(dsmem_addr(bar(kBarComb), owner));
            mbar_wait_acquire_cluster(bar(kBarComb), 0);
#else
            cluster_arrive();
            cluster_wait();
#endif
            if (threadIdx.x == 0) LG_TRACE(7);
            for (int task = threadIdx.x; task < rows_per * 8; task += kSoftmaxWarps * 32) {
                const int lr = task >> 3;  // owned row
                const int cp = task & 7;   // 8-column group of the 64 output columns
                const int grow = tile_m * kBlockM + split * rows_per + lr;
                if (grow >= n_q) continue;
                float ms[kMaxSplits], ls[kMaxSplits];
                float4 v0[kMaxSplits], v1[kMaxSplits];
                float m_glob = -INFINITY;
#pragma unroll
                for (int sp = 0; sp < kMaxSplits; ++sp) {
                    if (sp < num_splits) {
                        const int slot = sp * rows_per + lr;
                        const float2 st = ld_shared_v2f(smem + kOff + kSmemCombStat + slot * 8);
                        ms[sp] = st.x;
                        ls[sp] = st.y;
                        m_glob = fmaxf(m_glob, st.x);
                        if constexpr (kOutF32) {
                            const uint32_t src = recv_o + slot * 256;
                            v0[sp] = ld_shared_v4f(src + (((2 * cp) ^ (lr & 15)) << 4));
                            v1[sp] = ld_shared_v4f(src + (((2 * cp + 1) ^ (lr & 15)) << 4));
                        } else {
                            const uint4 hv = ld_shared_v4(recv_o + slot * 128 + ((cp ^ (lr & 7)) << 4));
                            const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&hv.x));
                            const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&hv.y));
                            const float2 f2 = __half22float2(*reinterpret_cast<const __half2*>(&hv.z));
                            const float2 f3 = __half22float2(*reinterpret_cast<const __half2*>(&hv.w));
                            v0[sp] = make_float4(f0.x, f0.y, f1.x, f1.y);
                            v1[sp] = make_float4(f2.x, f2.y, f3.x, f3.y);
                        }
                    }
                }
                float l_tot = 0.f;
                float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int sp = 0; sp < kMaxSplits; ++sp) {
                    if (sp < num_splits) {
                        float a = ex2_approx((ms[sp] - m_glob) * kScaleLog2);
                        l_tot = fmaf(ls[sp], a, l_tot);
                        if constexpr (!kOutF32) a *= ls[sp];   // FP16 partials arrive normalised by their own l
                        acc[0] = fmaf(v0[sp].x, a, acc[0]); acc[1] = fmaf(v0[sp].y, a, acc[1]);
                        acc[2] = fmaf(v0[sp].z, a, acc[2]); acc[3] = fmaf(v0[sp].w, a, acc[3]);
                        acc[4] = fmaf(v1[sp].x, a, acc[4]); acc[5] = fmaf(v1[sp].y, a, acc[5]);
                        acc[6] = fmaf(v1[sp].z, a, acc[6]); acc[7] = fmaf(v1[sp].w, a, acc[7]);
                    }
                }
                const float inv = 1.0f / l_tot;
                const long long off = bi * ol.sb + hi * ol.sh + grow * ol.sr + cp * 8;
                if constexpr (!kOutF32) {
                    uint4 v;
                    v.x = pack_half2(acc[0] * inv, acc[1] * inv);
                    v.y = pack_half2(acc[2] * inv, acc[3] * inv);
                    v.z = pack_half2(acc[4] * inv, acc[5] * inv);
                    v.w = pack_half2(acc[6] * inv, acc[7] * inv);
                    st_global_v4(reinterpret_cast<__half*>(out) + off, v);
                } else {
                    uint4 w0, w1;
                    w0.x = __float_as_uint(acc[0] * inv); w0.y = __float_as_uint(acc[1] * inv);
                    w0.z = __float_as_uint(acc[2] * inv); w0.w = __float_as_uint(acc[3] * inv);
                    w1.x = __float_as_uint(acc[4] * inv); w1.y = __float_as_uint(acc[5] * inv);
                    w1.z = __float_as_uint(acc[6] * inv); w1.w = __float_as_uint(acc[7] * inv);
                    st_global_v4(reinterpret_cast<float*>(out) + off, w0);
                    st_global_v4(reinterpret_cast<float*>(out) + off + 4, w1);
                }
            }
        } else if constexpr (!kOutF32) {
            const float inv_l = 1.0f / l_row;
            const uint32_t stage = smem + kSmemStage + warp * (32 * 64);  // 32 rows x 64 B per warp
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    w[e] = pack_half2(__uint_as_float(orow[8 * c + 2 * e]) * inv_l,
                                      __uint_as_float(orow[8 * c + 2 * e + 1]) * inv_l);
                st_shared_v4(stage + lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4), w[0], w[1], w[2], w[3]);
            }
            __syncwarp();
            __half* o = reinterpret_cast<__half*>(out) + (bi * ol.sb + hi * ol.sh) + half * 32;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = i * 8 + (lane >> 2);
                const int c = lane & 3;
                const uint4 v = ld_shared_v4(stage + r * 64 + ((c ^ ((r >> 1) & 3)) << 4));
                if (row0 + r < n_q) st_global_v4(o + (row0 + r) * ol.sr + c * 8, v);
            }
        } else {
            const float inv_l = 1.0f / l_row;
            const uint32_t stage = smem + kSmemStage + warp * (32 * 128);  // 32 rows x 128 B per warp
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                st_shared_v4(stage + lane * 128 + ((c ^ (lane & 7)) << 4),
                             __float_as_uint(__uint_as_float(orow[4 * c + 0]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 1]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 2]) * inv_l),
                             __float_as_uint(__uint_as_float(orow[4 * c + 3]) * inv_l));
            }
            __syncwarp();
            float* o = reinterpret_cast<float*>(out) + (bi * ol.sb + hi * ol.sh) + half * 32;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + (lane >> 3);
                const int c = lane & 7;
                const uint4 v = ld_shared_v4(stage + r * 128 + ((c ^ (r & 7)) << 4));
                if (row0 + r < n_q) st_global_v4(o + (row0 + r) * ol.sr + c * 4, v);
            }
        }
    }

    if (threadIdx.x == 0) LG_TRACE(8);
    tc_fence_before();
    __syncthreads();
    if (warp == kSoftmaxWarps + 2) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
    if (threadIdx.x == 0) LG_TRACE(9);
}

}  // namespace w8
}  // namespace lgattn
