// B200-native (sm_100a) fused multi-head attention for LightGlue, head_dim 64 — MULTI-Q-TILE persistent throughput kernel.
//
//   O = softmax(Q K^T / 8) V      per (batch, head); Q [B,H,N,64], K/V [B,H,M,64], O [B,H,N,64]
//
// Replaces the hot loop of the reference kernels
// (/root/reference/lightglue_attention_plugin/attention_headdim_64_fp16in_fp32out.cu:400-665 and the FP16-out twin).
//
// One CTA per SM owns ALL 512 TMEM columns and kT 128-row Q tiles of one (batch, head) at a time ("work item"); the kT tiles
// share every K/V tile that TMA brings in (one L2->smem transfer per kT*128 Q rows instead of one per 128), two MMA-issuing
// warps feed the in-order tensor pipe (two tiles each, fixed order), and kT softmax warpgroups (one per Q tile, one S row per
// thread) overlap one another's TMEM loads / stores, barrier round trips and MMA latency.  Two layouts are instantiated:
//
//   <kT=4, kBN=64,  alias=1>   S[64] (P aliased onto its first 32 columns) | O[64] per tile (128 columns): four softmax warps
//                              per scheduler; per tile the chain QK -> softmax -> PV -> QK is serial and the other three
//                              tiles fill the pipes meanwhile.  With P aliased the tensor pipe's issue order (PV(j) before
//                              QK(j+1), same thread) is what keeps QK(j+1) from overwriting P(j) early, and s_full(j) also
//                              tells the softmax group that PV(j-1) is complete (O final, rescale legal) — one barrier wait
//                              per step.  The default full-grid kernel.
//   <kT=2, kBN=128, alias=0>   S[128] | P[64] | O[64] per tile (256 columns): QK(j+1) is issued as soon as the softmax group has
//                              pulled S(j) into registers, so S(j+1) is ready before the group asks for it.  Used when the
//                              Q-tile count packs better into pairs.
//
// kSplit = 2 (tuning builds only; measured slower): kSplit softmax warps share a 32-row slab and take
// kBN / kSplit columns of every step each (16 softmax warps on the <2,128> layout).
//
// kLazyMax (alias layouts): the row max is only computed on the first KV step of a Q tile and when the step's row sum says
// that some probability reached 2^7 (the stale reference is then too small); S is still intact in TMEM at that point and
// the step is redone with the true max.  That removes the 0.5 FMNMX3 per score of the common step.
//
// Warps: [0, 4kT) softmax groups | 4kT: K/V TMA producer | 4kT+1: Q TMA producer + TMEM allocator | 4kT+2.. : MMA issuers.
// Service warps run converged and issue through elect.sync (see the comment at their entry).
// CTAs are persistent: grid = min(#SM, items), items walked blockIdx.x + i * gridDim.x; barrier phases are carried by
// running step / item counters; K/V stream across item boundaries; Q(i+1) is prefetched when the last QK of item i is done.
// Items have a variable number of active tiles (struct Sched): idle tiles only exchange barrier arrivals.
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

#ifndef LG_ATTN_SPIN_NS
#define LG_ATTN_SPIN_NS 64
#endif
#ifndef LG_ATTN_MQ_STAGES64
#define LG_ATTN_MQ_STAGES64 4   // K/V ring depth for 64-row tiles (8 KiB each)
#endif
#ifndef LG_ATTN_MQ_STAGES128
#define LG_ATTN_MQ_STAGES128 3  // K/V ring depth for 128-row tiles (16 KiB each)
#endif
#ifndef LG_ATTN_MQ_MMA_SPIN_NS
#define LG_ATTN_MQ_MMA_SPIN_NS 0   // poll interval of the MMA issuer (0: tight try_wait loop; its waits are on every tile's critical chain)
#endif
#ifndef LG_ATTN_MQ_FIRST_READY
#define LG_ATTN_MQ_FIRST_READY 0   // 1: MMA warps with two tiles serve the tile whose P is ready first (measured slower: 869 vs 899 TFLOP/s)
#endif
#ifndef LG_ATTN_MQ_STAGGER_NS
#define LG_ATTN_MQ_STAGGER_NS 0
#endif
#ifndef LG_ATTN_MQ_STRICT_ORDER
#define LG_ATTN_MQ_STRICT_ORDER 0  // 1: alias layouts wait for pv_done(j) before issuing QK(j+1) (no reliance on pipe order)
#endif

namespace lgattn {
namespace mq {

using namespace ptx;

// Optional phase tracing (tuning builds only, -DLG_ATTN_MQ_TRACE): clock64 stamps of warp 0 / lane 0 of every softmax group and of
// the MMA issuer, steps [0, kTraceSteps) of the first item of the first CTAs; read back through lg_attn_debug_read_mq_trace().
//   softmax events: 0 before s_full wait, 1 s_full seen, 2 S in registers, 3 exponentials done, 4 p_full arrived
//   MMA events:     5 p_full seen, 6 PV (+ next QK) issued and committed
#ifdef LG_ATTN_MQ_TRACE
constexpr int kTraceCtas = 8, kTraceSteps = 64, kTraceTiles = 4, kTraceEvents = 8;
__device__ long long g_mq_trace[kTraceCtas][kTraceSteps][kTraceTiles][kTraceEvents];
#define MQ_TRACE(cond, step, tile_, ev)                                                                  \
    do {                                                                                                \
        if ((cond) && blockIdx.x < kTraceCtas && (step) < kTraceSteps) g_mq_trace[blockIdx.x][(step)][(tile_)][(ev)] = clock64(); \
    } while (0)
#else
#define MQ_TRACE(cond, step, tile_, ev)
#endif

#ifndef LG_ATTN_MQ_MMA_WARPS
#define LG_ATTN_MQ_MMA_WARPS 2   // MMA-issuing warps; warp m serves tiles [m*kT/n, (m+1)*kT/n) in a fixed order (clamped to kT)
#endif

template <int kT, int kBN, bool kAlias, int kSplit = 1>
struct Cfg {
    static_assert(kBN == 64 || kBN == 128, "KV step");
    static_assert(kSplit == 1 || (kSplit == 2 && !kAlias), "column halves: S and P separate");
    static constexpr int kWarpsPerTile = 4 * kSplit;   // kSplit warps share a 32-row slab, each takes kBN / kSplit columns of it
    static constexpr int kSoftmaxWarps = kWarpsPerTile * kT;
    static constexpr int kMmaWarps = LG_ATTN_MQ_MMA_WARPS > kT ? kT : (kT % LG_ATTN_MQ_MMA_WARPS == 0 ? LG_ATTN_MQ_MMA_WARPS : 1);
    static constexpr int kServiceWarps = (2 + kMmaWarps + 3) / 4 * 4;   // K/V producer, Q producer (+ TMEM alloc), MMA issuers; setmaxnreg wants whole warpgroups
    static constexpr int kThreads = (kSoftmaxWarps + kServiceWarps) * 32;
    static constexpr int kStages = kBN == 64 ? LG_ATTN_MQ_STAGES64 : LG_ATTN_MQ_STAGES128;
    static constexpr uint32_t kQBytes = kBlockM * kHeadDim * 2;   // 16 KiB
    static constexpr uint32_t kKVBytes = kBN * kHeadDim * 2;      // 8 / 16 KiB
    static constexpr uint32_t kSmemQ = 0;
    static constexpr uint32_t kSmemK = kSmemQ + kT * kQBytes;
    static constexpr uint32_t kSmemV = kSmemK + kStages * kKVBytes;
    // FP16 output: one 128-row x 128-B staging tile per Q tile for the epilogue's TMA store (same 128-byte swizzle as the loads)
    static constexpr uint32_t kSmemO = kSmemV + kStages * kKVBytes;
    static constexpr uint32_t kSmemBar = kSmemO + kT * kQBytes;
    // barriers (8 bytes each): per tile 7, ring 4 * kStages
    static constexpr uint32_t kBarQFull = 0, kBarQEmpty = kT, kBarSFull = 2 * kT, kBarSEmpty = 3 * kT, kBarPFull = 4 * kT,
                              kBarPVDone = 5 * kT, kBarOEmpty = 6 * kT, kBarKFull = 7 * kT, kBarKEmpty = kBarKFull + kStages,
                              kBarVFull = kBarKEmpty + kStages, kBarVEmpty = kBarVFull + kStages,
                              kNumBars = kBarVEmpty + kStages;
    static constexpr uint32_t kSmemTmemSlot = kSmemBar + 8 * kNumBars;
    // kSplit == 2: row max / row sum exchange between the two column halves of a row, [tile][parity][half][row] floats
    static constexpr uint32_t kSmemExch = kSmemTmemSlot + 16;
    static constexpr uint32_t kSmemUsed = kSmemExch + (kSplit > 1 ? kT * 2 * kSplit * kBlockM * 4 : 0);
    // >= 114 KiB so that a second CTA can never become resident next to this one (it would spin in tcgen05.alloc)
    static constexpr uint32_t kSmemBytes = (kSmemUsed + 1024 > 120 * 1024) ? kSmemUsed + 1024 : 120 * 1024;
    static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
    // TMEM columns of one tile
    static constexpr uint32_t kColS = 0;
    static constexpr uint32_t kColP = kAlias ? 0 : kBN;
    static constexpr uint32_t kColO = kAlias ? kBN : kBN + kBN / 2;
    static constexpr uint32_t kTileCols = kColO + kHeadDim;
    static_assert(kT * kTileCols <= 512, "TMEM budget");
    // register split: pool = kThreads * launch registers (launch_bounds(kThreads, 1)), service warps shrink to 24
    // pool = kThreads * floor8(65536 / kThreads); service warps keep kRegsService, the softmax warps get the rest
    static constexpr int kRegsLaunch = (65536 / kThreads) / 8 * 8;
    static constexpr int kRegsService = 32;
    static constexpr int kRegsSoftmaxMax = ((kThreads * kRegsLaunch - kServiceWarps * 32 * kRegsService) / (kSoftmaxWarps * 32)) / 8 * 8;
    static constexpr int kRegsSoftmax = kRegsSoftmaxMax > 232 ? 232 : kRegsSoftmaxMax;
};

// Work items.  A full item = kT consecutive Q tiles of one (batch, head) ("group").  The persistent grid walks the items
// blockIdx.x + i * gridDim.x, so when the number of groups is not a multiple of the grid the last wave would leave most SMs idle
// for a whole item time (B16 N4096: 512 groups on 148 SMs = 3.46 waves).  The host therefore cuts the groups of that last partial
// wave into `parts` smaller items (kT / parts tiles each): more CTAs busy, each for a shorter time.  Tiles of an item past its
// n_active (split items, and groups that stick out past the last Q tile) take part in every barrier handshake but issue no
// loads, MMAs, arithmetic or stores.
struct Sched {
    int groups;       // groups per (batch, head) = ceil(tiles_m / kT)
    int tiles_m;      // Q tiles per (batch, head)
    int n_full;       // items [0, n_full) are whole groups
    int parts;        // items >= n_full: group n_full + (w - n_full) / parts, part (w - n_full) % parts
    int total_items;
};

struct Item {
    int tile_m0, n_active, hi, bi, bk;
};

template <int kT>
__device__ __forceinline__ Item item_coord(int w, const Sched& sc, const OutLayout& ol) {
    Item c;
    int grp = w, off = 0, cnt = kT;
    if (w >= sc.n_full) {
        const int idx = w - sc.n_full;
        grp = sc.n_full + idx / sc.parts;
        cnt = kT / sc.parts;
        off = (idx % sc.parts) * cnt;
    }
    c.tile_m0 = (grp % sc.groups) * kT + off;
    const int left = sc.tiles_m - c.tile_m0;
    c.n_active = left < cnt ? (left < 0 ? 0 : left) : cnt;
    const int bh = grp / sc.groups;
    c.hi = bh % ol.heads;
    c.bi = bh / ol.heads;
    c.bk = c.bi + ol.kv_shift >= ol.batch ? c.bi + ol.kv_shift - ol.batch : c.bi + ol.kv_shift;
    return c;
}

template <int kT, int kBN, bool kAlias, bool kLazyMax, bool kOutF32, int kSplit = 1>
__global__ void __launch_bounds__(Cfg<kT, kBN, kAlias, kSplit>::kThreads, 1)
lg_attn_hd64_fwd(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_o, void* __restrict__ out, int n_q,
                 int n_kv, const Sched sc, const OutLayout ol, int tma_store) {
    using C = Cfg<kT, kBN, kAlias, kSplit>;
    static_assert(!kLazyMax || kAlias, "the lazy row max re-reads S from TMEM: S must survive until P is stored");
    extern __shared__ uint8_t smem_raw[];
    uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;
    asm volatile("mov.u32 %0, %0;" : "+r"(smem));   // pin: ptxas otherwise re-derives the aligned base at every barrier op
    const uint32_t bars = smem + C::kSmemBar;
    auto bar = [&](uint32_t i) { return bars + 8u * i; };

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int nkv = (n_kv + kBN - 1) / kBN;
    const int first = blockIdx.x, stride = gridDim.x;
    const int total_items = sc.total_items;
    constexpr int kWarpKV = C::kSoftmaxWarps, kWarpQ = kWarpKV + 1, kWarpAlloc = kWarpQ, kWarpMMA = kWarpKV + 2;

    if (warp == kWarpKV && lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
#pragma unroll
        for (int t = 0; t < kT; ++t) {
            mbar_init(bar(C::kBarQFull + t), 1);
            mbar_init(bar(C::kBarQEmpty + t), 1);
            mbar_init(bar(C::kBarSFull + t), 1);
            mbar_init(bar(C::kBarSEmpty + t), C::kWarpsPerTile);   // one arrival per softmax warp of the tile
            mbar_init(bar(C::kBarPFull + t), C::kWarpsPerTile);
            mbar_init(bar(C::kBarPVDone + t), 1);
            mbar_init(bar(C::kBarOEmpty + t), C::kWarpsPerTile);
        }
#pragma unroll
        for (int s = 0; s < C::kStages; ++s) {
            mbar_init(bar(C::kBarKFull + s), 1);
            mbar_init(bar(C::kBarKEmpty + s), C::kMmaWarps);   // one commit per MMA warp, after its last tile
            mbar_init(bar(C::kBarVFull + s), 1);
            mbar_init(bar(C::kBarVEmpty + s), C::kMmaWarps);
        }
        fence_mbar_init();
    }
    if (warp == kWarpAlloc) {
        tmem_alloc<512>(smem + C::kSmemTmemSlot);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem + C::kSmemTmemSlot));

    if (warp >= C::kSoftmaxWarps) {
        setmaxnreg_dec<C::kRegsService>();
        // Service warps run CONVERGED: all 32 lanes execute the loops and the barrier waits, one elected lane (elect.sync) issues
        // the TMA / tcgen05 instructions.  Every operand of those instructions lives in uniform registers; in a `lane == 0`
        // branch ptxas cannot prove uniformity and wraps each UTCHMMA / UTMALDG / UTCBAR in an R2UR + ELECT + BRA.U.ANY waterfall
        // loop (~80 clk per MMA, measured with the phase trace: 650 clk to issue 8 MMAs).  For the same reason the base addresses
        // are re-derived here from provably uniform sources instead of the pinned / asm-loaded copies the softmax warps use.
        const uint32_t smem_u = (smem_u32(smem_raw) + 1023u) & ~1023u;
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t bars_u = smem_u + C::kSmemBar;
        auto ubar = [&](uint32_t i) { return bars_u + 8u * i; };
        if (warp == kWarpQ) {
            // ===================================================================== Q producer
            int it = 0;
            for (int w = first; w < total_items; w += stride, ++it) {
                const Item c = item_coord<kT>(w, sc, ol);
#pragma unroll 1
                for (int t = 0; t < kT; ++t) {
                    if (it > 0) mbar_wait_backoff(ubar(C::kBarQEmpty + t), (it - 1) & 1, LG_ATTN_SPIN_NS);  // last QK of the previous item done
                    if (elect_one()) {
                        if (t < c.n_active) {
                            mbar_arrive_expect_tx(ubar(C::kBarQFull + t), C::kQBytes);
                            tma_load_4d(smem_u + C::kSmemQ + t * C::kQBytes, &tm_q, ubar(C::kBarQFull + t), 0, (c.tile_m0 + t) * kBlockM, c.hi, c.bi);
                        } else {
                            mbar_arrive(ubar(C::kBarQFull + t));   // idle tile of this item: nothing to load, the phase still turns
                        }
                    }
                    __syncwarp();
                }
            }
        } else if (warp == kWarpKV) {
            // ===================================================================== K/V producer (K runs one step ahead of V)
            int g = 0;   // running KV-step counter: ring slots and phases continue across items
            auto load_k = [&](int gg, int j, const Item& c) {
                const int s = gg % C::kStages;
                if (gg >= C::kStages) mbar_wait_backoff(ubar(C::kBarKEmpty + s), ((gg / C::kStages) - 1) & 1, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(C::kBarKFull + s), C::kKVBytes);
                    tma_load_4d(smem_u + C::kSmemK + s * C::kKVBytes, &tm_k, ubar(C::kBarKFull + s), 0, j * kBN, c.hi, c.bk);
                }
                __syncwarp();
            };
            auto load_v = [&](int gg, int j, const Item& c) {
                const int s = gg % C::kStages;
                if (gg >= C::kStages) mbar_wait_backoff(ubar(C::kBarVEmpty + s), ((gg / C::kStages) - 1) & 1, LG_ATTN_SPIN_NS);
                if (elect_one()) {
                    mbar_arrive_expect_tx(ubar(C::kBarVFull + s), C::kKVBytes);
                    tma_load_4d(smem_u + C::kSmemV + s * C::kKVBytes, &tm_v, ubar(C::kBarVFull + s), 0, j * kBN, c.hi, c.bk);
                }
                __syncwarp();
            };
            for (int w = first; w < total_items; w += stride) {
                const Item c = item_coord<kT>(w, sc, ol);
                for (int j = 0; j < nkv; ++j, ++g) {
                    if (j == 0) load_k(g, 0, c);
                    if (j + 1 < nkv) load_k(g + 1, j + 1, c);
                    load_v(g, j, c);
                }
            }
        } else if (warp >= kWarpMMA && warp < kWarpMMA + C::kMmaWarps) {
            // ===================================================================== MMA issuers (each: its tiles in a fixed round-robin order)
            constexpr int kTilesPerWarp = kT / C::kMmaWarps;
            const int t_begin = (warp - kWarpMMA) * kTilesPerWarp, t_end = t_begin + kTilesPerWarp;
            auto mma_wait = [&](uint32_t b, uint32_t parity) {
                if (LG_ATTN_MQ_MMA_SPIN_NS > 0) mbar_wait_backoff(b, parity, LG_ATTN_MQ_MMA_SPIN_NS);
                else mbar_wait(b, parity);
            };
            const int n_work = first < total_items ? (total_items - first + stride - 1) / stride : 0;
#if LG_ATTN_MQ_STAGGER_NS > 0
            // offset the issue schedules of the MMA warps once: the tiles of one warp stay a service time apart by construction, and
            // without this the two warps' schedules coincide (phase trace: tiles 1|2 and 3|0 in lockstep)
            if (warp > kWarpMMA) __nanosleep(LG_ATTN_MQ_STAGGER_NS * (warp - kWarpMMA));
#endif

            bool skip_qk = false;
            int n_active = kT;   // of the item being issued (tiles >= n_active only exchange barrier arrivals)
            // (elected lane) S_t = Q_t K(g)^T.  The caller has made sure K(g) has landed and S_t may be overwritten.
            auto issue_qk = [&](int t, int g) {
                if (skip_qk || t >= n_active) return;
                umma_qk<kBN>(tmem_u + t * C::kTileCols + C::kColS, smem_u + C::kSmemQ + t * C::kQBytes,
                             smem_u + C::kSmemK + (g % C::kStages) * C::kKVBytes);
            };
            // (elected lane) O_t (+)= P_t V(g)
            auto issue_pv = [&](int t, int g, bool accumulate) {
                if (t >= n_active) return;
                umma_pv<kBN>(tmem_u + t * C::kTileCols + C::kColO, tmem_u + t * C::kTileCols + C::kColP,
                             smem_u + C::kSmemV + (g % C::kStages) * C::kKVBytes, C::kKVBytes, accumulate);
            };
            auto wait_k = [&](int g) {
                mma_wait(ubar(C::kBarKFull + g % C::kStages), (g / C::kStages) & 1);
                tc_fence_after();
            };
            // QK of step g for every tile (item prologue, and the non-alias steady state)
            auto qk_all = [&](int g, int it, bool first_of_item, bool last_of_item) {
                wait_k(g);
#pragma unroll 1
                for (int t = t_begin; t < t_end; ++t) {
                    if (first_of_item) mma_wait(ubar(C::kBarQFull + t), it & 1);
                    if (!kAlias && g > 0) mma_wait(ubar(C::kBarSEmpty + t), (g - 1) & 1);  // S(g-1) is in registers
#if LG_ATTN_MQ_STRICT_ORDER
                    if (kAlias && g > 0) mma_wait(ubar(C::kBarPVDone + t), (g - 1) & 1);
#endif
                    tc_fence_after();
                    if (elect_one()) {
                        issue_qk(t, g);
                        if (last_of_item) mma_commit(ubar(C::kBarQEmpty + t));
                        mma_commit(ubar(C::kBarSFull + t));
                        if (t == t_end - 1) mma_commit(ubar(C::kBarKEmpty + g % C::kStages));
                    }
                    __syncwarp();
                }
            };

            int g = 0;
            for (int it = 0; it < n_work; ++it) {
                n_active = item_coord<kT>(first + it * stride, sc, ol).n_active;
#ifdef LG_ATTN_WHATIF_F3
                // what-if build (WRONG results, timing only; tools/f3_whatif.py): row f3 of SURVEY section 8 — with both cross-attention
                // directions in one launch (kv_shift), the second direction could reuse S(0->1)^T instead of computing its own
                // Q K^T.  This drops the second direction's QK MMAs (commits stay) to bound what that could gain.
                skip_qk = ol.kv_shift > 0 && item_coord<kT>(first + it * stride, sc, ol).bi >= ol.batch / 2;
#endif
                qk_all(g, it, true, nkv == 1);
                for (int j = 0; j < nkv; ++j, ++g) {
                    const bool more = j + 1 < nkv;
                    if (!kAlias && more) qk_all(g + 1, it, false, j + 2 == nkv);
                    mma_wait(ubar(C::kBarVFull + g % C::kStages), (g / C::kStages) & 1);
                    if (kAlias && more) wait_k(g + 1);
                    // Service order inside a step: with two tiles per MMA warp, whichever tile's P arrives first is served first
                    // (a fixed order makes the earlier tile queue behind the later one: ~230 clk per step in the phase trace)
                    int swap = 0;
                    if (LG_ATTN_MQ_FIRST_READY && kTilesPerWarp == 2) {
                        const uint32_t par = static_cast<uint32_t>(g & 1);
                        while (true) {
                            if (mbar_test_wait(ubar(C::kBarPFull + t_begin), par)) break;
                            if (mbar_test_wait(ubar(C::kBarPFull + t_begin + 1), par)) { swap = 1; break; }
                        }
                        swap = __shfl_sync(0xffffffffu, swap, 0);   // one decision for the warp (probes of different lanes may differ)
                    }
#pragma unroll 1
                    for (int ti = 0; ti < kTilesPerWarp; ++ti) {
                        const int t = t_begin + (swap ? kTilesPerWarp - 1 - ti : ti);
                        const bool last_tile = ti == kTilesPerWarp - 1;
                        mma_wait(ubar(C::kBarPFull + t), g & 1);
                        MQ_TRACE(lane == 0, g, t, 5);
                        if (j == 0 && it > 0) mma_wait(ubar(C::kBarOEmpty + t), (it - 1) & 1);  // epilogue has read O
                        tc_fence_after();
#if LG_ATTN_MQ_STRICT_ORDER
                        if (elect_one()) {
                            issue_pv(t, g, j > 0);
                            mma_commit(ubar(C::kBarPVDone + t));
                            if (last_tile) mma_commit(ubar(C::kBarVEmpty + g % C::kStages));
                        }
                        __syncwarp();
                        if (kAlias && more) {
                            mma_wait(ubar(C::kBarPVDone + t), g & 1);
                            tc_fence_after();
                            if (elect_one()) {
                                issue_qk(t, g + 1);
                                if (j + 2 == nkv) mma_commit(ubar(C::kBarQEmpty + t));
                                mma_commit(ubar(C::kBarSFull + t));
                                if (last_tile) mma_commit(ubar(C::kBarKEmpty + (g + 1) % C::kStages));
                            }
                            __syncwarp();
                        }
#else
                        if (elect_one()) {
                            issue_pv(t, g, j > 0);
                            if (!kAlias || !more) mma_commit(ubar(C::kBarPVDone + t));   // alias: only the epilogue waits for it
                            if (last_tile) mma_commit(ubar(C::kBarVEmpty + g % C::kStages));
                            if (kAlias && more) {
                                issue_qk(t, g + 1);   // after PV_t(g) in the pipe: may overwrite P_t(g)
                                if (j + 2 == nkv) mma_commit(ubar(C::kBarQEmpty + t));
                                mma_commit(ubar(C::kBarSFull + t));
                                if (last_tile) mma_commit(ubar(C::kBarKEmpty + (g + 1) % C::kStages));
                            }
                        }
                        __syncwarp();
#endif
                        MQ_TRACE(lane == 0, g, t, 6);
                    }
                }
            }
        }
    } else {
        // ========================================================================= softmax warps (tile = warp / (4 * kSplit))
        // kSplit == 2: warps q and q + 4 of a tile own the same 32 rows (TMEM lanes 32 * (warp % 4) ..) and one half of the step's
        // columns each; they agree on the row max through shared memory + a 64-thread named barrier, and each keeps / rescales /
        // stores its half of the O columns.
        setmaxnreg_inc<C::kRegsSoftmax>();
        constexpr int kCols = kBN / kSplit, kOCols = kHeadDim / kSplit;
        const int tile = warp / C::kWarpsPerTile;
        const int half = (warp >> 2) % kSplit;
        const uint32_t lane_sel = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t t_s = tmem_base + lane_sel + tile * C::kTileCols + C::kColS + half * kCols;
        const uint32_t t_o = tmem_base + lane_sel + tile * C::kTileCols + C::kColO + half * kOCols;
        const uint32_t t_p = tmem_base + lane_sel + tile * C::kTileCols + C::kColP + half * (kCols / 2);
        const uint32_t exch = smem + C::kSmemExch + static_cast<uint32_t>(tile * 2 * kSplit * kBlockM + (warp & 3) * 32 + lane) * 4u;
        auto exch_addr = [&](int parity, int h) { return exch + static_cast<uint32_t>((parity * kSplit + h) * kBlockM) * 4u; };
        auto pair_sync = [&]() { named_bar_sync(1u + static_cast<uint32_t>(tile * 4 + (warp & 3)), 64u); };
        const uint32_t bar_s_full = bar(C::kBarSFull + tile), bar_s_empty = bar(C::kBarSEmpty + tile);
        const uint32_t bar_p_full = bar(C::kBarPFull + tile), bar_pv_done = bar(C::kBarPVDone + tile);
        const uint32_t bar_o_empty = bar(C::kBarOEmpty + tile);

        int g = 0, it = 0;
        for (int w = first; w < total_items; w += stride, ++it) {
            const Item c = item_coord<kT>(w, sc, ol);
            float m_used = -INFINITY;            // row max the exponent is referenced to (raw score units)
            float m_thr = INFINITY, neg_m = 0.f; // rescale threshold m_used + tau/c, and -m_used*c
            float l_sum = 0.f;
            int kv_left = n_kv;

            for (int j = 0; j < nkv; ++j, ++g) {
                MQ_TRACE((warp & 3) == 0 && lane == 0, g, tile, 0);
                if (kAlias || j == 0) {  // non-alias: s_full(g) of later steps was awaited under step g-1 (see below)
                    mbar_wait(bar_s_full, g & 1);
                    tc_fence_after();
                }
                MQ_TRACE((warp & 3) == 0 && lane == 0, g, tile, 1);
                if (tile >= c.n_active) {
                    // idle tile of this item: keep every barrier phase turning (the MMA warp commits without issuing), nothing else
                    __syncwarp();
                    if (!kAlias) {
                        if (lane == 0) mbar_arrive(bar_s_empty);
                        if (j + 1 < nkv) mbar_wait(bar_s_full, (g + 1) & 1);
                        else if (j > 0) mbar_wait(bar_pv_done, (g - 1) & 1);
                    }
                    if (lane == 0) mbar_arrive(bar_p_full);
                    continue;
                }
                const int valid = kv_left - half * kCols;
                kv_left -= kBN;

                uint32_t pk[kCols / 2];
                float tile_sum, alpha = 1.f;
                bool rescale = false;
                bool need_max = !kLazyMax || j == 0;
                while (true) {
                    uint32_t sr[kCols];
#pragma unroll
                    for (int cc = 0; cc < kCols / 32; ++cc) tmem_ld_x32(t_s + 32u * cc, sr + 32 * cc);
                    tmem_wait_ld();
                    MQ_TRACE((warp & 3) == 0 && lane == 0, g, tile, 2);
                    if (!kAlias) {   // QK(g+1) may overwrite S now
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_s_empty);
                    }
                    mask_kv_tail<kCols>(sr, valid);   // ragged KV tail: columns past n_kv get -inf
                    if (need_max) {
                        float mx = row_max<kCols>(sr);
                        if constexpr (kSplit > 1) {   // both halves continue with the row's max (identical rescale decisions)
                            st_shared_f32(exch_addr(g & 1, half), mx);
                            pair_sync();
                            mx = fmaxf(mx, ld_shared_f32(exch_addr(g & 1, half ^ 1)));
                        }
                        // lazy rescale decision (warp-uniform so the tcgen05.ld/st below stay aligned)
                        if (j == 0) {
                            m_used = mx;
                            m_thr = mx + kTau / kScaleLog2;
                            neg_m = -mx * kScaleLog2;
                        } else {
                            rescale = __any_sync(0xffffffffu, mx > m_thr) || kLazyMax;
                            if (rescale) {
                                const float m_new = fmaxf(m_used, mx);
                                alpha = ex2_approx((m_used - m_new) * kScaleLog2);
                                m_used = m_new;
                                m_thr = m_new + kTau / kScaleLog2;
                                neg_m = -m_new * kScaleLog2;
                            }
                        }
                    }
                    // P = exp2(s*c - m*c) as packed FP16 pairs + the FP32 row sum (softmax_math.cuh)
#if defined(LG_ATTN_WHATIF) && LG_ATTN_WHATIF >= 2
                    // what-if builds (WRONG results, timing decomposition only; LG_ATTN_TUNING_BUILD): 2 = no exponentials (one FFMA2 +
                    // pack per pair), 3 = no arithmetic at all ("skeleton": S load, P store, barriers, MMAs)
#pragma unroll
                    for (int cc = 0; cc < kCols; cc += 2) {
#if LG_ATTN_WHATIF == 2
                        float x0, x1;
                        unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(sr[cc]), __uint_as_float(sr[cc + 1])),
                                               pack_f32x2(kScaleLog2, kScaleLog2), pack_f32x2(neg_m, neg_m)), x0, x1);
                        pk[cc / 2] = pack_half2(x0, x1);
#else
                        pk[cc / 2] = sr[cc] ^ sr[cc + 1];
#endif
                    }
                    tile_sum = 1.0f;
#else
                    tile_sum = softmax_exp_pass<kCols>(sr, pk, neg_m);
#endif
                    if (kLazyMax && !need_max) {
                        // a probability >= 2^tau (the polynomial path saturates exactly there) shows in the row sum: the stale
                        // reference is too small for this step -> redo it with the true row max (S is intact: P not stored yet)
                        if (__any_sync(0xffffffffu, !(tile_sum < kLazyTrigger))) {
                            need_max = true;
                            continue;
                        }
                    }
                    break;
                }

                MQ_TRACE((warp & 3) == 0 && lane == 0, g, tile, 3);
                if (!kAlias) {
                    // one barrier round trip per step: QK(g+1) is issued after PV(g-1) by the same thread and tcgen05.commit covers
                    // all of that thread's earlier MMAs, so s_full(g+1) also means the P buffer is free and O(g-1) is final
                    if (j + 1 < nkv) {
                        mbar_wait(bar_s_full, (g + 1) & 1);
                        tc_fence_after();
                    } else if (j > 0) {
                        mbar_wait(bar_pv_done, (g - 1) & 1);
                        tc_fence_after();
                    }
                }
#pragma unroll
                for (int cc = 0; cc < kCols / 64; ++cc) tmem_st_x32(t_p + 32u * cc, pk + 32 * cc);
                if (rescale) {
                    rescale_o_tmem<kOCols>(t_o, alpha);
                    l_sum *= alpha;
                }
                l_sum += tile_sum;
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_p_full);
                MQ_TRACE((warp & 3) == 0 && lane == 0, g, tile, 4);
            }

            // --------------------------------------------------------------------------- epilogue
            mbar_wait(bar_pv_done, (kAlias && !LG_ATTN_MQ_STRICT_ORDER ? it : g - 1) & 1);
            tc_fence_after();
            if (tile >= c.n_active) {
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_o_empty);
                continue;
            }
            uint32_t orow[kOCols];
#pragma unroll
            for (int cc = 0; cc < kOCols / 32; ++cc) tmem_ld_x32(t_o + 32u * cc, orow + 32 * cc);
            tmem_wait_ld();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_o_empty);  // the next item's first PV may overwrite O now
            if constexpr (kSplit > 1) {   // row sum = both halves' (the parity slot last written two steps ago is free)
                st_shared_f32(exch_addr(g & 1, half), l_sum);
                pair_sync();
                l_sum += ld_shared_f32(exch_addr(g & 1, half ^ 1));
            }
            const float inv_l = 1.0f / l_sum;
            const int row = (c.tile_m0 + tile) * kBlockM + (warp & 3) * 32 + lane;
            if constexpr (!kOutF32 && kSplit == 1) {
                if (tma_store) {
                    // FP16 output through shared memory and ONE TMA store per Q tile.  A thread owns a whole output row, so direct
                    // stores are 16 B per lane at a 128-B stride: 32 LSU wavefronts per instruction, ~4 000 per item, with all 16
                    // softmax warps of the CTA in their epilogues at the same time (measured: 1.4 us per item, 12 % of a launch at
                    // N = M = 512).  The staging tile is written with the loads' 128-byte swizzle (4 wavefronts per instruction), the
                    // tile's four warps meet on a named barrier, one elected thread issues the store; rows past n_q are clipped
                    // by the tensor map.  The store of item i is waited for (source read) in front of the writes of item i + 1.
                    const bool issuer = (warp & 3) == 0;
                    if (issuer) {
                        if (elect_one()) bulk_wait_group_read0();
                        __syncwarp();
                    }
                    named_bar_sync(1u + static_cast<uint32_t>(tile), 128u);
                    const int r_in = (warp & 3) * 32 + lane;
                    const uint32_t stg = smem + C::kSmemO + static_cast<uint32_t>(tile) * C::kQBytes + static_cast<uint32_t>(r_in) * 128u;
#pragma unroll
                    for (int cc = 0; cc < 8; ++cc)
                        st_shared_v4(stg + (static_cast<uint32_t>(cc ^ (r_in & 7)) << 4),
                                     pack_half2(__uint_as_float(orow[8 * cc + 0]) * inv_l, __uint_as_float(orow[8 * cc + 1]) * inv_l),
                                     pack_half2(__uint_as_float(orow[8 * cc + 2]) * inv_l, __uint_as_float(orow[8 * cc + 3]) * inv_l),
                                     pack_half2(__uint_as_float(orow[8 * cc + 4]) * inv_l, __uint_as_float(orow[8 * cc + 5]) * inv_l),
                                     pack_half2(__uint_as_float(orow[8 * cc + 6]) * inv_l, __uint_as_float(orow[8 * cc + 7]) * inv_l));
                    fence_proxy_async_smem();
                    named_bar_sync(1u + static_cast<uint32_t>(tile), 128u);
                    if (issuer) {
                        if (elect_one()) {
                            tma_store_4d(&tm_o, smem + C::kSmemO + static_cast<uint32_t>(tile) * C::kQBytes, 0, (c.tile_m0 + tile) * kBlockM,
                                         c.hi, c.bi);
                            bulk_commit_group();
                        }
                        __syncwarp();
                    }
                    continue;
                }
            }
#ifdef LG_ATTN_WHATIF_NOSTORE   // what-if build (WRONG results, LG_ATTN_TUNING_BUILD): what the per-thread row stores of the epilogue cost
            if (row < n_q && inv_l == 123.f) {
#else
            if (row < n_q) {
#endif
                const long long off = c.bi * ol.sb + c.hi * ol.sh + row * ol.sr + half * kOCols;
                if constexpr (!kOutF32) store_o_row<false, kOCols>(reinterpret_cast<__half*>(out) + off, orow, inv_l);
                else store_o_row<true, kOCols>(reinterpret_cast<float*>(out) + off, orow, inv_l);
            }
        }
    }

    if constexpr (!kOutF32 && kSplit == 1) {
        // the staging tiles must outlive the stores that read them
        if (warp < C::kSoftmaxWarps && (warp & 3) == 0 && tma_store) {
            if (elect_one()) bulk_wait_group_read0();
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kWarpAlloc) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

}  // namespace mq
}  // namespace lgattn
