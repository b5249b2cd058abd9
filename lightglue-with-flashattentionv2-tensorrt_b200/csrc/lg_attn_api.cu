// C-ABI of the sm_100a attention path (declared in include/lg_attn.h).
//
// Host-side glue only: argument checks, per-device one-time setup, TMA tensor-map encoding
// (cuTensorMapEncodeTiled through cudaGetDriverEntryPoint — a pure host call, legal while a
// stream is being captured) and the kernel launch.  Mirrors what the reference's
// LightGlueAttentionPlugin::enqueue does around its launchers
// (/root/reference/lightglue_attention_plugin/lightglue_attention_plugin.cpp:165-267), minus the
// pad / unpad passes that the new kernel makes unnecessary.

#include "attn_sm100.h"
#include "attn_sm100_kernel.cuh"     // lgattn::w8 — latency shapes, split-KV clusters, PDL
#include "attn_sm100_kernel_w4.cuh"  // lgattn::w4 — full grids whose Q-tile count packs badly into 2- / 4-tile work items
#include "attn_sm100_kernel_mq.cuh"  // lgattn::mq — full grids: one persistent CTA per SM owning several Q tiles (throughput)

#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/lg_attn.h"

namespace {

using namespace lgattn;

constexpr int kMaxDevices = 64;
constexpr int kAbiVersion = 1;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

std::mutex g_init_mutex;
std::atomic<int> g_device_ready[kMaxDevices];  // 0 = not initialised, 1 = ready
std::atomic<EncodeTiledFn> g_encode{nullptr};
int g_sm_count[kMaxDevices];              // written once per device under g_init_mutex, before g_device_ready
int g_max_clusters[kMaxDevices][4];       // co-resident clusters of 1 << i CTAs (same publication rule)
int g_max_clusters_f32[kMaxDevices][4];   // same for the FP32-input instantiation (1 CTA per SM)
std::atomic<int> g_forced_splits{0};      // tuning override (LG_ATTN_SPLITS), 0 = automatic
std::atomic<int> g_mq_tma_store{1};       // LG_ATTN_MQ_TMA_STORE=0: the multi-Q-tile kernel's FP16 epilogue stores rows per thread (A/B runs)
std::atomic<int> g_use_pdl{1};            // LG_ATTN_PDL=0 disables programmatic dependent launch
std::atomic<int> g_full_grid_kernel{-1};  // full grids: -1 = automatic, 0 = w4, 1 = mq (4 Q tiles), 2 = mq (2 Q tiles)
                                          // (LG_ATTN_FULL_GRID_KERNEL, lg_attn_debug_set_throughput_kernel: tests, A/B runs)

// The multi-Q-tile kernel variants compiled into the library (lg_attn_debug_set_throughput_kernel modes 2, 3, ...).
template <int T, int BN, bool A, bool L, int S = 1>
struct MqVariant {
    static constexpr int kT = T, kBN = BN, kSplit = S;
    static constexpr bool kAlias = A, kLazy = L;
    using C = mq::Cfg<T, BN, A, S>;
};
constexpr int kMqModeBase = 1;
template <typename F>
void mq_for_each_variant(F&& f) {
    f(MqVariant<4, 64, true, true>{});      // mode 1: 4 Q tiles, 64-row KV steps, P aliased onto S, lazy row max
    f(MqVariant<2, 128, false, false>{});   // mode 2: 2 Q tiles, 128-row KV steps, S / P / O separate
#ifdef LG_ATTN_MQ_EXTRA_VARIANTS            // tuning builds: more layouts to A/B
    f(MqVariant<4, 64, true, false>{});     // mode 3
    f(MqVariant<3, 64, false, false>{});    // mode 4
    f(MqVariant<3, 64, true, true>{});      // mode 5
    // mode 6: mode 2's layout with two softmax warps per 32-row slab (64 of the step's 128 columns each, row max agreed through
    // shared memory every step).  No softmax warp ever waits for the tensor pipe in this layout (phase trace: wait 44 clk / step),
    // and it is still slower than mode 1 (879 vs 958 TFLOP/s at B64 N2048): the four softmax warps of a sub-partition already
    // saturate its issue / FMA / MUFU pipes at ~600 clk per 64-column row slice, so hiding the S round trip buys nothing and the
    // per-step row max (30 FMNMX3 + exchange) costs ~60 clk more.
    f(MqVariant<2, 128, false, false, 2>{});
#endif
}
#ifdef LG_ATTN_MQ_EXTRA_VARIANTS
constexpr int kMqVariants = 6;
#else
constexpr int kMqVariants = 2;
#endif

// How the multi-Q-tile kernel cuts a launch into work items (mq::Sched): whole groups of t Q tiles for every full wave of the
// persistent grid; the groups of a last partial wave that fills <= 1/2 (1/4) of the SMs are cut into 2 (4) smaller items.
mq::Sched mq_schedule(long long bh, long long tiles_m, int t, int sms) {
    mq::Sched sc;
    sc.groups = static_cast<int>((tiles_m + t - 1) / t);
    sc.tiles_m = static_cast<int>(tiles_m);
    const long long g = bh * sc.groups;
    const long long rem = g % sms;
    int parts = 1;
    if (g > sms && rem > 0) {
        if (4 * rem <= sms && t % 4 == 0) parts = 4;
        else if (2 * rem <= sms && t % 2 == 0) parts = 2;
    }
    sc.parts = parts;
    sc.n_full = static_cast<int>(parts > 1 ? g - rem : g);
    sc.total_items = static_cast<int>(parts > 1 ? g - rem + rem * parts : g);
    return sc;
}

// Relative time of an mq launch under that schedule, in units of one full 4-tile item of the 4-tile layout.  An item with fewer
// active tiles is cheaper, but not proportionally (fewer tiles hide less of each other's latency): about 0.85 / 0.65 / 0.50 of a
// full item for 3 / 2 / 1 of 4 tiles (B16 N4096: 818 -> 930 TFLOP/s with the last wave cut in two); the 2-tile layout's full
// item costs 2/4 / 0.90.  The grid walks items blockIdx.x + i * gridDim.x, so the launch takes as long as its busiest CTA: summed
// per CTA here exactly as the device decodes the items (a few thousand additions; beyond 64 items per CTA the average is exact
// enough).
double mq_time(long long bh, long long tiles_m, int t, int sms) {
    static const double kCost4[5] = {0.10, 0.50, 0.65, 0.85, 1.00};   // by active tiles, 4-tile layout
    static const double kCost2[3] = {0.06, 0.36, 0.556};              // 2-tile layout, same unit
    const mq::Sched sc = mq_schedule(bh, tiles_m, t, sms);
    auto item_cost = [&](long long w) {
        long long grp = w;
        int off = 0, cnt = t;
        if (w >= sc.n_full) {
            const long long idx = w - sc.n_full;
            grp = sc.n_full + idx / sc.parts;
            cnt = t / sc.parts;
            off = static_cast<int>(idx % sc.parts) * cnt;
        }
        const long long left = tiles_m - ((grp % sc.groups) * t + off);
        const int act = static_cast<int>(left < cnt ? (left < 0 ? 0 : left) : cnt);
        return t == 4 ? kCost4[act] : kCost2[act];
    };
    const long long items = sc.total_items;
    if (items > 64LL * sms) {   // long launches: per-(batch, head) average times items per CTA
        double per_bh = 0.0;
        for (int g = 0; g < sc.groups; ++g) per_bh += item_cost(g);
        return per_bh * static_cast<double>(bh) / sms;
    }
    double busy[256] = {0.0};
    const int n = sms < 256 ? sms : 256;
    for (long long w = 0; w < items; ++w) busy[w % n] += item_cost(w);
    double worst = 0.0;
    for (int i = 0; i < n; ++i) worst = busy[i] > worst ? busy[i] : worst;
    return worst;
}

// Which full-grid kernel is expected to finish first (same unit: one full item of the 4-tile layout = 4 Q tiles at relative speed
// 1.00; w4 runs two one-tile CTAs per SM at relative speed 0.85).  Relative speeds are same-box medians on square shapes, B200.
int choose_full_grid_kernel(long long bh, long long tiles_m, int sms) {
    const long long waves_w4 = (bh * tiles_m + 2LL * sms - 1) / (2LL * sms);
    const double t_w4 = static_cast<double>(waves_w4) * 2.0 * 0.25 / 0.85;
    const double t_mq4 = mq_time(bh, tiles_m, 4, sms);
    const double t_mq2 = mq_time(bh, tiles_m, 2, sms);
    if (t_mq4 <= t_mq2 && t_mq4 <= t_w4) return 1;
    return t_mq2 <= t_w4 ? 2 : 0;
}

int resolve_encode() {
    if (g_encode.load(std::memory_order_acquire) != nullptr) return 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
        (void)cudaGetLastError();
        return LG_ATTN_EDRIVER;
    }
    g_encode.store(reinterpret_cast<EncodeTiledFn>(fn), std::memory_order_release);
    return 0;
}

// 4-D map over a [batch, heads, rows, 64] half tensor with arbitrary (16-byte aligned) batch / head /
// row strides: dims (64, rows, heads, batch), box (64, 128, 1, 1), 128-byte swizzle.  Rows past `rows`
// inside a (batch, head) slice are zero-filled by the TMA unit, which is what makes ragged N / M work
// without the reference's padded workspace copy; the strides are what lets packed projection outputs
// ([B, N, 3*256] etc.) feed the kernel without a head-split transpose.
struct RowStrides {
    long long batch, head, row;  // elements
};

RowStrides contiguous_strides(int heads, int rows) {
    return RowStrides{static_cast<long long>(heads) * rows * kHeadDim, static_cast<long long>(rows) * kHeadDim, kHeadDim};
}

bool strides_ok(const RowStrides& st, int batch, int heads, int rows, int elem_bytes) {
    const long long a = 16 / elem_bytes;
    auto ok = [&](long long v, int extent) { return extent == 1 || (v > 0 && v % a == 0 && v * elem_bytes < (1LL << 40)); };
    return ok(st.batch, batch) && ok(st.head, heads) && ok(st.row, rows) && (rows == 1 || st.row >= kHeadDim);
}

int encode_rows_map(CUtensorMap* map, const void* base, int rows, int heads, int batch, const RowStrides& st,
                    int box_rows = kBlockN) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(kHeadDim), static_cast<cuuint64_t>(rows),
                                static_cast<cuuint64_t>(heads), static_cast<cuuint64_t>(batch)};
    // extent-1 dimensions never advance; give them any legal stride
    const cuuint64_t srow = static_cast<cuuint64_t>(rows > 1 ? st.row : kHeadDim) * 2;
    const cuuint64_t shead = heads > 1 ? static_cast<cuuint64_t>(st.head) * 2 : srow;
    const cuuint64_t sbatch = batch > 1 ? static_cast<cuuint64_t>(st.batch) * 2 : srow;
    const cuuint64_t strides[3] = {srow, shead, sbatch};
    const cuuint32_t box[4] = {static_cast<cuuint32_t>(kHeadDim), static_cast<cuuint32_t>(box_rows), 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = g_encode.load(std::memory_order_acquire)(
        map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : LG_ATTN_EDRIVER;
}

// FP32 Q / K / V of the plugin's kFLOAT branch, read directly by the latency kernel's FP32-input instantiation: same dims,
// box (32 floats = 128 B, 128 rows, 1, 1), no swizzle (the converter warps write the swizzled FP16 operand tile themselves).
int encode_rows_map_f32(CUtensorMap* map, const void* base, int rows, int heads, int batch, const RowStrides& st) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(kHeadDim), static_cast<cuuint64_t>(rows),
                                static_cast<cuuint64_t>(heads), static_cast<cuuint64_t>(batch)};
    const cuuint64_t srow = static_cast<cuuint64_t>(rows > 1 ? st.row : kHeadDim) * 4;
    const cuuint64_t shead = heads > 1 ? static_cast<cuuint64_t>(st.head) * 4 : srow;
    const cuuint64_t sbatch = batch > 1 ? static_cast<cuuint64_t>(st.batch) * 4 : srow;
    const cuuint64_t strides[3] = {srow, shead, sbatch};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(kBlockN), 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = g_encode.load(std::memory_order_acquire)(
        map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : LG_ATTN_EDRIVER;
}

int init_device_locked(int device) {
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (prop.major != 10) return LG_ATTN_EARCH;
    g_sm_count[device] = prop.multiProcessorCount;
    if (const char* env = getenv("LG_ATTN_SPLITS")) {
        const int v = atoi(env);
        if (v == 1 || v == 2 || v == 4 || v == 8) g_forced_splits.store(v, std::memory_order_relaxed);
    }
    if (const char* env = getenv("LG_ATTN_MQ_TMA_STORE")) g_mq_tma_store.store(atoi(env) != 0 ? 1 : 0, std::memory_order_relaxed);
    if (const char* env = getenv("LG_ATTN_PDL")) g_use_pdl.store(atoi(env) != 0 ? 1 : 0, std::memory_order_relaxed);
    if (const char* env = getenv("LG_ATTN_FULL_GRID_KERNEL")) {
        const int v = atoi(env);
        g_full_grid_kernel.store(v < 0 || v >= kMqModeBase + kMqVariants ? -1 : v, std::memory_order_relaxed);
    }
    int rc = resolve_encode();
    if (rc != 0) return rc;
    int prev = -1;
    e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (prev != device) {
        e = cudaSetDevice(device);
        if (e != cudaSuccess) return static_cast<int>(e);
    }
    // opt in to > 48 KiB dynamic shared memory (per device; the attribute is per context)
    auto opt_in = [&](const void* fn, uint32_t bytes) {
        if (e == cudaSuccess) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    };
    e = cudaSuccess;
    opt_in(reinterpret_cast<const void*>(w4::lg_attn_hd64_fwd<false>), w4::kSmemBytes);
    opt_in(reinterpret_cast<const void*>(w4::lg_attn_hd64_fwd<true>), w4::kSmemBytes);
    mq_for_each_variant([&](auto v) {
        using V = decltype(v);
        opt_in(reinterpret_cast<const void*>(mq::lg_attn_hd64_fwd<V::kT, V::kBN, V::kAlias, V::kLazy, false, V::kSplit>), V::C::kSmemBytes);
        opt_in(reinterpret_cast<const void*>(mq::lg_attn_hd64_fwd<V::kT, V::kBN, V::kAlias, V::kLazy, true, V::kSplit>), V::C::kSmemBytes);
    });
    opt_in(reinterpret_cast<const void*>(w8::lg_attn_hd64_fwd<false, false>), w8::kSmemBytes);
    opt_in(reinterpret_cast<const void*>(w8::lg_attn_hd64_fwd<true, false>), w8::kSmemBytes);
    opt_in(reinterpret_cast<const void*>(w8::lg_attn_hd64_fwd<false, true>), w8::kSmemBytes + w8::kSmemSplitExtra);
    opt_in(reinterpret_cast<const void*>(w8::lg_attn_hd64_fwd<true, true>), w8::kSmemBytes + w8::kSmemSplitExtra);
    opt_in(reinterpret_cast<const void*>(w8::lg_attn_hd64_fwd<true, true, true>), w8::kSmemBytesF32);
    // how many clusters of 2 / 4 / 8 split-KV CTAs are co-resident on this device
    for (int lg = 1; lg <= 3 && e == cudaSuccess; ++lg) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(64u << lg);
        cfg.blockDim = dim3(w8::kThreads);
        cfg.dynamicSmemBytes = w8::kSmemBytes + w8::kSmemSplitExtra;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 1u << lg;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, w8::lg_attn_hd64_fwd<false, true>, &cfg) != cudaSuccess) {
            (void)cudaGetLastError();
            n = 0;  // no split launches on this device
        }
        g_max_clusters[device][lg] = n;
        cfg.blockDim = dim3(w8::kThreadsF32);
        cfg.dynamicSmemBytes = w8::kSmemBytesF32;
        n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, w8::lg_attn_hd64_fwd<true, true, true>, &cfg) != cudaSuccess) {
            (void)cudaGetLastError();
            n = 0;
        }
        g_max_clusters_f32[device][lg] = n;
    }
    if (prev != device && prev >= 0) (void)cudaSetDevice(prev);
    return static_cast<int>(e);
}

size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

bool valid_dtype(int d) { return d == LG_ATTN_F16_F16 || d == LG_ATTN_F32_F32 || d == LG_ATTN_F16_F32; }

}  // namespace

#ifdef LG_ATTN_W4_TRACE
extern "C" int lg_attn_debug_read_w4_trace(long long* dst, unsigned* smid) {
    cudaError_t e = cudaMemcpyFromSymbol(dst, lgattn::w4::g_w4_trace, sizeof(lgattn::w4::g_w4_trace));
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(smid, lgattn::w4::g_w4_smid, sizeof(lgattn::w4::g_w4_smid));
    return static_cast<int>(e);
}
#endif
#ifdef LG_ATTN_MQ_TRACE
extern "C" int lg_attn_debug_read_mq_trace(long long* dst) {
    return static_cast<int>(cudaMemcpyFromSymbol(dst, lgattn::mq::g_mq_trace, sizeof(lgattn::mq::g_mq_trace)));
}
#endif
#ifdef LG_ATTN_TRACE
static std::atomic<int> g_trace_seq{0};
#define LG_TRACE_SEQ , g_trace_seq.fetch_add(1)
extern "C" int lg_attn_debug_max_clusters(int device, int log2_splits) {
    return (device >= 0 && device < kMaxDevices && log2_splits >= 1 && log2_splits <= 3) ? g_max_clusters[device][log2_splits] : -1;
}
extern "C" int lg_attn_debug_read_trace(unsigned long long* dst, int reset_seq) {
    if (reset_seq) g_trace_seq.store(0);
    return static_cast<int>(cudaMemcpyFromSymbol(dst, lgattn::w8::g_trace, sizeof(lgattn::w8::g_trace)));
}
#else
#define LG_TRACE_SEQ
#endif

// Shared tail of both public launchers: tensor maps + kernel choice + launch.  FP16 Q/K/V only.
static int launch_core(const void* qh, const void* kh, const void* vh, void* o, int batch, int heads, int n_q,
                       int n_kv, bool out_f32, const RowStrides& sq, const RowStrides& sk, const RowStrides& sv,
                       const RowStrides& so, int kv_shift, int device, cudaStream_t stream, bool in_f32 = false) {
    CUtensorMap tm_q, tm_k, tm_v;
    int rc;
    if (in_f32) {
        if ((rc = encode_rows_map_f32(&tm_q, qh, n_q, heads, batch, sq)) != 0) return rc;
        if ((rc = encode_rows_map_f32(&tm_k, kh, n_kv, heads, batch, sk)) != 0) return rc;
        if ((rc = encode_rows_map_f32(&tm_v, vh, n_kv, heads, batch, sv)) != 0) return rc;
    } else {
        if ((rc = encode_rows_map(&tm_q, qh, n_q, heads, batch, sq)) != 0) return rc;
        if ((rc = encode_rows_map(&tm_k, kh, n_kv, heads, batch, sk)) != 0) return rc;
        if ((rc = encode_rows_map(&tm_v, vh, n_kv, heads, batch, sv)) != 0) return rc;
    }
    const OutLayout ol{so.batch, so.head, so.row, heads, batch, kv_shift};
    const long long bh_ll = static_cast<long long>(batch) * heads;
    const long long tiles_m_ll = (static_cast<long long>(n_q) + kBlockM - 1) / kBlockM;
    const int tiles_m = static_cast<int>(tiles_m_ll);
    cudaError_t e = cudaSuccess;

    // Full grids (>= 2 Q tiles per SM) take a throughput kernel — the multi-Q-tile persistent kernel (mq) in its 4- or 2-tile
    // layout, or the one-tile-per-CTA kernel (w4) when the Q-tile count packs badly into work items (choose_full_grid_kernel);
    // grids that cannot fill the GPU take the 8-softmax-warp kernel (w8), which splits each
    // Q tile's KV range over a cluster of `splits` CTAs (DSMEM combine) and uses programmatic
    // dependent launch so its prologue overlaps the previous kernel's tail.  splits = the largest
    // power of two <= min(8, KV tiles) whose clusters are all co-resident (one wave), from
    // cudaOccupancyMaxActiveClusters at device init (B200: 74 / 33 / 15 clusters of 2 / 4 / 8).
    const int kv_tiles = (n_kv + kBlockN - 1) / kBlockN;
    const long long base = bh_ll * tiles_m_ll;
    const int forced = g_forced_splits.load(std::memory_order_relaxed);
    // FP32 inputs (the plugin's kFLOAT branch: B = 1 engines) always take the latency kernel, whose FP32-input instantiation
    // converts on the way into shared memory: one launch, no workspace, at the price of one CTA per SM on large grids
    const bool full_grid = base >= 2LL * g_sm_count[device] && forced == 0 && !in_f32;

    cudaLaunchConfig_t cfg = {};
    cfg.stream = stream;
    cudaLaunchAttribute attrs[2];
    unsigned n_attrs = 0;
    int fg_kernel = g_full_grid_kernel.load(std::memory_order_relaxed);
    if (full_grid && fg_kernel < 0) fg_kernel = choose_full_grid_kernel(bh_ll, tiles_m_ll, g_sm_count[device]);
    if (full_grid && fg_kernel >= kMqModeBase) {
        // one persistent CTA per SM, kT Q tiles per work item, K/V tiles of kBN rows shared by the kT tiles
        int idx = kMqModeBase;
        mq_for_each_variant([&](auto v) {
            using V = decltype(v);
            if (idx++ != fg_kernel || e != cudaSuccess || rc != 0) return;
            if (V::kBN != kBlockN) {
                if ((rc = encode_rows_map(&tm_k, kh, n_kv, heads, batch, sk, V::kBN)) != 0) return;
                if ((rc = encode_rows_map(&tm_v, vh, n_kv, heads, batch, sv, V::kBN)) != 0) return;
            }
            const mq::Sched sc = mq_schedule(bh_ll, tiles_m_ll, V::kT, g_sm_count[device]);
            const long long items = sc.total_items;
            // FP16 output leaves through a TMA store per Q tile (same map shape as Q's: rows past n_q are clipped); a base or
            // strides the tensor map cannot express fall back to per-thread row stores
            CUtensorMap tm_o = tm_q;
            int tma_store = 0;
            if (!out_f32 && V::kSplit == 1 && g_mq_tma_store.load(std::memory_order_relaxed) != 0 &&
                (reinterpret_cast<uintptr_t>(o) & 15u) == 0 && encode_rows_map(&tm_o, o, n_q, heads, batch, so, kBlockM) == 0)
                tma_store = 1;
            cfg.gridDim = dim3(static_cast<unsigned>(items < g_sm_count[device] ? items : g_sm_count[device]));
            cfg.blockDim = dim3(V::C::kThreads);
            cfg.dynamicSmemBytes = V::C::kSmemBytes;
            cfg.attrs = attrs;
            cfg.numAttrs = 0;
            e = cudaLaunchKernelEx(&cfg,
                                   !out_f32 ? mq::lg_attn_hd64_fwd<V::kT, V::kBN, V::kAlias, V::kLazy, false, V::kSplit>
                                            : mq::lg_attn_hd64_fwd<V::kT, V::kBN, V::kAlias, V::kLazy, true, V::kSplit>,
                                   tm_q, tm_k, tm_v, tm_o, o, n_q, n_kv, sc, ol, tma_store);
        });
        if (rc != 0) return rc;
    } else if (full_grid) {
        cfg.gridDim = dim3(static_cast<unsigned>(base));
        cfg.blockDim = dim3(w4::kThreads);
        cfg.dynamicSmemBytes = w4::kSmemBytes;
        cfg.attrs = attrs;
        cfg.numAttrs = 0;
        e = cudaLaunchKernelEx(&cfg, !out_f32 ? w4::lg_attn_hd64_fwd<false> : w4::lg_attn_hd64_fwd<true>, tm_q,
                               tm_k, tm_v, o, n_q, n_kv, tiles_m, ol);
    } else {
        int splits = 1;
        if (forced > 0) {
            splits = forced;
        } else if (base < g_sm_count[device]) {
            const int* max_clusters = in_f32 ? g_max_clusters_f32[device] : g_max_clusters[device];
            for (int lg = 1; lg <= 3; ++lg)
                if ((1 << lg) <= kv_tiles && base <= max_clusters[lg]) splits = 1 << lg;
        }
        while (splits > 1 && splits > kv_tiles) splits /= 2;
        if (base * splits > 2147483647LL) return LG_ATTN_EINVAL;
        int flags = 0;
        if (base * splits <= 2LL * g_sm_count[device]) flags |= w8::kFlagEarlyTrigger;
        cfg.gridDim = dim3(static_cast<unsigned>(base) * static_cast<unsigned>(splits));
        cfg.blockDim = dim3(in_f32 ? w8::kThreadsF32 : w8::kThreads);
        cfg.dynamicSmemBytes = in_f32 ? w8::kSmemBytesF32 : w8::kSmemBytes + (splits > 1 ? w8::kSmemSplitExtra : 0);
        if (splits > 1) {
            attrs[n_attrs].id = cudaLaunchAttributeClusterDimension;
            attrs[n_attrs].val.clusterDim.x = static_cast<unsigned>(splits);
            attrs[n_attrs].val.clusterDim.y = 1;
            attrs[n_attrs].val.clusterDim.z = 1;
            ++n_attrs;
        }
        if (g_use_pdl.load(std::memory_order_relaxed)) {
            // this kernel's prologue may overlap the previous kernel's tail (it still waits, at
            // griddepcontrol.wait, for that kernel to complete before touching memory)
            attrs[n_attrs].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attrs[n_attrs].val.programmaticStreamSerializationAllowed = 1;
            ++n_attrs;
        }
        cfg.attrs = attrs;
        cfg.numAttrs = n_attrs;
        const bool f32 = out_f32;
        if (in_f32) {
            e = cudaLaunchKernelEx(&cfg, w8::lg_attn_hd64_fwd<true, true, true>, tm_q, tm_k, tm_v, o, n_q, n_kv, tiles_m, splits, flags,
                                   ol LG_TRACE_SEQ);
        } else if (splits > 1) {
            e = cudaLaunchKernelEx(&cfg, f32 ? w8::lg_attn_hd64_fwd<true, true> : w8::lg_attn_hd64_fwd<false, true>, tm_q, tm_k,
                                   tm_v, o, n_q, n_kv, tiles_m, splits, flags, ol LG_TRACE_SEQ);
        } else {
            e = cudaLaunchKernelEx(&cfg, f32 ? w8::lg_attn_hd64_fwd<true, false> : w8::lg_attn_hd64_fwd<false, false>, tm_q, tm_k,
                                   tm_v, o, n_q, n_kv, tiles_m, splits, flags, ol LG_TRACE_SEQ);
        }
    }
    if (e != cudaSuccess) return static_cast<int>(e);
    return static_cast<int>(cudaGetLastError());
}

extern "C" {

int lg_attn_version(void) { return kAbiVersion; }

int lg_attn_build_flags(void) {
    int f = 0;
#ifdef LG_ATTN_TUNING_BUILD
    f |= 1;
#endif
#if defined(LG_ATTN_WHATIF) || defined(LG_ATTN_WHATIF_F3)
    f |= 2;
#endif
#ifdef LG_ATTN_EXPERIMENT_NOSUM
    f |= 4;
#endif
#if defined(LG_ATTN_TRACE) || defined(LG_ATTN_W4_TRACE) || defined(LG_ATTN_MQ_TRACE)
    f |= 8;
#endif
    return f;
}

int lg_attn_debug_set_throughput_kernel(int mode) {
    if (mode < -1 || mode >= kMqModeBase + kMqVariants) return LG_ATTN_EINVAL;
    g_full_grid_kernel.store(mode, std::memory_order_relaxed);
    return 0;
}

int lg_attn_debug_set_pdl(int enabled) {
    g_use_pdl.store(enabled != 0 ? 1 : 0, std::memory_order_relaxed);
    return 0;
}

int lg_attn_debug_describe_launch(int batch, int heads, int n_q, int n_kv, char* buf, size_t buf_bytes) {
    if (buf == nullptr || buf_bytes == 0 || batch <= 0 || heads <= 0 || n_q <= 0 || n_kv <= 0) return LG_ATTN_EINVAL;
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_init_device(device);
    if (rc != 0) return rc;
    const long long bh = static_cast<long long>(batch) * heads, tiles_m = (static_cast<long long>(n_q) + kBlockM - 1) / kBlockM;
    const long long base = bh * tiles_m;
    const int kv_tiles = (n_kv + kBlockN - 1) / kBlockN, sms = g_sm_count[device];
    const int forced = g_forced_splits.load(std::memory_order_relaxed);
    if (base >= 2LL * sms && forced == 0) {
        int k = g_full_grid_kernel.load(std::memory_order_relaxed);
        if (k < 0) k = choose_full_grid_kernel(bh, tiles_m, sms);
        if (k == 0) {
            snprintf(buf, buf_bytes, "lgattn::w4 (one 128-row Q tile per CTA, 2 CTAs/SM), grid %lld", base);
        } else {
            const int t = k == 1 ? 4 : 2;
            const mq::Sched sc = mq_schedule(bh, tiles_m, t, sms);
            const long long items = sc.total_items;
            snprintf(buf, buf_bytes,
                     "lgattn::mq<%d Q tiles, %d-row KV steps> (persistent, 1 CTA/SM), grid %lld over %lld work items (last wave cut in %d)",
                     t, k == 1 ? 64 : 128, items < sms ? items : static_cast<long long>(sms), items, sc.parts);
        }
    } else {
        int splits = 1;
        if (forced > 0) {
            splits = forced;
        } else if (base < sms) {
            for (int lg = 1; lg <= 3; ++lg)
                if ((1 << lg) <= kv_tiles && base <= g_max_clusters[device][lg]) splits = 1 << lg;
        }
        while (splits > 1 && splits > kv_tiles) splits /= 2;
        snprintf(buf, buf_bytes, "lgattn::w8 (latency kernel, %d-CTA split-KV cluster%s, PDL %s), grid %lld", splits, splits > 1 ? "s" : "",
                 g_use_pdl.load(std::memory_order_relaxed) ? "on" : "off", base * splits);
    }
    return 0;
}

int lg_attn_hd64_init_device(int device) {
    if (device < 0 || device >= kMaxDevices) return LG_ATTN_EINVAL;
    if (g_device_ready[device].load(std::memory_order_acquire) == 1) return 0;
    std::lock_guard<std::mutex> lock(g_init_mutex);
    if (g_device_ready[device].load(std::memory_order_acquire) == 1) return 0;
    int rc = init_device_locked(device);
    if (rc == 0) g_device_ready[device].store(1, std::memory_order_release);
    return rc;
}

size_t lg_attn_hd64_workspace_bytes(int batch, int heads, int n_q, int n_kv, int io_dtype) {
    (void)batch; (void)heads; (void)n_q; (void)n_kv; (void)io_dtype;
    return 0;   // every mode is one launch without scratch (the kFLOAT conversion happens on the way into shared memory)
}

int lg_attn_hd64_launches_per_call(int io_dtype) {
    if (!valid_dtype(io_dtype)) return 0;
    return 1;
}

int lg_attn_hd64_launch(const void* q, const void* k, const void* v, void* o, int batch, int heads,
                        int n_q, int n_kv, int io_dtype, void* workspace, size_t workspace_bytes,
                        void* stream_v) {
    if (q == nullptr || k == nullptr || v == nullptr || o == nullptr) return LG_ATTN_EINVAL;
    if (batch <= 0 || heads <= 0 || n_q <= 0 || n_kv <= 0 || !valid_dtype(io_dtype)) return LG_ATTN_EINVAL;
    const long long bh_ll = static_cast<long long>(batch) * heads;
    const long long tiles_m_ll = (static_cast<long long>(n_q) + kBlockM - 1) / kBlockM;
    if (bh_ll * tiles_m_ll > 2147483647LL) return LG_ATTN_EINVAL;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);

    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_init_device(device);
    if (rc != 0) return rc;

    const int bh = static_cast<int>(bh_ll);
    const void* qh = q;
    const void* kh = k;
    const void* vh = v;

    (void)workspace;        // kept in the signature for ABI stability; no mode needs scratch any more
    (void)workspace_bytes;
    (void)bh;
    const RowStrides cq = contiguous_strides(heads, n_q), ckv = contiguous_strides(heads, n_kv);
    return launch_core(qh, kh, vh, o, batch, heads, n_q, n_kv, io_dtype != LG_ATTN_F16_F16, cq, ckv, ckv, cq, 0, device,
                       stream, io_dtype == LG_ATTN_F32_F32);
}

int lg_attn_hd64_launch_strided(const void* q, const void* k, const void* v, void* o, int batch, int heads,
                                int n_q, int n_kv, int out_f32, const lg_attn_strides* sq, const lg_attn_strides* sk,
                                const lg_attn_strides* sv, const lg_attn_strides* so, int kv_batch_shift, void* stream_v) {
    if (q == nullptr || k == nullptr || v == nullptr || o == nullptr) return LG_ATTN_EINVAL;
    if (sq == nullptr || sk == nullptr || sv == nullptr || so == nullptr) return LG_ATTN_EINVAL;
    if (batch <= 0 || heads <= 0 || n_q <= 0 || n_kv <= 0 || (out_f32 != 0 && out_f32 != 1)) return LG_ATTN_EINVAL;
    if (kv_batch_shift < 0 || kv_batch_shift >= batch) return LG_ATTN_EINVAL;
    const long long tiles = (static_cast<long long>(n_q) + kBlockM - 1) / kBlockM;
    if (static_cast<long long>(batch) * heads * tiles > 2147483647LL) return LG_ATTN_EINVAL;
    const RowStrides rq{sq->batch, sq->head, sq->row}, rk{sk->batch, sk->head, sk->row};
    const RowStrides rv{sv->batch, sv->head, sv->row}, ro{so->batch, so->head, so->row};
    auto aligned16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(o)) return LG_ATTN_EINVAL;
    if (!strides_ok(rq, batch, heads, n_q, 2) || !strides_ok(rk, batch, heads, n_kv, 2) ||
        !strides_ok(rv, batch, heads, n_kv, 2) || !strides_ok(ro, batch, heads, n_q, out_f32 ? 4 : 2))
        return LG_ATTN_EINVAL;
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_init_device(device);
    if (rc != 0) return rc;
    return launch_core(q, k, v, o, batch, heads, n_q, n_kv, out_f32 != 0, rq, rk, rv, ro, kv_batch_shift, device,
                       static_cast<cudaStream_t>(stream_v));
}

int lg_attn_hd64_run_host(const void* q_host, const void* k_host, const void* v_host, void* o_host,
                          int batch, int heads, int n_q, int n_kv, int io_dtype, int device) {
    if (q_host == nullptr || k_host == nullptr || v_host == nullptr || o_host == nullptr) return LG_ATTN_EINVAL;
    if (batch <= 0 || heads <= 0 || n_q <= 0 || n_kv <= 0 || !valid_dtype(io_dtype)) return LG_ATTN_EINVAL;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_cast<int>(e);
    const size_t in_es = (io_dtype == LG_ATTN_F32_F32) ? 4 : 2;
    const size_t out_es = (io_dtype == LG_ATTN_F16_F16) ? 2 : 4;
    const size_t bh = static_cast<size_t>(batch) * heads;
    const size_t q_bytes = bh * n_q * kHeadDim * in_es;
    const size_t kv_bytes = bh * n_kv * kHeadDim * in_es;
    const size_t o_bytes = bh * n_q * kHeadDim * out_es;
    const size_t ws_bytes = lg_attn_hd64_workspace_bytes(batch, heads, n_q, n_kv, io_dtype);
    uint8_t* dev = nullptr;
    const size_t total = align256(q_bytes) + 2 * align256(kv_bytes) + align256(o_bytes) + ws_bytes;
    e = cudaMalloc(&dev, total);
    if (e != cudaSuccess) return static_cast<int>(e);
    uint8_t* dq = dev;
    uint8_t* dk = dq + align256(q_bytes);
    uint8_t* dv = dk + align256(kv_bytes);
    uint8_t* dout = dv + align256(kv_bytes);
    uint8_t* dws = dout + align256(o_bytes);
    cudaStream_t stream = nullptr;
    int rc = 0;
    e = cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dq, q_host, q_bytes, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dk, k_host, kv_bytes, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dv, v_host, kv_bytes, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) {
        rc = lg_attn_hd64_launch(dq, dk, dv, dout, batch, heads, n_q, n_kv, io_dtype, ws_bytes ? dws : nullptr,
                                 ws_bytes, stream);
    }
    if (e == cudaSuccess && rc == 0) e = cudaMemcpyAsync(o_host, dout, o_bytes, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess && rc == 0) e = cudaStreamSynchronize(stream);
    if (stream != nullptr) (void)cudaStreamDestroy(stream);
    (void)cudaFree(dev);
    if (rc != 0) return rc;
    return static_cast<int>(e);
}

// Host-buffer session.  Three streams — H2D copies, launches, D2H copies — tied together per slot by events, so that each copy
// engine and the SMs each see one back-to-back queue: the input copy of problem k+1.. runs under the kernel of problem k and the
// output copy of problem k-1 (one stream per slot, round 1, left the H2D engine idle whenever a slot's own D2H had not finished:
// 25 of the 30 TFLOP/s-equivalent a copy-only loop reaches on the same box, tools/copy_only_probe.py).
struct lg_attn_session {
    int device = 0, max_batch = 0, heads = 0, max_n_q = 0, max_n_kv = 0, io_dtype = 0, n_slots = 0;
    size_t q_bytes = 0, kv_bytes = 0, o_bytes = 0;
    uint8_t* pool = nullptr;
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    cudaEvent_t* ev_in = nullptr;    // [slot] inputs have landed
    cudaEvent_t* ev_run = nullptr;   // [slot] kernel done: inputs consumed, output written
    cudaEvent_t* ev_out = nullptr;   // [slot] output copied out: the slot's output buffer may be rewritten
    unsigned long long next = 0;
    // pool layout: [n_slots x (q | k | v)] then [n_slots x o] — consecutive slots are contiguous on the device, so a packed group of
    // problems moves with ONE copy each way (lg_attn_session_submit_many)
    size_t in_bytes() const { return q_bytes + 2 * kv_bytes; }
    uint8_t* slot_in(int i) const { return pool + in_bytes() * static_cast<size_t>(i); }
    uint8_t* slot_out(int i) const { return pool + in_bytes() * static_cast<size_t>(n_slots) + o_bytes * static_cast<size_t>(i); }
};

int lg_attn_session_create(lg_attn_session** out, int device, int max_batch, int heads, int max_n_q, int max_n_kv,
                           int io_dtype, int n_slots) {
    if (out == nullptr || max_batch <= 0 || heads <= 0 || max_n_q <= 0 || max_n_kv <= 0 || n_slots <= 0 ||
        n_slots > 64 || !valid_dtype(io_dtype))
        return LG_ATTN_EINVAL;
    *out = nullptr;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_init_device(device);
    if (rc != 0) return rc;
    auto* s = new lg_attn_session();
    s->device = device; s->max_batch = max_batch; s->heads = heads; s->max_n_q = max_n_q; s->max_n_kv = max_n_kv;
    s->io_dtype = io_dtype; s->n_slots = n_slots;
    const size_t in_es = (io_dtype == LG_ATTN_F32_F32) ? 4 : 2;
    const size_t out_es = (io_dtype == LG_ATTN_F16_F16) ? 2 : 4;
    const size_t bh = static_cast<size_t>(max_batch) * heads;
    s->q_bytes = align256(bh * max_n_q * kHeadDim * in_es);
    s->kv_bytes = align256(bh * max_n_kv * kHeadDim * in_es);
    s->o_bytes = align256(bh * max_n_q * kHeadDim * out_es);
    const size_t slot = s->q_bytes + 2 * s->kv_bytes + s->o_bytes;
    e = cudaMalloc(&s->pool, slot * n_slots);
    if (e != cudaSuccess) { delete s; return static_cast<int>(e); }
    s->ev_in = new cudaEvent_t[n_slots]();
    s->ev_run = new cudaEvent_t[n_slots]();
    s->ev_out = new cudaEvent_t[n_slots]();
    e = cudaStreamCreateWithFlags(&s->s_in, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->s_run, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->s_out, cudaStreamNonBlocking);
    for (int i = 0; i < n_slots && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&s->ev_in[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_run[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s->ev_out[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) { lg_attn_session_destroy(s); return static_cast<int>(e); }
    *out = s;
    return 0;
}

int lg_attn_session_submit(lg_attn_session* s, const void* q_host, const void* k_host, const void* v_host,
                           void* o_host, int batch, int n_q, int n_kv) {
    if (s == nullptr || q_host == nullptr || k_host == nullptr || v_host == nullptr || o_host == nullptr)
        return LG_ATTN_EINVAL;
    if (batch <= 0 || batch > s->max_batch || n_q <= 0 || n_q > s->max_n_q || n_kv <= 0 || n_kv > s->max_n_kv)
        return LG_ATTN_EINVAL;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return static_cast<int>(e);
    const unsigned long long k = s->next++;
    const int i = static_cast<int>(k % static_cast<unsigned long long>(s->n_slots));
    const bool reused = k >= static_cast<unsigned long long>(s->n_slots);
    uint8_t* dq = s->slot_in(i);
    uint8_t* dk = dq + s->q_bytes;
    uint8_t* dv = dk + s->kv_bytes;
    uint8_t* dout = s->slot_out(i);
    const size_t in_es = (s->io_dtype == LG_ATTN_F32_F32) ? 4 : 2;
    const size_t out_es = (s->io_dtype == LG_ATTN_F16_F16) ? 2 : 4;
    const size_t bh = static_cast<size_t>(batch) * s->heads;
    const size_t qb = bh * n_q * kHeadDim * in_es, kvb = bh * n_kv * kHeadDim * in_es;
    const uint8_t* qh8 = static_cast<const uint8_t*>(q_host);
    // ---- input copies: the slot's previous kernel must have consumed its inputs
    if (reused) e = cudaStreamWaitEvent(s->s_in, s->ev_run[i], 0);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (static_cast<const uint8_t*>(k_host) == qh8 + qb && static_cast<const uint8_t*>(v_host) == qh8 + qb + kvb &&
        qb == s->q_bytes && kvb == s->kv_bytes) {
        // q | k | v adjacent in host memory and the slot is exactly this shape: one 3x larger PCIe transfer
        // (0.5 MiB copies reach ~2/3 of the link rate of 1.5 MiB ones on this box)
        e = cudaMemcpyAsync(dq, q_host, qb + 2 * kvb, cudaMemcpyHostToDevice, s->s_in);
    } else {
        e = cudaMemcpyAsync(dq, q_host, qb, cudaMemcpyHostToDevice, s->s_in);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dk, k_host, kvb, cudaMemcpyHostToDevice, s->s_in);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dv, v_host, kvb, cudaMemcpyHostToDevice, s->s_in);
    }
    if (e == cudaSuccess) e = cudaEventRecord(s->ev_in[i], s->s_in);
    // ---- launch: inputs landed, and the slot's previous output has been copied out
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s->s_run, s->ev_in[i], 0);
    if (e == cudaSuccess && reused) e = cudaStreamWaitEvent(s->s_run, s->ev_out[i], 0);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_launch(dq, dk, dv, dout, batch, s->heads, n_q, n_kv, s->io_dtype, nullptr, 0, s->s_run);
    if (rc != 0) return rc;
    e = cudaEventRecord(s->ev_run[i], s->s_run);
    // ---- output copy
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s->s_out, s->ev_run[i], 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(o_host, dout, bh * n_q * kHeadDim * out_es, cudaMemcpyDeviceToHost, s->s_out);
    if (e == cudaSuccess) e = cudaEventRecord(s->ev_out[i], s->s_out);
    return static_cast<int>(e);
}

int lg_attn_session_submit_many(lg_attn_session* s, const void* q_host, const void* k_host, const void* v_host, void* o_host,
                                int count, size_t q_stride_bytes, size_t kv_stride_bytes, size_t out_stride_bytes, int group, int batch,
                                int n_q, int n_kv) {
    if (s == nullptr || q_host == nullptr || k_host == nullptr || v_host == nullptr || o_host == nullptr || count <= 0)
        return LG_ATTN_EINVAL;
    if (batch <= 0 || batch > s->max_batch || n_q <= 0 || n_q > s->max_n_q || n_kv <= 0 || n_kv > s->max_n_kv)
        return LG_ATTN_EINVAL;
    const size_t in_es = (s->io_dtype == LG_ATTN_F32_F32) ? 4 : 2;
    const size_t out_es = (s->io_dtype == LG_ATTN_F16_F16) ? 2 : 4;
    const size_t bh = static_cast<size_t>(batch) * s->heads;
    const size_t qb = bh * n_q * kHeadDim * in_es, kvb = bh * n_kv * kHeadDim * in_es, ob = bh * n_q * kHeadDim * out_es;
    const uint8_t* q8 = static_cast<const uint8_t*>(q_host);
    const uint8_t* k8 = static_cast<const uint8_t*>(k_host);
    const uint8_t* v8 = static_cast<const uint8_t*>(v_host);
    uint8_t* o8 = static_cast<uint8_t*>(o_host);
    // packed = the host arrays are laid out exactly like `count` consecutive device slots: [q | k | v] per problem, outputs dense
    const size_t in_stride_bytes = q_stride_bytes;
    const bool packed = k8 == q8 + qb && v8 == k8 + kvb && q_stride_bytes == qb + 2 * kvb && kv_stride_bytes == q_stride_bytes && out_stride_bytes == ob &&
                        qb == s->q_bytes && kvb == s->kv_bytes && ob == s->o_bytes;
    if (group <= 0) group = 4;
    if (group > s->n_slots / 2) group = s->n_slots / 2;   // at least two groups of slots in flight
    if (!packed || group <= 1) {
        for (int p = 0; p < count; ++p) {
            const int rc = lg_attn_session_submit(s, q8 + p * q_stride_bytes, k8 + p * kv_stride_bytes, v8 + p * kv_stride_bytes,
                                                  o8 + p * out_stride_bytes, batch, n_q, n_kv);
            if (rc != 0) return rc;
        }
        return 0;
    }
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) return static_cast<int>(e);
    for (int done = 0; done < count;) {
        const int i0 = static_cast<int>(s->next % static_cast<unsigned long long>(s->n_slots));
        int g = group < count - done ? group : count - done;
        if (g > s->n_slots - i0) g = s->n_slots - i0;   // a group never wraps around the pool
        // ---- ONE H2D copy for the group: every slot's previous kernel must have consumed its inputs
        for (int j = 0; j < g && e == cudaSuccess; ++j)
            if (s->next + j >= static_cast<unsigned long long>(s->n_slots)) e = cudaStreamWaitEvent(s->s_in, s->ev_run[i0 + j], 0);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(s->slot_in(i0), q8 + done * in_stride_bytes, g * in_stride_bytes, cudaMemcpyHostToDevice, s->s_in);
        if (e == cudaSuccess) e = cudaEventRecord(s->ev_in[i0], s->s_in);
        // ---- g launches (one problem each, as the plugin would enqueue them): inputs landed, previous outputs copied out
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s->s_run, s->ev_in[i0], 0);
        for (int j = 0; j < g && e == cudaSuccess; ++j)
            if (s->next + j >= static_cast<unsigned long long>(s->n_slots)) e = cudaStreamWaitEvent(s->s_run, s->ev_out[i0 + j], 0);
        if (e != cudaSuccess) return static_cast<int>(e);
        for (int j = 0; j < g; ++j) {
            uint8_t* dq = s->slot_in(i0 + j);
            const int rc = lg_attn_hd64_launch(dq, dq + s->q_bytes, dq + s->q_bytes + s->kv_bytes, s->slot_out(i0 + j), batch, s->heads, n_q,
                                               n_kv, s->io_dtype, nullptr, 0, s->s_run);
            if (rc != 0) return rc;
        }
        for (int j = 0; j < g && e == cudaSuccess; ++j) e = cudaEventRecord(s->ev_run[i0 + j], s->s_run);
        // ---- ONE D2H copy for the group
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s->s_out, s->ev_run[i0 + g - 1], 0);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(o8 + done * out_stride_bytes, s->slot_out(i0), g * out_stride_bytes, cudaMemcpyDeviceToHost, s->s_out);
        for (int j = 0; j < g && e == cudaSuccess; ++j) e = cudaEventRecord(s->ev_out[i0 + j], s->s_out);
        if (e != cudaSuccess) return static_cast<int>(e);
        s->next += static_cast<unsigned long long>(g);
        done += g;
    }
    return 0;
}

int lg_attn_session_sync(lg_attn_session* s) {
    if (s == nullptr) return LG_ATTN_EINVAL;
    cudaError_t e = cudaStreamSynchronize(s->s_in);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->s_run);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->s_out);
    return static_cast<int>(e);
}

void lg_attn_session_destroy(lg_attn_session* s) {
    if (s == nullptr) return;
    (void)cudaSetDevice(s->device);
    for (cudaStream_t st : {s->s_in, s->s_run, s->s_out})
        if (st != nullptr) { (void)cudaStreamSynchronize(st); (void)cudaStreamDestroy(st); }
    for (cudaEvent_t* arr : {s->ev_in, s->ev_run, s->ev_out}) {
        if (arr == nullptr) continue;
        for (int i = 0; i < s->n_slots; ++i)
            if (arr[i] != nullptr) (void)cudaEventDestroy(arr[i]);
        delete[] arr;
    }
    if (s->pool != nullptr) (void)cudaFree(s->pool);
    delete s;
}

const char* lg_attn_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case LG_ATTN_EINVAL: return "lg_attn: invalid argument";
        case LG_ATTN_EWORKSPACE: return "lg_attn: workspace missing or too small";
        case LG_ATTN_EDRIVER: return "lg_attn: cuTensorMapEncodeTiled unavailable or failed";
        case LG_ATTN_EARCH: return "lg_attn: device is not compute capability 10.x (sm_100a required)";
        case 100010: return "lg_matcher: a cuBLASLt call failed";
        case 100011: return "lg_matcher: malformed weight blob";
        case 100012: return "lg_match_assign_static: shape needs more scratch than LG_MATCH_ASSIGN_STATIC_MAX_BYTES (4 MiB); use lg_match_assign";
        default: return cudaGetErrorString(static_cast<cudaError_t>(code));
    }
}

}  // extern "C"
