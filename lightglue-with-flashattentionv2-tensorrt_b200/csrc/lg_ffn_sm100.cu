// C-ABI of the matcher's hand-written GEMM kernels — fused ffn-up (ffn_sm100_kernel.cuh) and input projection + rotary
// (proj_sm100_kernel.cuh) — declared in include/lg_match.h: host-side glue only — argument checks, per-device one-time
// set-up, three TMA tensor maps per call (cuTensorMapEncodeTiled through cudaGetDriverEntryPoint: a pure host call, legal while a
// stream is being captured) and one launch.  Never allocates, never synchronises.

#include "ffn_sm100_kernel.cuh"
#include "proj_sm100_kernel.cuh"

#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdlib>
#include <mutex>

#include "../../include/lg_attn.h"
#include "../../include/lg_match.h"

namespace {

using namespace lgffn;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int kMaxDevices = 64;
std::mutex g_mutex;
std::atomic<int> g_ready[kMaxDevices];
std::atomic<EncodeTiledFn> g_encode{nullptr};
int g_sms[kMaxDevices];
int g_small_clusters[kMaxDevices];   // co-resident 4-CTA clusters of the few-rows kernel
std::atomic<int> g_small_enabled{1}; // LG_MATCH_FFN_SMALL=0: no few-rows kernel (A/B runs)
std::atomic<int> g_pdl{1};           // LG_ATTN_PDL=0: no programmatic dependent launch
std::atomic<int> g_pair{-1};         // persistent ffn kernel as CTA pairs (cta_group::2): -1 = by size (>= 3 waves of tiles), 0 / 1 = forced
                                     // (LG_MATCH_FFN_PAIR, lg_match_debug_set_ffn_pair)

int init_locked(int device) {
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (prop.major != 10) return LG_ATTN_EARCH;
    if (g_encode.load(std::memory_order_acquire) == nullptr) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return LG_ATTN_EDRIVER;
        }
        g_encode.store(reinterpret_cast<EncodeTiledFn>(fn), std::memory_order_release);
    }
    int prev = 0;
    if ((e = cudaGetDevice(&prev)) != cudaSuccess) return static_cast<int>(e);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return static_cast<int>(e);
    e = cudaFuncSetAttribute(ffn1_ln_gelu_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBytes));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(ffn1_ln_gelu_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemBytes));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(small::ffn1_ln_gelu_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(small::kSmemBytesS));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(lgproj::proj_rotary_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(lgproj::Cfg<256>::kSmemBytes));
    if (e == cudaSuccess)
        e = cudaFuncSetAttribute(lgproj::proj_rotary_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 static_cast<int>(lgproj::Cfg<128>::kSmemBytes));
    if (e == cudaSuccess) {
        // how many 4-CTA clusters of the few-rows kernel are resident at once (one tile each, no persistence: a launch must fit)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(static_cast<unsigned>(prop.multiProcessorCount / small::kSplit * small::kSplit));
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = small::kSmemBytesS;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = small::kSplit;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, small::ffn1_ln_gelu_small_kernel, &cfg) != cudaSuccess) {
            (void)cudaGetLastError();
            n = 0;
        }
        g_small_clusters[device] = n;
    }
    (void)cudaSetDevice(prev);
    if (e != cudaSuccess) return static_cast<int>(e);
    g_sms[device] = prop.multiProcessorCount;
    if (const char* env = getenv("LG_MATCH_FFN_SMALL")) g_small_enabled.store(atoi(env) != 0 ? 1 : 0, std::memory_order_relaxed);
    if (const char* env = getenv("LG_ATTN_PDL")) g_pdl.store(atoi(env) != 0 ? 1 : 0, std::memory_order_relaxed);
    if (const char* env = getenv("LG_MATCH_FFN_PAIR")) g_pair.store(atoi(env) != 0 ? 1 : 0, std::memory_order_relaxed);
    return 0;
}

// 2-D row-major FP16 matrix [rows, cols] with row stride `ld` elements, as a 4-D map (cols, rows, 1, 1) so that the kernels'
// tma_load_4d / tma_store_4d wrappers serve it; rows past `rows` read as zeros / are not written.
int encode_matrix(CUtensorMap* map, const void* base, long long rows, int cols, long long ld, int box_cols, int box_rows,
                  CUtensorMapSwizzle swizzle) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows), 1, 1};
    const cuuint64_t srow = static_cast<cuuint64_t>(ld) * 2;
    const cuuint64_t strides[3] = {srow, srow, srow};   // extent-1 dimensions never advance; any legal stride
    const cuuint32_t box[4] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows), 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = g_encode.load(std::memory_order_acquire)(
        map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
        swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : LG_ATTN_EDRIVER;
}

bool use_small(int device, long long tiles) {
    return device >= 0 && device < kMaxDevices && g_ready[device].load(std::memory_order_acquire) &&
           g_small_enabled.load(std::memory_order_relaxed) && tiles <= g_small_clusters[device];
}

}  // namespace

extern "C" {

int lg_match_ffn_init_device(int device) {
    if (device < 0 || device >= kMaxDevices) return LG_ATTN_EINVAL;
    if (g_ready[device].load(std::memory_order_acquire)) return 0;
    std::lock_guard<std::mutex> lock(g_mutex);
    if (g_ready[device].load(std::memory_order_relaxed)) return 0;
    const int rc = init_locked(device);
    if (rc == 0) g_ready[device].store(1, std::memory_order_release);
    return rc;
}

int lg_match_debug_set_ffn_pair(int enabled) {
    g_pair.store(enabled < 0 ? -1 : (enabled != 0 ? 1 : 0), std::memory_order_relaxed);
    return 0;
}

int lg_match_ffn1_prefer_fused(long long rows) {
    // Tiles are 128 rows x all 512 columns (a row's LayerNorm never leaves the CTA), one tile per CTA at a time: small row counts
    // leave most SMs idle (a single 1024-keypoint pair is 16 tiles) and the library GEMM + lg_match_layernorm_gelu pair, which
    // spreads the same work over every SM, finishes first.  Measured crossover on B200 (tools/pairs_sweep.py): between 64 and 128 tiles.
    int device = 0, sms = 148;
    if (cudaGetDevice(&device) == cudaSuccess && device >= 0 && device < kMaxDevices && g_ready[device].load(std::memory_order_acquire))
        sms = g_sms[device];
    const long long tiles = (rows + kBM - 1) / kBM;
    if (tiles * 3 >= static_cast<long long>(sms) * 2) return 1;
    // very few rows (a single pair): the 4-CTA-cluster kernel, one tile per cluster, when every cluster is resident at once
    return use_small(device, tiles) ? 1 : 0;
}

int lg_match_ffn1_ln_gelu(const void* x_half, long long rows, int ldx, const void* w_nk_half, const void* bias_half,
                          const float* gamma, const float* beta, float eps, void* hid_half, int ldh, void* stream) {
    if (x_half == nullptr || w_nk_half == nullptr || bias_half == nullptr || gamma == nullptr || beta == nullptr ||
        hid_half == nullptr || rows <= 0 || rows > (1LL << 31) - kBM)
        return LG_ATTN_EINVAL;
    if (ldx < kK || ldh < kN || (ldx & 7) != 0 || (ldh & 7) != 0) return LG_ATTN_EINVAL;
    if (((reinterpret_cast<uintptr_t>(x_half) | reinterpret_cast<uintptr_t>(w_nk_half) | reinterpret_cast<uintptr_t>(hid_half)) & 15u) != 0)
        return LG_ATTN_EINVAL;
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_match_ffn_init_device(device);
    if (rc != 0) return rc;

    const int tiles = static_cast<int>((rows + kBM - 1) / kBM);
    const bool small_shape = use_small(device, tiles);
    CUtensorMap tm_x, tm_w, tm_h;
    if ((rc = encode_matrix(&tm_x, x_half, rows, kK, ldx, kBK, kBM, CU_TENSOR_MAP_SWIZZLE_128B)) != 0) return rc;
    if ((rc = encode_matrix(&tm_w, w_nk_half, kN, kK, kK, kBK, small_shape ? small::kNC : 256, CU_TENSOR_MAP_SWIZZLE_128B)) != 0) return rc;
    if ((rc = encode_matrix(&tm_h, hid_half, rows, kN, ldh, kRound, 32, CU_TENSOR_MAP_SWIZZLE_64B)) != 0) return rc;
    if (small_shape) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(static_cast<unsigned>(tiles) * small::kSplit);
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = small::kSmemBytesS;
        cfg.stream = static_cast<cudaStream_t>(stream);
        cudaLaunchAttribute attrs[2];
        int n_attrs = 0;
        attrs[n_attrs].id = cudaLaunchAttributeClusterDimension;
        attrs[n_attrs].val.clusterDim.x = small::kSplit;
        attrs[n_attrs].val.clusterDim.y = 1;
        attrs[n_attrs].val.clusterDim.z = 1;
        ++n_attrs;
        if (g_pdl.load(std::memory_order_relaxed)) {
            // the prologue (barriers, TMEM, parameter loads) may overlap the previous kernel's tail; the kernel waits at
            // griddepcontrol.wait before it touches an activation
            attrs[n_attrs].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attrs[n_attrs].val.programmaticStreamSerializationAllowed = 1;
            ++n_attrs;
        }
        cfg.attrs = attrs;
        cfg.numAttrs = static_cast<unsigned>(n_attrs);
        e = cudaLaunchKernelEx(&cfg, small::ffn1_ln_gelu_small_kernel, tm_x, tm_w, tm_h, static_cast<const __half*>(bias_half), gamma,
                               beta, eps);
        return static_cast<int>(e);
    }
    // CTA pairs win once the launch is many waves long (262 144 rows: 154 vs 160 us, 65 536: 47.7 vs 49.8; their double-buffered output staging takes the
    // TMA-store wait out of every round) and lose a little on short launches (32 768 rows: 29.3 vs 27.3 us)
    const int pair_mode = g_pair.load(std::memory_order_relaxed);
    if (pair_mode > 0 || (pair_mode < 0 && tiles >= 3 * g_sms[device])) {
        // CTA pairs (cta_group::2): W1' tiles of 128 rows per CTA, clusters of 2, a pair per 256-row super-tile
        if ((rc = encode_matrix(&tm_w, w_nk_half, kN, kK, kK, kBK, 128, CU_TENSOR_MAP_SWIZZLE_128B)) != 0) return rc;
        const int pairs_max = g_sms[device] / 2, super_tiles = (tiles + 1) / 2;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2u * static_cast<unsigned>(super_tiles < pairs_max ? super_tiles : pairs_max));
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = kSmemBytes;
        cfg.stream = static_cast<cudaStream_t>(stream);
        cudaLaunchAttribute attrs[1];
        attrs[0].id = cudaLaunchAttributeClusterDimension;
        attrs[0].val.clusterDim.x = 2;
        attrs[0].val.clusterDim.y = 1;
        attrs[0].val.clusterDim.z = 1;
        cfg.attrs = attrs;
        cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, ffn1_ln_gelu_kernel<true>, tm_x, tm_w, tm_h, static_cast<const __half*>(bias_half), gamma, beta,
                               eps, tiles);
        return static_cast<int>(e);
    }
    const int grid = tiles < g_sms[device] ? tiles : g_sms[device];
    ffn1_ln_gelu_kernel<false><<<grid, kThreads, kSmemBytes, static_cast<cudaStream_t>(stream)>>>(
        tm_x, tm_w, tm_h, static_cast<const __half*>(bias_half), gamma, beta, eps, tiles);
    e = cudaGetLastError();
    return static_cast<int>(e);
}

int lg_match_proj_prefer_fused(long long rows, int n_blocks, int n_rot_blocks) {
    // Items are (128-row tile, 256-column block), one per CTA at a time.  Measured inside the matcher's CUDA graph (B200,
    // tools/pairs_sweep.py, LG_MATCH_FUSED_PROJ=0/1): at 32 pairs x 1024 / 2048 keypoints the kernel gains 2 % of a forward (the
    // rotary pass is gone, the GEMM itself runs at the library's speed: both sit at the ~4.3 TB/s a write-heavy stream reaches); for
    // a single pair (16 - 32 row tiles) the library GEMM + lg_match_rotary pair is faster (0.478 vs 0.498 ms per 1024-keypoint pair):
    // same rule as the ffn kernel, row tiles >= two thirds of the SM count.
    (void)n_blocks;
    (void)n_rot_blocks;
    int device = 0, sms = 148;
    if (cudaGetDevice(&device) == cudaSuccess && device >= 0 && device < kMaxDevices && g_ready[device].load(std::memory_order_acquire))
        sms = g_sms[device];
    if (const char* env = getenv("LG_MATCH_FUSED_PROJ")) {
        const int v = atoi(env);
        if (v == 0) return 0;
        if (v == 2) return 1;   // always (A/B runs)
    }
    const long long tiles = (rows + lgproj::kBM - 1) / lgproj::kBM;
    return tiles * 3 >= static_cast<long long>(sms) * 2 ? 1 : 0;
}

int lg_match_proj_rotary(const void* x_half, long long rows, int ldx, const void* w_nk_half, const void* bias_half, int n_blocks,
                         int n_rot_blocks, const float* cs, void* out_half, int ldo, void* stream) {
    if (x_half == nullptr || w_nk_half == nullptr || bias_half == nullptr || out_half == nullptr || rows <= 0 ||
        rows > (1LL << 31) - lgproj::kBM || n_blocks < 1 || n_blocks > lgproj::kMaxBlocks || n_rot_blocks < 0 || n_rot_blocks > n_blocks ||
        (n_rot_blocks > 0 && cs == nullptr))
        return LG_ATTN_EINVAL;
    const int n_cols = n_blocks * 256;
    if (ldx < lgproj::kK || ldo < n_cols || (ldx & 7) != 0 || (ldo & 7) != 0) return LG_ATTN_EINVAL;
    if (((reinterpret_cast<uintptr_t>(x_half) | reinterpret_cast<uintptr_t>(w_nk_half) | reinterpret_cast<uintptr_t>(out_half) |
          reinterpret_cast<uintptr_t>(cs)) & 15u) != 0)
        return LG_ATTN_EINVAL;
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_match_ffn_init_device(device);
    if (rc != 0) return rc;
    // few rows (a single 1024-keypoint pair: 16 row tiles): items of 128 columns while they still fit one wave, so that more SMs
    // work and an item's four K slabs are all in flight.  (Inside the matcher's graph this is as fast as library GEMM +
    // lg_match_rotary, 0.483 vs 0.476 ms per pair, not faster: lg_match_proj_prefer_fused keeps the library pair there.)
    const long long row_tiles = (rows + lgproj::kBM - 1) / lgproj::kBM;
    const bool narrow = row_tiles * n_blocks * 2 <= g_sms[device];
    const int bn = narrow ? 128 : 256;
    CUtensorMap tm_x, tm_w, tm_o;
    if ((rc = encode_matrix(&tm_x, x_half, rows, lgproj::kK, ldx, lgproj::kBK, lgproj::kBM, CU_TENSOR_MAP_SWIZZLE_128B)) != 0) return rc;
    if ((rc = encode_matrix(&tm_w, w_nk_half, n_cols, lgproj::kK, lgproj::kK, lgproj::kBK, bn, CU_TENSOR_MAP_SWIZZLE_128B)) != 0) return rc;
    if ((rc = encode_matrix(&tm_o, out_half, rows, n_cols, ldo, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) != 0) return rc;
    const long long items = row_tiles * (n_cols / bn);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(items < g_sms[device] ? items : g_sms[device]));
    cfg.blockDim = dim3(lgproj::kThreads);
    cfg.dynamicSmemBytes = narrow ? lgproj::Cfg<128>::kSmemBytes : lgproj::Cfg<256>::kSmemBytes;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = g_pdl.load(std::memory_order_relaxed) ? 1u : 0u;
    const int early_trigger = 1;   // the next kernel (attention, a programmatic dependent launch) may set up under this one
    if (narrow)
        e = cudaLaunchKernelEx(&cfg, lgproj::proj_rotary_kernel<128>, tm_x, tm_w, tm_o, static_cast<const __half*>(bias_half),
                               reinterpret_cast<const float4*>(cs), rows, n_blocks, n_rot_blocks, static_cast<int>(items), early_trigger);
    else
        e = cudaLaunchKernelEx(&cfg, lgproj::proj_rotary_kernel<256>, tm_x, tm_w, tm_o, static_cast<const __half*>(bias_half),
                               reinterpret_cast<const float4*>(cs), rows, n_blocks, n_rot_blocks, static_cast<int>(items), early_trigger);
    return static_cast<int>(e);
}

}  // extern "C"
