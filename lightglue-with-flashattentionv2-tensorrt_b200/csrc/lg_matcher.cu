// Native pair-matcher engine (C-ABI in include/lg_matcher.h): the role of the reference's compiled host
// demo/lightglue_trt.cpp (LightGlueTRT: engine + buffers + captured CUDA graph + PostProcess), written
// directly against this repo's sm_100a kernels and cuBLASLt.  Host-side C++ only — every device
// function it launches lives in lg_attn_api.cu / lg_match_kernels.cu (or in cuBLASLt for plain GEMMs).
//
// Forward = LightGlue.forward (/root/reference/lightglue_pytorch_with_plugin/lightglue.py:333-359) +
// filter_matches (:253-276) on the packed layout of DESIGN.md §8:
//   XC   [R, 512]  FP16   cols 0..255 residual stream x | cols 256..511 attention output (= cat(x, ctx))
//   PROJ [R, 768]  FP16   q|k|v (self) or qk|v (cross) projection, q/k rotated in place
//   HID  [R, 512]  FP16   ffn hidden, LayerNorm+GELU in place
//   MD   [R, 264]  FP16   final projection / 4th-root scale | matchability logit
//   SIM  [B, N0, N1] FP32
// with R = B*(N0+N1), image-0 rows first.

#include <cublasLt.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <utility>
#include <vector>

#include "../../include/lg_attn.h"
#include "../../include/lg_match.h"
#include "../../include/lg_matcher.h"

namespace {

constexpr int kDim = 256, kHeads = 4, kHeadDim = 64, kFfn = 512, kMdCols = 264;
constexpr size_t kLtWorkspace = 32u << 20;

size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// ---- weight blob: header + arrays, each padded to 256 bytes, in this order -------------------------------
struct BlobHeader {
    char magic[8];  // "LGMATCH1"
    int32_t n_layers;
    int32_t reserved;
};

struct BlockW {           // device pointers into the blob copy
    const __half* w_in_t; // [256, n_in]   n_in = 768 (self: q|k|v) or 512 (cross: qk|v)
    const __half* b_in;   // [n_in]
    const __half* w1_t;   // [512, 512]
    const __half* b1;     // [512]
    const float* gamma;   // [512]
    const float* beta;    // [512]
    const __half* w2_t;   // [512, 256]
    int n_in;
    const __half* w1_nk;  // [512 out, 512 in]: w1_t transposed at create time (outside the blob) — the fused ffn kernel's B operand
    const __half* w_in_nk; // [n_in out, 256 in]: w_in_t transposed likewise — the fused projection kernel's B operand
};

size_t block_bytes(int n_in) {
    return align256(static_cast<size_t>(kDim) * n_in * 2) + align256(static_cast<size_t>(n_in) * 2) +
           align256(static_cast<size_t>(kFfn) * kFfn * 2) + align256(kFfn * 2) + 2 * align256(kFfn * 4) +
           align256(static_cast<size_t>(kFfn) * kDim * 2);
}

size_t blob_bytes(int n_layers) {
    return align256(sizeof(BlobHeader)) + align256(32 * 2 * 4) +
           static_cast<size_t>(n_layers) * (block_bytes(3 * kDim) + block_bytes(2 * kDim)) +
           align256(static_cast<size_t>(kDim) * kMdCols * 2) + align256(kMdCols * 2) + align256(kDim * 2);
}

// ---- one cuBLASLt GEMM: C[rows, n] (row-major FP16 or FP32) = A[rows, k] . Wt[k, n] (+ bias | + C) --------
// Row-major operands are handed to cuBLASLt as their column-major transposes: C^T[n, rows] = Wt^T[n, k] . A^T[k, rows].
struct Gemm {
    cublasLtMatmulDesc_t op = nullptr;
    cublasLtMatrixLayout_t a = nullptr, b = nullptr, c = nullptr;
    cublasLtMatmulAlgo_t algo{};
    bool have_algo = false;
    float beta = 0.f;

    void destroy() {
        if (a) cublasLtMatrixLayoutDestroy(a);
        if (b) cublasLtMatrixLayoutDestroy(b);
        if (c) cublasLtMatrixLayoutDestroy(c);
        if (op) cublasLtMatmulDescDestroy(op);
        a = b = c = nullptr;
        op = nullptr;
    }
};

#define LT_CHECK(x)                                   \
    do {                                              \
        if ((x) != CUBLAS_STATUS_SUCCESS) return LG_MATCHER_ECUBLAS; \
    } while (0)

// rows x k (lda) times k x n (ldw) -> rows x n (ldc).  bias != null: epilogue bias (beta = 0); accumulate: beta = 1.
int make_gemm(cublasLtHandle_t lt, Gemm& g, long long rows, int n, int k, long long lda, long long ldw, long long ldc,
              bool with_bias, bool accumulate, size_t ws_bytes) {
    LT_CHECK(cublasLtMatmulDescCreate(&g.op, CUBLAS_COMPUTE_32F, CUDA_R_32F));
    const cublasOperation_t nt = CUBLAS_OP_N;
    LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_TRANSA, &nt, sizeof(nt)));
    LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_TRANSB, &nt, sizeof(nt)));
    if (with_bias) {
        const cublasLtEpilogue_t epi = CUBLASLT_EPILOGUE_BIAS;
        LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_EPILOGUE, &epi, sizeof(epi)));
        const cudaDataType_t bt = CUDA_R_16F;
        LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_BIAS_DATA_TYPE, &bt, sizeof(bt)));
    }
    // column-major views: "A" = Wt^T [n, k] ld ldw ; "B" = A^T [k, rows] ld lda ; "C" = C^T [n, rows] ld ldc
    LT_CHECK(cublasLtMatrixLayoutCreate(&g.a, CUDA_R_16F, n, k, ldw));
    LT_CHECK(cublasLtMatrixLayoutCreate(&g.b, CUDA_R_16F, k, rows, lda));
    LT_CHECK(cublasLtMatrixLayoutCreate(&g.c, CUDA_R_16F, n, rows, ldc));
    g.beta = accumulate ? 1.f : 0.f;
    cublasLtMatmulPreference_t pref = nullptr;
    LT_CHECK(cublasLtMatmulPreferenceCreate(&pref));
    cublasLtMatmulPreferenceSetAttribute(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &ws_bytes, sizeof(ws_bytes));
    cublasLtMatmulHeuristicResult_t res{};
    int found = 0;
    const cublasStatus_t st = cublasLtMatmulAlgoGetHeuristic(lt, g.op, g.a, g.b, g.c, g.c, pref, 1, &res, &found);
    cublasLtMatmulPreferenceDestroy(pref);
    if (st != CUBLAS_STATUS_SUCCESS || found == 0) return LG_MATCHER_ECUBLAS;
    g.algo = res.algo;
    g.have_algo = true;
    return 0;
}

int run_gemm(cublasLtHandle_t lt, const Gemm& g, const void* act, const void* wt, void* c, const void* bias, void* ws,
             size_t ws_bytes, cudaStream_t stream) {
    if (bias != nullptr) {
        // the descriptor is shared by every layer's GEMM of this shape: the bias pointer is set per call (host-side)
        if (cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_BIAS_POINTER, &bias, sizeof(bias)) != CUBLAS_STATUS_SUCCESS)
            return LG_MATCHER_ECUBLAS;
    }
    const float alpha = 1.f;
    LT_CHECK(cublasLtMatmul(lt, g.op, &alpha, wt, g.a, act, g.b, &g.beta, c, g.c, c, g.c, &g.algo, ws, ws_bytes, stream));
    return 0;
}

}  // namespace

// Everything that depends on the (pairs, n0, n1) shape: device buffers, cuBLASLt plans, the captured graph.
struct ShapePlan {
    int pairs = 0, n0 = 0, n1 = 0;
    long long rows = 0;
    uint8_t* pool = nullptr;
    float *kpts = nullptr, *cs = nullptr, *sim = nullptr, *mscores = nullptr;
    __half *xc = nullptr, *proj = nullptr, *hid = nullptr, *md = nullptr;
    int* matches = nullptr;
    void *assign_ws = nullptr, *lt_ws = nullptr;
    size_t assign_ws_bytes = 0;
    Gemm g_qkv, g_cross_in, g_ffn1, g_ffn2, g_final, g_sim;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int launches = 0;
    // asynchronous host-buffer path: device staging slots (slot layout: kpts0 | kpts1 | desc0 | desc1)
    uint8_t* stage[2] = {nullptr, nullptr};
    unsigned long long last_use = 0;
};

// The reference engine is dynamic (1..2048 keypoints per image, a different count for every pair — steps.txt:152 profiles);
// a shape change here costs cudaMalloc + a warm-up forward + graph capture, so the engine keeps the active plan plus a few
// parked ones (least recently used is evicted) instead of re-planning every time two shapes alternate.
constexpr size_t kMaxParkedPlans = 3;

struct lg_matcher : ShapePlan {   // the inherited members are the ACTIVE plan
    int device = 0, n_layers = 0;
    float threshold = 0.1f;
    cublasLtHandle_t lt = nullptr;
    uint8_t* blob = nullptr;  // device copy of the weight blob
    __half* w1_nk = nullptr;  // [2 * n_layers][512][512]: every block's W1' in nn.Linear layout (see BlockW::w1_nk)
    bool fused_ffn = true;    // LG_MATCH_FUSED_FFN=0: library GEMM + lg_match_layernorm_gelu instead (A/B runs)
    __half* w_in_nk = nullptr; // [n_layers][(768 + 512)][256]: every block's input projection in nn.Linear layout
    bool fused_proj = true;   // LG_MATCH_FUSED_PROJ=0: library GEMM (+ lg_match_rotary) instead of lg_match_proj_rotary
    const float* wr = nullptr;
    std::vector<BlockW> self_w, cross_w;
    const __half *wf_t = nullptr, *bf = nullptr;
    cudaStream_t stream = nullptr;  // capture / host-buffer stream
    // asynchronous host-buffer path: a copy stream fills one of two staging slots while the compute stream runs the previous
    // sub-batch
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_free[2] = {nullptr, nullptr};
    unsigned long long submits = 0;
    std::vector<ShapePlan> parked;
    unsigned long long tick = 0;
    ShapePlan& active() { return *this; }
};

namespace {

void free_plan(ShapePlan* m) {
    if (m->exec) cudaGraphExecDestroy(m->exec);
    if (m->graph) cudaGraphDestroy(m->graph);
    m->exec = nullptr;
    m->graph = nullptr;
    for (Gemm* g : {&m->g_qkv, &m->g_cross_in, &m->g_ffn1, &m->g_ffn2, &m->g_final, &m->g_sim}) g->destroy();
    if (m->pool) cudaFree(m->pool);
    m->pool = nullptr;
    for (int i = 0; i < 2; ++i) {
        if (m->stage[i]) cudaFree(m->stage[i]);
        m->stage[i] = nullptr;
    }
    m->pairs = m->n0 = m->n1 = 0;
    m->launches = 0;
}

lg_attn_strides heads_strides(long long rows_per_batch, long long ld) { return lg_attn_strides{rows_per_batch * ld, kHeadDim, ld}; }

// One forward on `st` (eager or under capture).  Returns 0 or an error code; counts launches.
int forward(lg_matcher* m, cudaStream_t st, int* launches_out) {
    const int B = m->pairs, n0 = m->n0, n1 = m->n1;
    const long long R = m->rows, r0 = static_cast<long long>(B) * n0;
    int rc, launches = 0;
    __half* x = m->xc;  // [R, 512], cols 0..255
    if ((rc = lg_match_posenc(m->kpts, m->wr, m->cs, R, st)) != 0) return rc;
    ++launches;

    auto attention = [&](const __half* q, const __half* k, const __half* v, long long ld_in, __half* o, int batch, int nq, int nkv,
                         int shift) {
        const lg_attn_strides sq = heads_strides(nq, ld_in), sk = heads_strides(nkv, ld_in), so = heads_strides(nq, kFfn);
        ++launches;
        return lg_attn_hd64_launch_strided(q, k, v, o, batch, kHeads, nq, nkv, 0, &sq, &sk, &sk, &so, shift, st);
    };
    auto ffn = [&](const BlockW& w) {
        int e;
        if (m->fused_ffn && lg_match_ffn1_prefer_fused(R)) {
            // cat(x, ctx) @ W1'^T + b1', LayerNorm, GELU in one tcgen05 kernel (LayerNorm + GELU in the GEMM's epilogue)
            e = lg_match_ffn1_ln_gelu(m->xc, R, kFfn, w.w1_nk, w.b1, w.gamma, w.beta, 1e-5f, m->hid, kFfn, st);
            launches += 2;
        } else {
            e = run_gemm(m->lt, m->g_ffn1, m->xc, w.w1_t, m->hid, w.b1, m->lt_ws, kLtWorkspace, st);
            if (e == 0) e = lg_match_layernorm_gelu(m->hid, R, kFfn, w.gamma, w.beta, 1e-5f, st);
            launches += 3;
        }
        if (e == 0) e = run_gemm(m->lt, m->g_ffn2, m->hid, w.w2_t, x, nullptr, m->lt_ws, kLtWorkspace, st);  // x += hid @ W2^T
        return e;
    };

    for (int l = 0; l < m->n_layers; ++l) {
        // ---- self attention on both images (lightglue.py:137-151)
        const BlockW& sw = m->self_w[l];
        if (m->fused_proj && lg_match_proj_prefer_fused(R, 3, 2)) {
            // Wqkv with the rotary of q, k in the GEMM epilogue: one launch, no second pass over q | k
            if ((rc = lg_match_proj_rotary(x, R, kFfn, sw.w_in_nk, sw.b_in, 3, 2, m->cs, m->proj, 3 * kDim, st)) != 0) return rc;
            ++launches;
        } else {
            if ((rc = run_gemm(m->lt, m->g_qkv, x, sw.w_in_t, m->proj, sw.b_in, m->lt_ws, kLtWorkspace, st)) != 0) return rc;
            if ((rc = lg_match_rotary(m->proj, R, 3 * kDim, 2, m->cs, st)) != 0) return rc;
            launches += 2;
        }
        const long long ldp = 3 * kDim;
        if (n0 == n1) {
            rc = attention(m->proj, m->proj + kDim, m->proj + 2 * kDim, ldp, m->xc + kDim, 2 * B, n0, n0, 0);
        } else {
            rc = attention(m->proj, m->proj + kDim, m->proj + 2 * kDim, ldp, m->xc + kDim, B, n0, n0, 0);
            if (rc == 0) {
                const __half* p1 = m->proj + r0 * ldp;
                rc = attention(p1, p1 + kDim, p1 + 2 * kDim, ldp, m->xc + r0 * kFfn + kDim, B, n1, n1, 0);
            }
        }
        if (rc != 0 || (rc = ffn(sw)) != 0) return rc;
        // ---- cross attention, both directions (lightglue.py:188-206); PROJ reused as [R, 512]: qk | v
        const BlockW& cw = m->cross_w[l];
        if (m->fused_proj && lg_match_proj_prefer_fused(R, 2, 0))
            rc = lg_match_proj_rotary(x, R, kFfn, cw.w_in_nk, cw.b_in, 2, 0, nullptr, m->proj, 2 * kDim, st);
        else
            rc = run_gemm(m->lt, m->g_cross_in, x, cw.w_in_t, m->proj, cw.b_in, m->lt_ws, kLtWorkspace, st);
        if (rc != 0) return rc;
        ++launches;
        const long long ldc = 2 * kDim;
        if (n0 == n1) {
            rc = attention(m->proj, m->proj, m->proj + kDim, ldc, m->xc + kDim, 2 * B, n0, n0, B);
        } else {
            const __half* p1 = m->proj + r0 * ldc;
            rc = attention(m->proj, p1, p1 + kDim, ldc, m->xc + kDim, B, n0, n1, 0);
            if (rc == 0) rc = attention(p1, m->proj, m->proj + kDim, ldc, m->xc + r0 * kFfn + kDim, B, n1, n0, 0);
        }
        if (rc != 0 || (rc = ffn(cw)) != 0) return rc;
    }
    // ---- assignment (lightglue.py:236-246) + filter (:253-276)
    if ((rc = run_gemm(m->lt, m->g_final, x, m->wf_t, m->md, m->bf, m->lt_ws, kLtWorkspace, st)) != 0) return rc;
    {
        // sim[b] = md0[b] . md1[b]^T  — column-major: sim^T[n1, n0] = (md1^T)^T[n1, 256] . md0^T[256, n0]
        const float alpha = 1.f, beta = 0.f;
        if (cublasLtMatmul(m->lt, m->g_sim.op, &alpha, m->md + r0 * kMdCols, m->g_sim.a, m->md, m->g_sim.b, &beta, m->sim,
                           m->g_sim.c, m->sim, m->g_sim.c, &m->g_sim.algo, m->lt_ws, kLtWorkspace, st) != CUBLAS_STATUS_SUCCESS)
            return LG_MATCHER_ECUBLAS;
    }
    rc = lg_match_assign(m->sim, m->md + kDim, m->md + r0 * kMdCols + kDim, kMdCols, B, n0, n1, m->threshold, m->matches,
                         m->mscores, nullptr, m->assign_ws, m->assign_ws_bytes, st);
    launches += 2 + lg_match_assign_launches(0);
    if (launches_out) *launches_out = launches;
    return rc;
}

}  // namespace

namespace {
// Descriptor rows [rows,256] FP16 (contiguous) -> left half of XC (row pitch 1024 B).  A pitched copy-engine
// transfer of 512 B rows runs far below HBM speed; 16 B per thread, one row per warp, is bandwidth-bound.
__global__ void __launch_bounds__(256) place_desc_kernel(const uint4* __restrict__ src, uint4* __restrict__ xc, long long n_vec) {
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n_vec; i += static_cast<long long>(gridDim.x) * 256)
        xc[(i >> 5) * 64 + (i & 31)] = src[i];
}

cudaError_t place_desc(const void* src, __half* xc, long long rows, cudaStream_t st) {
    const long long n_vec = rows * 32;
    if (n_vec == 0) return cudaSuccess;
    const long long blocks = (n_vec + 255) / 256;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    const long long cap = 16LL * sms;   // grid-stride beyond 16 CTAs per SM
    place_desc_kernel<<<static_cast<unsigned>(blocks < cap ? blocks : cap), 256, 0, st>>>(
        static_cast<const uint4*>(src), reinterpret_cast<uint4*>(xc), n_vec);
    return cudaGetLastError();
}
}  // namespace

extern "C" {

size_t lg_matcher_blob_bytes(int n_layers) { return n_layers > 0 ? blob_bytes(n_layers) : 0; }

int lg_matcher_create(lg_matcher** out, int device, int n_layers, const void* weights_host, size_t weights_bytes,
                      float threshold) {
    if (out == nullptr || weights_host == nullptr || n_layers <= 0 || n_layers > 64) return LG_ATTN_EINVAL;
    *out = nullptr;
    if (weights_bytes != blob_bytes(n_layers)) return LG_MATCHER_EBLOB;
    BlobHeader hdr;
    memcpy(&hdr, weights_host, sizeof(hdr));
    if (memcmp(hdr.magic, "LGMATCH1", 8) != 0 || hdr.n_layers != n_layers) return LG_MATCHER_EBLOB;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_cast<int>(e);
    int rc = lg_attn_hd64_init_device(device);
    if (rc == 0) rc = lg_match_ffn_init_device(device);
    if (rc != 0) return rc;
    lg_matcher* m = new (std::nothrow) lg_matcher();
    if (m == nullptr) return static_cast<int>(cudaErrorMemoryAllocation);
    m->device = device;
    m->n_layers = n_layers;
    m->threshold = threshold;
    if (cublasLtCreate(&m->lt) != CUBLAS_STATUS_SUCCESS) { delete m; return LG_MATCHER_ECUBLAS; }
    e = cudaMalloc(&m->blob, weights_bytes);
    if (e == cudaSuccess) e = cudaMemcpy(m->blob, weights_host, weights_bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&m->ev_in[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_free[i], cudaEventDisableTiming);
    }
    if (const char* env = getenv("LG_MATCH_FUSED_FFN")) m->fused_ffn = atoi(env) != 0;
    if (const char* env = getenv("LG_MATCH_FUSED_PROJ")) m->fused_proj = atoi(env) != 0;
    const size_t w1_elems = static_cast<size_t>(kFfn) * kFfn;
    if (e == cudaSuccess) e = cudaMalloc(&m->w1_nk, 2 * static_cast<size_t>(n_layers) * w1_elems * 2);
    if (e == cudaSuccess) e = cudaMalloc(&m->w_in_nk, static_cast<size_t>(n_layers) * 5 * kDim * kDim * 2);
    if (e != cudaSuccess) { lg_matcher_destroy(m); return static_cast<int>(e); }
    // carve the device copy
    uint8_t* p = m->blob + align256(sizeof(BlobHeader));
    auto take = [&](size_t bytes) { uint8_t* q = p; p += align256(bytes); return q; };
    std::vector<uint16_t> w1_host(w1_elems);
    int n_blocks = 0;
    size_t w_in_used = 0;   // elements of m->w_in_nk handed out
    m->wr = reinterpret_cast<const float*>(take(32 * 2 * 4));
    auto take_block = [&](int n_in) {
        BlockW w{};
        w.n_in = n_in;
        w.w_in_t = reinterpret_cast<const __half*>(take(static_cast<size_t>(kDim) * n_in * 2));
        w.b_in = reinterpret_cast<const __half*>(take(static_cast<size_t>(n_in) * 2));
        w.w1_t = reinterpret_cast<const __half*>(take(static_cast<size_t>(kFfn) * kFfn * 2));
        w.b1 = reinterpret_cast<const __half*>(take(kFfn * 2));
        w.gamma = reinterpret_cast<const float*>(take(kFfn * 4));
        w.beta = reinterpret_cast<const float*>(take(kFfn * 4));
        w.w2_t = reinterpret_cast<const __half*>(take(static_cast<size_t>(kFfn) * kDim * 2));
        // W1' the other way round ([out, in], rows = output features), transposed on the host from the caller's blob
        const uint16_t* src = reinterpret_cast<const uint16_t*>(static_cast<const uint8_t*>(weights_host) +
                                                                (reinterpret_cast<const uint8_t*>(w.w1_t) - m->blob));
        for (int k = 0; k < kFfn; ++k)
            for (int n = 0; n < kFfn; ++n) w1_host[static_cast<size_t>(n) * kFfn + k] = src[static_cast<size_t>(k) * kFfn + n];
        __half* dst = m->w1_nk + static_cast<size_t>(n_blocks++) * w1_elems;
        if (e == cudaSuccess) e = cudaMemcpy(dst, w1_host.data(), w1_elems * 2, cudaMemcpyHostToDevice);
        w.w1_nk = dst;
        const uint16_t* src_in = reinterpret_cast<const uint16_t*>(static_cast<const uint8_t*>(weights_host) +
                                                                   (reinterpret_cast<const uint8_t*>(w.w_in_t) - m->blob));
        for (int k = 0; k < kDim; ++k)
            for (int n = 0; n < n_in; ++n) w1_host[static_cast<size_t>(n) * kDim + k] = src_in[static_cast<size_t>(k) * n_in + n];
        __half* dst_in = m->w_in_nk + w_in_used;
        w_in_used += static_cast<size_t>(n_in) * kDim;
        if (e == cudaSuccess) e = cudaMemcpy(dst_in, w1_host.data(), static_cast<size_t>(n_in) * kDim * 2, cudaMemcpyHostToDevice);
        w.w_in_nk = dst_in;
        return w;
    };
    for (int l = 0; l < n_layers; ++l) {
        m->self_w.push_back(take_block(3 * kDim));
        m->cross_w.push_back(take_block(2 * kDim));
    }
    m->wf_t = reinterpret_cast<const __half*>(take(static_cast<size_t>(kDim) * kMdCols * 2));
    m->bf = reinterpret_cast<const __half*>(take(kMdCols * 2));
    if (e != cudaSuccess) { lg_matcher_destroy(m); return static_cast<int>(e); }
    *out = m;
    return 0;
}

static int build_plan(lg_matcher* m, int pairs, int n0, int n1) {
    cudaError_t e;
    const long long R = static_cast<long long>(pairs) * (n0 + n1);
    const size_t b_kpts = align256(R * 2 * 4), b_cs = align256(R * 64 * 4), b_xc = align256(R * kFfn * 2),
                 b_proj = align256(R * 3 * kDim * 2), b_hid = align256(R * kFfn * 2), b_md = align256(R * kMdCols * 2),
                 b_sim = align256(static_cast<size_t>(pairs) * n0 * n1 * 4), b_m = align256(static_cast<size_t>(pairs) * n0 * 4),
                 b_aws = align256(lg_match_assign_workspace_bytes(pairs, n0, n1));
    const size_t total = b_kpts + b_cs + b_xc + b_proj + b_hid + b_md + b_sim + 2 * b_m + b_aws + kLtWorkspace;
    e = cudaMalloc(&m->pool, total);
    if (e != cudaSuccess) { m->pool = nullptr; return static_cast<int>(e); }
    uint8_t* p = m->pool;
    auto take = [&](size_t bytes) { uint8_t* q = p; p += bytes; return q; };
    m->kpts = reinterpret_cast<float*>(take(b_kpts));
    m->cs = reinterpret_cast<float*>(take(b_cs));
    m->xc = reinterpret_cast<__half*>(take(b_xc));
    m->proj = reinterpret_cast<__half*>(take(b_proj));
    m->hid = reinterpret_cast<__half*>(take(b_hid));
    m->md = reinterpret_cast<__half*>(take(b_md));
    m->sim = reinterpret_cast<float*>(take(b_sim));
    m->matches = reinterpret_cast<int*>(take(b_m));
    m->mscores = reinterpret_cast<float*>(take(b_m));
    m->assign_ws = take(b_aws);
    m->assign_ws_bytes = b_aws;
    m->lt_ws = take(kLtWorkspace);
    m->pairs = pairs; m->n0 = n0; m->n1 = n1; m->rows = R;
    for (int i = 0; i < 2; ++i) {
        e = cudaMalloc(&m->stage[i], align256(R * 8) + static_cast<size_t>(R) * kDim * 2);
        if (e != cudaSuccess) { m->stage[i] = nullptr; return static_cast<int>(e); }
    }
    e = cudaMemsetAsync(m->pool, 0, total, m->stream);   // finite values everywhere (XC's right half is read by the first ffn)
    if (e != cudaSuccess) return static_cast<int>(e);

    // ---- GEMM plans (one per distinct shape; weights / bias pointers are per call)
    int rc;
    if ((rc = make_gemm(m->lt, m->g_qkv, R, 3 * kDim, kDim, kFfn, 3 * kDim, 3 * kDim, true, false, kLtWorkspace)) != 0) return rc;
    if ((rc = make_gemm(m->lt, m->g_cross_in, R, 2 * kDim, kDim, kFfn, 2 * kDim, 2 * kDim, true, false, kLtWorkspace)) != 0) return rc;
    if ((rc = make_gemm(m->lt, m->g_ffn1, R, kFfn, kFfn, kFfn, kFfn, kFfn, true, false, kLtWorkspace)) != 0) return rc;
    if ((rc = make_gemm(m->lt, m->g_ffn2, R, kDim, kFfn, kFfn, kDim, kFfn, false, true, kLtWorkspace)) != 0) return rc;
    if ((rc = make_gemm(m->lt, m->g_final, R, kMdCols, kDim, kFfn, kMdCols, kMdCols, true, false, kLtWorkspace)) != 0) return rc;
    {   // sim^T[n1, n0] (FP32, ld n1) = op_T(md1^T [256, n1] ld 264) . md0^T [256, n0] ld 264, strided over pairs
        Gemm& g = m->g_sim;
        LT_CHECK(cublasLtMatmulDescCreate(&g.op, CUBLAS_COMPUTE_32F, CUDA_R_32F));
        const cublasOperation_t tr = CUBLAS_OP_T, nt = CUBLAS_OP_N;
        LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_TRANSA, &tr, sizeof(tr)));
        LT_CHECK(cublasLtMatmulDescSetAttribute(g.op, CUBLASLT_MATMUL_DESC_TRANSB, &nt, sizeof(nt)));
        LT_CHECK(cublasLtMatrixLayoutCreate(&g.a, CUDA_R_16F, kDim, n1, kMdCols));
        LT_CHECK(cublasLtMatrixLayoutCreate(&g.b, CUDA_R_16F, kDim, n0, kMdCols));
        LT_CHECK(cublasLtMatrixLayoutCreate(&g.c, CUDA_R_32F, n1, n0, n1));
        const int32_t bc = pairs;
        const int64_t sa = static_cast<int64_t>(n1) * kMdCols, sb = static_cast<int64_t>(n0) * kMdCols, sc = static_cast<int64_t>(n0) * n1;
        for (auto pr : {std::make_pair(g.a, sa), std::make_pair(g.b, sb), std::make_pair(g.c, sc)}) {
            LT_CHECK(cublasLtMatrixLayoutSetAttribute(pr.first, CUBLASLT_MATRIX_LAYOUT_BATCH_COUNT, &bc, sizeof(bc)));
            LT_CHECK(cublasLtMatrixLayoutSetAttribute(pr.first, CUBLASLT_MATRIX_LAYOUT_STRIDED_BATCH_OFFSET, &pr.second, sizeof(pr.second)));
        }
        cublasLtMatmulPreference_t pref = nullptr;
        LT_CHECK(cublasLtMatmulPreferenceCreate(&pref));
        const size_t wsb = kLtWorkspace;
        cublasLtMatmulPreferenceSetAttribute(pref, CUBLASLT_MATMUL_PREF_MAX_WORKSPACE_BYTES, &wsb, sizeof(wsb));
        cublasLtMatmulHeuristicResult_t res{};
        int found = 0;
        const cublasStatus_t st = cublasLtMatmulAlgoGetHeuristic(m->lt, g.op, g.a, g.b, g.c, g.c, pref, 1, &res, &found);
        cublasLtMatmulPreferenceDestroy(pref);
        if (st != CUBLAS_STATUS_SUCCESS || found == 0) return LG_MATCHER_ECUBLAS;
        g.algo = res.algo;
        g.have_algo = true;
    }
    // ---- warm-up (eager) then capture
    if ((rc = forward(m, m->stream, nullptr)) != 0) return rc;
    e = cudaStreamSynchronize(m->stream);
    if (e != cudaSuccess) return static_cast<int>(e);
    e = cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeRelaxed);
    if (e != cudaSuccess) return static_cast<int>(e);
    rc = forward(m, m->stream, &m->launches);
    e = cudaStreamEndCapture(m->stream, &m->graph);
    if (rc != 0) return rc;
    if (e != cudaSuccess) return static_cast<int>(e);
    e = cudaGraphInstantiate(&m->exec, m->graph, 0);
    return static_cast<int>(e);
}

int lg_matcher_plan(lg_matcher* m, int pairs, int n0, int n1) {
    if (m == nullptr || pairs <= 0 || n0 <= 0 || n1 <= 0) return LG_ATTN_EINVAL;
    m->last_use = ++m->tick;
    if (m->exec != nullptr && m->pairs == pairs && m->n0 == n0 && m->n1 == n1) return 0;
    cudaError_t e = cudaSetDevice(m->device);
    if (e != cudaSuccess) return static_cast<int>(e);
    // a parked plan of this shape: swap it in.  Nothing is freed, so forwards still in flight on the old plan's buffers are safe.
    for (ShapePlan& q : m->parked) {
        if (q.exec != nullptr && q.pairs == pairs && q.n0 == n0 && q.n1 == n1) {
            std::swap(m->active(), q);
            m->last_use = m->tick;
            return 0;
        }
    }
    if (m->exec != nullptr) {   // park the active plan; evict the least recently used one if the shelf is full
        if (m->parked.size() >= kMaxParkedPlans) {
            size_t lru = 0;
            for (size_t i = 1; i < m->parked.size(); ++i)
                if (m->parked[i].last_use < m->parked[lru].last_use) lru = i;
            // the evicted plan's buffers may still be in use by a forward enqueued on a caller's stream
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) return static_cast<int>(e);
            free_plan(&m->parked[lru]);
            m->parked.erase(m->parked.begin() + static_cast<long>(lru));
        }
        m->parked.push_back(m->active());
        m->active() = ShapePlan{};
    } else {
        free_plan(m);   // a half-initialised active plan (previous build failed) holds nothing, but keep the invariant
    }
    m->last_use = m->tick;
    const int rc = build_plan(m, pairs, n0, n1);
    if (rc != 0) {          // never leave a half-built plan behind: the next call must re-plan, not launch it
        (void)cudaStreamSynchronize(m->stream);
        free_plan(m);
        (void)cudaGetLastError();   // a failed allocation must not surface later as a launch error
    }
    return rc;
}

int lg_matcher_plans_resident(const lg_matcher* m) {
    if (m == nullptr) return 0;
    return static_cast<int>(m->parked.size()) + (m->exec != nullptr ? 1 : 0);
}

int lg_matcher_launches(const lg_matcher* m) { return m ? m->launches : 0; }

static int enqueue(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0, const void* desc1, int pairs,
                   int n0, int n1, int* matches0, float* mscores0, cudaStream_t st, cudaMemcpyKind in_kind,
                   cudaMemcpyKind out_kind) {
    const long long r0 = static_cast<long long>(pairs) * n0, r1 = static_cast<long long>(pairs) * n1;
    cudaError_t e = cudaMemcpyAsync(m->kpts, kpts0, r0 * 8, in_kind, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->kpts + r0 * 2, kpts1, r1 * 8, in_kind, st);
    // descriptors land in the left half of XC (row pitch 1024 B); input_proj is the identity for 256-d descriptors
    const bool aligned16 = ((reinterpret_cast<uintptr_t>(desc0) | reinterpret_cast<uintptr_t>(desc1)) & 15) == 0;
    if (in_kind == cudaMemcpyDeviceToDevice && aligned16) {
        if (e == cudaSuccess) e = place_desc(desc0, m->xc, r0, st);
        if (e == cudaSuccess) e = place_desc(desc1, m->xc + r0 * kFfn, r1, st);
    } else {
        if (e == cudaSuccess) e = cudaMemcpy2DAsync(m->xc, kFfn * 2, desc0, kDim * 2, kDim * 2, r0, in_kind, st);
        if (e == cudaSuccess) e = cudaMemcpy2DAsync(m->xc + r0 * kFfn, kFfn * 2, desc1, kDim * 2, kDim * 2, r1, in_kind, st);
    }
    if (e == cudaSuccess) e = cudaGraphLaunch(m->exec, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(matches0, m->matches, r0 * 4, out_kind, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(mscores0, m->mscores, r0 * 4, out_kind, st);
    return static_cast<int>(e);
}

int lg_matcher_match(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half, const void* desc1_half,
                     int pairs, int n0, int n1, int* matches0, float* mscores0, void* stream) {
    if (m == nullptr || kpts0 == nullptr || kpts1 == nullptr || desc0_half == nullptr || desc1_half == nullptr ||
        matches0 == nullptr || mscores0 == nullptr)
        return LG_ATTN_EINVAL;
    int rc = lg_matcher_plan(m, pairs, n0, n1);
    if (rc != 0) return rc;
    return enqueue(m, kpts0, kpts1, desc0_half, desc1_half, pairs, n0, n1, matches0, mscores0, static_cast<cudaStream_t>(stream),
                   cudaMemcpyDeviceToDevice, cudaMemcpyDeviceToDevice);
}

int lg_matcher_match_host(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half,
                          const void* desc1_half, int pairs, int n0, int n1, int* matches0, float* mscores0) {
    if (m == nullptr || kpts0 == nullptr || kpts1 == nullptr || desc0_half == nullptr || desc1_half == nullptr ||
        matches0 == nullptr || mscores0 == nullptr)
        return LG_ATTN_EINVAL;
    int rc = lg_matcher_plan(m, pairs, n0, n1);
    if (rc != 0) return rc;
    rc = enqueue(m, kpts0, kpts1, desc0_half, desc1_half, pairs, n0, n1, matches0, mscores0, m->stream, cudaMemcpyHostToDevice,
                 cudaMemcpyDeviceToHost);
    if (rc != 0) return rc;
    return static_cast<int>(cudaStreamSynchronize(m->stream));
}

int lg_matcher_submit_host(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half,
                           const void* desc1_half, int pairs, int n0, int n1, int* matches0, float* mscores0) {
    if (m == nullptr || kpts0 == nullptr || kpts1 == nullptr || desc0_half == nullptr || desc1_half == nullptr ||
        matches0 == nullptr || mscores0 == nullptr)
        return LG_ATTN_EINVAL;
    if (m->exec == nullptr || m->pairs != pairs || m->n0 != n0 || m->n1 != n1) {
        bool parked = false;
        for (const ShapePlan& q : m->parked) parked = parked || (q.exec != nullptr && q.pairs == pairs && q.n0 == n0 && q.n1 == n1);
        int rc = parked ? 0 : lg_matcher_sync(m);   // building a plan drains the pipeline; swapping a parked one in does not
        if (rc == 0) rc = lg_matcher_plan(m, pairs, n0, n1);
        if (rc != 0) return rc;
    }
    cudaError_t e = cudaSetDevice(m->device);
    if (e != cudaSuccess) return static_cast<int>(e);
    const int slot = static_cast<int>(m->submits++ & 1);
    const long long r0 = static_cast<long long>(pairs) * n0, r1 = static_cast<long long>(pairs) * n1;
    uint8_t* sk = m->stage[slot];
    uint8_t* sd = sk + align256((r0 + r1) * 8);
    // copy stream: wait until the forward that last read this slot has consumed it, then stage the inputs
    e = cudaStreamWaitEvent(m->copy_stream, m->ev_free[slot], 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sk, kpts0, r0 * 8, cudaMemcpyHostToDevice, m->copy_stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sk + r0 * 8, kpts1, r1 * 8, cudaMemcpyHostToDevice, m->copy_stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sd, desc0_half, r0 * kDim * 2, cudaMemcpyHostToDevice, m->copy_stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(sd + r0 * kDim * 2, desc1_half, r1 * kDim * 2, cudaMemcpyHostToDevice, m->copy_stream);
    if (e == cudaSuccess) e = cudaEventRecord(m->ev_in[slot], m->copy_stream);
    // compute stream: staged inputs -> plan buffers (on-device), forward, results to the host
    if (e == cudaSuccess) e = cudaStreamWaitEvent(m->stream, m->ev_in[slot], 0);
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->kpts, sk, (r0 + r1) * 8, cudaMemcpyDeviceToDevice, m->stream);
    if (e == cudaSuccess) e = place_desc(sd, m->xc, r0 + r1, m->stream);
    if (e == cudaSuccess) e = cudaEventRecord(m->ev_free[slot], m->stream);
    if (e == cudaSuccess) e = cudaGraphLaunch(m->exec, m->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(matches0, m->matches, r0 * 4, cudaMemcpyDeviceToHost, m->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(mscores0, m->mscores, r0 * 4, cudaMemcpyDeviceToHost, m->stream);
    return static_cast<int>(e);
}

int lg_matcher_sync(lg_matcher* m) {
    if (m == nullptr) return LG_ATTN_EINVAL;
    cudaError_t e = cudaStreamSynchronize(m->copy_stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
    return static_cast<int>(e);
}

void lg_matcher_destroy(lg_matcher* m) {
    if (m == nullptr) return;
    (void)cudaSetDevice(m->device);
    (void)cudaDeviceSynchronize();   // parked plans may have forwards in flight on callers' streams
    free_plan(m);
    for (ShapePlan& q : m->parked) free_plan(&q);
    m->parked.clear();
    if (m->stream) (void)cudaStreamDestroy(m->stream);
    if (m->copy_stream) (void)cudaStreamDestroy(m->copy_stream);
    for (int i = 0; i < 2; ++i) {
        if (m->ev_in[i]) (void)cudaEventDestroy(m->ev_in[i]);
        if (m->ev_free[i]) (void)cudaEventDestroy(m->ev_free[i]);
    }
    if (m->blob) (void)cudaFree(m->blob);
    if (m->w1_nk) (void)cudaFree(m->w1_nk);
    if (m->w_in_nk) (void)cudaFree(m->w_in_nk);
    if (m->lt) cublasLtDestroy(m->lt);
    delete m;
}

}  // extern "C"
