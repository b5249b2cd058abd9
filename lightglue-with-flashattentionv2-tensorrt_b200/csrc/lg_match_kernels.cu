// Memory-bound glue either side of the attention hot path (C-ABI in include/lg_match.h).
//
// HBM-bound elementwise / reduction work: no tensor cores, 128-bit coalesced accesses, grids sized
// in multiples of the SM count (grid-stride where the row count is large), FP32 arithmetic on FP16
// storage.  Each kernel replaces a run of eager / TensorRT elementwise launches of the reference
// (file:line under /root/reference/lightglue_pytorch_with_plugin/):
//   posenc_kernel           lightglue.py:77-86    Wr(x) -> cos / sin
//   rotary_kernel           lightglue.py:143-144, 154-165   rotate_half + cached rotary on q, k
//   layernorm_gelu_kernel   lightglue.py:131-136 (ffn[1], ffn[2])
//   assignment kernels      lightglue.py:216-224 (sigmoid_log_double_softmax), :253-276 (filter_matches)

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <atomic>
#include <cstdlib>
#include <mutex>

#include "../../include/lg_attn.h"
#include "../../include/lg_match.h"
#include "glue_math.cuh"

namespace lgmatch {

// SM count of the current device (queried once per device; the glue kernels size their grid-stride grids from it)
static int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}
constexpr int kFfnCols = 512;   // 2 * embed_dim (lightglue.py:131)
constexpr int kAngles = 32;     // head_dim / 2

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Programmatic dependent launch (the single-pair forward is a chain of ~100 short kernels, launch-to-launch latency is a large
// part of it): every glue kernel is launched with the programmatic-stream-serialization attribute (launch_glue below), so it may
// be scheduled while its predecessor is still running; it must therefore wait for the predecessor's completion before touching
// memory, and it lets its own successor be scheduled at once.  Completion is transitive along the chain: the predecessor did the
// same wait before its first access.  Without the attribute (LG_ATTN_PDL=0) both instructions are no-ops.
#define LG_GLUE_PDL_ENTRY()                                        \
    do {                                                           \
        asm volatile("griddepcontrol.wait;" ::: "memory");         \
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); \
    } while (0)

// ------------------------------------------------------------------------------------------ posenc
__global__ void __launch_bounds__(256) posenc_kernel(const float2* __restrict__ kpts, const float2* __restrict__ wr,
                                                     float2* __restrict__ cs, long long total) {
    LG_GLUE_PDL_ENTRY();
    for (long long t = blockIdx.x * 256LL + threadIdx.x; t < total; t += gridDim.x * 256LL) {
        const long long r = t >> 5;
        const int i = static_cast<int>(t & 31);
        const float2 p = kpts[r];
        const float2 w = wr[i];
        const float a = fmaf(p.y, w.y, p.x * w.x);
        float s, c;
        sincosf(a, &s, &c);
        cs[t] = make_float2(c, s);
    }
}

// ------------------------------------------------------------------------------------------ rotary
// One thread = 8 halves (4 rotary pairs) of one head of one block of one row.
__global__ void __launch_bounds__(256) rotary_kernel(__half* __restrict__ buf, long long rows, int ld, int chunks_per_row,
                                                     const float4* __restrict__ cs) {
    // (the attention launch that follows is a programmatic dependent launch too: its CTAs set up — barriers, TMEM, tensor-map
    // prefetch — while this kernel runs, and wait for this grid to complete before they read q / k)
    LG_GLUE_PDL_ENTRY();
    const long long total = rows * chunks_per_row;
    for (long long t = blockIdx.x * 256LL + threadIdx.x; t < total; t += gridDim.x * 256LL) {
        const long long r = t / chunks_per_row;
        const int c = static_cast<int>(t - r * chunks_per_row);  // 16-byte chunk inside the row's rotary span
        const int a4 = c & 7;                                    // which 4 angles (chunk inside the 64-wide head)
        uint4* p = reinterpret_cast<uint4*>(buf + r * ld) + c;
        uint4 v = *p;
        const float4 cs0 = cs[(r * kAngles + a4 * 4) / 2];      // (cos0, sin0, cos1, sin1)
        const float4 cs1 = cs[(r * kAngles + a4 * 4) / 2 + 1];  // (cos2, sin2, cos3, sin3)
        auto rot = [](uint32_t w, float co, float si) {
            const float2 ab = __half22float2(*reinterpret_cast<const __half2*>(&w));
            const __half2 o = __floats2half2_rn(fmaf(ab.x, co, -ab.y * si), fmaf(ab.y, co, ab.x * si));
            return *reinterpret_cast<const uint32_t*>(&o);
        };
        v.x = rot(v.x, cs0.x, cs0.y);
        v.y = rot(v.y, cs0.z, cs0.w);
        v.z = rot(v.z, cs1.x, cs1.y);
        v.w = rot(v.w, cs1.z, cs1.w);
        *p = v;
    }
}

// ------------------------------------------------------------------------------------------ LN + GELU
// erf-form GELU: gelu_erf() in glue_math.cuh (shared with the fused ffn kernel).
// One warp per row of 512: lane owns columns [8*lane, 8*lane+8) and [256 + 8*lane, ...).
__global__ void __launch_bounds__(256) layernorm_gelu_kernel(__half* __restrict__ h, long long rows, int ld,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps) {
    LG_GLUE_PDL_ENTRY();
    const int lane = threadIdx.x & 31;
    const long long warp0 = blockIdx.x * 8LL + (threadIdx.x >> 5);
    float g[16], b[16];   // gamma / 2, beta / 2: the affine step yields y / 2, which gelu_erf_of_half takes (two multiplies per element fewer)
#pragma unroll
    for (int half_i = 0; half_i < 2; ++half_i) {
        const float4* gp = reinterpret_cast<const float4*>(gamma + half_i * 256 + lane * 8);
        const float4* bp = reinterpret_cast<const float4*>(beta + half_i * 256 + lane * 8);
        const float4 g0 = gp[0], g1 = gp[1], b0 = bp[0], b1 = bp[1];
        g[8 * half_i + 0] = g0.x; g[8 * half_i + 1] = g0.y; g[8 * half_i + 2] = g0.z; g[8 * half_i + 3] = g0.w;
        g[8 * half_i + 4] = g1.x; g[8 * half_i + 5] = g1.y; g[8 * half_i + 6] = g1.z; g[8 * half_i + 7] = g1.w;
        b[8 * half_i + 0] = b0.x; b[8 * half_i + 1] = b0.y; b[8 * half_i + 2] = b0.z; b[8 * half_i + 3] = b0.w;
        b[8 * half_i + 4] = b1.x; b[8 * half_i + 5] = b1.y; b[8 * half_i + 6] = b1.z; b[8 * half_i + 7] = b1.w;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        g[i] *= 0.5f;
        b[i] *= 0.5f;
    }
    // Two rows per warp iteration: both rows' loads are issued before any arithmetic.  Measured (B200, 131072 rows): 4.2 TB/s of
    // read+write, the same as one row per iteration, as a register-double-buffered prefetch of the next two rows (4.0) and as round
    // 1's five-coefficient erf: the kernel is neither op-count- nor latency-bound in any simple sense — ncu shows 358 warp
    // instructions per row at 63-69 % issue utilisation (XU 51 %, FMA 54 %, 21 of 64 warps resident), 3.2 TB/s of actual DRAM
    // traffic (part of the in-place writes stays in L2).
    const long long stride = gridDim.x * 8LL;
    for (long long r = warp0; r < rows; r += 2 * stride) {
        const long long r1 = r + stride;
        const bool two = r1 < rows;
        uint4* p0 = reinterpret_cast<uint4*>(h + r * ld);
        uint4* p1 = reinterpret_cast<uint4*>(h + (two ? r1 : r) * ld);
        uint4 raw[2][2] = {{p0[lane], p0[32 + lane]}, {p1[lane], p1[32 + lane]}};
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            float x[16];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const __half2* hp = reinterpret_cast<const __half2*>(&raw[w][i]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = __half22float2(hp[e]);
                    x[8 * i + 2 * e] = f.x;
                    x[8 * i + 2 * e + 1] = f.y;
                }
            }
            // mean and variance in one pass over the registers (two independent shuffle reductions in flight): FP32 sums of 512
            // FP16 values; the variance is taken around lane 0's first element (shifted data) so that a large common offset does
            // not cancel
            const float shift = __shfl_sync(0xffffffffu, x[0], 0);
            float s = 0.f, q = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float d = x[i] - shift;
                s += d;
                q = fmaf(d, d, q);
            }
            s = warp_sum(s);
            q = warp_sum(q);
            const float dm = s * (1.0f / kFfnCols);                       // mean - shift
            const float rstd = rsqrtf(fmaxf(fmaf(-dm, dm, q * (1.0f / kFfnCols)), 0.f) + eps);
            const float nmr = -(shift + dm) * rstd;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                __half2* hp = reinterpret_cast<__half2*>(&raw[w][i]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int k = 8 * i + 2 * e;
                    const float y0 = gelu_erf_of_half(fmaf(fmaf(x[k], rstd, nmr), g[k], b[k]));
                    const float y1 = gelu_erf_of_half(fmaf(fmaf(x[k + 1], rstd, nmr), g[k + 1], b[k + 1]));
                    hp[e] = __floats2half2_rn(y0, y1);
                }
            }
        }
        p0[lane] = raw[0][0];
        p0[32 + lane] = raw[0][1];
        if (two) {
            p1[lane] = raw[1][0];
            p1[32 + lane] = raw[1][1];
        }
    }
}

// ------------------------------------------------------------------------------------------ assignment
// Four streaming passes over sim [B, n0, n1] FP32 (row lse, column lse, row best, column best), each a
// pure read at L2 / HBM speed: 128-bit loads when n1 % 4 == 0 (VEC = 4), rows unrolled for memory-level
// parallelism, column passes split over row blocks (partials combined by a tiny second kernel) so the
// grid fills every SM at any batch size.
__device__ __forceinline__ float log_sigmoid(float z) { return fminf(z, 0.f) - log1pf(expf(-fabsf(z))); }

// (max, sum exp(x - max)) online merge
__device__ __forceinline__ void lse_merge(float& m, float& s, float m2, float s2) {
    const float mn = fmaxf(m, m2);
    if (mn == -INFINITY) { m = mn; s = 0.f; return; }
    s = s * __expf(m - mn) + s2 * __expf(m2 - mn);
    m = mn;
}
__device__ __forceinline__ void lse_push(float& m, float& s, float x) {
    // x == -inf (a masked / overflowed similarity) adds nothing; with m still -inf, exp(x - m) would be exp(NaN) and poison the column
    if (x > m) { s = s * __expf(m - x) + 1.f; m = x; } else if (x != -INFINITY) { s += __expf(x - m); }
}

template <int VEC>
__device__ __forceinline__ void load_vec(const float* p, float (&v)[VEC]) {
    if constexpr (VEC == 4) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
        v[0] = *p;
    }
}

constexpr int kColRowsPerCta = 256;  // rows of sim one column-pass CTA walks (8 row groups x 32 rows)

// Warp per row.  MODE 0: row_term[r] = logsigmoid(z0[r]) - logsumexp_j sim[r, j]
//                MODE 1: m0[r] = argmax_j (2 sim[r, j] + col_term[j]), score0[r] = that max + row_term[r]
template <int MODE, int VEC>
__global__ void __launch_bounds__(256) row_pass_kernel(const float* __restrict__ sim, int n0, int n1, long long rows,
                                                       const __half* __restrict__ z0, long long z_stride,
                                                       const float* __restrict__ col_term, float* __restrict__ row_term,
                                                       int* __restrict__ m0, float* __restrict__ score0) {
    LG_GLUE_PDL_ENTRY();
    const int lane = threadIdx.x & 31;
    for (long long r = blockIdx.x * 8LL + (threadIdx.x >> 5); r < rows; r += gridDim.x * 8LL) {
        const float* p = sim + r * n1;
        const float* ct = MODE == 1 ? col_term + (r / n0) * n1 : nullptr;
        float m = -INFINITY, s = 0.f;
        int bj = 0x7fffffff;
#pragma unroll 4
        for (int j = lane * VEC; j < n1; j += 32 * VEC) {
            float x[VEC];
            load_vec<VEC>(p + j, x);
            if (MODE == 0) {
                float mx = x[0];
#pragma unroll
                for (int e = 1; e < VEC; ++e) mx = fmaxf(mx, x[e]);
                if (mx > m) { s *= __expf(m - mx); m = mx; }
#pragma unroll
                for (int e = 0; e < VEC; ++e) s += __expf(x[e] - m);
            } else {
                float c[VEC];
                load_vec<VEC>(ct + j, c);
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const float v = fmaf(2.f, x[e], c[e]);
                    if (v > m) { m = v; bj = j + e; }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
            if (MODE == 0) {
                const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
                lse_merge(m, s, m2, s2);
            } else {
                const int j2 = __shfl_xor_sync(0xffffffffu, bj, o);
                if (m2 > m || (m2 == m && j2 < bj)) { m = m2; bj = j2; }
            }
        }
        if (lane == 0) {
            if (MODE == 0) {
                row_term[r] = log_sigmoid(__half2float(z0[r * z_stride])) - (m + logf(s));
            } else {
                m0[r] = bj;
                score0[r] = m + row_term[r];
            }
        }
    }
}

// Column pass, partial over one block of kColRowsPerCta rows: block = 32 column lanes (VEC columns each)
// x 8 row groups; grid (ceil(n1 / (32 VEC)), row blocks, batch).
// MODE 0: (max, sum exp) of sim[:, j] over the row block           -> part_a, part_b [B, S, n1]
// MODE 1: best (value, row) of 2 sim[i, j] + row_term[i] over it   -> part_a, part_b (row as int bits)
template <int MODE, int VEC>
__global__ void __launch_bounds__(256) col_partial_kernel(const float* __restrict__ sim, int n0, int n1,
                                                          const float* __restrict__ row_term,
                                                          float* __restrict__ part_a, float* __restrict__ part_b) {
    LG_GLUE_PDL_ENTRY();
    __shared__ float sh_a[8][32 * VEC + 1];
    __shared__ float sh_b[8][32 * VEC + 1];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int b = blockIdx.z, sp = blockIdx.y;
    const int j0 = (blockIdx.x * 32 + cx) * VEC;
    const int r_begin = sp * kColRowsPerCta;
    const int r_end = min(n0, r_begin + kColRowsPerCta);
    const float* p = sim + static_cast<long long>(b) * n0 * n1 + j0;
    const float* rt = row_term + static_cast<long long>(b) * n0;
    float m[VEC], s[VEC];
    int best[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) { m[e] = -INFINITY; s[e] = 0.f; best[e] = 0x7fffffff; }
    if (j0 < n1) {
#pragma unroll 4
        for (int i = r_begin + ry; i < r_end; i += 8) {
            float x[VEC];
            load_vec<VEC>(p + static_cast<long long>(i) * n1, x);
            if (MODE == 0) {
#pragma unroll
                for (int e = 0; e < VEC; ++e) lse_push(m[e], s[e], x[e]);
            } else {
                const float t = rt[i];
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const float v = fmaf(2.f, x[e], t);
                    if (v > m[e]) { m[e] = v; best[e] = i; }
                }
            }
        }
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        sh_a[ry][cx * VEC + e] = m[e];
        sh_b[ry][cx * VEC + e] = MODE == 0 ? s[e] : __int_as_float(best[e]);
    }
    __syncthreads();
    // 32 * VEC columns, one per thread of the first VEC warps
    const int col = threadIdx.x;
    const int j = blockIdx.x * 32 * VEC + col;
    if (col < 32 * VEC && j < n1) {
        float ma = sh_a[0][col], sb = sh_b[0][col];
        int bi = __float_as_int(sb);
        for (int g = 1; g < 8; ++g) {
            if (MODE == 0) {
                lse_merge(ma, sb, sh_a[g][col], sh_b[g][col]);
            } else {
                const float v = sh_a[g][col];
                const int i2 = __float_as_int(sh_b[g][col]);
                if (v > ma || (v == ma && i2 < bi)) { ma = v; bi = i2; }
            }
        }
        const long long o = (static_cast<long long>(b) * gridDim.y + sp) * n1 + j;
        part_a[o] = ma;
        part_b[o] = MODE == 0 ? sb : __int_as_float(bi);
    }
}

// Combine the row-block partials of a column.  MODE 0: col_term[b, j] = logsigmoid(z1) - logsumexp_i sim[i, j];
// MODE 1: m1[b, j] = best row.
template <int MODE>
__global__ void __launch_bounds__(256) col_combine_kernel(const float* __restrict__ part_a, const float* __restrict__ part_b,
                                                          int n_parts, int n1, long long cols,
                                                          const __half* __restrict__ z1, long long z_stride,
                                                          float* __restrict__ col_term, int* __restrict__ m1) {
    LG_GLUE_PDL_ENTRY();
    for (long long t = blockIdx.x * 256LL + threadIdx.x; t < cols; t += gridDim.x * 256LL) {
        const long long b = t / n1;
        const long long base = b * n_parts * n1 + (t - b * n1);
        float ma = part_a[base], sb = part_b[base];
        int bi = __float_as_int(sb);
        for (int g = 1; g < n_parts; ++g) {
            const float a2 = part_a[base + static_cast<long long>(g) * n1], b2 = part_b[base + static_cast<long long>(g) * n1];
            if (MODE == 0) {
                lse_merge(ma, sb, a2, b2);
            } else {
                const int i2 = __float_as_int(b2);
                if (a2 > ma || (a2 == ma && i2 < bi)) { ma = a2; bi = i2; }
            }
        }
        if (MODE == 0) col_term[t] = log_sigmoid(__half2float(z1[t * z_stride])) - (ma + logf(sb));
        else m1[t] = bi;
    }
}

// filter_matches (lightglue.py:253-276): mutual check, exp, threshold.
__global__ void __launch_bounds__(256) mutual_kernel(const int* __restrict__ m0, const int* __restrict__ m1,
                                                     const float* __restrict__ score0, int batch, int n0, int n1,
                                                     float threshold, int* __restrict__ matches0,
                                                     float* __restrict__ mscores0) {
    LG_GLUE_PDL_ENTRY();
    const long long rows = static_cast<long long>(batch) * n0;
    for (long long r = blockIdx.x * 256LL + threadIdx.x; r < rows; r += gridDim.x * 256LL) {
        const long long b = r / n0;
        const int i = static_cast<int>(r - b * n0);
        // a row of NaN / -inf similarities (FP16 overflow upstream) has no argmax: its index stays at the INT_MAX sentinel,
        // which must neither be dereferenced nor reported as a match
        const int j = m0[r];
        const bool valid = static_cast<unsigned>(j) < static_cast<unsigned>(n1);
        const bool mutual = valid && m1[b * n1 + j] == i;
        const float ms = mutual ? expf(score0[r]) : 0.f;
        mscores0[r] = ms;
        matches0[r] = ms > threshold ? j : -1;
    }
}

// scores[b,i,j] = 2 sim - row_lse_i - col_lse_j + lsz0_i + lsz1_j  (lightglue.py:216-224)
__global__ void __launch_bounds__(256) scores_kernel(const float4* __restrict__ sim, long long rows, int n0, int n1,
                                                     const float* __restrict__ row_term,
                                                     const float* __restrict__ col_term, float4* __restrict__ scores) {
    LG_GLUE_PDL_ENTRY();
    const int q = n1 >> 2;  // n1 % 4 == 0 on this path
    const long long total = rows * q;
    for (long long t = blockIdx.x * 256LL + threadIdx.x; t < total; t += gridDim.x * 256LL) {
        const long long r = t / q;
        const int j4 = static_cast<int>(t - r * q);
        const long long b = r / n0;
        const float4 x = sim[t];
        const float4 c = reinterpret_cast<const float4*>(col_term + b * n1)[j4];
        const float rt = row_term[r];
        scores[t] = make_float4(2.f * x.x + rt + c.x, 2.f * x.y + rt + c.y, 2.f * x.z + rt + c.z, 2.f * x.w + rt + c.w);
    }
}
__global__ void __launch_bounds__(256) scores_kernel_scalar(const float* __restrict__ sim, long long rows, int n0, int n1,
                                                            const float* __restrict__ row_term,
                                                            const float* __restrict__ col_term,
                                                            float* __restrict__ scores) {
    LG_GLUE_PDL_ENTRY();
    const long long total = rows * n1;
    for (long long t = blockIdx.x * 256LL + threadIdx.x; t < total; t += gridDim.x * 256LL) {
        const long long r = t / n1;
        const int j = static_cast<int>(t - r * n1);
        const long long b = r / n0;
        scores[t] = 2.f * sim[t] + row_term[r] + col_term[b * n1 + j];
    }
}

unsigned grid_for(long long work_items, int per_block, int waves = 8) {
    long long blocks = (work_items + per_block - 1) / per_block;
    const long long cap = static_cast<long long>(sm_count()) * waves;  // grid-stride beyond `waves` CTAs per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<unsigned>(blocks);
}

size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// LG_ATTN_PDL=0 turns programmatic dependent launch off for the glue kernels as well (same switch as the attention launcher)
bool glue_pdl() {
    static const bool on = [] {
        const char* env = getenv("LG_ATTN_PDL");
        return env == nullptr || atoi(env) != 0;
    }();
    return on;
}

// <<<grid, 256, 0, stream>>> with the programmatic-stream-serialization attribute (see LG_GLUE_PDL_ENTRY)
template <typename... KArgs, typename... Args>
void launch_glue(void (*kernel)(KArgs...), dim3 grid, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = glue_pdl() ? 1u : 0u;
    (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);   // errors surface through last_error()
}

int last_error() { return static_cast<int>(cudaGetLastError()); }

}  // namespace lgmatch

using namespace lgmatch;

extern "C" {

int lg_match_posenc(const float* kpts, const float* wr, float* cs, long long rows, void* stream) {
    if (kpts == nullptr || wr == nullptr || cs == nullptr || rows <= 0) return LG_ATTN_EINVAL;
    const long long total = rows * kAngles;
    launch_glue(posenc_kernel, dim3(grid_for(total, 256)), static_cast<cudaStream_t>(stream), 
        reinterpret_cast<const float2*>(kpts), reinterpret_cast<const float2*>(wr), reinterpret_cast<float2*>(cs), total);
    return last_error();
}

int lg_match_rotary(void* packed_half, long long rows, int ld, int n_blocks, const float* cs, void* stream) {
    if (packed_half == nullptr || cs == nullptr || rows <= 0 || n_blocks <= 0) return LG_ATTN_EINVAL;
    if (ld < n_blocks * 256 || (ld & 7) != 0 || (reinterpret_cast<uintptr_t>(packed_half) & 15u) != 0) return LG_ATTN_EINVAL;
    const int chunks = n_blocks * 32;  // 16-byte chunks per row that carry rotary data
    launch_glue(rotary_kernel, dim3(grid_for(rows * chunks, 256)), static_cast<cudaStream_t>(stream), 
        static_cast<__half*>(packed_half), rows, ld, chunks, reinterpret_cast<const float4*>(cs));
    return last_error();
}

int lg_match_layernorm_gelu(void* h_half, long long rows, int ld, const float* gamma, const float* beta, float eps,
                            void* stream) {
    if (h_half == nullptr || gamma == nullptr || beta == nullptr || rows <= 0) return LG_ATTN_EINVAL;
    if (ld < kFfnCols || (ld & 7) != 0 || (reinterpret_cast<uintptr_t>(h_half) & 15u) != 0) return LG_ATTN_EINVAL;
    launch_glue(layernorm_gelu_kernel, dim3(grid_for(rows, 8)), static_cast<cudaStream_t>(stream), 
        static_cast<__half*>(h_half), rows, ld, gamma, beta, eps);
    return last_error();
}

size_t lg_match_assign_workspace_bytes(int batch, int n0, int n1) {
    if (batch <= 0 || n0 <= 0 || n1 <= 0) return 0;
    const size_t r0 = align256(static_cast<size_t>(batch) * n0 * 4), r1 = align256(static_cast<size_t>(batch) * n1 * 4);
    const size_t parts = (static_cast<size_t>(n0) + kColRowsPerCta - 1) / kColRowsPerCta;
    // row_term, m0, score0 | col_term, m1 | two partial planes [batch, row blocks, n1]
    return 3 * r0 + 2 * r1 + 2 * align256(static_cast<size_t>(batch) * parts * n1 * 4);
}

int lg_match_assign_launches(int with_scores) { return with_scores ? 8 : 7; }

// ---- lg_match_assign_static: the same assignment with its scratch in a file-scope device array (no allocation anywhere) --------
// Every GPU that loads the library holds these LG_MATCH_ASSIGN_STATIC_MAX_BYTES until exit.  In host code `g_assign_scratch` is a
// host-side stand-in for the symbol, NOT a device address: the kernels get the address cudaGetSymbolAddress returns, resolved once
// per device (with that device current) and cached.
static __device__ __align__(256) unsigned char g_assign_scratch[LG_MATCH_ASSIGN_STATIC_MAX_BYTES];

namespace {
constexpr int kMaxScratchDevices = 64;
void* g_assign_scratch_ptr[kMaxScratchDevices];                 // device address of g_assign_scratch on device i
std::atomic<int> g_assign_scratch_state[kMaxScratchDevices];    // 0 = unresolved, 1 = resolved (release / acquire publishes the pointer)
std::mutex g_assign_scratch_mutex;

// Device address of the scratch array on the CURRENT device.  First call per device: cudaGetSymbolAddress (a lookup, not an
// allocation: the array is part of the loaded module's image).
int assign_scratch_address(void** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (dev < 0 || dev >= kMaxScratchDevices) return LG_ATTN_EINVAL;
    if (g_assign_scratch_state[dev].load(std::memory_order_acquire) == 0) {
        std::lock_guard<std::mutex> lock(g_assign_scratch_mutex);
        if (g_assign_scratch_state[dev].load(std::memory_order_relaxed) == 0) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_assign_scratch);
            if (e != cudaSuccess) return static_cast<int>(e);
            if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) return LG_ATTN_EINVAL;
            g_assign_scratch_ptr[dev] = p;
            g_assign_scratch_state[dev].store(1, std::memory_order_release);
        }
    }
    *out = g_assign_scratch_ptr[dev];
    return 0;
}
}  // namespace

int lg_match_assign_static(const float* sim, const void* z0_half, const void* z1_half, long long z_stride, int batch, int n0,
                           int n1, float threshold, int* matches0, float* mscores0, float* scores, void* stream) {
    if (batch <= 0 || n0 <= 0 || n1 <= 0) return LG_ATTN_EINVAL;
    // the size check comes first and needs no device: shapes beyond the fixed array are refused, never truncated
    const size_t need = lg_match_assign_workspace_bytes(batch, n0, n1);
    if (need > static_cast<size_t>(LG_MATCH_ASSIGN_STATIC_MAX_BYTES)) return LG_MATCH_ESTATIC_TOO_LARGE;
    void* scratch = nullptr;
    const int rc = assign_scratch_address(&scratch);
    if (rc != 0) return rc;
    // Nothing in the scratch has to start at zero: within one call every word (row / column terms, arg-max indices, the column
    // partial planes) is written by an earlier kernel of the call before a later one reads it, and no kernel accumulates into it —
    // so the previous call's contents are simply overwritten.  Two calls that overlap on the GPU (different streams) race on it.
    return lg_match_assign(sim, z0_half, z1_half, z_stride, batch, n0, n1, threshold, matches0, mscores0, scores, scratch, need, stream);
}

int lg_match_assign(const float* sim, const void* z0_half, const void* z1_half, long long z_stride, int batch, int n0,
                    int n1, float threshold, int* matches0, float* mscores0, float* scores, void* workspace,
                    size_t workspace_bytes, void* stream_v) {
    if (sim == nullptr || z0_half == nullptr || z1_half == nullptr || matches0 == nullptr || mscores0 == nullptr)
        return LG_ATTN_EINVAL;
    if (batch <= 0 || n0 <= 0 || n1 <= 0 || z_stride <= 0 || batch > 65535) return LG_ATTN_EINVAL;
    if (workspace == nullptr || workspace_bytes < lg_match_assign_workspace_bytes(batch, n0, n1)) return LG_ATTN_EWORKSPACE;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    const long long rows0 = static_cast<long long>(batch) * n0, rows1 = static_cast<long long>(batch) * n1;
    const size_t r0 = align256(static_cast<size_t>(rows0) * 4), r1 = align256(static_cast<size_t>(rows1) * 4);
    const int parts = (n0 + kColRowsPerCta - 1) / kColRowsPerCta;
    if (parts > 65535) return LG_ATTN_EINVAL;
    const size_t pp = align256(static_cast<size_t>(batch) * parts * n1 * 4);
    uint8_t* ws = static_cast<uint8_t*>(workspace);
    float* row_term = reinterpret_cast<float*>(ws);
    int* m0 = reinterpret_cast<int*>(ws + r0);
    float* score0 = reinterpret_cast<float*>(ws + 2 * r0);
    float* col_term = reinterpret_cast<float*>(ws + 3 * r0);
    int* m1 = reinterpret_cast<int*>(ws + 3 * r0 + r1);
    float* part_a = reinterpret_cast<float*>(ws + 3 * r0 + 2 * r1);
    float* part_b = reinterpret_cast<float*>(ws + 3 * r0 + 2 * r1 + pp);
    const __half* z0 = static_cast<const __half*>(z0_half);
    const __half* z1 = static_cast<const __half*>(z1_half);
    const bool vec = (n1 & 3) == 0 && (reinterpret_cast<uintptr_t>(sim) & 15u) == 0;
    const int cols_per_cta = vec ? 128 : 32;
    const dim3 col_grid(static_cast<unsigned>((n1 + cols_per_cta - 1) / cols_per_cta), static_cast<unsigned>(parts),
                        static_cast<unsigned>(batch));
    const unsigned row_grid = grid_for(rows0, 8, 16), cmb_grid = grid_for(rows1, 256);

    if (vec) {
        launch_glue(row_pass_kernel<0, 4>, dim3(row_grid), stream, sim, n0, n1, rows0, z0, z_stride, nullptr, row_term, nullptr, nullptr);
        launch_glue(col_partial_kernel<0, 4>, dim3(col_grid), stream, sim, n0, n1, row_term, part_a, part_b);
    } else {
        launch_glue(row_pass_kernel<0, 1>, dim3(row_grid), stream, sim, n0, n1, rows0, z0, z_stride, nullptr, row_term, nullptr, nullptr);
        launch_glue(col_partial_kernel<0, 1>, dim3(col_grid), stream, sim, n0, n1, row_term, part_a, part_b);
    }
    launch_glue(col_combine_kernel<0>, dim3(cmb_grid), stream, part_a, part_b, parts, n1, rows1, z1, z_stride, col_term, nullptr);
    if (vec) {
        launch_glue(row_pass_kernel<1, 4>, dim3(row_grid), stream, sim, n0, n1, rows0, z0, z_stride, col_term, row_term, m0, score0);
        launch_glue(col_partial_kernel<1, 4>, dim3(col_grid), stream, sim, n0, n1, row_term, part_a, part_b);
    } else {
        launch_glue(row_pass_kernel<1, 1>, dim3(row_grid), stream, sim, n0, n1, rows0, z0, z_stride, col_term, row_term, m0, score0);
        launch_glue(col_partial_kernel<1, 1>, dim3(col_grid), stream, sim, n0, n1, row_term, part_a, part_b);
    }
    launch_glue(col_combine_kernel<1>, dim3(cmb_grid), stream, part_a, part_b, parts, n1, rows1, z1, z_stride, nullptr, m1);
    launch_glue(mutual_kernel, dim3(grid_for(rows0, 256)), stream, m0, m1, score0, batch, n0, n1, threshold, matches0, mscores0);
    if (scores != nullptr) {
        if (vec && (reinterpret_cast<uintptr_t>(scores) & 15u) == 0)
            launch_glue(scores_kernel, dim3(grid_for(rows0 * (n1 >> 2), 256)), stream, 
                reinterpret_cast<const float4*>(sim), rows0, n0, n1, row_term, col_term, reinterpret_cast<float4*>(scores));
        else
            launch_glue(scores_kernel_scalar, dim3(grid_for(rows0 * n1, 256)), stream, sim, rows0, n0, n1, row_term, col_term, scores);
    }
    return last_error();
}

}  // extern "C"
