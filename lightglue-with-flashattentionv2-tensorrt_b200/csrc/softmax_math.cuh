// Per-thread softmax arithmetic of the attention kernel: one S row slice in registers ->
// P = exp2(s*c - m*c) packed to FP16 pairs + the FP32 row sum.
//
// d = 64 attention does one exponential per 256 tensor FLOPs, so this file — not the MMAs — sets
// the kernel's speed.  Measured issue costs on B200 per SM sub-partition (tools/microbench/pipe_rates.cu):
//   MUFU.EX2 8 cycles | FFMA/FADD ~1-1.4 (heavy + lite pipes) | FFMA2/FADD2 ~3.5 (heavy pipe only)
//   FMNMX / FMNMX3 / LEA / IMAD 2 (16-lane pipes)
// so a fraction of the exponentials is moved from the MUFU pipe to the FMA pipes (Cody-Waite +
// degree-3 polynomial), and packed f32x2 forms are used only where they save issue slots without
// overloading the heavy pipe.  Compile-time knobs (defaults = best measured):
//   LG_ATTN_POLY_OF_8    how many of every 8 columns take the polynomial path
//   LG_ATTN_PACKED_SCALE 1: the s*c - m*c FFMA runs two columns per FFMA2 on the MUFU-path columns
//   LG_ATTN_PACKED_SUM   1: row sums accumulate with FADD2
//   LG_ATTN_PACKED_POLY  1: polynomial columns in pairs on FFMA2
#pragma once
#include "ptx_sm100.cuh"

// Build switches that produce WRONG results by design (timing decompositions only) must be asked for twice, so that a stray -D
// can never ship a kernel that fails parity: csrc/build.py's product build does not define LG_ATTN_TUNING_BUILD, and
// tests/test_boundary_cpu.py checks the shipped library was built without any of them.
#if (defined(LG_ATTN_WHATIF) || defined(LG_ATTN_EXPERIMENT_NOSUM) || defined(LG_ATTN_WHATIF_F3) || defined(LG_ATTN_WHATIF_NOSTORE)) && \
    !defined(LG_ATTN_TUNING_BUILD)
#error "LG_ATTN_WHATIF / LG_ATTN_WHATIF_F3 / LG_ATTN_EXPERIMENT_NOSUM give wrong results: tuning builds only (-DLG_ATTN_TUNING_BUILD)"
#endif

#ifndef LG_ATTN_POLY_OF_8
#define LG_ATTN_POLY_OF_8 2
#endif
#ifndef LG_ATTN_PACKED_SCALE
#define LG_ATTN_PACKED_SCALE 1
#endif
#ifndef LG_ATTN_PACKED_POLY
#define LG_ATTN_PACKED_POLY 0   // 1: polynomial columns are processed in pairs with FFMA2 (ex2_poly_pair)
#endif
#ifndef LG_ATTN_PACKED_SUM
#define LG_ATTN_PACKED_SUM 1
#endif

namespace lgattn {

using namespace ptx;

// softmax(x/8): exp(0.125*(s-m)) = exp2((s-m) * 0.125 * log2(e))
constexpr float kScaleLog2 = 0.125f * 1.4426950408889634f;
// lazy rescale: keep the stale row max while the new one is < 2^kTau larger (P stays < 2^8 in fp16)
#ifndef LG_ATTN_TAU
#define LG_ATTN_TAU 8.0f
#endif
constexpr float kTau = LG_ATTN_TAU;
// lazy row max (attn_sm100_kernel_mq.cuh): a step whose row sum reaches this is redone with the true row max.  2^(tau-1): the
// polynomial path saturates at 2^tau * 0.99993, so the trigger has to sit below that.
constexpr float kLazyTrigger = 128.0f;
static_assert(LG_ATTN_TAU == 8.0f, "kLazyTrigger assumes tau = 8");

// ---- polynomial exp2 on the FMA pipes -----------------------------------------------------------
// x = s*c - m*c lies in (-inf, kTau].  y = sat((x + 120) / 128) in [0, 1] comes out of ONE saturating
// FFMA (the clamp is free), x' = 128*y = x + 120 in [0, 128]; n = round(x') via the 1.5*2^23 magic
// add, f = x' - n in [-0.5, 0.5]; 2^f by a degree-3 minimax polynomial (relative error 7.5e-5, below
// the FP16 rounding of P) whose coefficients carry the 2^-120 bias; the integer part is added to
// the exponent field with one shift-add.  Columns masked to -inf give 2^-120 (0 in FP16).
constexpr float kPolyBias = 120.f;
constexpr float kPolyInvSpan = 1.f / 128.f;
constexpr float kMagic = 12582912.f;  // 1.5 * 2^23
constexpr float kTwoM120 = 7.52316384526264e-37f;  // 2^-120
constexpr float kC0 = 0.9999280572f * kTwoM120;
constexpr float kC1 = 0.6932609677f * kTwoM120;
constexpr float kC2 = 0.2426111251f * kTwoM120;
constexpr float kC3 = 0.0551716648f * kTwoM120;

struct PolyConsts {
    float scale;  // c / 128
    float bias;   // (120 - m*c) / 128
};

__device__ __forceinline__ PolyConsts make_poly_consts(float neg_m_scaled) {
    return {kScaleLog2 * kPolyInvSpan, (neg_m_scaled + kPolyBias) * kPolyInvSpan};
}

__device__ __forceinline__ float ex2_poly(float s, const PolyConsts& pc) {
    float y;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(s), "f"(pc.scale), "f"(pc.bias));
    const float t = fmaf(y, 128.f, kMagic);      // integer part of x' in the low mantissa bits
    const float nn = kMagic - t;                 // -round(x')
    const float f = fmaf(y, 128.f, nn);          // x' - round(x') in [-0.5, 0.5]
    float p = fmaf(f, kC3, kC2);
    p = fmaf(p, f, kC1);
    p = fmaf(p, f, kC0);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// Two columns at once: the clamp stays scalar (fma.sat has no f32x2 form), range reduction and Horner steps run as FFMA2 / FADD2
// (one issue slot for two columns; 5 instead of 8 issue slots per column).
__device__ __forceinline__ void ex2_poly_pair(float s0, float s1, const PolyConsts& pc, float& p0, float& p1) {
    float y0, y1;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(y0) : "f"(s0), "f"(pc.scale), "f"(pc.bias));
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(y1) : "f"(s1), "f"(pc.scale), "f"(pc.bias));
    const uint64_t y = pack_f32x2(y0, y1);
    const uint64_t k128 = pack_f32x2(128.f, 128.f), magic = pack_f32x2(kMagic, kMagic);
    const uint64_t t = fma_f32x2(y, k128, magic);   // integer part of x' in the low mantissa bits
    const uint64_t nn = sub_f32x2(magic, t);        // -round(x')
    const uint64_t f = fma_f32x2(y, k128, nn);      // x' - round(x') in [-0.5, 0.5]
    uint64_t p = fma_f32x2(f, pack_f32x2(kC3, kC3), pack_f32x2(kC2, kC2));
    p = fma_f32x2(p, f, pack_f32x2(kC1, kC1));
    p = fma_f32x2(p, f, pack_f32x2(kC0, kC0));
    float q0, q1, t0, t1;
    unpack_f32x2(p, q0, q1);
    unpack_f32x2(t, t0, t1);
    p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
    p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}

// ---- one row slice ------------------------------------------------------------------------------
// sr[kCols]: raw scores (bit patterns); pk[kCols/2]: FP16 pairs (column 2i in the low half).
// Returns the FP32 sum of the (unrounded) probabilities.
template <int kCols>
__device__ __forceinline__ float softmax_exp_pass(const uint32_t (&sr)[kCols], uint32_t (&pk)[kCols / 2],
                                                  float neg_m_scaled) {
    const PolyConsts pc = make_poly_consts(neg_m_scaled);
    float p[kCols];
#pragma unroll
    for (int c = 0; c < kCols; c += 2) {
        const bool poly0 = ((c % 8) < LG_ATTN_POLY_OF_8);
        const bool poly1 = (((c + 1) % 8) < LG_ATTN_POLY_OF_8);
        if (poly0 && poly1 && LG_ATTN_PACKED_POLY) {
            ex2_poly_pair(__uint_as_float(sr[c]), __uint_as_float(sr[c + 1]), pc, p[c], p[c + 1]);
        } else if (!poly0 && !poly1 && LG_ATTN_PACKED_SCALE) {
            float x0, x1;
            unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(sr[c]), __uint_as_float(sr[c + 1])),
                                   pack_f32x2(kScaleLog2, kScaleLog2), pack_f32x2(neg_m_scaled, neg_m_scaled)),
                         x0, x1);
            p[c] = ex2_approx(x0);
            p[c + 1] = ex2_approx(x1);
        } else {
            p[c] = poly0 ? ex2_poly(__uint_as_float(sr[c]), pc)
                         : ex2_approx(fmaf(__uint_as_float(sr[c]), kScaleLog2, neg_m_scaled));
            p[c + 1] = poly1 ? ex2_poly(__uint_as_float(sr[c + 1]), pc)
                             : ex2_approx(fmaf(__uint_as_float(sr[c + 1]), kScaleLog2, neg_m_scaled));
        }
        pk[c / 2] = pack_half2(p[c], p[c + 1]);
    }
    float sum;
#ifdef LG_ATTN_EXPERIMENT_NOSUM
    // what-if build (WRONG results): drops the row-sum adds to bound what moving them to the tensor core could gain
    sum = p[0] + p[kCols - 1];
#else
    if (LG_ATTN_PACKED_SUM) {
        uint64_t acc[4] = {0ull, 0ull, 0ull, 0ull};
#pragma unroll
        for (int c = 0; c < kCols; c += 2) acc[(c / 2) % 4] = add_f32x2(acc[(c / 2) % 4], pack_f32x2(p[c], p[c + 1]));
        float s0, s1;
        unpack_f32x2(add_f32x2(add_f32x2(acc[0], acc[1]), add_f32x2(acc[2], acc[3])), s0, s1);
        sum = s0 + s1;
    } else {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int c = 0; c < kCols; c += 4) {
            a0 += p[c];
            a1 += p[c + 1];
            a2 += p[c + 2];
            a3 += p[c + 3];
        }
        sum = (a0 + a1) + (a2 + a3);
    }
#endif
    return sum;
}


// row max of a slice with 3-input FMNMX chains
#ifndef LG_ATTN_MAX_CHAINS
#define LG_ATTN_MAX_CHAINS 4
#endif
template <int kCols>
__device__ __forceinline__ float row_max(const uint32_t (&sr)[kCols]) {
    constexpr int kCh = LG_ATTN_MAX_CHAINS;
    float mx[kCh];
#pragma unroll
    for (int i = 0; i < kCh; ++i) mx[i] = __uint_as_float(sr[i]);
#pragma unroll
    for (int c = kCh; c < kCols; c += kCh) {
#pragma unroll
        for (int i = 0; i < kCh; ++i) mx[i] = fmaxf(mx[i], __uint_as_float(sr[c + i]));
    }
#pragma unroll
    for (int w = kCh / 2; w > 0; w /= 2) {
#pragma unroll
        for (int i = 0; i < w; ++i) mx[i] = fmaxf(mx[i], mx[i + w]);
    }
    return mx[0];
}

}  // namespace lgattn
