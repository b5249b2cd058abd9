// Scalar arithmetic shared by the glue kernels (lg_match_kernels.cu) and the fused ffn kernel (ffn_sm100_kernel.cuh).
#pragma once

namespace lgmatch {

// erf-form GELU, 0.5 x (1 + erf(x / sqrt 2)) = 0.5 x + 0.5 |x| erf(|x| / sqrt 2), with erf by Abramowitz-Stegun 7.1.25
// (three coefficients, |err| <= 2.5e-5: below half an FP16 ulp of the result everywhere) — 10 FP32 ops + 2 MUFU (rcp, ex2) per
// element.  The kernels that use it are bound by FP32 issue, not by HBM (~0.7 element-ops / clk / scheduler,
// tools/microbench/pipe_rates.cu), so every op per element counts: round 1's 7.1.26 form (five coefficients) cost two more FMAs
// for accuracy FP16 cannot hold.
__device__ __forceinline__ float gelu_erf(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.47047f, z, 1.0f)));
    float p = fmaf(0.7478556f, t, -0.0958798f);
    p = fmaf(p, t, 0.3480242f);
    p *= t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"((z * -1.4426950408889634f) * z));
    const float erf_abs = fmaf(-p, e, 1.0f);
    const float hx = 0.5f * x;
    return fmaf(fabsf(hx), erf_abs, hx);
}

// The same function of hx = x / 2 (callers that produce x by an affine step halve its coefficients instead: one multiply less),
// with the exponent's argument folded into the scale of z (z' = |x| / sqrt 2 * sqrt(log2 e), so exp(-(x / sqrt 2)^2) = 2^(-z'^2):
// another multiply less): 8 FP32 ops + 2 MUFU.  Used by the fused ffn kernels, whose epilogue is bound by instruction issue.
__device__ __forceinline__ float gelu_erf_of_half(float hx) {
    constexpr float kSqrtLog2e = 1.2011224087864498f;
    const float z = fabsf(hx) * (1.4142135623730951f * kSqrtLog2e);
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.47047f / kSqrtLog2e, z, 1.0f)));
    float p = fmaf(0.7478556f, t, -0.0958798f);
    p = fmaf(p, t, 0.3480242f);
    p *= t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-z * z));
    return fmaf(fabsf(hx), fmaf(-p, e, 1.0f), hx);
}

}  // namespace lgmatch
