// TEST / BASELINE INFRASTRUCTURE — not product code.
// extern "C" shim over the REFERENCE's own CUDA launchers (compiled unmodified from
// /root/reference/lightglue_attention_plugin/*.cu by oracle/build_ref.py) so that tests and the
// bench can run the reference kernels on a B200 next to the new kernel.  The call order reproduces
// LightGlueAttentionPlugin::enqueue (lightglue_attention_plugin.cpp:165-267):
//   kHALF : pad-copy (always) -> attention fp16->fp16 (padded operands only where needed) -> unpad iff N%64
//   kFLOAT: convert+pad (always) -> attention fp16->fp32 -> unpad iff N%64
// Workspace layout follows lightglue_attention_plugin.h:107-120 generalised to any batch/length:
// Q | K | V (half, padded) | O (float, padded).
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "attention_headdim_64_fp16in_fp16out.cuh"
#include "attention_headdim_64_fp16in_fp32out.cuh"

static inline int pad64(int n) { return ((n + 63) / 64) * 64; }

extern "C" size_t ref_attn_workspace_bytes(int batch, int heads, int n_q, int n_kv) {
    size_t bh = (size_t)batch * heads;
    return bh * pad64(n_q) * 64 * 2 + 2 * bh * pad64(n_kv) * 64 * 2 + bh * pad64(n_q) * 64 * 4;
}

// is_float: 0 = kHALF branch, 1 = kFLOAT branch. Returns number of kernel launches, <0 on error.
extern "C" int ref_attn_enqueue(const void* q, const void* k, const void* v, void* o, int batch, int heads, int n_q,
                                int n_kv, int is_float, void* workspace, void* stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    const int nqp = pad64(n_q), nkp = pad64(n_kv);
    const size_t bh = (size_t)batch * heads;
    uint8_t* ws = (uint8_t*)workspace;
    half* qw = (half*)ws;
    half* kw = (half*)(ws + bh * nqp * 64 * 2);
    half* vw = (half*)(ws + bh * nqp * 64 * 2 + bh * nkp * 64 * 2);
    void* ow = ws + bh * nqp * 64 * 2 + 2 * bh * nkp * 64 * 2;
    int launches = 0;
    if (!is_float) {
        AttentionHeadDim64::launch_QKV_move_from_unpadded_to_padded_fp16((const half*)q, (const half*)k, (const half*)v, qw,
                                                                          kw, vw, batch, heads, n_q, n_kv, nqp, nkp, stream);
        ++launches;
        AttentionHeadDim64::launch_attention_kernel_headdim_64_no_remainder_fp16in_fp16out(
            n_q != nqp ? qw : (const half*)q, n_kv != nkp ? kw : (const half*)k, n_kv != nkp ? vw : (const half*)v,
            n_q != nqp ? (half*)ow : (half*)o, batch, heads, nqp, nkp, n_q, n_kv, stream);
        ++launches;
        if (n_q != nqp) {
            AttentionHeadDim64::launch_O_move_from_padded_to_unpadded_fp16((const half*)ow, (half*)o, batch, heads, n_q, nqp,
                                                                            stream);
            ++launches;
        }
    } else {
        AttentionHeadDim64::launch_QKV_convert_from_fp32_to_fp16((const float*)q, (const float*)k, (const float*)v, qw, kw, vw,
                                                                  batch, heads, n_q, n_kv, nqp, nkp, stream);
        ++launches;
        AttentionHeadDim64::launch_attention_kernel_headdim_64_no_remainder_fp16in_fp32out(
            qw, kw, vw, n_q != nqp ? (float*)ow : (float*)o, batch, heads, nqp, nkp, n_q, n_kv, stream);
        ++launches;
        if (n_q != nqp) {
            AttentionHeadDim64::launch_O_move_from_padded_to_unpadded_fp32((const float*)ow, (float*)o, batch, heads, n_q, nqp,
                                                                            stream);
            ++launches;
        }
    }
    return cudaGetLastError() == cudaSuccess ? launches : -1;
}

// The reference's internal half-in / float-out kernel alone, on pre-padded operands (no pad kernels).
extern "C" int ref_attn_f16_f32_padded(const void* q, const void* k, const void* v, void* o, int batch, int heads,
                                       int n_q_padded, int n_kv_padded, int n_q, int n_kv, void* stream_v) {
    AttentionHeadDim64::launch_attention_kernel_headdim_64_no_remainder_fp16in_fp32out(
        (const half*)q, (const half*)k, (const half*)v, (float*)o, batch, heads, n_q_padded, n_kv_padded, n_q, n_kv,
        (cudaStream_t)stream_v);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}
