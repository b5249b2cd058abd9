/* lg_match.h — C-ABI of the memory-bound glue either side of the attention hot path
 * (widened only after rows a–e were parity-green and measured).
 *
 * The reference runs these as many small eager / TensorRT-fused elementwise launches around each
 * plugin call.  Here each is ONE hand-written sm_100a kernel over packed FP16 activation buffers, so
 * that the strided attention entry (lg_attn_hd64_launch_strided) never needs a layout copy:
 *
 *   lg_match_posenc          LearnableFourierPositionalEncoding.forward
 *                            (lightglue_pytorch_with_plugin/lightglue.py:77-86)
 *   lg_match_rotary          SelfBlock.rotate_half / apply_cached_rotary_emb on q and k
 *                            (lightglue.py:143-144, 154-165), in place inside the packed projection
 *   lg_match_layernorm_gelu  the LayerNorm(512) + GELU in the middle of both ffn stacks
 *                            (lightglue.py:131-136, 181-186)
 *   lg_match_ffn1_ln_gelu    ffn[0..2] of both stacks as one tcgen05 GEMM with the LayerNorm + GELU in its epilogue
 *                            (lightglue.py:131-136, 181-186)
 *   lg_match_proj_rotary     Wqkv / to_qk|to_v projection with the rotary in the GEMM epilogue (lightglue.py:139-147, 190-193)
 *   lg_match_assign          sigmoid_log_double_softmax + filter_matches
 *                            (lightglue.py:216-224, 253-276; demo/lightglue_trt.cpp:291-345)
 *
 * All pointers are DEVICE pointers; every function only enqueues on `stream` (a cudaStream_t as
 * void*): no allocation, no synchronisation, legal under stream capture.  Returns 0 or a
 * cudaError_t / LG_ATTN_E* code (see lg_attn.h).
 */
#ifndef LG_MATCH_H_
#define LG_MATCH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cs[r][i] = (cos, sin)(kpts[r] . wr[i]) for i < 32: kpts [rows,2] float (normalised keypoints),
 * wr [32,2] float (posenc.Wr.weight), cs [rows,32,2] float. */
int lg_match_posenc(const float* kpts, const float* wr, float* cs, long long rows, void* stream);

/* In-place rotary on `n_blocks` consecutive 256-column blocks (4 heads x 64) at the start of every
 * row of a packed FP16 buffer with leading dimension `ld` (elements): for each pair (2i, 2i+1) of a
 * head, (a, b) -> (a cos_i - b sin_i, b cos_i + a sin_i), the same 32 angles for every head. */
int lg_match_rotary(void* packed_half, long long rows, int ld, int n_blocks, const float* cs, void* stream);

/* In place on h [rows, 512] FP16 (row stride `ld` elements): y = GELU(LayerNorm(h) * gamma + beta),
 * erf-form GELU, statistics in FP32.  gamma, beta: float[512]. */
int lg_match_layernorm_gelu(void* h_half, long long rows, int ld, const float* gamma, const float* beta,
                            float eps, void* stream);

/* The first half of an ffn stack as ONE launch (lightglue.py:131-136, 181-186: ffn[0] Linear(512, 512), ffn[1] LayerNorm(512),
 * ffn[2] GELU): hid = GELU(LayerNorm(x @ w^T + bias) * gamma + beta).  A hand-written tcgen05 GEMM (FP32 accumulate) whose epilogue
 * does the LayerNorm statistics, the affine and the erf-form GELU from registers under the next tile's MMAs — replaces a library
 * GEMM + lg_match_layernorm_gelu (one pass over [rows, 512] less, and the FP32-issue-bound arithmetic hidden).
 *   x     [rows, 512] FP16, row stride `ldx` elements (>= 512, multiple of 8): cat(x, context)
 *   w_nk  [512 (out), 512 (in)] FP16 contiguous — nn.Linear.weight layout (row n = output feature n)
 *   bias  FP16[512];  gamma, beta: float[512]
 *   hid   [rows, 512] FP16, row stride `ldh` elements; must not overlap x
 * x, w_nk, hid 16-byte aligned.  Enqueues on `stream` only (three host-side tensor-map encodes, one launch). */
int lg_match_ffn1_ln_gelu(const void* x_half, long long rows, int ldx, const void* w_nk_half, const void* bias_half,
                          const float* gamma, const float* beta, float eps, void* hid_half, int ldh, void* stream);
/* Per-device one-time set-up of the kernel above (shared-memory opt-in, driver entry point); idempotent, thread-safe, also done
 * lazily by the first lg_match_ffn1_ln_gelu call on a device. */
int lg_match_ffn_init_device(int device);
/* 1 when the fused kernel is expected to beat a library GEMM + lg_match_layernorm_gelu at this row count on the current device,
 * else 0: its tiles are whole rows (128 x 512), so few rows leave most SMs idle (crossover on B200: about two thirds of the SM
 * count in tiles).  The matcher's callers ask this once per shape plan. */
int lg_match_ffn1_prefer_fused(long long rows);
/* Tuning / tests: the persistent variant of lg_match_ffn1_ln_gelu also exists as CTA pairs (clusters of 2, tcgen05.mma
 * cta_group::2 with M = 256, half of every weight tile per CTA, double-buffered output staging).  Same results bit for bit; faster
 * once a launch is several waves of tiles long, slightly slower below, so the launcher picks by size (>= 3 tiles per SM).
 * enabled: 1 / 0 force it on / off, -1 = by size (default).  Also LG_MATCH_FFN_PAIR=0/1 in the environment. */
int lg_match_debug_set_ffn_pair(int enabled);

/* Input projection of a transformer block with the rotary embedding in its epilogue, ONE launch (lightglue.py:139-147: Wqkv +
 * apply_cached_rotary_emb on q, k; :190-193: to_qk | to_v): out[:, 256 t ..] = x @ w[256 t .., :]^T + bias for t < n_blocks, then
 * blocks t < n_rot_blocks are rotated like lg_match_rotary does.  A hand-written tcgen05 GEMM (K = 256) with double-buffered TMEM
 * accumulators; replaces a library GEMM + lg_match_rotary (no second pass over q | k, rotation applied before the FP16 rounding).
 *   x    [rows, 256] FP16, row stride `ldx` elements;  w_nk [256 n_blocks (out), 256 (in)] FP16 contiguous (nn.Linear.weight layout)
 *   bias FP16[256 n_blocks];  cs [rows, 32, 2] float from lg_match_posenc (may be NULL when n_rot_blocks == 0)
 *   out  [rows, 256 n_blocks] FP16, row stride `ldo` elements;  n_blocks <= 3;  x, w_nk, out, cs 16-byte aligned. */
int lg_match_proj_rotary(const void* x_half, long long rows, int ldx, const void* w_nk_half, const void* bias_half, int n_blocks,
                         int n_rot_blocks, const float* cs, void* out_half, int ldo, void* stream);
/* 1 when lg_match_proj_rotary is expected to beat a library GEMM (+ lg_match_rotary) for this shape on the current device. */
int lg_match_proj_prefer_fused(long long rows, int n_blocks, int n_rot_blocks);

/* Assignment + mutual-nearest-neighbour filter for `batch` pairs.
 *   sim  [batch, n0, n1] float  : mdesc0 . mdesc1^T (already divided by dim^0.25 on both sides)
 *   z0   matchability logits of image 0, FP16, element (b, i) at z0[(b*n0 + i) * z_stride]; z1 likewise
 *   matches0 [batch, n0] int32  : index into image 1, or -1         (LightGlue `matches0`)
 *   mscores0 [batch, n0] float  : exp(score) where mutual, else 0   (LightGlue `mscores0`)
 *   scores   [batch, n0, n1] float or NULL: the full log-assignment matrix (`lightglue_scores`,
 *            export_onnx_with_plugin.py:197) — only materialised when asked for.
 * workspace: lg_match_assign_workspace_bytes() bytes. */
size_t lg_match_assign_workspace_bytes(int batch, int n0, int n1);
int lg_match_assign(const float* sim, const void* z0_half, const void* z1_half, long long z_stride,
                    int batch, int n0, int n1, float threshold, int* matches0, float* mscores0,
                    float* scores, void* workspace, size_t workspace_bytes, void* stream);

/* lg_match_assign for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and kernels as lg_match_assign; the scratch is a file-scope device array of
 * LG_MATCH_ASSIGN_STATIC_MAX_BYTES (4 MiB) inside the library — every GPU that loads the library holds it until exit.  The call
 * allocates and frees nothing (first call per device: one cudaGetSymbolAddress lookup, cached).
 * Limit: lg_match_assign_workspace_bytes(batch, n0, n1) <= LG_MATCH_ASSIGN_STATIC_MAX_BYTES (fits e.g. 64 pairs of 1024 x 1024
 * keypoints, 16 of 2048 x 2048, 4 of 4096 x 4096); larger shapes return LG_MATCH_ESTATIC_TOO_LARGE and launch nothing — use lg_match_assign with caller memory.
 * All calls on a device share the one array: calls that may overlap on the GPU (different streams without an ordering between
 * them) race on it and must be serialised by the caller; nothing in it needs clearing between calls. */
#define LG_MATCH_ASSIGN_STATIC_MAX_BYTES 4194304
#define LG_MATCH_ESTATIC_TOO_LARGE 100012 /* shape needs more scratch than LG_MATCH_ASSIGN_STATIC_MAX_BYTES */
int lg_match_assign_static(const float* sim, const void* z0_half, const void* z1_half, long long z_stride,
                           int batch, int n0, int n1, float threshold, int* matches0, float* mscores0,
                           float* scores, void* stream);

/* Number of kernels each entry enqueues (for the bench's gpu_launches accounting). */
int lg_match_assign_launches(int with_scores);

#ifdef __cplusplus
}
#endif
#endif /* LG_MATCH_H_ */
