/* lg_attn.h — C-ABI of the B200-native (sm_100a) LightGlue fused attention (head_dim 64).
 *
 * Drop-in boundary for the hot path of qdLMF/LightGlue-with-FlashAttentionV2-TensorRT.
 * The reference exports no extern "C" symbol: its TensorRT plugin `enqueue`
 * (lightglue_attention_plugin/lightglue_attention_plugin.cpp:114-270) calls C++-mangled
 * launchers in namespace AttentionHeadDim64
 * (attention_headdim_64_fp16in_fp16out.cuh:17-57, attention_headdim_64_fp16in_fp32out.cuh:17-75).
 * The functions below are what a C / ctypes / cgo binding of that path binds instead; the
 * plugin translation unit in this repo calls them from `enqueue`.
 *
 * Conventions
 *   - all tensors are contiguous row-major [batch, heads, rows, 64] (TensorRT kLINEAR),
 *     device pointers unless a function says "host";
 *   - O = softmax(Q K^T / 8) V per (batch, head)  (lightglue_pytorch_no_plugin/lightglue.py:82-85);
 *   - n_q (rows of Q and O) and n_kv (rows of K and V) are independent, any value >= 1;
 *   - functions taking a stream only enqueue work: no allocation, no synchronisation,
 *     legal inside CUDA stream capture;
 *   - return value: 0 on success, otherwise a cudaError_t / LG_ATTN_E* code (never aborts).
 */
#ifndef LG_ATTN_H_
#define LG_ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LG_ATTN_HEAD_DIM 64

/* io_dtype: what the caller's buffers hold. Compute is always FP16 operands, FP32 accumulate. */
enum lg_attn_io_dtype {
    LG_ATTN_F16_F16 = 0, /* Q,K,V half  -> O half   : plugin kHALF branch  (…plugin.cpp:177-221; kernel …fp16out.cu:253-733) */
    LG_ATTN_F32_F32 = 1, /* Q,K,V float -> O float  : plugin kFLOAT branch (…plugin.cpp:222-267; convert …fp32out.cu:706-804) */
    LG_ATTN_F16_F32 = 2  /* Q,K,V half  -> O float  : launch_attention_kernel_…fp16in_fp32out (…fp32out.cu:962-989) */
};

/* error codes beyond cudaError_t (chosen above the cudaError_t range) */
#define LG_ATTN_EINVAL      100001 /* bad argument (null pointer, non-positive size, unknown dtype) */
#define LG_ATTN_EWORKSPACE  100002 /* workspace missing or smaller than lg_attn_hd64_workspace_bytes() */
#define LG_ATTN_EDRIVER     100003 /* cuTensorMapEncodeTiled entry point unavailable / failed */
#define LG_ATTN_EARCH       100004 /* device is not compute capability 10.x */

/* ABI version of this header (bumped on any signature change). */
int lg_attn_version(void);

/* Bit mask of the tuning switches this library was compiled with; 0 for a product build.  bit 0 LG_ATTN_TUNING_BUILD,
 * bit 1 LG_ATTN_WHATIF (wrong results by design), bit 2 LG_ATTN_EXPERIMENT_NOSUM (wrong results), bit 3 a phase-trace build
 * (LG_ATTN_TRACE / LG_ATTN_W4_TRACE / LG_ATTN_MQ_TRACE).  tests/test_boundary_cpu.py requires 0 of the shipped .so. */
int lg_attn_build_flags(void);

/* Test / tuning hook (no reference counterpart): which full-grid (throughput) kernel lg_attn_hd64_launch* uses.
 * -1 automatic (default: whichever packs the launch best, see choose_full_grid_kernel in lg_attn_api.cu),
 * 0 one 128-row Q tile per CTA, 1 persistent CTAs owning four Q tiles (64-row KV steps), 2 persistent CTAs owning two Q
 * tiles (128-row KV steps).  All meet the same tolerances; also settable as LG_ATTN_FULL_GRID_KERNEL. */
int lg_attn_debug_set_throughput_kernel(int mode);

/* Measurement hooks (no reference counterpart).  set_pdl: programmatic dependent launch of the latency kernel on / off (also
 * LG_ATTN_PDL) so that a bench can report a serialised single-call latency both ways.  describe_launch: writes which kernel,
 * grid and split count lg_attn_hd64_launch would use for this shape on the current device into buf (NUL-terminated). */
int lg_attn_debug_set_pdl(int enabled);
int lg_attn_debug_describe_launch(int batch, int heads, int n_q, int n_kv, char* buf, size_t buf_bytes);

/* Per-device one-time setup (shared-memory opt-in, driver entry point lookup). Idempotent and
 * thread-safe; lg_attn_hd64_launch() calls it lazily, but call it before stream capture.
 * Replaces the static-initialiser cudaFuncSetAttribute of the reference (…fp16out.cu:242-250). */
int lg_attn_hd64_init_device(int device);

/* Scratch bytes the launcher needs for this problem: 0 in every mode (LG_ATTN_F32_F32 converts to FP16 on the way into shared
 * memory).  Replaces the fixed 5 MiB of getWorkspaceSize (…plugin.cpp:96-112); kept so that callers keep asking. */
size_t lg_attn_hd64_workspace_bytes(int batch, int heads, int n_q, int n_kv, int io_dtype);

/* The hot path: enqueue one fused attention on `stream` (a cudaStream_t passed as void*).
 * Replaces the pad -> attention -> unpad launch sequence of enqueue (…plugin.cpp:177-267). */
int lg_attn_hd64_launch(const void* q, const void* k, const void* v, void* o,
                        int batch, int heads, int n_q, int n_kv, int io_dtype,
                        void* workspace, size_t workspace_bytes, void* stream);

/* Strided variant of the hot path for callers that hold Q/K/V inside packed projection buffers
 * (the reshape / transpose / .contiguous() glue around every plugin call in the reference:
 * lightglue_pytorch_with_plugin/lightglue.py:139-147, 154-156, 190-199).  Q, K, V are FP16; element
 * (b, h, r, c) of a tensor lives at base + b*batch + h*head + r*row + c (strides in ELEMENTS, the 64
 * head-dim elements contiguous).  Strides must be multiples of 16 bytes, pointers 16-byte aligned.
 * out_f32: 0 -> FP16 output, 1 -> FP32 output.  kv_batch_shift (0 <= shift < batch): batch entry b of Q
 * attends to batch entry (b + shift) mod batch of K and V; with two image sets stacked along the batch
 * dimension, shift = batch / 2 runs both directions of CrossBlock.forward (lightglue.py:188-199) in one
 * launch.  No workspace.  Same guarantees as lg_attn_hd64_launch. */
typedef struct lg_attn_strides {
    long long batch, head, row;
} lg_attn_strides;
int lg_attn_hd64_launch_strided(const void* q, const void* k, const void* v, void* o,
                                int batch, int heads, int n_q, int n_kv, int out_f32,
                                const lg_attn_strides* sq, const lg_attn_strides* sk,
                                const lg_attn_strides* sv, const lg_attn_strides* so,
                                int kv_batch_shift, void* stream);

/* Convenience for non-CUDA hosts (ctypes / cgo): HOST buffers in, HOST buffer out.
 * Allocates device scratch, copies H2D, launches, copies D2H, synchronises.  Not for the
 * TensorRT path; used by the end-to-end (`e2e`) measurement and by bindings without a CUDA runtime. */
int lg_attn_hd64_run_host(const void* q_host, const void* k_host, const void* v_host, void* o_host,
                          int batch, int heads, int n_q, int n_kv, int io_dtype, int device);

/* Host-buffer session: persistent device buffers + streams for callers that hold HOST tensors
 * (bindings without a CUDA runtime, the bench's end-to-end leg).  `n_slots` independent
 * (device buffers, stream) sets are used round-robin, so copies of one call overlap the kernel of
 * another when the host buffers are pinned.  submit() is asynchronous: it enqueues
 * H2D(q,k,v) -> launch -> D2H(o) on the slot's stream and returns; o_host is valid after sync().
 * Shapes up to the create-time maxima; any n_q, n_kv >= 1 below them. */
typedef struct lg_attn_session lg_attn_session;
int lg_attn_session_create(lg_attn_session** out, int device, int max_batch, int heads, int max_n_q, int max_n_kv,
                           int io_dtype, int n_slots);
int lg_attn_session_submit(lg_attn_session* s, const void* q_host, const void* k_host, const void* v_host,
                           void* o_host, int batch, int n_q, int n_kv);
/* `count` problems of one shape in one call.  Problem p reads q at q_host + p * q_stride_bytes, k / v at their pointers +
 * p * kv_stride_bytes, and writes o at o_host + p * out_stride_bytes.  When the host arrays are packed like the session's device slots — [q | k | v] per problem back to
 * back, outputs dense, shape = the create-time maxima — `group` consecutive problems (default 4, at most n_slots / 2) move with ONE
 * H2D and ONE D2H transfer (1.5 MiB copies reach ~42 GB/s of this box's PCIe link beside the return traffic, 6 MiB copies ~49) and
 * are launched one by one; any other layout falls back to lg_attn_session_submit() per problem.  Asynchronous like submit(). */
int lg_attn_session_submit_many(lg_attn_session* s, const void* q_host, const void* k_host, const void* v_host, void* o_host,
                                int count, size_t q_stride_bytes, size_t kv_stride_bytes, size_t out_stride_bytes, int group, int batch,
                                int n_q, int n_kv);
int lg_attn_session_sync(lg_attn_session* s);
void lg_attn_session_destroy(lg_attn_session* s);

/* Number of kernels lg_attn_hd64_launch() enqueues for a dtype: 1 in every mode (the reference enqueues 2-3: pad / convert,
 * attention, unpad — …plugin.cpp:177-267). */
int lg_attn_hd64_launches_per_call(int io_dtype);

/* Human-readable name of a return code (static storage). */
const char* lg_attn_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* LG_ATTN_H_ */
