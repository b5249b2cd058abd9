/* lg_matcher.h — C-ABI of the native (C++ / CUDA / cuBLASLt) LightGlue pair-matcher engine.
 *
 * What the reference's compiled host does for this path: demo/lightglue_trt.cpp — `LightGlueTRT`
 * owns a TensorRT engine, device buffers and a captured CUDA graph (:347-366), takes two images'
 * keypoints + descriptors and returns matches (`Forward` :200-289, `PostProcess` :291-345).  TensorRT is
 * unavailable on this box, so the same role is filled by a small engine written directly against the
 * sm_100a kernels of this repo (include/lg_attn.h, include/lg_match.h) and cuBLASLt for the plain GEMMs:
 * no Python, no torch, no TensorRT in the call path.  Layout and weight packing are those of
 * `matcher.py` (one packed FP16 activation matrix, folded projections — see DESIGN.md §8).
 *
 * All functions return 0 or a cudaError_t / cublasStatus_t-derived / LG_ATTN_E* code; none aborts.
 * An engine owns ONE set of plan buffers: calls on one engine are not thread-safe, and the stream path
 * (lg_matcher_match) must not be mixed with in-flight lg_matcher_submit_host work without a sync between them.
 * Create one engine per GPU / per concurrent stream of work.  A shape change re-plans (device-wide
 * synchronisation, allocation, graph capture): keep shapes stable on the hot path.
 */
#ifndef LG_MATCHER_H_
#define LG_MATCHER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lg_matcher lg_matcher;

#define LG_MATCHER_ECUBLAS 100010 /* a cuBLASLt call failed */
#define LG_MATCHER_EBLOB   100011 /* weight blob malformed (magic / size / layer count) */

/* Create an engine on `device` from a packed weight blob in HOST memory (layout: lg_matcher_blob_bytes()
 * documents the size; `MatcherWeights.to_blob()` in matcher.py writes it).  `threshold` is LightGlue's
 * filter_threshold (lightglue.py:284, default 0.1). */
int lg_matcher_create(lg_matcher** out, int device, int n_layers, const void* weights_host, size_t weights_bytes,
                      float threshold);

/* Size of the weight blob for `n_layers` layers. */
size_t lg_matcher_blob_bytes(int n_layers);

/* Allocate the buffers for `pairs` pairs of (n0, n1) keypoints and capture the forward into a CUDA graph.
 * Called implicitly by the match functions when the shape changes; call it up front to keep the first
 * match free of allocation.  One shape is resident at a time. */
int lg_matcher_plan(lg_matcher* m, int pairs, int n0, int n1);

/* Match `pairs` image pairs.  DEVICE pointers: kpts0 [pairs,n0,2] / kpts1 [pairs,n1,2] float (normalised
 * keypoints), desc0 [pairs,n0,256] / desc1 [pairs,n1,256] FP16; outputs matches0 [pairs,n0] int32 (index into
 * image 1 or -1) and mscores0 [pairs,n0] float.  Everything is enqueued on `stream` (cudaStream_t as void*);
 * no synchronisation. */
int lg_matcher_match(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half,
                     const void* desc1_half, int pairs, int n0, int n1, int* matches0, float* mscores0, void* stream);

/* Same with HOST buffers (pinned for overlap): H2D, match, D2H on the engine's own stream, then synchronise. */
int lg_matcher_match_host(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half,
                          const void* desc1_half, int pairs, int n0, int n1, int* matches0, float* mscores0);

/* Asynchronous host-buffer path for streams of sub-batches: submit() stages the (pinned) host inputs on a copy
 * stream into one of two device slots while the previous forward still runs, then enqueues forward + D2H and
 * returns; outputs are valid after lg_matcher_sync().  Host buffers must stay untouched until then. */
int lg_matcher_submit_host(lg_matcher* m, const float* kpts0, const float* kpts1, const void* desc0_half,
                           const void* desc1_half, int pairs, int n0, int n1, int* matches0, float* mscores0);
int lg_matcher_sync(lg_matcher* m);

/* Kernel launches inside one captured forward of the current plan (0 before planning). */
int lg_matcher_launches(const lg_matcher* m);

/* How many shape plans (device buffers + captured graph) the engine currently holds: the active one plus the parked ones
 * (at most 4; the least recently used is evicted).  Shapes that alternate are swapped in without re-planning. */
int lg_matcher_plans_resident(const lg_matcher* m);

void lg_matcher_destroy(lg_matcher* m);

#ifdef __cplusplus
}
#endif
#endif /* LG_MATCHER_H_ */
