// Compile-time geometry of the sm_100a attention kernel (shared by kernel and launcher).
#pragma once
#include <stdint.h>

namespace lgattn {

constexpr int kHeadDim = 64;    // ATTENTION_HEAD_DIM (reference lightglue_attention_plugin.h:22)
constexpr int kBlockM = 128;    // Q rows per CTA  (= tcgen05.mma M, one TMEM lane per row)
constexpr int kBlockN = 128;    // KV rows per pipeline step (= tcgen05.mma N of S = Q K^T)
constexpr int kStages = 2;      // K/V shared-memory ring depth
#ifndef LG_ATTN_SOFTMAX_WARPS
#define LG_ATTN_SOFTMAX_WARPS 8
#endif
constexpr int kSoftmaxWarps = LG_ATTN_SOFTMAX_WARPS;  // 8: two column-half warpgroups over the same 128 rows; 4: whole rows
constexpr int kThreads = 32 * (kSoftmaxWarps + 4);    // softmax warps + TMA + MMA + TMEM-alloc + spare
// Q tile + 2x(K,V) tiles (16 KiB each) + barriers and row-stat exchange, + 1 KiB slack for manual 1024-B alignment
constexpr uint32_t kSmemBytes = (1 + 2 * kStages) * 16384 + 5120 + 1024;

}  // namespace lgattn
