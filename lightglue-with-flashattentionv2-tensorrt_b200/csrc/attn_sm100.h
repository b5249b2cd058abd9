// Compile-time geometry of the sm_100a attention kernel (shared by kernel and launcher).
#pragma once
#include <stdint.h>

namespace lgattn {

constexpr int kHeadDim = 64;    // ATTENTION_HEAD_DIM (reference lightglue_attention_plugin.h:22)
constexpr int kBlockM = 128;    // Q rows per CTA  (= tcgen05.mma M, one TMEM lane per row)
constexpr int kBlockN = 128;    // KV rows per pipeline step (= tcgen05.mma N of S = Q K^T)
constexpr int kStages = 2;      // K/V shared-memory ring depth
constexpr int kThreads = 256;   // 4 softmax warps + TMA + MMA + TMEM-alloc + spare
// Q tile + 2x(K,V) tiles (16 KiB each) + barriers, + 1 KiB slack for manual 1024-B alignment
constexpr uint32_t kSmemBytes = (1 + 2 * kStages) * 16384 + 1024 + 1024;

}  // namespace lgattn
