// Compile-time geometry shared by the sm_100a attention kernels and the launcher.
#pragma once
#include <stdint.h>

namespace lgattn {

constexpr int kHeadDim = 64;    // ATTENTION_HEAD_DIM (reference lightglue_attention_plugin.h:22)
constexpr int kBlockM = 128;    // Q rows per CTA  (= tcgen05.mma M, one TMEM lane per row)
constexpr int kBlockN = 128;    // KV rows per pipeline step (= tcgen05.mma N of S = Q K^T)
constexpr int kStages = 2;      // K/V shared-memory ring depth

// Where a (batch, head) slice of O starts and how far apart its rows are, in ELEMENTS of the output
// type.  Contiguous [B,H,N,64] (the plugin layout): sb = H*N*64, sh = N*64, sr = 64.
// kv_shift: K/V of batch entry b are read from batch entry (b + kv_shift) mod batch — lets both directions
// of a cross-attention between two stacked image sets run as ONE launch (kv_shift = batch / 2).
struct OutLayout {
    long long sb, sh, sr;
    int heads;
    int batch;
    int kv_shift;
};

}  // namespace lgattn
