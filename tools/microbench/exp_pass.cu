// Saturated cost of the softmax arithmetic of one 64-column row slice (softmax_math.cuh, the product header) per SM
// sub-partition, without TMEM / barriers / MMAs: W warps per sub-partition loop over softmax_exp_pass<64> on register-resident
// scores.  Tells how much of the throughput kernel's ~600 clk per (warp, 64-column step) is arithmetic.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I lightglue-with-flashattentionv2-tensorrt_b200/csrc \
//        [-DLG_ATTN_POLY_OF_8=..] [-DLG_ATTN_PACKED_SCALE=..] [-DLG_ATTN_PACKED_SUM=..] [-DLG_ATTN_PACKED_POLY=..] [-DWITH_MAX=1] \
//        -o exp_pass tools/microbench/exp_pass.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "softmax_math.cuh"

#ifndef WITH_MAX
#define WITH_MAX 0
#endif
#ifndef ITERS
#define ITERS 256
#endif

using namespace lgattn;

#ifndef LB_THREADS
#define LB_THREADS 512
#endif
__global__ void __launch_bounds__(LB_THREADS, 1) k(uint64_t* out_cycles, float* sink, const float* in, float neg_m0) {
    uint32_t sr[64];
#pragma unroll
    for (int c = 0; c < 64; ++c) sr[c] = __float_as_uint(in[(threadIdx.x * 64 + c) & 4095]);
    float neg_m = neg_m0, acc = 0.f;
    uint32_t x = 0;
    __syncthreads();
    const uint64_t t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        uint32_t pk[32];
        if (WITH_MAX) neg_m = -row_max<64>(sr) * kScaleLog2 + neg_m0 * it;
        acc += softmax_exp_pass<64>(sr, pk, neg_m);
#pragma unroll
        for (int c = 0; c < 32; c += 2) x ^= pk[c] ^ pk[c + 1];   // P is "stored": 16 LOP3 stand in for the tcgen05.st
        neg_m -= 1e-3f;
        // keep the scores opaque across iterations without touching them (the kernel gets fresh ones from TMEM)
#pragma unroll
        for (int c = 0; c < 64; ++c) asm volatile("" : "+r"(sr[c]));
    }
    const uint64_t t1 = clock64();
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + __uint_as_float(x);
    if (threadIdx.x == 0) out_cycles[blockIdx.x] = t1 - t0;
}

int main() {
    uint64_t* d_c;
    float *d_s, *d_in;
    cudaMalloc(&d_c, 1024);
    cudaMalloc(&d_s, 1 << 20);
    cudaMalloc(&d_in, 4096 * 4);
    float h[4096];
    for (int i = 0; i < 4096; ++i) h[i] = float((i * 2654435761u) % 2001) * 0.01f - 10.f;   // scores in [-10, 10]
    cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice);
    printf("POLY_OF_8=%d PACKED_SCALE=%d PACKED_SUM=%d PACKED_POLY=%d WITH_MAX=%d:", LG_ATTN_POLY_OF_8, LG_ATTN_PACKED_SCALE, LG_ATTN_PACKED_SUM,
           LG_ATTN_PACKED_POLY, WITH_MAX);
    for (int warps : {4, 8, 12, 16}) {
        k<<<1, warps * 32>>>(d_c, d_s, d_in, -1.5f);
        k<<<1, warps * 32>>>(d_c, d_s, d_in, -1.5f);
        uint64_t c;
        if (cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
        printf("  W=%d: %.0f clk/pass/warp = %.0f clk/pass/SMSP", warps / 4, c / double(ITERS), c / double(ITERS) / (warps / 4.0));
    }
    printf("\n");
    return 0;
}
