// Micro-benchmark (GPU box): issue rates of the softmax-side instructions on sm_100a, per SM
// sub-partition, for 1/2/4 warps per scheduler.  Not product code; informs the kernel's op mix.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu && ./pipe_rates
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHAINS 8
#define ITERS 512

enum Op { FFMA, FFMA2, FADD2, FMNMX, FMNMX3, F2FP, MUFU, IMADSHL, LEA_, FADD, MIX_FFMA2_MUFU, MIX_FFMA2_ALU, MIX_FFMA_ALU,
          MIX_ALL, MIX_FFMA2_FFMA, MIX_MUFU_ALU, NOPS };
static const char* names[] = {"FFMA", "FFMA2", "FADD2", "FMNMX", "FMNMX3", "F2FP", "MUFU.EX2", "IMAD(shl-add)", "LEA", "FADD",
                              "FFMA2+MUFU (1:1)", "FFMA2+FMNMX (1:1)", "FFMA+FMNMX (1:1)", "FFMA2+FADD2+MUFU+FMNMX+F2FP",
                              "FFMA2+FFMA (1:1)", "MUFU+FMNMX (1:1)"};
static const int per_iter[] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 5, 2, 2};

template <int OP>
__global__ void k(uint64_t* out_cycles, float* sink, float a, float b) {
    float x[CHAINS];
    uint64_t y[CHAINS];
    uint32_t u[CHAINS];
    for (int i = 0; i < CHAINS; ++i) {
        x[i] = a + i + threadIdx.x;
        asm("mov.b64 %0, {%1, %2};" : "=l"(y[i]) : "f"(x[i]), "f"(x[i] + 1.f));
        u[i] = threadIdx.x + i;
    }
    uint64_t ab, bb;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ab) : "f"(a), "f"(a));
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b), "f"(b));
    __syncthreads();
    const uint64_t t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (OP == FFMA || OP == MIX_FFMA_ALU || OP == MIX_FFMA2_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
            if (OP == FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(b));
            if (OP == FFMA2 || OP == MIX_FFMA2_MUFU || OP == MIX_FFMA2_ALU || OP == MIX_ALL || OP == MIX_FFMA2_FFMA)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(y[i]) : "l"(ab), "l"(bb));
            if (OP == FADD2 || OP == MIX_ALL) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(y[i]) : "l"(bb));
            if (OP == FMNMX || OP == MIX_FFMA2_ALU || OP == MIX_FFMA_ALU || OP == MIX_ALL || OP == MIX_MUFU_ALU)
                asm volatile("max.f32 %0, %0, %1;" : "+f"(x[(i + 1) % CHAINS]) : "f"(b));
            if (OP == FMNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
            if (OP == F2FP || OP == MIX_ALL) asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(u[i]) : "f"(x[i]), "f"(x[(i + 3) % CHAINS]));
            if (OP == MUFU || OP == MIX_FFMA2_MUFU || OP == MIX_ALL || OP == MIX_MUFU_ALU)
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[(i + 2) % CHAINS]));
            if (OP == IMADSHL) asm volatile("mad.lo.u32 %0, %0, 8388608, %1;" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]));
            if (OP == LEA_) asm volatile("{.reg .u32 t; shl.b32 t, %0, 23; add.u32 %0, t, %1;}" : "+r"(u[i]) : "r"(u[(i + 1) % CHAINS]));
        }
    }
    const uint64_t t1 = clock64();
    float s = 0;
    for (int i = 0; i < CHAINS; ++i) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(y[i]));
        s += x[i] + lo + hi + __uint_as_float(u[i]);
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) out_cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(uint64_t* d_c, float* d_s) {
    for (int warps : {4, 8, 16, 32}) {
        k<OP><<<1, warps * 32>>>(d_c, d_s, 1.0001f, 0.5f);
        k<OP><<<1, warps * 32>>>(d_c, d_s, 1.0001f, 0.5f);
        uint64_t c;
        cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
        const double instr_per_smsp = double(ITERS) * CHAINS * per_iter[OP] * (warps / 4.0);
        printf("%-32s warps/SMSP=%d  cycles/warp-instr/SMSP = %.3f\n", names[OP], warps / 4, c / instr_per_smsp);
    }
}

int main() {
    uint64_t* d_c;
    float* d_s;
    cudaMalloc(&d_c, 1024);
    cudaMalloc(&d_s, 1 << 20);
    run<FFMA>(d_c, d_s); run<FADD>(d_c, d_s); run<FFMA2>(d_c, d_s); run<FADD2>(d_c, d_s); run<FMNMX>(d_c, d_s);
    run<FMNMX3>(d_c, d_s); run<F2FP>(d_c, d_s); run<MUFU>(d_c, d_s); run<IMADSHL>(d_c, d_s); run<LEA_>(d_c, d_s);
    run<MIX_FFMA2_MUFU>(d_c, d_s); run<MIX_FFMA2_ALU>(d_c, d_s); run<MIX_FFMA_ALU>(d_c, d_s); run<MIX_ALL>(d_c, d_s);
    run<MIX_FFMA2_FFMA>(d_c, d_s); run<MIX_MUFU_ALU>(d_c, d_s);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
