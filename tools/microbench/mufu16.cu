#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CHAINS 8
#define ITERS 512
template <int OP>
__global__ void k(uint64_t* out_cycles, uint32_t* sink, uint32_t a) {
    uint32_t u[CHAINS]; float x[CHAINS];
    for (int i = 0; i < CHAINS; ++i) { u[i] = a + threadIdx.x + i; x[i] = a * 0.001f + i; }
    __syncthreads();
    const uint64_t t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(u[i]));
            if (OP == 2) asm volatile("{.reg .b16 lo, hi; mov.b32 {lo, hi}, %0; ex2.approx.f16 lo, lo; mov.b32 %0, {lo, hi};}" : "+r"(u[i]));
            if (OP == 3) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 4) asm volatile("fma.rn.f16x2 %0, %0, %0, %0;" : "+r"(u[i]));
            if (OP == 5) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        }
    }
    const uint64_t t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < CHAINS; ++i) s += u[i] + __float_as_uint(x[i]);
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) out_cycles[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, uint64_t* d_c, uint32_t* d_s) {
    for (int warps : {4, 8, 16, 32}) {
        k<OP><<<1, warps * 32>>>(d_c, d_s, 3);
        k<OP><<<1, warps * 32>>>(d_c, d_s, 3);
        uint64_t c; cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
        printf("%-28s warps/SMSP=%d  cycles/warp-instr(PTX)/SMSP = %.3f\n", name, warps / 4, c / (double(ITERS) * CHAINS * (warps / 4.0)));
    }
}
int main() {
    uint64_t* d_c; uint32_t* d_s; cudaMalloc(&d_c, 1024); cudaMalloc(&d_s, 1 << 20);
    run<0>("ex2.f32", d_c, d_s); run<1>("ex2.f16x2 (2 MUFU.F16+PRMT)", d_c, d_s); run<2>("ex2.f16 scalar", d_c, d_s);
    run<3>("tanh.f32", d_c, d_s); run<4>("fma.f16x2", d_c, d_s); run<5>("rcp.f32", d_c, d_s);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
