// Micro-probe (GPU box): which kernel feature drops the occupancy of a 384-thread / 80-reg / 88 KB CTA to 1 per SM?
#include <cstdio>
#include <cuda_runtime.h>
template <int FEAT>
__global__ void __launch_bounds__(384, 2) kern(float* p, int n) {
    extern __shared__ float sm[];
    float v[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = p[threadIdx.x + i * 32];
    if (FEAT & 1) {  // cluster barrier + mapa
        if (n > 1000) {
            unsigned r, a = 0;
            asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(r));
            asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
            asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
            v[0] += a;
        }
    }
    if (FEAT & 2) {  // named barriers 1..5
        asm volatile("bar.sync 1, 64;" ::: "memory");
        asm volatile("bar.sync 5, 256;" ::: "memory");
    }
    if (FEAT & 4) {  // setmaxnreg
        if (threadIdx.x >= 256) asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
    }
    if (FEAT & 8) {  // tcgen05 alloc / dealloc
        if (threadIdx.x < 32) {
            unsigned addr = 0;
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(addr) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(addr) : "memory");
        }
    }
    if (FEAT & 16) {  // griddepcontrol
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    }
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = v[i] * v[(i + 1) % 64] + sm[i];
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 64; ++i) s += v[i];
    p[threadIdx.x] = s;
}
template <typename K>
void probe(const char* name, K k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 88064);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k);
    int nb = -1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, 384, 88064);
    printf("%-28s regs=%3d -> blocks/SM %d\n", name, fa.numRegs, nb);
}
int main() {
    probe("plain", kern<0>);
    probe("cluster barrier + mapa", kern<1>);
    probe("named barriers", kern<2>);
    probe("setmaxnreg", kern<4>);
    probe("tcgen05 alloc", kern<8>);
    probe("griddepcontrol", kern<16>);
    probe("all", kern<31>);
    return 0;
}
