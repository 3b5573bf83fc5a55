// dsmem_bench.cu — throughput of random atomicOr over the shared memory of a whole thread-block cluster
// (would a table slice spread over 8 or 16 SMs be countable at a useful rate?)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/dsmem_bench tools/dsmem_bench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// MODE 0: atom (returns old), 1: red (no return), 2: local only (baseline, atom)
template <int MODE, int ILP>
__global__ void __launch_bounds__(1024, 1) k_dsmem(int words_per_cta, int iters, uint32_t* sink)
{
    extern __shared__ uint32_t s[];
    cg::cluster_group cl = cg::this_cluster();
    const uint32_t nr = cl.num_blocks();
    for (int i = threadIdx.x; i < words_per_cta; i += blockDim.x) s[i] = 0;
    cl.sync();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(s);
    uint32_t x = mix(blockIdx.x * 1024u + threadIdx.x + 1u), acc = 0;
    for (int it = 0; it < iters; it++) {
        uint32_t addr[ILP], old[ILP];
#pragma unroll
        for (int k = 0; k < ILP; k++) {
            x = mix(x + k + 1);
            const uint32_t word = x % (uint32_t)words_per_cta, rank = MODE == 2 ? cl.block_rank() : (x >> 20) % nr;
            uint32_t la = base + word * 4, ra;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(rank));
            addr[k] = ra;
        }
#pragma unroll
        for (int k = 0; k < ILP; k++) {
            if (MODE == 1) { asm volatile("red.shared::cluster.or.b32 [%0], %1;" :: "r"(addr[k]), "r"(1u << (x & 31)) : "memory"); old[k] = 0; }
            else asm volatile("atom.shared::cluster.or.b32 %0, [%1], %2;" : "=r"(old[k]) : "r"(addr[k]), "r"(1u << (x & 31)) : "memory");
        }
#pragma unroll
        for (int k = 0; k < ILP; k++) acc += old[k];
    }
    cl.sync();
    if (acc == 0x12345678u) *sink = acc;
}

template <int MODE, int ILP>
void run(const char* name, int csize, int kb, uint32_t* sink)
{
    const int iters = 256;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    CK(cudaFuncSetAttribute(k_dsmem<MODE, ILP>, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024));
    if (csize > 8) CK(cudaFuncSetAttribute(k_dsmem<MODE, ILP>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    int nclusters = 0;
    cfg.gridDim = dim3(csize * 64); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = kb * 1024; cfg.attrs = at; cfg.numAttrs = 1;
    if (cudaOccupancyMaxActiveClusters(&nclusters, k_dsmem<MODE, ILP>, &cfg) != cudaSuccess) { cudaGetLastError(); nclusters = -1; }
    cfg.gridDim = dim3(csize * (nclusters > 0 ? nclusters : 8));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float ms = 0;
    for (int rep = 0; rep < 2; rep++) {
        CK(cudaEventRecord(a));
        cudaError_t e = cudaLaunchKernelEx(&cfg, k_dsmem<MODE, ILP>, kb * 256, iters, sink);
        if (e != cudaSuccess) { printf("  %-28s cluster %2d: launch failed: %s\n", name, csize, cudaGetErrorString(e)); cudaGetLastError(); return; }
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b));
    }
    const double ops = (double)cfg.gridDim.x * 1024 * iters * ILP;
    printf("  %-28s cluster %2d (%3d resident clusters, %3d CTAs) ILP %d: %8.1f Gop/s\n", name, csize, nclusters, cfg.gridDim.x, ILP, ops / ms / 1e6);
}

int main()
{
    uint32_t* sink; CK(cudaMalloc(&sink, 4));
    for (int csize : {1, 2, 4, 8, 16}) {
        run<0, 4>("atom.shared::cluster.or", csize, 192, sink);
        run<1, 4>("red.shared::cluster.or", csize, 192, sink);
        run<2, 4>("atom, own CTA only", csize, 192, sink);
    }
    run<0, 8>("atom.shared::cluster.or", 16, 192, sink);
    run<0, 1>("atom.shared::cluster.or", 16, 192, sink);
    return 0;
}
