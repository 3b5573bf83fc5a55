// probe_bench.cu — how fast can a B200 answer random 4-byte probes of a bitmap, as a function of the bitmap's size
// and of the load flavour?  (k_probe_full is bound by exactly this; ncu shows ~100 B of DRAM traffic per probe.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/probe_bench tools/probe_bench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE> __device__ __forceinline__ uint32_t load(const uint32_t* p)
{
    uint32_t v;
    if (MODE == 0) v = __ldg(p);
    else if (MODE == 1) asm volatile("ld.global.nc.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (MODE == 2) asm volatile("ld.global.nc.L2::128B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (MODE == 4) asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (MODE == 5) asm volatile("ld.global.cv.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (MODE == 6) asm volatile("ld.global.L1::evict_first.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    }
    return v;
}

template <int MODE, int ILP>
__global__ void __launch_bounds__(256) k_probe(const uint32_t* tab, uint32_t words_mask, int iters, uint32_t* sink)
{
    uint32_t x = mix(blockIdx.x * 256u + threadIdx.x + 1u), acc = 0;
    for (int it = 0; it < iters; it++) {
        uint32_t a[ILP];
#pragma unroll
        for (int k = 0; k < ILP; k++) { x = mix(x + k + 1); a[k] = x & words_mask; }
#pragma unroll
        for (int k = 0; k < ILP; k++) acc += load<MODE>(tab + a[k]);
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <class F> float time_ms(F f)
{
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f(); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms;
}

template <int MODE> void run(const char* name, const uint32_t* tab, uint32_t wm, uint32_t* sink)
{
    const int iters = 64, grid = 148 * 16;
    float ms = time_ms([&] { k_probe<MODE, 8><<<grid, 256>>>(tab, wm, iters, sink); });
    printf("    %-34s %7.1f Gprobe/s\n", name, (double)grid * 256 * iters * 8 / ms / 1e6);
}

int main()
{
    uint32_t *tab, *sink;
    const size_t bytes = (size_t)1 << 31;
    CK(cudaMalloc(&tab, bytes)); CK(cudaMemset(tab, 0, bytes)); CK(cudaMalloc(&sink, 4));
    for (int mib : {16, 32, 64, 128, 256, 512, 2048}) {
        const uint32_t wm = (uint32_t)((size_t)mib << 18) - 1;
        printf("bitmap of %d MiB\n", mib);
        run<0>("__ldg (ld.global.nc)", tab, wm, sink);
        run<1>("ld.global.nc.L2::64B", tab, wm, sink);
        run<2>("ld.global.nc.L2::128B", tab, wm, sink);
        run<3>("ld.global.nc.L1::no_allocate", tab, wm, sink);
        run<4>("ld.global.cg", tab, wm, sink);
        run<5>("ld.global.cv", tab, wm, sink);
        run<6>("ld.global.L1::evict_first", tab, wm, sink);
        run<7>("ld.global.nc + L2 evict_first hint", tab, wm, sink);
    }
    for (size_t g : {32, 64, 128}) {
        CK(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g));
        printf("cudaLimitMaxL2FetchGranularity = %zu, 512 MiB\n", g);
        run<0>("__ldg (ld.global.nc)", tab, (uint32_t)((size_t)512 << 18) - 1, sink);
    }
    return 0;
}
