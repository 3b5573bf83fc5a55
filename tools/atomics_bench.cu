// Development microbenchmark: throughput of random 32-bit atomics on B200.
//   global: atomicOr with the old value used (ATOMG), atomicOr without (REDG), plain loads — over an L2-resident region
//   shared: atomicOr with the old value used (ATOMS) over a 64 KiB / 192 KiB shared-memory table
// usage: atomics_bench [region_MiB=48]
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int MODE, int ILP>
__global__ void __launch_bounds__(256) k_global(uint32_t* tab, uint32_t words_mask, int iters, uint32_t* sink)
{
    uint32_t seed = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0;
    for (int it = 0; it < iters; it++) {
        uint32_t a[ILP], o[ILP];
#pragma unroll
        for (int q = 0; q < ILP; q++) { seed = mix(seed + q + 1); a[q] = seed; }
#pragma unroll
        for (int q = 0; q < ILP; q++) {
            uint32_t* p = tab + ((a[q] >> 5) & words_mask);
            if (MODE == 0) o[q] = atomicOr(p, 1u << (a[q] & 31));
            else if (MODE == 1) { atomicOr(p, 1u << (a[q] & 31)); o[q] = 0; }
            else o[q] = __ldcg(p);
        }
#pragma unroll
        for (int q = 0; q < ILP; q++) acc += o[q];
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_shared(int words, int iters, uint32_t* sink)
{
    extern __shared__ uint32_t s[];
    for (int i = threadIdx.x; i < words; i += blockDim.x) s[i] = 0;
    __syncthreads();
    uint32_t seed = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 777u;
    uint32_t acc = 0;
    for (int it = 0; it < iters; it++) {
        uint32_t a[ILP], o[ILP];
#pragma unroll
        for (int q = 0; q < ILP; q++) { seed = mix(seed + q + 1); a[q] = seed; }
#pragma unroll
        for (int q = 0; q < ILP; q++) o[q] = atomicOr(&s[(a[q] >> 5) % (uint32_t)words], 1u << (a[q] & 31));
#pragma unroll
        for (int q = 0; q < ILP; q++) acc += o[q];
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <class F> float time_ms(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main(int argc, char** argv)
{
    int mib = argc > 1 ? atoi(argv[1]) : 48;
    size_t words = (size_t)mib << 18;
    uint32_t wm = 1; while ((size_t)wm * 2 <= words) wm *= 2; wm -= 1;
    uint32_t *tab, *sink;
    CK(cudaMalloc(&tab, (size_t)(wm + 1) * 4)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(tab, 0, (size_t)(wm + 1) * 4));
    printf("region %.0f MiB\n", (wm + 1) * 4.0 / (1 << 20));
    const char* names[3] = {"ATOMG.OR (return used)", "REDG.OR (no return)", "LDG.cg random"};
    for (int ctas = 1; ctas <= 8; ctas *= 2) {
        const int grid = 148 * ctas, iters = 256 / ctas * 8;
#define RUN(MODE, ILP) { float ms = time_ms([&] { k_global<MODE, ILP><<<grid, 256>>>(tab, wm, iters, sink); }); \
            printf("  %-24s ctas/SM=%d ILP=%2d : %7.1f Gop/s\n", names[MODE], ctas, ILP, (double)grid * 256 * iters * ILP / ms / 1e6); }
        RUN(0, 1) RUN(0, 4) RUN(0, 8) RUN(0, 16) RUN(1, 4) RUN(1, 16) RUN(2, 4) RUN(2, 16)
    }
    for (int kb = 64; kb <= 192; kb += 128) {
        CK(cudaFuncSetAttribute(k_shared<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024));
        CK(cudaFuncSetAttribute(k_shared<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024));
        CK(cudaFuncSetAttribute(k_shared<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kb * 1024));
        const int per_sm = kb == 64 ? 3 : 1;
        const int grid = 148 * per_sm, iters = 4096;
        for (int threads = 256; threads <= 1024; threads *= 2) {
            float m1 = time_ms([&] { k_shared<1><<<grid, threads, kb * 1024>>>(kb * 256, iters, sink); });
            float m4 = time_ms([&] { k_shared<4><<<grid, threads, kb * 1024>>>(kb * 256, iters / 4, sink); });
            float m8 = time_ms([&] { k_shared<8><<<grid, threads, kb * 1024>>>(kb * 256, iters / 8, sink); });
            printf("  ATOMS.OR %3d KiB table, %d CTA/SM x %4d thr: ILP1 %7.1f  ILP4 %7.1f  ILP8 %7.1f Gop/s\n", kb, per_sm, threads,
                   (double)grid * threads * iters / m1 / 1e6, (double)grid * threads * iters / m4 / 1e6, (double)grid * threads * iters / m8 / 1e6);
        }
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
