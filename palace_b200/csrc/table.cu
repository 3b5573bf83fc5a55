// table.cu — the query k-mer table (north_star parts 2+3).
//
// The reference keeps one saturating byte counter per 32-bit code, shared by the three code
// streams (bin/extract_ref.cpp:25-26, :995-997), and only ever asks "count == 3" (:531).  In HBM
// that is three 2^32-bit bitmaps, level L meaning "seen at least L times".  An occurrence sets its
// bit in level 1; if it was already set it moves on to level 2, then level 3.  Of n occurrences of a
// code exactly one finds level 1 clear and exactly one of the rest finds level 2 clear, so level 3
// ends up set iff n >= 3, whatever the order the atomics land in: the build is deterministic.
#include "eref_internal.cuh"

namespace {

constexpr int BATCH = 4;     // window starts whose 12 level-1 atomics are issued back to back

__device__ __forceinline__ void count_code(uint32_t code, uint32_t old1, uint32_t* __restrict__ bm2,
                                           uint32_t* __restrict__ bm3)
{
    const uint32_t bit = 1u << (code & 31);
    if (old1 & bit) {
        uint32_t old2 = atomicOr(bm2 + (code >> 5), bit);
        if (old2 & bit) atomicOr(bm3 + (code >> 5), bit);
    }
}

// One thread per plane word = 32 window starts.  Only starts whose 32 positions are all valid
// bases are visited (valid_windows), so FASTQ headers/qualities cost one mask computation.
__global__ void __launch_bounds__(256)
k_count_kmers(Planes pl, CoderMasks m, uint32_t* __restrict__ bm1, uint32_t* __restrict__ bm2,
              uint32_t* __restrict__ bm3)
{
    uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= pl.nwords) return;
    const uint32_t ok0 = pl.ok[w], ok1 = pl.ok[w + 1];          // pad word makes w+1 readable
    uint32_t vm = valid_windows(ok0, ok1);
    if (!vm) return;
    const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
    while (vm) {
        uint32_t code[BATCH * 3], old[BATCH * 3];
        int nb = 0;
#pragma unroll
        for (int q = 0; q < BATCH; q++) {
            if (vm) {
                int s = __clz(vm);
                vm &= ~(0x80000000u >> s);
                WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
#pragma unroll
                for (int i = 0; i < 3; i++) code[q * 3 + i] = stream_code(m, b, i);
                nb = q + 1;
            }
        }
#pragma unroll
        for (int q = 0; q < BATCH * 3; q++)
            if (q < nb * 3) old[q] = atomicOr(bm1 + (code[q] >> 5), 1u << (code[q] & 31));
#pragma unroll
        for (int q = 0; q < BATCH * 3; q++)
            if (q < nb * 3) count_code(code[q], old[q], bm2, bm3);
    }
}

__global__ void k_popcount(const uint32_t* __restrict__ bm, uint64_t nwords, unsigned long long* out)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned long long c = 0;
    for (; i < nwords; i += stride) c += __popc(bm[i]);
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

} // namespace

cudaError_t launch_count_kmers(Planes pl, CoderMasks m, uint32_t* bm1, uint32_t* bm2, uint32_t* bm3,
                               cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_count_kmers<<<(unsigned)((pl.nwords + 255) / 256), 256, 0, st>>>(pl, m, bm1, bm2, bm3);
    return cudaGetLastError();
}

cudaError_t launch_popcount(const uint32_t* bm, uint64_t nwords, unsigned long long* out, cudaStream_t st)
{
    k_popcount<<<148 * 8, 256, 0, st>>>(bm, nwords, out);
    return cudaGetLastError();
}
