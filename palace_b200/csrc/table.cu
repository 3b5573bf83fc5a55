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

// ---- bucketed build --------------------------------------------------------------------------------
// Random atomics over three 512 MiB bitmaps miss L2 on every access (ncu: 4.6 % hit rate, ~80 B of DRAM
// traffic per atomic).  The codes are therefore first partitioned by their top BUCKET_BITS bits; the
// counting kernel then walks the buckets in order, so the 3 x 16 MiB of table slices it is updating at
// any moment stay resident in the 126 MB L2 and DRAM sees each table sector about once.
// Atomics commute, so the resulting bitmaps are identical to the direct kernel's.
constexpr int BUCKET_BITS = 5;
constexpr int NBUCKETS = 1 << BUCKET_BITS;
constexpr int BUCKET_SHIFT = 32 - BUCKET_BITS;

template <class F>
__device__ __forceinline__ void for_each_code(const Planes& pl, const CoderMasks& m, uint64_t w, F&& f)
{
    uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]);
    if (!vm) return;
    const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
    while (vm) {
        int s = __clz(vm);
        vm &= ~(0x80000000u >> s);
        WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
#pragma unroll
        for (int i = 0; i < 3; i++) f(stream_code(m, b, i));
    }
}

// SCATTER=false: global histogram of codes per bucket.  SCATTER=true: the CTA counts its codes per
// bucket, reserves one contiguous range per bucket with 32 global atomics, then writes its codes there.
template <bool SCATTER>
__global__ void __launch_bounds__(256)
k_bucket_codes(Planes pl, CoderMasks m, unsigned long long* __restrict__ g_hist,
               const unsigned long long* __restrict__ bucket_off, unsigned long long* __restrict__ g_tail,
               uint32_t* __restrict__ codes_out)
{
    __shared__ uint32_t s_hist[8][NBUCKETS];
    __shared__ unsigned long long s_base[NBUCKETS];
    __shared__ uint32_t s_cursor[NBUCKETS];
    const int warp = threadIdx.x >> 5;
    s_hist[threadIdx.x >> 5][threadIdx.x & 31] = 0;
    __syncthreads();
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = w < pl.nwords;
    if (live) for_each_code(pl, m, w, [&](uint32_t code) { atomicAdd(&s_hist[warp][code >> BUCKET_SHIFT], 1u); });
    __syncthreads();
    if (threadIdx.x < NBUCKETS) {
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) c += s_hist[k][threadIdx.x];
        if (!SCATTER) {
            if (c) atomicAdd(g_hist + threadIdx.x, (unsigned long long)c);
        } else {
            s_base[threadIdx.x] = bucket_off[threadIdx.x] + (c ? atomicAdd(g_tail + threadIdx.x, (unsigned long long)c) : 0ull);
            s_cursor[threadIdx.x] = 0;
        }
    }
    if (!SCATTER) return;
    __syncthreads();
    if (live) for_each_code(pl, m, w, [&](uint32_t code) {
        const uint32_t b = code >> BUCKET_SHIFT;
        codes_out[s_base[b] + atomicAdd(&s_cursor[b], 1u)] = code;
    });
}

// Counts one bucket of up to two bucket-ordered code lists (the two FASTQ files).  A thread owns U
// 16-byte units, strided by the grid so that every LDG.128 of a warp is one contiguous 512 B run; all
// level-1 atomics of a thread are issued before the first dependent level-2 atomic.
template <int U>
__global__ void __launch_bounds__(256)
k_count_codes(const uint32_t* __restrict__ c0, uint64_t n0, const uint32_t* __restrict__ c1, uint64_t n1,
              uint32_t* __restrict__ bm1, uint32_t* __restrict__ bm2, uint32_t* __restrict__ bm3)
{
    const uint64_t units0 = (n0 + 3) >> 2, total = units0 + ((n1 + 3) >> 2);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t first = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t code[4 * U], old[4 * U];
    uint32_t live = 0;                                   // bit q set: code[q] is a real code
#pragma unroll
    for (int j = 0; j < U; j++) {
        const uint64_t u = first + (uint64_t)j * stride;
        if (u < total) {
            const bool second = u >= units0;
            const uint64_t lu = second ? u - units0 : u;
            const uint32_t* base = second ? c1 : c0;
            const uint64_t n = second ? n1 : n0;
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(base) + lu);   // bucket starts are 16-byte aligned,
            code[4 * j] = v.x; code[4 * j + 1] = v.y; code[4 * j + 2] = v.z; code[4 * j + 3] = v.w;   // padded to 4 codes
            const uint64_t left = n - (lu << 2);
            live |= (left >= 4 ? 0xFu : ((1u << left) - 1u)) << (4 * j);
        }
    }
#pragma unroll
    for (int q = 0; q < 4 * U; q++)
        if ((live >> q) & 1u) old[q] = atomicOr(bm1 + (code[q] >> 5), 1u << (code[q] & 31));
#pragma unroll
    for (int q = 0; q < 4 * U; q++)
        if ((live >> q) & 1u) count_code(code[q], old[q], bm2, bm3);
}

// summary bit = OR of 16 consecutive level-3 bits: a 32 MiB filter that stays in L2 during the scan
__global__ void __launch_bounds__(256)
k_build_summary(const uint32_t* __restrict__ bm3, uint32_t* __restrict__ summary)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;          // summary word: 16 table words
    const uint4* p = reinterpret_cast<const uint4*>(bm3) + (size_t)i * 4;
    uint32_t out = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint4 v = __ldg(p + q);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int word = q * 4 + k;                                // table word inside the group of 16
            out |= (uint32_t)((w[k] & 0x0000FFFFu) != 0) << (word * 2);
            out |= (uint32_t)((w[k] & 0xFFFF0000u) != 0) << (word * 2 + 1);
        }
    }
    summary[i] = out;
}

// zero one bucket's slice of all three levels; the lines stay dirty in L2 for the atomics that follow
__global__ void __launch_bounds__(256)
k_zero_slices(uint4* __restrict__ a, uint4* __restrict__ b, uint4* __restrict__ c, uint32_t nvec)
{
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += gridDim.x * blockDim.x) {
        a[i] = z; b[i] = z; c[i] = z;
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

cudaError_t launch_bucket_hist(Planes pl, CoderMasks m, unsigned long long* g_hist, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_bucket_codes<false><<<(unsigned)((pl.nwords + 255) / 256), 256, 0, st>>>(pl, m, g_hist, nullptr, nullptr, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_bucket_scatter(Planes pl, CoderMasks m, const unsigned long long* bucket_off,
                                  unsigned long long* g_tail, uint32_t* codes_out, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_bucket_codes<true><<<(unsigned)((pl.nwords + 255) / 256), 256, 0, st>>>(pl, m, nullptr, bucket_off, g_tail, codes_out);
    return cudaGetLastError();
}

cudaError_t launch_count_codes(const uint32_t* c0, uint64_t n0, const uint32_t* c1, uint64_t n1, int units,
                               uint32_t* bm1, uint32_t* bm2, uint32_t* bm3, cudaStream_t st)
{
    const uint64_t total = ((n0 + 3) >> 2) + ((n1 + 3) >> 2);
    if (!total) return cudaSuccess;
    if (units >= 4) k_count_codes<4><<<(unsigned)((total + 1023) / 1024), 256, 0, st>>>(c0, n0, c1, n1, bm1, bm2, bm3);
    else if (units == 2) k_count_codes<2><<<(unsigned)((total + 511) / 512), 256, 0, st>>>(c0, n0, c1, n1, bm1, bm2, bm3);
    else k_count_codes<1><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(c0, n0, c1, n1, bm1, bm2, bm3);
    return cudaGetLastError();
}

cudaError_t launch_zero_slices(uint32_t* a, uint32_t* b, uint32_t* c, uint32_t nwords, cudaStream_t st)
{
    k_zero_slices<<<148 * 8, 256, 0, st>>>(reinterpret_cast<uint4*>(a), reinterpret_cast<uint4*>(b),
                                          reinterpret_cast<uint4*>(c), nwords / 4);
    return cudaGetLastError();
}

cudaError_t launch_build_summary(const uint32_t* bm3, uint32_t* summary, cudaStream_t st)
{
    k_build_summary<<<SUMMARY_WORDS / 256, 256, 0, st>>>(bm3, summary);
    return cudaGetLastError();
}

cudaError_t launch_popcount(const uint32_t* bm, uint64_t nwords, unsigned long long* out, cudaStream_t st)
{
    k_popcount<<<148 * 8, 256, 0, st>>>(bm, nwords, out);
    return cudaGetLastError();
}
