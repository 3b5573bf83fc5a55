// noalloc.cu — eref_census_noalloc: the FASTQ census (cal_sam_ratio's pass, bin/extract_ref.cpp:1124-1148) for callers
// that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference (include/eref_gpu.h).
//
// The existing entry point (eref_query_census_device) keeps its per-tile newline counts, their prefix and the per-tile
// kept-byte counts in buffers that grow with the text (DevBuf::ensure -> cudaMalloc) and scans them with cub (temp
// storage, cudaMalloc).  Here the same per-tile scratch lives in ONE file-scope __device__ array of fixed size, and the
// prefix over the tiles is a single-CTA scan — nothing is allocated, nothing waits for the host, every launch goes to the
// caller's stream.  The array exists once per GPU that loads the library and is shared by every call: calls that overlap
// on the GPU (two streams, two threads) race on it; calls on one stream are ordered and safe.
#include "eref_internal.cuh"
#include <mutex>

namespace {

constexpr uint64_t NOALLOC_MAX_BYTES = EREF_NOALLOC_MAX_BYTES;                         // 256 MiB of FASTQ text
constexpr uint32_t NOALLOC_MAX_TILES = (uint32_t)(NOALLOC_MAX_BYTES / TILE_BYTES);     // 32768 tiles of 8 KiB

struct CensusScratch {                                    // 512 KiB + 16 B per GPU, held until the process exits
    uint32_t           tile_counts[NOALLOC_MAX_TILES + 1];     // newlines per tile
    uint32_t           tile_prefix[NOALLOC_MAX_TILES + 1];     // newlines before the tile; [ntiles] = all
    unsigned long long tile_kept[NOALLOC_MAX_TILES + 1];       // k_fastq_kept's per-tile plane positions (not reported)
};

static __device__ __align__(256) unsigned char g_census_scratch[sizeof(CensusScratch)];

// In host code `g_census_scratch` is only a stand-in for the symbol: its device address is asked for once per device,
// with that device current, and handed to the kernels.
constexpr int MAX_DEVICES = 64;
std::mutex g_addr_mu;
CensusScratch* g_scratch_addr[MAX_DEVICES] = {};

int scratch_for_current_device(CensusScratch** out)
{
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return eref_fail(EREF_ERR_LIMIT, "eref_census_noalloc: device %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_addr_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_census_scratch));
        if (!p) return eref_fail(EREF_ERR_CUDA, "eref_census_noalloc: no device address for the scratch array");
        g_scratch_addr[dev] = static_cast<CensusScratch*>(p);
    }
    *out = g_scratch_addr[dev];
    return 0;
}

// Exclusive prefix of counts[0 .. ntiles) into prefix[0 .. ntiles], prefix[ntiles] = total; ONE CTA of 1024 threads,
// each owning a contiguous run of tiles (at most 32).  Also the line count of cal_sam_ratio: out[1] = sequence lines.
__global__ void __launch_bounds__(1024)
k_tiles_scan_noalloc(const uint32_t* __restrict__ counts, uint32_t* __restrict__ prefix, uint32_t ntiles,
                     const uint8_t* __restrict__ text, uint64_t n, unsigned long long* __restrict__ out)
{
    __shared__ uint32_t s_warp[32];
    const uint32_t per = (ntiles + 1023u) / 1024u;
    const uint32_t lo = min(threadIdx.x * per, ntiles), hi = min(lo + per, ntiles);
    uint32_t sum = 0;
    for (uint32_t t = lo; t < hi; t++) sum += counts[t];
    uint32_t inc = sum;                                   // inclusive scan over the 1024 run sums
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31) >= d) inc += v;
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t w = s_warp[threadIdx.x];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, w, d);
            if (threadIdx.x >= d) w += v;
        }
        s_warp[threadIdx.x] = w;
    }
    __syncthreads();
    uint32_t run = inc - sum + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0u);
    for (uint32_t t = lo; t < hi; t++) { prefix[t] = run; run += counts[t]; }
    if (threadIdx.x == 1023) {
        const uint32_t newlines = s_warp[31];
        prefix[ntiles] = newlines;
        // getline yields one line per newline plus an unterminated last one; lines 1, 5, 9, ... are sequence lines (:947)
        const unsigned long long lines = (unsigned long long)newlines + ((n && text[n - 1] != '\n') ? 1ull : 0ull);
        out[1] = (lines + 2ull) / 4ull;
    }
}

} // namespace

extern "C" int eref_census_noalloc(const void* fastq_dev, size_t n, uint64_t* out_dev, void* cuda_stream)
{
    if (!out_dev || (n && !fastq_dev)) return eref_fail(EREF_ERR_ARG, "eref_census_noalloc: null argument");
    if ((uint64_t)n > NOALLOC_MAX_BYTES)
        return eref_fail(EREF_ERR_LIMIT, "eref_census_noalloc: text of %zu bytes exceeds the fixed scratch (EREF_NOALLOC_MAX_BYTES = %llu); "
                                         "use eref_query_census_device", n, (unsigned long long)NOALLOC_MAX_BYTES);
    if (((uintptr_t)fastq_dev & 15) != 0) return eref_fail(EREF_ERR_ARG, "eref_census_noalloc: device text must be 16-byte aligned");
    if (((uintptr_t)out_dev & 7) != 0) return eref_fail(EREF_ERR_ARG, "eref_census_noalloc: out_dev must be 8-byte aligned");
    CensusScratch* s = nullptr;
    int rc = scratch_for_current_device(&s);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const uint8_t* text = static_cast<const uint8_t*>(fastq_dev);
    unsigned long long* out = reinterpret_cast<unsigned long long*>(out_dev);
    const uint64_t ntiles = ((uint64_t)n + TILE_BYTES - 1) / TILE_BYTES;
    // the array keeps the last call's contents: every word read below is written first by this call (counts and prefix
    // for tiles [0, ntiles), prefix[ntiles]); the caller's two output words are the only thing that must start at zero
    CUDA_TRY(cudaMemsetAsync(out, 0, 16, st));
    CUDA_TRY(launch_count_newlines(text, n, s->tile_counts, ntiles, st));
    k_tiles_scan_noalloc<<<1, 1024, 0, st>>>(s->tile_counts, s->tile_prefix, (uint32_t)ntiles, text, n, out);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(launch_fastq_kept(text, n, s->tile_prefix, nullptr, s->tile_kept, out, ntiles, st));
    return 0;
}
