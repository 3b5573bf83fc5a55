// table.cu — the query k-mer table (north_star parts 2+3).
//
// The reference keeps one saturating byte counter per 32-bit code, shared by the three code
// streams (bin/extract_ref.cpp:25-26, :995-997), and only ever asks "count == 3" (:531).  Here that is
// three bitmaps, level L meaning "seen at least L times".  An occurrence sets its bit in level 1; if it
// was already set it moves on to level 2, then level 3.  Of n occurrences of a code exactly one finds
// level 1 clear and exactly one of the rest finds level 2 clear, so level 3 ends up set iff n >= 3,
// whatever the order the atomics land in: the build is deterministic.  Only level 3 is ever probed.
//
// Measured on B200 (tools/atomics_bench.cu): random L2-resident atomics that return the old value top out
// at ~129 G/s whatever the occupancy, shared-memory atomics at ~1 T/s.  The build therefore never touches
// the table with global atomics: the codes are radix-partitioned by their top 13 bits (a coarse scatter
// from the bit-planes, then a fine partition staged through shared memory so that the writes stay
// coalesced), and one CTA per 2^19-slot table slice counts its codes in a 192 KiB shared-memory copy of the
// three levels and writes the finished level-3 slice (and its summary) to HBM with plain stores.
#include "eref_internal.cuh"
#include <cstdlib>
#include <algorithm>

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

// Direct build (EREF_FLAG_DIRECT_COUNT, kept as the cross-check of the partitioned build): global atomics
// straight into three 2^32-bit levels.  One thread per plane word = 32 window starts; only starts whose 32
// positions are all valid bases are visited (valid_windows).
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

// ---- partitioned build -------------------------------------------------------------------------------

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

// P0: histogram of the codes of one file over the NSLICES table slices (top 13 code bits).  Persistent grid:
// a CTA accumulates in shared memory over all its tiles and flushes once.  (Two 1024-thread CTAs per SM — a third
// of the flush atomics — measured the same 55 us per file: the flush is not what bounds it.)
constexpr int HIST_THREADS = NSLICES * 4 > 48 * 1024 ? 512 : 256;      // 64 KiB of bins: three CTAs of 512 per SM

// s0/ns: the table slices this GPU counts (sharded build, api.cu); codes of other slices are some other GPU's work.
template <bool FILTER>
__global__ void __launch_bounds__(HIST_THREADS)
k_slice_hist(Planes pl, const unsigned long long* __restrict__ plane_bits, CoderMasks m,
             unsigned long long* __restrict__ g_hist, uint32_t s0, uint32_t ns)
{
    extern __shared__ uint32_t s_hist[];                   // NSLICES
    for (int i = threadIdx.x; i < NSLICES; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const uint64_t nwords = (*plane_bits + 31) >> 5;      // the host has not seen the packed size yet
    for (uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (uint64_t)gridDim.x * blockDim.x)
        for_each_code(pl, m, w, [&](uint32_t code) {
            const uint32_t s = code >> SLICE_BITS;
            if (!FILTER || s - s0 < ns) atomicAdd(&s_hist[s], 1u);
        });
    __syncthreads();
    for (int i = threadIdx.x; i < NSLICES; i += blockDim.x) {
        const uint32_t v = s_hist[i];
        if (v) atomicAdd(g_hist + i, (unsigned long long)v);
    }
}

// P1: coarse scatter by the top `32 - shift` bits.  The CTA counts its codes per bucket, reserves one
// contiguous range per bucket with one global atomic each, then writes its codes there.  slice_off is the
// slice offset table: coarse bucket b starts where its first slice starts.
__global__ void __launch_bounds__(256)
k_coarse_scatter(Planes pl, CoderMasks m, int shift, const unsigned long long* __restrict__ slice_off,
                 unsigned long long* __restrict__ g_tail, uint32_t* __restrict__ codes_out, uint32_t s0, uint32_t ns)
{
    __shared__ uint32_t s_hist[8][MAX_BUCKETS];
    __shared__ unsigned long long s_base[MAX_BUCKETS];
    __shared__ uint32_t s_cursor[MAX_BUCKETS];
    const int warp = threadIdx.x >> 5;
    const int nb = 1 << (32 - shift);
    for (int i = threadIdx.x; i < 8 * nb; i += blockDim.x) s_hist[i / nb][i % nb] = 0;
    __syncthreads();
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = w < pl.nwords;
    if (live) for_each_code(pl, m, w, [&](uint32_t code) { if ((code >> SLICE_BITS) - s0 < ns) atomicAdd(&s_hist[warp][code >> shift], 1u); });
    __syncthreads();
    if (threadIdx.x < nb) {
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) c += s_hist[k][threadIdx.x];
        s_base[threadIdx.x] = slice_off[(size_t)threadIdx.x << (shift - SLICE_BITS)] +
                              (c ? atomicAdd(g_tail + threadIdx.x, (unsigned long long)c) : 0ull);
        s_cursor[threadIdx.x] = 0;
    }
    __syncthreads();
    if (live) for_each_code(pl, m, w, [&](uint32_t code) {
        const uint32_t b = code >> shift;
        if ((code >> SLICE_BITS) - s0 < ns) codes_out[s_base[b] + atomicAdd(&s_cursor[b], 1u)] = code;
    });
}

// P1 (staged): the same coarse scatter with the CTA's codes sorted by bucket in shared memory first, so that the
// global stores are consecutive (k_coarse_scatter issues one 4-byte store per code to up to 32 different places per
// warp, and runs at the one-sector-per-clock-per-SM store rate: 0.19 ms for 52 M codes; staged: 0.135 ms).  A CTA
// of 256 threads owns 128 plane words (a thread: 16 window starts), at most 12288 codes = 48 KiB of staging.
constexpr int CS_WORDS = 128, CS_THREADS = 256, CS_CODES = CS_WORDS * 32 * 3;

template <class F>
__device__ __forceinline__ void for_each_code_half(const Planes& pl, const CoderMasks& m, uint64_t w, int half, F&& f)
{
    uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]) & (half ? 0x0000FFFFu : 0xFFFF0000u);
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

template <bool FILTER>
__global__ void __launch_bounds__(CS_THREADS)
k_coarse_scatter_staged(Planes pl, const unsigned long long* __restrict__ plane_bits, const uint32_t* __restrict__ skip,
                        CoderMasks m, int shift, const unsigned long long* __restrict__ slice_off,
                        unsigned long long* __restrict__ g_tail, uint32_t* __restrict__ codes_out, uint32_t s0, uint32_t ns,
                        unsigned long long out_cap)
{
    // plane_bits: the packed size, when the host has not seen it (the grid then covers an upper bound); skip: set by
    // k_slice_offsets when the codes would not fit the buffers (api.cu: optimistic sizes)
    const uint64_t nwords = plane_bits ? (*plane_bits + 31) >> 5 : pl.nwords;
    if ((uint64_t)blockIdx.x * CS_WORDS >= nwords || (skip && *skip)) return;
    extern __shared__ uint32_t s_codes[];                  // CS_CODES
    __shared__ uint32_t s_cur[CS_THREADS / 32][MAX_BUCKETS];   // per warp: counts, then cursors into s_codes
    __shared__ uint32_t s_start[MAX_BUCKETS + 1];
    __shared__ unsigned long long s_base[MAX_BUCKETS];
    __shared__ uint32_t s_wsum[CS_THREADS / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = 1 << (32 - shift);
    for (int i = threadIdx.x; i < (CS_THREADS / 32) * MAX_BUCKETS; i += CS_THREADS) (&s_cur[0][0])[i] = 0;
    __syncthreads();
    const uint64_t w = (uint64_t)blockIdx.x * CS_WORDS + (threadIdx.x >> 1);
    const int half = threadIdx.x & 1;
    const bool live = w < nwords;
    if (live) for_each_code_half(pl, m, w, half, [&](uint32_t code) { if (!FILTER || (code >> SLICE_BITS) - s0 < ns) atomicAdd(&s_cur[warp][code >> shift], 1u); });
    __syncthreads();
    // bucket totals -> exclusive prefix (nb <= 256: one bucket per thread), global reservation, per-warp cursors
    uint32_t c = 0;
    if (threadIdx.x < nb) {
#pragma unroll
        for (int k = 0; k < CS_THREADS / 32; k++) c += s_cur[k][threadIdx.x];
    }
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int k = 0; k < CS_THREADS / 32; k++) if (k < warp) base += s_wsum[k];
    if (threadIdx.x < nb) {
        uint32_t run = base + inc - c;
        s_start[threadIdx.x] = run;
        s_base[threadIdx.x] = slice_off[(size_t)threadIdx.x << (shift - SLICE_BITS)] +
                              (c ? atomicAdd(g_tail + threadIdx.x, (unsigned long long)c) : 0ull);
#pragma unroll
        for (int k = 0; k < CS_THREADS / 32; k++) { const uint32_t v = s_cur[k][threadIdx.x]; s_cur[k][threadIdx.x] = run; run += v; }
    }
    if (threadIdx.x == CS_THREADS - 1) s_start[nb] = base + inc;
    __syncthreads();
    if (live) for_each_code_half(pl, m, w, half, [&](uint32_t code) {
        if (!FILTER || (code >> SLICE_BITS) - s0 < ns) s_codes[atomicAdd(&s_cur[warp][code >> shift], 1u)] = code;
    });
    __syncthreads();
    const uint32_t total = s_start[nb];
    EREF_ASSERT(total <= (uint32_t)CS_CODES);
    if (threadIdx.x < nb) s_base[threadIdx.x] -= s_start[threadIdx.x];     // element i of the sorted tile goes to s_base[b] + i
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += CS_THREADS) {
        const uint32_t cd = s_codes[i];
        EREF_ASSERT(s_base[cd >> shift] + i < out_cap);
        codes_out[s_base[cd >> shift] + i] = cd;
    }
}

// P1 (staged, codes kept in registers; EREF_COARSE_STAGED=2, a measured variant that LOST): the staged scatter above
// forms every code twice — once to count the CTA's codes per bucket, once to place them.  Here a thread owns a quarter
// word (8 window starts, 24 codes), forms its codes once into registers (fully unrolled, invalid starts predicated off)
// and uses them in both passes.  512 threads per CTA over the same 128 plane words, two CTAs per SM: 9 % slower.
constexpr int CR_THREADS = 512, CR_STARTS = 8;

template <bool FILTER>
__global__ void __launch_bounds__(CR_THREADS, 2)
k_coarse_scatter_reg(Planes pl, const unsigned long long* __restrict__ plane_bits, const uint32_t* __restrict__ skip,
                     CoderMasks m, int shift, const unsigned long long* __restrict__ slice_off,
                     unsigned long long* __restrict__ g_tail, uint32_t* __restrict__ codes_out, uint32_t s0, uint32_t ns,
                     unsigned long long out_cap)
{
    extern __shared__ uint32_t s_codes[];                  // CS_CODES
    __shared__ uint32_t s_cur[CR_THREADS / 32][MAX_BUCKETS];   // per warp: counts, then cursors into s_codes
    __shared__ uint32_t s_start[MAX_BUCKETS + 1];
    __shared__ unsigned long long s_base[MAX_BUCKETS];
    __shared__ uint32_t s_wsum[CR_THREADS / 32];
    const uint64_t nwords = plane_bits ? (*plane_bits + 31) >> 5 : pl.nwords;
    if ((uint64_t)blockIdx.x * CS_WORDS >= nwords || (skip && *skip)) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = 1 << (32 - shift);
    for (int i = threadIdx.x; i < (CR_THREADS / 32) * MAX_BUCKETS; i += CR_THREADS) (&s_cur[0][0])[i] = 0;
    __syncthreads();
    const uint64_t w = (uint64_t)blockIdx.x * CS_WORDS + (threadIdx.x >> 2);
    const int first = (threadIdx.x & 3) * CR_STARTS;
    // this thread's codes: code[3 j + i] for start first + j, stream i; bit j of vm = that start is a valid window
    uint32_t code[3 * CR_STARTS];
    uint32_t keep = 0;                                     // bit 3 j + i: the code is counted (valid, and in this GPU's slices)
    if (w < nwords) {
        const uint32_t vm = (valid_windows(pl.ok[w], pl.ok[w + 1]) >> (32 - CR_STARTS - first)) & ((1u << CR_STARTS) - 1u);
        if (vm) {
            const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
#pragma unroll
            for (int j = 0; j < CR_STARTS; j++) {
                const int s = first + j;
                const WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
                const bool v = (vm >> (CR_STARTS - 1 - j)) & 1u;
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const uint32_t c = stream_code(m, b, i);
                    code[3 * j + i] = c;
                    if (v && (!FILTER || (c >> SLICE_BITS) - s0 < ns)) keep |= 1u << (3 * j + i);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 3 * CR_STARTS; q++)
        if ((keep >> q) & 1u) atomicAdd(&s_cur[warp][code[q] >> shift], 1u);
    __syncthreads();
    // bucket totals -> exclusive prefix (nb <= 256: one bucket per thread), global reservation, per-warp cursors
    uint32_t c = 0;
    if (threadIdx.x < nb) {
#pragma unroll
        for (int k = 0; k < CR_THREADS / 32; k++) c += s_cur[k][threadIdx.x];
    }
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int k = 0; k < CR_THREADS / 32; k++) if (k < warp) base += s_wsum[k];
    if (threadIdx.x < nb) {
        uint32_t run = base + inc - c;
        s_start[threadIdx.x] = run;
        s_base[threadIdx.x] = slice_off[(size_t)threadIdx.x << (shift - SLICE_BITS)] +
                              (c ? atomicAdd(g_tail + threadIdx.x, (unsigned long long)c) : 0ull);
#pragma unroll
        for (int k = 0; k < CR_THREADS / 32; k++) { const uint32_t v = s_cur[k][threadIdx.x]; s_cur[k][threadIdx.x] = run; run += v; }
    }
    // (threads nb .. 255 of the first 8 warps hold c = 0: the running sum of the last of them is the CTA's total)
    if (threadIdx.x == MAX_BUCKETS - 1) s_start[nb] = base + inc;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 3 * CR_STARTS; q++)
        if ((keep >> q) & 1u) s_codes[atomicAdd(&s_cur[warp][code[q] >> shift], 1u)] = code[q];
    __syncthreads();
    const uint32_t total = s_start[nb];
    EREF_ASSERT(total <= (uint32_t)CS_CODES);
    if (threadIdx.x < nb) s_base[threadIdx.x] -= s_start[threadIdx.x];     // element i of the sorted tile goes to s_base[b] + i
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += CR_THREADS) {
        const uint32_t cd = s_codes[i];
        EREF_ASSERT(s_base[cd >> shift] + i < out_cap);
        codes_out[s_base[cd >> shift] + i] = cd;
    }
}

// Slice offsets of one file from its slice histogram, on the device (no host round trip, no H2D copy that
// would queue behind the database upload): off[s] = start of slice s (padded to 4 codes so that every slice
// starts 16-byte aligned), cnt = copy of the histogram, coarse_n[b] = codes of coarse bucket b,
// chunk_start[b] = first fine-partition chunk of bucket b, totals = {codes, padded size, chunks}.

__global__ void __launch_bounds__(1024)
k_slice_offsets(const unsigned long long* __restrict__ hist, int cbits, unsigned long long* __restrict__ off,
                unsigned long long* __restrict__ cnt, unsigned long long* __restrict__ coarse_n,
                uint32_t* __restrict__ chunk_start, unsigned long long* __restrict__ totals,
                unsigned long long cap_codes, uint32_t* __restrict__ overflow)
{
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_padded;
    __shared__ unsigned long long s_coarse[MAX_BUCKETS];
    constexpr int PER = NSLICES / 1024;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long v[PER], sum = 0;
#pragma unroll
    for (int k = 0; k < PER; k++) { v[k] = hist[threadIdx.x * PER + k]; cnt[threadIdx.x * PER + k] = v[k]; sum += (v[k] + 3) & ~3ull; }
    unsigned long long inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { unsigned long long t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned long long base = 0;
    for (int k = 0; k < warp; k++) base += s_warp[k];
    unsigned long long run = base + inc - sum;
#pragma unroll
    for (int k = 0; k < PER; k++) { off[threadIdx.x * PER + k] = run; run += (v[k] + 3) & ~3ull; }
    if (threadIdx.x == 1023) { off[NSLICES] = run; s_padded = run; }
    // codes per coarse bucket: a bucket's 2^fbits slices sit in 2^fbits / PER neighbouring threads (at most one warp for
    // cbits >= 32 - SLICE_BITS - 8), so it is a segmented shuffle sum of what the threads already hold
    const int fbits = 32 - SLICE_BITS - cbits, NC = 1 << cbits;
    const int tpb = (1 << fbits) / PER;                    // threads per bucket: a power of two, 1..32
    unsigned long long raw = 0;
#pragma unroll
    for (int k = 0; k < PER; k++) raw += v[k];
    for (int d = 1; d < tpb; d <<= 1) raw += __shfl_xor_sync(0xffffffffu, raw, d);
    if ((threadIdx.x & (tpb - 1)) == 0) {
        coarse_n[threadIdx.x / tpb] = raw;
        s_coarse[threadIdx.x / tpb] = raw;
    }
    __syncthreads();
    // cap_codes: what the code buffers hold when they were sized before this count was known.  A file that needs more
    // (or follows one that did) is not partitioned at all — no chunks, the scatter skips, its slices count as empty —
    // and the host is told.
    const bool over = cap_codes && (s_padded > cap_codes || (overflow && *overflow));
    if (over) {
#pragma unroll
        for (int k = 0; k < PER; k++) cnt[threadIdx.x * PER + k] = 0;
    }
    if (threadIdx.x == 0) {
        unsigned long long total = 0;
        uint32_t chunks = 0;
        for (int b = 0; b < NC; b++) {
            chunk_start[b] = chunks;
            chunks += (uint32_t)((s_coarse[b] + FINE_CHUNK - 1) / FINE_CHUNK);
            total += s_coarse[b];
        }
        if (over) { chunks = 0; if (overflow) *overflow = 1u; }
        chunk_start[NC] = chunks;
        totals[0] = total; totals[2] = chunks;
    }
    if (threadIdx.x == 1023) totals[1] = run;
}

// P2: fine partition of one coarse bucket's codes into its 2^fbits slices.  A CTA takes a chunk of
// FINE_CHUNK codes, ranks them per slice with shared-memory atomics, reserves one range per slice, sorts the
// chunk by slice in shared memory and writes it out, consecutive threads to consecutive addresses.
template <int FINE_THREADS>
__global__ void __launch_bounds__(FINE_THREADS, 2048 / FINE_THREADS >= 3 ? 3 : 2)
k_fine_partition(const uint32_t* __restrict__ in, const unsigned long long* __restrict__ slice_off,
                 const unsigned long long* __restrict__ coarse_n, const uint32_t* __restrict__ chunk_start,
                 int cbits, unsigned long long* __restrict__ g_tail, uint32_t* __restrict__ out, unsigned long long out_cap)
{
    __shared__ uint32_t s_stage[FINE_CHUNK];
    __shared__ uint32_t s_cnt[MAX_BUCKETS], s_off[MAX_BUCKETS + 1];
    __shared__ uint32_t s_delta[MAX_BUCKETS];
    __shared__ uint32_t s_warp[8];
    __shared__ int s_bucket;
    constexpr int FINE_PER_THREAD = FINE_CHUNK / FINE_THREADS;
    const int fbits = 32 - SLICE_BITS - cbits, nf = 1 << fbits, ncoarse = 1 << cbits;
    if (blockIdx.x >= chunk_start[ncoarse]) return;                         // the grid is sized from an upper bound
    // which coarse bucket does this chunk belong to?  chunk_start is non-decreasing: b = #{i : chunk_start[i] <= chunk} - 1,
    // counted by the first warp with one load per lane and round and handed over at the barrier that follows the
    // zeroing of the counters (a binary search here was five dependent loads in every thread before the first code load)
    if (threadIdx.x < 32) {
        int below = 0;
        for (int i0 = 0; i0 < ncoarse; i0 += 32) {
            const int i = i0 + (int)threadIdx.x;
            below += __popc(__ballot_sync(0xffffffffu, i < ncoarse && chunk_start[i] <= blockIdx.x));
        }
        if (threadIdx.x == 0) s_bucket = below - 1;
    }
    if (threadIdx.x < nf) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int b = s_bucket;
    const uint64_t first = (uint64_t)(blockIdx.x - chunk_start[b]) * FINE_CHUNK;
    const uint64_t nb = coarse_n[b];
    if (first >= nb) return;
    const uint32_t cnt = (uint32_t)min((uint64_t)FINE_CHUNK, nb - first);
    const uint32_t* src = in + slice_off[(size_t)b << fbits] + first;       // 16-byte aligned: slice starts and chunks are
    uint32_t code[FINE_PER_THREAD];
    uint16_t rank[FINE_PER_THREAD];
#pragma unroll
    for (int j = 0; j < FINE_PER_THREAD / 4; j++) {
        const uint32_t u = (uint32_t)(j * FINE_THREADS + threadIdx.x);      // 16-byte unit inside the chunk
        uint4 v = make_uint4(0, 0, 0, 0);
        if (u * 4 < cnt) v = __ldcs(reinterpret_cast<const uint4*>(src) + u);
        code[4 * j] = v.x; code[4 * j + 1] = v.y; code[4 * j + 2] = v.z; code[4 * j + 3] = v.w;
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (u * 4 + k < cnt) rank[4 * j + k] = (uint16_t)atomicAdd(&s_cnt[(code[4 * j + k] >> SLICE_BITS) & (nf - 1)], 1u);
    }
    __syncthreads();
    // exclusive scan of the per-slice counts (nf <= 256: one value per thread of the first 8 warps) + global reservation
    if (threadIdx.x < 256) {
        const uint32_t v = threadIdx.x < nf ? s_cnt[threadIdx.x] : 0u;
        uint32_t inc = v;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) s_warp[warp] = inc;
        asm volatile("bar.sync 1, 256;");                                   // the first 8 warps only
        uint32_t base = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) if (k < warp) base += s_warp[k];
        if (threadIdx.x < nf) {
            s_off[threadIdx.x] = base + inc - v;
            const size_t slice = ((size_t)b << fbits) + threadIdx.x;
            // where element i of the sorted chunk goes, relative to the bucket's first code: s_delta[f] + i
            s_delta[threadIdx.x] = (uint32_t)(slice_off[slice] - slice_off[(size_t)b << fbits] +
                                              (v ? atomicAdd(g_tail + slice, (unsigned long long)v) : 0ull)) - (base + inc - v);
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < FINE_PER_THREAD / 4; j++) {
        const uint32_t u = (uint32_t)(j * FINE_THREADS + threadIdx.x);
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (u * 4 + k < cnt) {
                EREF_ASSERT(s_off[(code[4 * j + k] >> SLICE_BITS) & (nf - 1)] + rank[4 * j + k] < (uint32_t)FINE_CHUNK);
                s_stage[s_off[(code[4 * j + k] >> SLICE_BITS) & (nf - 1)] + rank[4 * j + k]] = code[4 * j + k];
            }
    }
    __syncthreads();
    uint32_t* out_b = out + slice_off[(size_t)b << fbits];
    for (uint32_t i = threadIdx.x; i < cnt; i += FINE_THREADS) {
        const uint32_t c = s_stage[i];
        EREF_ASSERT(slice_off[(size_t)b << fbits] + s_delta[(c >> SLICE_BITS) & (nf - 1)] + i < out_cap);
        out_b[s_delta[(c >> SLICE_BITS) & (nf - 1)] + i] = c;
    }
}

// P3: one CTA per table slice.  The three levels of the slice live in shared memory (3 x 64 KiB); the codes
// of the slice from every pending file are counted with shared-memory atomics; the finished level-3 slice
// and its part of the summary (1 bit per 16 slots) go to HBM with plain, coalesced stores.  Also
// accumulates the number of codes at each level (eref_query_popcount).
constexpr int COUNT_THREADS = SLICE_BITS >= 19 ? 1024 : 512;     // 192 KiB of levels: one CTA per SM; 96 KiB: two
constexpr int COUNT_CTAS = SLICE_BITS >= 19 ? 1 : 2;
constexpr int COUNT_PRE = 2;                 // 16-byte units per thread and list fetched before the levels are zeroed

__device__ __forceinline__ void count_unit(const uint4& v, uint32_t left, uint32_t* L1, uint32_t* L2, uint32_t* L3)
{
    const uint32_t code[4] = {v.x, v.y, v.z, v.w};
    uint32_t old[4];
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (k < left) old[k] = atomicOr(&L1[(code[k] & SLICE_MASK) >> 5], 1u << (code[k] & 31));
#pragma unroll
    for (int k = 0; k < 4; k++)
        if (k < left) {
            const uint32_t bit = 1u << (code[k] & 31), w = (code[k] & SLICE_MASK) >> 5;
            if (old[k] & bit) {
                const uint32_t old2 = atomicOr(&L2[w], bit);
                if (old2 & bit) atomicOr(&L3[w], bit);
            }
        }
}

template <bool POP>
__global__ void __launch_bounds__(COUNT_THREADS, 1)
k_count_slices(CountLists lists, uint32_t* __restrict__ bm3, uint32_t* __restrict__ summary,
               unsigned long long* __restrict__ level_pop, uint32_t s0)
{
    extern __shared__ uint4 s_raw[];
    uint32_t* L1 = reinterpret_cast<uint32_t*>(s_raw);
    uint32_t* L2 = L1 + SLICE_WORDS;
    uint32_t* L3 = L2 + SLICE_WORDS;
    const uint32_t slice = s0 + blockIdx.x;
    // the first codes of the first two lists are requested before the levels are zeroed, so that the HBM round
    // trip overlaps the shared-memory stores
    uint4 pre[2][COUNT_PRE];
    uint64_t n_pre[2] = {0, 0};
#pragma unroll
    for (int l = 0; l < 2; l++)
        if (l < lists.n) {
            n_pre[l] = lists.cnt[l][slice];
            const uint4* src = reinterpret_cast<const uint4*>(lists.codes[l] + lists.off[l][slice]);   // 16-byte aligned, padded to 4
#pragma unroll
            for (int j = 0; j < COUNT_PRE; j++) {
                const uint64_t u = threadIdx.x + (uint64_t)j * COUNT_THREADS;
                pre[l][j] = u * 4 < n_pre[l] ? __ldcs(src + u) : make_uint4(0, 0, 0, 0);
            }
        }
    for (int i = threadIdx.x; i < 3 * SLICE_WORDS / 4; i += COUNT_THREADS) s_raw[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
#pragma unroll
    for (int l = 0; l < 2; l++)
#pragma unroll
        for (int j = 0; j < COUNT_PRE; j++) {
            const uint64_t u = threadIdx.x + (uint64_t)j * COUNT_THREADS;
            if (u * 4 < n_pre[l]) count_unit(pre[l][j], (uint32_t)min((uint64_t)4, n_pre[l] - u * 4), L1, L2, L3);
        }
    for (int l = 0; l < lists.n; l++) {
        const uint64_t n = lists.cnt[l][slice];
        const uint4* src = reinterpret_cast<const uint4*>(lists.codes[l] + lists.off[l][slice]);
        for (uint64_t u = threadIdx.x + (l < 2 ? (uint64_t)COUNT_PRE * COUNT_THREADS : 0ull); u * 4 < n; u += COUNT_THREADS)
            count_unit(__ldcs(src + u), (uint32_t)min((uint64_t)4, n - u * 4), L1, L2, L3);
    }
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(bm3 + (size_t)slice * SLICE_WORDS);
    const uint4* l3 = reinterpret_cast<const uint4*>(L3);
    for (int i = threadIdx.x; i < SLICE_WORDS / 4; i += COUNT_THREADS) __stcs(dst + i, l3[i]);
    // summary word j of the slice covers table words 16j .. 16j+15
    for (int j = threadIdx.x; j < SLICE_WORDS / 16; j += COUNT_THREADS) {
        uint32_t out = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint4 v = l3[j * 4 + q];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int word = q * 4 + k;
                out |= (uint32_t)((w[k] & 0x0000FFFFu) != 0) << (word * 2);
                out |= (uint32_t)((w[k] & 0xFFFF0000u) != 0) << (word * 2 + 1);
            }
        }
        summary[(size_t)slice * (SLICE_WORDS / 16) + j] = out;
    }
    if (POP) {
        uint32_t p1 = 0, p2 = 0, p3 = 0;
        for (int i = threadIdx.x; i < SLICE_WORDS; i += COUNT_THREADS) { p1 += __popc(L1[i]); p2 += __popc(L2[i]); p3 += __popc(L3[i]); }
        p1 = __reduce_add_sync(0xffffffffu, p1); p2 = __reduce_add_sync(0xffffffffu, p2); p3 = __reduce_add_sync(0xffffffffu, p3);
        if ((threadIdx.x & 31) == 0) {
            if (p1) atomicAdd(level_pop + 0, (unsigned long long)p1);
            if (p2) atomicAdd(level_pop + 1, (unsigned long long)p2);
            if (p3) atomicAdd(level_pop + 2, (unsigned long long)p3);
        }
    }
}

// P3 (persistent): one CTA per SM walks the slices blockIdx.x, blockIdx.x + gridDim.x, ...  While a slice is being
// counted and written out, the first codes of the NEXT slice are already on their way into registers, so the HBM
// round trip and the CTA turn-around of k_count_slices (about a third of its 8.8 us per slice) disappear.
//
// Sharded build (TableShare, eref_internal.cuh): the CTAs walk only this GPU's slices [s0, s0 + ns) and store every
// finished level-3 slice and its summary into the table of EVERY GPU of the build — plain 16-byte stores through
// peer pointers, i.e. the all-gather of the table rides on the copy-out over NVLink, slice by slice, while the
// other SMs are still counting.  The last CTA to finish publishes the build's epoch in every GPU's flag word
// (release at system scope); k_table_wait on the consumer's stream acquires all of them before the first probe.
template <bool POP, bool SHARED>
__global__ void __launch_bounds__(COUNT_THREADS, COUNT_CTAS)
k_count_slices_persistent(CountLists lists, TableShare share, unsigned long long* __restrict__ level_pop)
{
    extern __shared__ uint4 s_raw[];
    uint32_t* L1 = reinterpret_cast<uint32_t*>(s_raw);
    uint32_t* L2 = L1 + SLICE_WORDS;
    uint32_t* L3 = L2 + SLICE_WORDS;
    uint4 pre[2][COUNT_PRE];
    uint64_t n_pre[2] = {0, 0};
    const uint32_t slice_end = share.s0 + share.ns;
    auto fetch = [&](uint32_t slice) {
#pragma unroll
        for (int l = 0; l < 2; l++) {
            n_pre[l] = 0;
            if (l < lists.n && slice < slice_end) {
                n_pre[l] = lists.cnt[l][slice];
                EREF_ASSERT(n_pre[l] == 0 || lists.off[l][slice] + n_pre[l] <= lists.cap[l]);
                const uint4* src = reinterpret_cast<const uint4*>(lists.codes[l] + lists.off[l][slice]);
#pragma unroll
                for (int j = 0; j < COUNT_PRE; j++) {
                    const uint64_t u = threadIdx.x + (uint64_t)j * COUNT_THREADS;
                    pre[l][j] = u * 4 < n_pre[l] ? __ldcs(src + u) : make_uint4(0, 0, 0, 0);
                }
            }
        }
    };
    uint32_t p1 = 0, p2 = 0, p3 = 0;
    fetch(share.s0 + blockIdx.x);
    for (int i = threadIdx.x; i < 3 * SLICE_WORDS / 4; i += COUNT_THREADS) s_raw[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    for (uint32_t slice = share.s0 + blockIdx.x; slice < slice_end; slice += gridDim.x) {
        // the prefetched units of this slice, then whatever the registers could not hold
        uint4 cur[2][COUNT_PRE];
        uint64_t n_cur[2];
#pragma unroll
        for (int l = 0; l < 2; l++) {
            n_cur[l] = n_pre[l];
#pragma unroll
            for (int j = 0; j < COUNT_PRE; j++) cur[l][j] = pre[l][j];
        }
        fetch(slice + gridDim.x);                          // next slice: in flight during the atomics and the write-out
#pragma unroll
        for (int l = 0; l < 2; l++)
#pragma unroll
            for (int j = 0; j < COUNT_PRE; j++) {
                const uint64_t u = threadIdx.x + (uint64_t)j * COUNT_THREADS;
                if (u * 4 < n_cur[l]) count_unit(cur[l][j], (uint32_t)min((uint64_t)4, n_cur[l] - u * 4), L1, L2, L3);
            }
        for (int l = 0; l < lists.n; l++) {
            const uint64_t n = lists.cnt[l][slice];
            const uint4* src = reinterpret_cast<const uint4*>(lists.codes[l] + lists.off[l][slice]);
            for (uint64_t u = threadIdx.x + (l < 2 ? (uint64_t)COUNT_PRE * COUNT_THREADS : 0ull); u * 4 < n; u += COUNT_THREADS)
                count_unit(__ldcs(src + u), (uint32_t)min((uint64_t)4, n - u * 4), L1, L2, L3);
        }
        __syncthreads();
        if (POP) {
            for (int i = threadIdx.x; i < SLICE_WORDS; i += COUNT_THREADS) { p1 += __popc(L1[i]); p2 += __popc(L2[i]); p3 += __popc(L3[i]); }
            __syncthreads();                               // the copy-out below clears words other threads count here
        }
        // level 3 out, one 16-byte unit per thread and round; summary word j covers table words 16j .. 16j+15 = four
        // consecutive units, i.e. four neighbouring lanes: 8 summary bits each, combined with two shuffles
        const size_t dst_unit = (size_t)slice * (SLICE_WORDS / 4);
        const uint4* l3 = reinterpret_cast<const uint4*>(L3);
        uint4* l3w = reinterpret_cast<uint4*>(L3);
        uint4* l1w = reinterpret_cast<uint4*>(L1);
        uint4* l2w = reinterpret_cast<uint4*>(L2);
        for (int i = threadIdx.x; i < SLICE_WORDS / 4; i += COUNT_THREADS) {
            const uint4 v = l3[i];
            __stcs(reinterpret_cast<uint4*>(share.bm3[0]) + dst_unit + i, v);
            if (SHARED) for (int p = 1; p < share.n; p++) __stcs(reinterpret_cast<uint4*>(share.bm3[p]) + dst_unit + i, v);
            // the three levels are zeroed for the next slice as they are read out (a thread clears what it alone read)
            l3w[i] = make_uint4(0, 0, 0, 0); l1w[i] = make_uint4(0, 0, 0, 0); l2w[i] = make_uint4(0, 0, 0, 0);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            // "half-word != 0" lands in bits 15 and 31 of t (no carry crosses a half), word k's pair goes to bits 2k and 16+2k
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const uint32_t t = (((w[k] & 0x7FFF7FFFu) + 0x7FFF7FFFu) | w[k]) & 0x80008000u;
                acc |= t >> (15 - 2 * k);
            }
            uint32_t out = (acc | (acc >> 15)) & 0xFFu;
            out <<= 8 * (i & 3);
            out |= __shfl_xor_sync(0xffffffffu, out, 1);
            out |= __shfl_xor_sync(0xffffffffu, out, 2);
            if ((i & 3) == 0) {
                share.summary[0][(size_t)slice * (SLICE_WORDS / 16) + (i >> 2)] = out;
                if (SHARED) for (int p = 1; p < share.n; p++) share.summary[p][(size_t)slice * (SLICE_WORDS / 16) + (i >> 2)] = out;
            }
        }
        __syncthreads();
    }
    if (SHARED && share.n > 1) {
        // publish: every thread's peer stores are ordered before the CTA's arrival; the last CTA to arrive raises this
        // part's flag on every GPU
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            if (atomicAdd(share.done, 1u) == gridDim.x - 1) {
                __threadfence_system();
                for (int p = 0; p < share.n; p++)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(share.flags[p] + share.part), "r"(share.epoch) : "memory");
            }
        }
    }
    if (POP) {
        p1 = __reduce_add_sync(0xffffffffu, p1); p2 = __reduce_add_sync(0xffffffffu, p2); p3 = __reduce_add_sync(0xffffffffu, p3);
        if ((threadIdx.x & 31) == 0) {
            if (p1) atomicAdd(level_pop + 0, (unsigned long long)p1);
            if (p2) atomicAdd(level_pop + 1, (unsigned long long)p2);
            if (p3) atomicAdd(level_pop + 2, (unsigned long long)p3);
        }
    }
}

// Consumer side of the sharded build: lane q waits until part q of build `epoch` has landed in this GPU's table
// (flag >= epoch, acquire at system scope).  Bounded: after ~4 s without progress the error word is set and the
// kernel returns, so a peer that died cannot hang the stream.
__global__ void k_table_wait(const uint32_t* __restrict__ flags, int n, uint32_t epoch, uint32_t* __restrict__ err)
{
    if ((int)threadIdx.x >= n) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
        if ((int32_t)(v - epoch) >= 0) return;
        __nanosleep(200);
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 4000000000ull) { atomicExch(err, 1u + threadIdx.x); return; }
    }
}

// ---- shared reads: every GPU partitions its share of the reads; a build pulls its slices out of every peer's lists ----
// "My lists of build `epoch` are complete": raise this GPU's word in every GPU's lists-ready flags (release, system scope).
__global__ void k_lists_signal(TableShare share, uint32_t flag0, uint32_t epoch)
{
    if (threadIdx.x >= (unsigned)share.n) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(share.flags[threadIdx.x] + flag0 + share.part), "r"(epoch) : "memory");
}

// One peer list (blockIdx.y) -> local staging: the codes of the slices [s0, s0 + ns) are ONE contiguous range of a
// slice-ordered list (every slice starts 16-byte aligned), copied with 16-byte loads over NVLink; the offsets are rebased
// onto the staging copy, the counts copied.  Lists use the arena layout off[NSLICES + 1] | cnt[NSLICES] | codes[cap].
__global__ void __launch_bounds__(256)
k_pull_lists(PullLists pl)
{
    const int slot = blockIdx.y;
    const uint8_t* src = pl.src[slot];
    uint8_t* dst = pl.dst + (size_t)slot * pl.list_stride;
    const unsigned long long* soff = reinterpret_cast<const unsigned long long*>(src);
    const unsigned long long* scnt = soff + NSLICES + 1;
    const uint4* scodes = reinterpret_cast<const uint4*>(src + ((size_t)2 * NSLICES + 2) * 8);
    unsigned long long* doff = reinterpret_cast<unsigned long long*>(dst);
    unsigned long long* dcnt = doff + NSLICES + 1;
    uint4* dcodes = reinterpret_cast<uint4*>(dst + ((size_t)2 * NSLICES + 2) * 8);
    const unsigned long long lo = soff[pl.s0], hi = soff[pl.s0 + pl.ns];
    EREF_ASSERT((lo & 3) == 0 && hi >= lo && hi - lo <= pl.cap_codes);
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nthreads = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long s = tid; s <= pl.ns; s += nthreads) {
        doff[pl.s0 + s] = soff[pl.s0 + s] - lo;
        if (s < pl.ns) dcnt[pl.s0 + s] = scnt[pl.s0 + s];
    }
    const unsigned long long units = (hi - lo + 3) >> 2;
    for (unsigned long long u = tid; u < units; u += nthreads) dcodes[u] = scodes[(lo >> 2) + u];
}

// summary bit = OR of 16 consecutive level-3 bits: a 32 MiB filter that stays in L2 during the scan
// (direct build only; the partitioned build writes the summary from shared memory)
__global__ void __launch_bounds__(256)
k_build_summary(const uint32_t* __restrict__ bm3, uint32_t* __restrict__ summary, uint32_t summary_words)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;          // summary word: 16 table words
    if (i >= summary_words) return;
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

cudaError_t launch_slice_hist(Planes pl, const unsigned long long* plane_bits, CoderMasks m, unsigned long long* g_hist,
                              uint32_t s0, uint32_t ns, cudaStream_t st)
{
    const int smem = NSLICES * 4;
    cudaError_t e = cudaFuncSetAttribute(k_slice_hist<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    const bool filter = !(s0 == 0 && ns >= (uint32_t)NSLICES);
    cudaError_t e2 = filter ? cudaFuncSetAttribute(k_slice_hist<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) : cudaSuccess;
    if (e2 != cudaSuccess) return e2;
    if (filter) k_slice_hist<true><<<148 * 1536 / HIST_THREADS, HIST_THREADS, smem, st>>>(pl, plane_bits, m, g_hist, s0, ns);
    else k_slice_hist<false><<<148 * 1536 / HIST_THREADS, HIST_THREADS, smem, st>>>(pl, plane_bits, m, g_hist, s0, ns);
    return cudaGetLastError();
}

cudaError_t launch_coarse_scatter(Planes pl, const unsigned long long* plane_bits, const uint32_t* skip, CoderMasks m, int cbits,
                                  const unsigned long long* slice_off, unsigned long long* g_tail, uint32_t* codes_out,
                                  unsigned long long out_cap, uint32_t s0, uint32_t ns, cudaStream_t st)
{
    // pl.nwords: the exact word count, or (plane_bits given) an upper bound that sizes the grid
    if (!pl.nwords) return cudaSuccess;
    // 1 (default): staged scatter; 2: the register-cached variant — measured SLOWER (B200, bench workload: bucket phases
    // 0.896 ms against 0.823 ms; configs[3]: 8.05 against 7.31 ms): 64 registers and two 512-thread CTAs per SM lose more
    // than forming each code once saves; 0: unstaged
    static const int staged = [] { const char* v = getenv("EREF_COARSE_STAGED"); return v ? atoi(v) : 1; }();
    if (staged == 2) {                                    // codes formed once, kept in registers between the two passes
        const int smem = CS_CODES * 4;
        const bool filter = !(s0 == 0 && ns >= (uint32_t)NSLICES);
        cudaError_t e = filter ? cudaFuncSetAttribute(k_coarse_scatter_reg<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                               : cudaFuncSetAttribute(k_coarse_scatter_reg<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        const unsigned grid = (unsigned)((pl.nwords + CS_WORDS - 1) / CS_WORDS);
        if (filter) k_coarse_scatter_reg<true><<<grid, CR_THREADS, smem, st>>>(pl, plane_bits, skip, m, 32 - cbits, slice_off, g_tail, codes_out, s0, ns, out_cap);
        else k_coarse_scatter_reg<false><<<grid, CR_THREADS, smem, st>>>(pl, plane_bits, skip, m, 32 - cbits, slice_off, g_tail, codes_out, s0, ns, out_cap);
        return cudaGetLastError();
    }
    if (staged) {
        const int smem = CS_CODES * 4;
        const bool filter = !(s0 == 0 && ns >= (uint32_t)NSLICES);
        cudaError_t e = filter ? cudaFuncSetAttribute(k_coarse_scatter_staged<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                               : cudaFuncSetAttribute(k_coarse_scatter_staged<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        const unsigned grid = (unsigned)((pl.nwords + CS_WORDS - 1) / CS_WORDS);
        if (filter) k_coarse_scatter_staged<true><<<grid, CS_THREADS, smem, st>>>(pl, plane_bits, skip, m, 32 - cbits, slice_off, g_tail, codes_out, s0, ns, out_cap);
        else k_coarse_scatter_staged<false><<<grid, CS_THREADS, smem, st>>>(pl, plane_bits, skip, m, 32 - cbits, slice_off, g_tail, codes_out, s0, ns, out_cap);
    } else {
        if (plane_bits) return cudaErrorNotSupported;     // the unstaged variant needs the exact size
        k_coarse_scatter<<<(unsigned)((pl.nwords + 255) / 256), 256, 0, st>>>(pl, m, 32 - cbits, slice_off, g_tail, codes_out, s0, ns);
    }
    return cudaGetLastError();
}

cudaError_t launch_slice_offsets(const unsigned long long* hist, int cbits, unsigned long long* off, unsigned long long* cnt,
                                 unsigned long long* coarse_n, uint32_t* chunk_start, unsigned long long* totals,
                                 unsigned long long cap_codes, uint32_t* overflow, cudaStream_t st)
{
    k_slice_offsets<<<1, 1024, 0, st>>>(hist, cbits, off, cnt, coarse_n, chunk_start, totals, cap_codes, overflow);
    return cudaGetLastError();
}

cudaError_t launch_fine_partition(const uint32_t* in, const unsigned long long* slice_off, const unsigned long long* coarse_n,
                                  const uint32_t* chunk_start, uint32_t nchunks, int cbits, unsigned long long* g_tail,
                                  uint32_t* out, unsigned long long out_cap, cudaStream_t st)
{
    if (!nchunks) return cudaSuccess;
    static const int threads = [] { const char* v = getenv("EREF_FINE_THREADS"); return v ? atoi(v) : 1024; }();
    if (threads == 512) k_fine_partition<512><<<nchunks, 512, 0, st>>>(in, slice_off, coarse_n, chunk_start, cbits, g_tail, out, out_cap);
    else if (threads == 256) k_fine_partition<256><<<nchunks, 256, 0, st>>>(in, slice_off, coarse_n, chunk_start, cbits, g_tail, out, out_cap);
    else k_fine_partition<1024><<<nchunks, 1024, 0, st>>>(in, slice_off, coarse_n, chunk_start, cbits, g_tail, out, out_cap);
    return cudaGetLastError();
}

cudaError_t launch_count_slices(const CountLists& lists, const TableShare& share, unsigned long long* level_pop,
                                cudaStream_t st)
{
    const int smem = 3 * SLICE_WORDS * 4;
    static const int persistent = [] { const char* v = getenv("EREF_COUNT_PERSISTENT"); return v ? atoi(v) : 1; }();
    if (!share.ns) return cudaSuccess;
    if (persistent || share.n > 1) {
        static int sms = 0;
        if (!sms) {
            int dev = 0;
            sms = 148;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        }
        const bool shared = share.n > 1;
        void (*kern)(CountLists, TableShare, unsigned long long*) =
            level_pop ? (shared ? k_count_slices_persistent<true, true> : k_count_slices_persistent<true, false>)
                      : (shared ? k_count_slices_persistent<false, true> : k_count_slices_persistent<false, false>);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (shared) { e = cudaMemsetAsync(share.done, 0, 4, st); if (e != cudaSuccess) return e; }
        const unsigned grid = (unsigned)std::min<uint32_t>((uint32_t)(sms * COUNT_CTAS), share.ns);
        kern<<<grid, COUNT_THREADS, smem, st>>>(lists, share, level_pop);
        return cudaGetLastError();
    }
    cudaError_t e = level_pop ? cudaFuncSetAttribute(k_count_slices<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                              : cudaFuncSetAttribute(k_count_slices<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    if (level_pop) k_count_slices<true><<<share.ns, COUNT_THREADS, smem, st>>>(lists, share.bm3[0], share.summary[0], level_pop, share.s0);
    else k_count_slices<false><<<share.ns, COUNT_THREADS, smem, st>>>(lists, share.bm3[0], share.summary[0], nullptr, share.s0);
    return cudaGetLastError();
}

// With lazy module loading (the CUDA 12 default) a kernel is loaded at its first launch, which may synchronise the
// context: a first launch issued while k_table_wait is already spinning on this GPU would wait for the spin, and the
// spin for a peer.  Loading the kernels of the shared build up front keeps the host out of that wait.
cudaError_t preload_share_kernels()
{
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, k_table_wait);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_lists_signal);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_pull_lists);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_count_slices_persistent<false, true>);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_count_slices_persistent<true, true>);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_count_slices_persistent<false, false>);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, k_count_slices_persistent<true, false>);
    return e;
}

cudaError_t launch_lists_signal(const TableShare& share, uint32_t flag0, uint32_t epoch, cudaStream_t st)
{
    k_lists_signal<<<1, 32, 0, st>>>(share, flag0, epoch);
    return cudaGetLastError();
}

cudaError_t launch_pull_lists(const PullLists& pl, cudaStream_t st)
{
    if (!pl.n || !pl.ns) return cudaSuccess;
    k_pull_lists<<<dim3(148 * 2, (unsigned)pl.n), 256, 0, st>>>(pl);
    return cudaGetLastError();
}

cudaError_t launch_table_wait(const uint32_t* flags, int n, uint32_t epoch, uint32_t* err, cudaStream_t st)
{
    k_table_wait<<<1, 32, 0, st>>>(flags, n, epoch, err);
    return cudaGetLastError();
}

cudaError_t launch_build_summary(const uint32_t* bm3, uint32_t* summary, uint32_t summary_words, cudaStream_t st)
{
    k_build_summary<<<(summary_words + 255) / 256, 256, 0, st>>>(bm3, summary, summary_words);
    return cudaGetLastError();
}

cudaError_t launch_popcount(const uint32_t* bm, uint64_t nwords, unsigned long long* out, cudaStream_t st)
{
    k_popcount<<<148 * 8, 256, 0, st>>>(bm, nwords, out);
    return cudaGetLastError();
}
