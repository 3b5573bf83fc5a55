// scan.cu — the streaming reference scan (north_star part 4): probe the query table at every
// reference position, apply the 500-position window rule, and reduce each record to
// {ordinal, frag, el, ref_len}.  Replaces read_index (:813-903) + slide_window (:504-624).
//
// Exact candidate filter.  A window is "good" only if at least three_min of its 500 positions are
// hit by ALL three code streams (:561), hence by stream 0.  Probing stream 0 at every 4th position
// bounds that count from above: trio(window) <= sampled_hits(window) + 375.  Words whose bound stays
// below three_min can never be good, so the full three-stream probe runs only on the words that can
// influence a window that still might be.  The selected records and their numbers are unchanged;
// random-sector probes drop from 3 per base to ~0.25 per base plus 3 per base on candidate regions.
#include "eref_internal.cuh"

namespace {

__device__ __forceinline__ uint32_t probe(const uint32_t* __restrict__ bm3, uint32_t code)
{
    // code 0 is the index's "no k-mer" sentinel (:733-735, :861-866): never a hit
    return code ? (__ldg(bm3 + (code >> 5)) >> (code & 31)) & 1u : 0u;
}

// ---- P1: stream 0 at every STRIDE-th plane position -------------------------------------------------
// Sampled positions are the multiples of STRIDE in plane coordinates (a record is contiguous in the planes, so
// any 500 consecutive positions of a record hold at least floor(500/STRIDE) of them).  The 32 MiB level-3
// summary (1 bit per 16 slots, table.cu) answers most probes of an absent reference from L2; only a set
// summary bit costs the DRAM sector of the real table.  Bit q of sample_hits[w] = q-th sampled position of word w.
template <int STRIDE>
__global__ void __launch_bounds__(256)
k_probe_sampled(Planes pl, CoderMasks m, const uint32_t* __restrict__ bm3, const uint32_t* __restrict__ summary,
                uint8_t* __restrict__ sample_hits)
{
    constexpr int NS = (32 + STRIDE - 1) / STRIDE;       // sampled positions per word, at most
    uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= pl.nwords) return;
    const uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]);
    const int first = (int)((STRIDE - (int)((w * 32) % STRIDE)) % STRIDE);
    uint32_t out = 0;
    if (vm) {
        const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
        uint32_t code[NS];
#pragma unroll
        for (int q = 0; q < NS; q++) {
            const int s = first + q * STRIDE;
            code[q] = 0u;
            if (s < 32 && ((vm >> (31 - s)) & 1u)) {
                WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
                code[q] = stream_code(m, b, 0);
            }
        }
        uint32_t maybe = 0;                          // first the summary (L2), all loads in flight
#pragma unroll
        for (int q = 0; q < NS; q++)
            maybe |= (code[q] ? (__ldg(summary + (code[q] >> 9)) >> ((code[q] >> 4) & 31)) & 1u : 0u) << q;
#pragma unroll
        for (int q = 0; q < NS; q++)
            if ((maybe >> q) & 1u) out |= ((__ldg(bm3 + (code[q] >> 5)) >> (code[q] & 31)) & 1u) << q;
    }
    sample_hits[w] = (uint8_t)out;
}

// ---- P2: candidate words and the words they depend on --------------------------------------------
// cand[w]: the sampled hits in words w-16..w of the record (a superset of every window ending in word w)
// reach min_sample_hits = three_min - (500 - floor(500/stride)).  need[w]: some word in w..w+16 of the same record is a
// candidate, i.e. a position of word w lies inside a window that may be good.  One CTA produces need[] for
// NEED_TILE words from a halo of 16 words on either side; both sliding sums are differences of block-wide
// prefix sums.
constexpr int NEED_TILE = 256;
constexpr int NEED_THREADS = NEED_TILE + 2 * WIN_WORDS;      // one thread per word of tile + halo (9 warps)

__device__ __forceinline__ uint32_t block_inclusive_scan_288(uint32_t v, uint32_t* s_warp)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int k = 0; k < NEED_THREADS / 32; k++) if (k < warp) base += s_warp[k];
    __syncthreads();
    return base + inc;
}

__global__ void __launch_bounds__(NEED_THREADS)
k_need(const uint8_t* __restrict__ sample_hits, const uint32_t* __restrict__ word_genome,
       const uint32_t* __restrict__ g_word, uint64_t nwords, int min_sample_hits, uint8_t* __restrict__ need,
       unsigned long long* __restrict__ need_count)
{
    __shared__ uint32_t s_P[NEED_THREADS + 1], s_C[NEED_THREADS + 1], s_warp[NEED_THREADS / 32];
    const int e = threadIdx.x;
    const int64_t first = (int64_t)blockIdx.x * NEED_TILE - WIN_WORDS;          // word of entry 0
    const int64_t w = first + e;
    const bool valid = w >= 0 && (uint64_t)w < nwords;
    uint32_t pop = 0;
    int64_t gs = 0, ge = 0;
    if (valid) {
        pop = __popc((uint32_t)sample_hits[w]);
        const uint32_t g = word_genome[w];
        gs = g_word[g]; ge = g_word[g + 1];
    }
    const uint32_t P = block_inclusive_scan_288(pop, s_warp);
    s_P[e + 1] = P;
    if (e == 0) { s_P[0] = 0; s_C[0] = 0; }
    __syncthreads();
    uint32_t cand = 0;
    if (valid && e >= WIN_WORDS) {
        const int64_t lo = max(w - WIN_WORDS, gs);                              // first word of the look-back, inside the record
        cand = (int)(P - s_P[lo - first]) >= min_sample_hits;
    }
    const uint32_t C = block_inclusive_scan_288(cand, s_warp);
    s_C[e + 1] = C;
    __syncthreads();
    uint32_t v = 0;
    if (valid && e >= WIN_WORDS && e < WIN_WORDS + NEED_TILE) {
        const int64_t hi = min(w + WIN_WORDS, ge - 1);                          // last word of the look-ahead, inside the record
        v = s_C[hi - first + 1] - s_C[e] > 0;
        need[w] = (uint8_t)v;
    }
    const uint32_t b = __ballot_sync(0xffffffffu, v != 0);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(need_count, (unsigned long long)__popc(b));
}

// ---- P3: all three streams at every position of the needed words ----------------------------------
// single bit = at least one stream hit, trio bit = all three hit (:528-546).
__global__ void __launch_bounds__(256)
k_probe_full(Planes pl, CoderMasks m, const uint32_t* __restrict__ bm3, const uint8_t* __restrict__ need,
             uint32_t* __restrict__ single, uint32_t* __restrict__ trio)
{
    uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= pl.nwords) return;
    uint32_t one = 0, three = 0;
    if (!need || need[w]) {
        const uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]);
        if (vm) {
            const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
#pragma unroll
            for (int s0 = 0; s0 < 32; s0 += 8) {
                uint32_t code[24];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const int s = s0 + q;
                    WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
                    const bool v = (vm >> (31 - s)) & 1u;
#pragma unroll
                    for (int i = 0; i < 3; i++) code[q * 3 + i] = v ? stream_code(m, b, i) : 0u;
                }
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    uint32_t h0 = probe(bm3, code[q * 3]), h1 = probe(bm3, code[q * 3 + 1]),
                             h2 = probe(bm3, code[q * 3 + 2]);
                    one |= (h0 | h1 | h2) << (31 - s0 - q);
                    three |= (h0 & h1 & h2) << (31 - s0 - q);
                }
            }
        }
    }
    single[w] = one; trio[w] = three;
}

// ---- P4: window counts -> good bit per position (:548-566) -----------------------------------------
// Position p of a record looks back over [max(0,p-499), p].  For the first position of word w that is
// bit 31 of word w, words w-1..w-15 and the low 19 bits of word w-16; each further position adds the
// next bit of word w and drops the next bit of the string (word w-16 : word w-15) from index 13 on.
__device__ __forceinline__ void window_load(const uint32_t* __restrict__ plane, uint64_t w, uint64_t gs,
                                            uint32_t& cur, uint32_t& body, uint64_t& leave)
{
    cur = plane[w];
    body = 0;
#pragma unroll
    for (int k = 1; k < WIN_WORDS; k++)
        if (w >= gs + k) body += __popc(plane[w - k]);
    uint32_t w16 = w >= gs + WIN_WORDS ? plane[w - WIN_WORDS] : 0u;
    uint32_t w15 = w >= gs + WIN_WORDS - 1 ? plane[w - (WIN_WORDS - 1)] : 0u;
    body += __popc(w16 & 0x7FFFFu);
    leave = ((uint64_t)w16 << 32) | w15;
}

__global__ void __launch_bounds__(256)
k_window(const uint32_t* __restrict__ single, const uint32_t* __restrict__ trio, const uint8_t* __restrict__ need,
         const uint32_t* __restrict__ word_genome, const uint32_t* __restrict__ g_word,
         const uint32_t* __restrict__ g_len, uint64_t nwords, int one_min, int three_min,
         uint32_t* __restrict__ good)
{
    uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    uint32_t out = 0;
    if (!need || need[w]) {
        const uint32_t g = word_genome[w];
        const uint64_t gs = g_word[g];
        const uint32_t len = g_len[g];
        const uint32_t p0 = (uint32_t)(w - gs) << 5;
        uint32_t s_cur, s_body, t_cur, t_body;
        uint64_t s_leave, t_leave;
        window_load(single, w, gs, s_cur, s_body, s_leave);
        window_load(trio, w, gs, t_cur, t_body, t_leave);
        int c1 = (int)s_body + (int)(s_cur >> 31);
        int c3 = (int)t_body + (int)(t_cur >> 31);
#pragma unroll
        for (int s = 0; s < 32; s++) {
            if (c1 >= one_min && c3 >= three_min && p0 + s < len) out |= 0x80000000u >> s;
            if (s < 31) {
                c1 += (int)((s_cur >> (30 - s)) & 1u) - (int)((s_leave >> (50 - s)) & 1ull);
                c3 += (int)((t_cur >> (30 - s)) & 1u) - (int)((t_leave >> (50 - s)) & 1ull);
            }
        }
    }
    good[w] = out;
}

// ---- P5: good runs -> padded, merged intervals -> {ordinal, frag, el, ref_len} (:568-617) ------------
// One warp per record walks its good words in order; the (rare) rising/falling edges are replayed by
// every lane in lock step, so the state machine stays warp-uniform.
__global__ void __launch_bounds__(128)
k_intervals(const uint32_t* __restrict__ good, const uint32_t* __restrict__ g_word,
            const uint32_t* __restrict__ g_len, uint32_t n_genomes, uint32_t* __restrict__ results)
{
    const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= n_genomes) return;
    const uint64_t gs = g_word[g];
    const int len = (int)g_len[g];
    const uint32_t nw = ((uint32_t)len + 31u) >> 5;
    int start = 0, last_end = 0, frag = 0, el = 0;
    bool open = false;
    uint32_t carry = 0;                               // good bit of the previous position
    for (uint32_t base = 0; base < nw; base += 32) {
        const uint32_t idx = base + lane;
        uint32_t x = idx < nw ? good[gs + idx] : 0u;
        uint32_t vmask = 0;
        if (idx < nw) {
            uint32_t rem = (uint32_t)len - (idx << 5);
            vmask = rem >= 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> rem);
        }
        x &= vmask;
        uint32_t pb = __shfl_up_sync(0xffffffffu, x & 1u, 1);
        if (lane == 0) pb = carry;
        const uint32_t prev = (x >> 1) | (pb << 31);
        const uint32_t rise = x & ~prev & vmask, fall = ~x & prev & vmask;
        uint32_t pending = __ballot_sync(0xffffffffu, (rise | fall) != 0);
        while (pending) {
            const int l = __ffs(pending) - 1;
            pending &= pending - 1;
            const uint32_t r = __shfl_sync(0xffffffffu, rise, l), f = __shfl_sync(0xffffffffu, fall, l);
            uint32_t e = r | f;
            while (e) {
                const int s = __clz(e);
                const uint32_t bit = 0x80000000u >> s;
                e &= ~bit;
                const int j = (int)((base + l) << 5) + s;
                if (r & bit) {                        // :568-574
                    start = j - 2 * WINDOW; if (start < 1) start = 1;
                    open = true;
                } else {                              // :575-589
                    int end = j + 2 * WINDOW; if (end > len) end = len;
                    if (frag > 0 && start - last_end < WINDOW) el += end - last_end;
                    else { frag++; el += end - start; }
                    last_end = end; open = false;
                }
            }
        }
        carry = __shfl_sync(0xffffffffu, x & 1u, 31);
    }
    if (open) {                                       // :599-609
        if (frag > 0 && start - last_end < WINDOW) el += len - last_end;
        else { frag++; el += len - start; }
    }
    if (lane == 0) {
        uint4 rec = make_uint4(g + 1, (uint32_t)frag, (uint32_t)el, (uint32_t)len);
        reinterpret_cast<uint4*>(results)[g] = rec;
    }
}

// ---- index body of one record (:711-738): 3 codes per window start, 0 = invalid --------------------
__global__ void k_record_codes(Planes pl, CoderMasks m, uint64_t first_word, uint32_t len,
                               uint32_t* __restrict__ codes)
{
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j + EREF_K > len) return;
    const uint64_t w = first_word + (j >> 5);
    const int s = j & 31;
    const bool v = window32(pl.ok[w], pl.ok[w + 1], s) == 0xFFFFFFFFu;
    WindowBits b = window_bits(window32(pl.hi[w], pl.hi[w + 1], s), window32(pl.lo[w], pl.lo[w + 1], s));
#pragma unroll
    for (int i = 0; i < 3; i++) codes[(size_t)j * 3 + i] = v ? stream_code(m, b, i) : 0u;
}

// ---- index body of a run of records, in the layout of <db>.k32.index.dat (:698, :711-738) -----------
// Per kept record: one word holding its length, then 3 codes per window start.  A "cell" is the length word
// or one window start; cell_pre[g] = cells of the records before g.  One thread per cell.
__global__ void __launch_bounds__(256)
k_index_body(Planes pl, CoderMasks m, const uint32_t* __restrict__ g_word, const uint32_t* __restrict__ g_len,
             const unsigned long long* __restrict__ cell_pre, uint32_t g0, uint32_t g1, unsigned long long ncells,
             uint32_t* __restrict__ out)
{
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells) return;
    const unsigned long long cell0 = cell_pre[g0], cell = cell0 + t;
    uint32_t lo = g0, hi = g1;                            // cell_pre[lo] <= cell < cell_pre[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (cell_pre[mid] <= cell) lo = mid; else hi = mid; }
    const uint32_t g = lo;
    const unsigned long long local = cell - cell_pre[g];
    uint32_t* dst = out + 3 * (cell_pre[g] - cell0) - 2ull * (g - g0);
    if (local == 0) { dst[0] = g_len[g]; return; }
    const uint32_t j = (uint32_t)(local - 1);
    const uint64_t w = (uint64_t)g_word[g] + (j >> 5);
    const int s = j & 31;
    const bool v = window32(pl.ok[w], pl.ok[w + 1], s) == 0xFFFFFFFFu;
    WindowBits b = window_bits(window32(pl.hi[w], pl.hi[w + 1], s), window32(pl.lo[w], pl.lo[w + 1], s));
#pragma unroll
    for (int i = 0; i < 3; i++) dst[1 + (size_t)j * 3 + i] = v ? stream_code(m, b, i) : 0u;
}

} // namespace

static inline unsigned grid_for(uint64_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

cudaError_t launch_probe_sampled(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* summary, int stride,
                                 uint8_t* sample_hits, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    const unsigned grid = grid_for(pl.nwords, 256);
    switch (stride) {
    case 4: k_probe_sampled<4><<<grid, 256, 0, st>>>(pl, m, bm3, summary, sample_hits); break;
    case 5: k_probe_sampled<5><<<grid, 256, 0, st>>>(pl, m, bm3, summary, sample_hits); break;
    case 6: k_probe_sampled<6><<<grid, 256, 0, st>>>(pl, m, bm3, summary, sample_hits); break;
    case 7: k_probe_sampled<7><<<grid, 256, 0, st>>>(pl, m, bm3, summary, sample_hits); break;
    case 8: k_probe_sampled<8><<<grid, 256, 0, st>>>(pl, m, bm3, summary, sample_hits); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_need(const uint8_t* sample_hits, const uint32_t* word_genome, const uint32_t* g_word, uint64_t nwords,
                        int min_sample_hits, uint8_t* need, unsigned long long* need_count, cudaStream_t st)
{
    if (!nwords) return cudaSuccess;
    k_need<<<grid_for(nwords, NEED_TILE), NEED_THREADS, 0, st>>>(sample_hits, word_genome, g_word, nwords, min_sample_hits,
                                                              need, need_count);
    return cudaGetLastError();
}

cudaError_t launch_probe_full(Planes pl, CoderMasks m, const uint32_t* bm3, const uint8_t* need,
                              uint32_t* single, uint32_t* trio, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_probe_full<<<grid_for(pl.nwords, 256), 256, 0, st>>>(pl, m, bm3, need, single, trio);
    return cudaGetLastError();
}

cudaError_t launch_window(const uint32_t* single, const uint32_t* trio, const uint8_t* need,
                          const uint32_t* word_genome, const uint32_t* g_word, const uint32_t* g_len,
                          uint64_t nwords, int one_min, int three_min, uint32_t* good, cudaStream_t st)
{
    if (!nwords) return cudaSuccess;
    k_window<<<grid_for(nwords, 256), 256, 0, st>>>(single, trio, need, word_genome, g_word, g_len, nwords,
                                                    one_min, three_min, good);
    return cudaGetLastError();
}

cudaError_t launch_intervals(const uint32_t* good, const uint32_t* g_word, const uint32_t* g_len,
                             uint32_t n_genomes, uint32_t* results, cudaStream_t st)
{
    if (!n_genomes) return cudaSuccess;
    k_intervals<<<grid_for((uint64_t)n_genomes * 32, 128), 128, 0, st>>>(good, g_word, g_len, n_genomes, results);
    return cudaGetLastError();
}

cudaError_t launch_record_codes(Planes pl, CoderMasks m, uint64_t first_word, uint32_t len, uint32_t* codes,
                                cudaStream_t st)
{
    if (len < (uint32_t)EREF_K) return cudaSuccess;
    k_record_codes<<<grid_for(len - EREF_K + 1, 256), 256, 0, st>>>(pl, m, first_word, len, codes);
    return cudaGetLastError();
}

cudaError_t launch_index_body(Planes pl, CoderMasks m, const uint32_t* g_word, const uint32_t* g_len,
                              const unsigned long long* cell_pre, uint32_t g0, uint32_t g1, unsigned long long ncells,
                              uint32_t* out, cudaStream_t st)
{
    if (!ncells) return cudaSuccess;
    k_index_body<<<grid_for(ncells, 256), 256, 0, st>>>(pl, m, g_word, g_len, cell_pre, g0, g1, ncells, out);
    return cudaGetLastError();
}
