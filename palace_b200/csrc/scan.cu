// scan.cu — the streaming reference scan (north_star part 4): probe the query table at every
// reference position, apply the 500-position window rule, and reduce each record to
// {ordinal, frag, el, ref_len}.  Replaces read_index (:813-903) + slide_window (:504-624).
//
// Exact candidate filter.  A window is "good" only if at least three_min of its 500 positions are
// hit by ALL three code streams (:561), hence by stream 0.  Probing stream 0 at every s-th position
// bounds that count from above: trio(window) <= sampled_hits(window) + 500 - floor(500/s).  Words whose bound
// stays below three_min can never be good, so the full three-stream probe and the window rule run only on the
// (listed) words that can influence a window that still might be.  The selected records and their numbers are
// unchanged; random-sector probes drop from 3 per base to 1/s per base plus 3 per base on candidate regions.
#include "eref_internal.cuh"
#include <algorithm>
#include <cstdlib>

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
// cand[x]: the sampled hits in words x-16..x of the record (a superset of every window ending in word x)
// reach min_sample_hits = three_min - (500 - floor(500/stride)).  need[w]: some word in w..w+16 of the same record
// is a candidate, i.e. a position of word w lies inside a window that may be good.
// A thread owns NEED_SPAN = 8 consecutive words and reads the 40 hit bytes w0-16 .. w0+23 around them (five
// 8-byte loads); both sliding sums run in registers.  When the 40 words belong to one record (two loads of
// word_genome tell) no clipping is needed; otherwise every term is guarded by "same record".  Needed words are
// appended to `list` (one atomicAdd per warp), and the thread zeroes single/trio/good of its words: the probe
// and window kernels then touch only the listed words (elsewhere single = trio = good = 0, as in the reference
// for positions no window can use).
constexpr int NEED_SPAN = 8;
constexpr int NEED_HALO = WIN_WORDS;                       // 16 words either side
constexpr int NEED_N = NEED_SPAN + 2 * NEED_HALO;          // 40 entries

__device__ __forceinline__ uint64_t byte_popcounts(uint64_t x)
{
    x = x - ((x >> 1) & 0x5555555555555555ull);
    x = (x & 0x3333333333333333ull) + ((x >> 2) & 0x3333333333333333ull);
    return (x + (x >> 4)) & 0x0F0F0F0F0F0F0F0Full;
}

// byte i of the 40-byte string held in five 64-bit words (static i after unrolling)
#define NEED_P(i) ((int)((pc[(i) >> 3] >> (8 * ((i) & 7))) & 0xFFull))

__global__ void __launch_bounds__(256)
k_need(const uint8_t* __restrict__ sample_hits, const uint32_t* __restrict__ word_genome, uint64_t nwords,
       int min_sample_hits, uint8_t* __restrict__ need, uint32_t* __restrict__ list,
       unsigned long long* __restrict__ need_count, uint32_t* __restrict__ single, uint32_t* __restrict__ trio,
       uint32_t* __restrict__ good, uint32_t* __restrict__ genome_listed)
{
    const uint64_t w0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * NEED_SPAN;
    uint32_t bits = 0;                                     // bit j: word w0+j is needed
    if (w0 < nwords) {
        // entry i = word w0 - 16 + i, i = 0..39
        uint64_t pc[NEED_N / 8];                           // per-entry hit counts, one byte each
        const bool inside = w0 >= NEED_HALO && w0 + NEED_SPAN + NEED_HALO <= nwords;
        bool one_record = false;
        if (inside) {
            const uint64_t* src = reinterpret_cast<const uint64_t*>(sample_hits + (w0 - NEED_HALO));
#pragma unroll
            for (int q = 0; q < NEED_N / 8; q++) pc[q] = byte_popcounts(src[q]);
            one_record = word_genome[w0 - NEED_HALO] == word_genome[w0 + NEED_SPAN + NEED_HALO - 1];
        }
        // sliding sums over full look-backs: Sfull(k) = entries k..k+16, for candidate word w0+k
        if (one_record) {
            int S = 0;
#pragma unroll
            for (int i = 0; i <= NEED_HALO; i++) S += NEED_P(i);
            uint32_t cm = 0;                               // bit k: word w0+k is a candidate, k = 0..23
#pragma unroll
            for (int k = 0; k < NEED_SPAN + NEED_HALO; k++) {
                cm |= (uint32_t)(S >= min_sample_hits) << k;
                if (k + 1 < NEED_SPAN + NEED_HALO) S += NEED_P(NEED_HALO + k + 1) - NEED_P(k);
            }
#pragma unroll
            for (int j = 0; j < NEED_SPAN; j++) bits |= (uint32_t)(((cm >> j) & 0x1FFFFu) != 0) << j;
        } else {
            // general case: a record boundary (or the end of the planes) inside the 40 entries.
            // bnd bit i: entry i starts a segment (first word of a record, or outside the planes); vld bit i: inside.
            uint64_t bnd = 0, vld = 0;
            if (inside) {
                vld = (1ull << NEED_N) - 1;
                const uint4* gsrc = reinterpret_cast<const uint4*>(word_genome + (w0 - NEED_HALO));
                uint32_t prev = 0;
#pragma unroll
                for (int q = 0; q < NEED_N / 4; q++) {
                    const uint4 g = gsrc[q];
                    if (q) bnd |= (uint64_t)(g.x != prev) << (4 * q);
                    bnd |= (uint64_t)(g.y != g.x) << (4 * q + 1);
                    bnd |= (uint64_t)(g.z != g.y) << (4 * q + 2);
                    bnd |= (uint64_t)(g.w != g.z) << (4 * q + 3);
                    prev = g.w;
                }
            } else {
                uint64_t raw[NEED_N / 8];
#pragma unroll
                for (int q = 0; q < NEED_N / 8; q++) raw[q] = 0;
                uint32_t prev = 0xFFFFFFFFu;
#pragma unroll
                for (int i = 0; i < NEED_N; i++) {
                    const int64_t x = (int64_t)w0 - NEED_HALO + i;
                    const bool in = x >= 0 && (uint64_t)x < nwords;
                    const uint32_t g = in ? word_genome[x] : 0xFFFFFFFFu;
                    raw[i >> 3] |= (uint64_t)(in ? sample_hits[x] : (uint8_t)0) << (8 * (i & 7));
                    vld |= (uint64_t)in << i;
                    bnd |= (uint64_t)(!in || g != prev) << i;
                    prev = g;
                }
#pragma unroll
                for (int q = 0; q < NEED_N / 8; q++) pc[q] = byte_popcounts(raw[q]);
            }
            int S = 0;
#pragma unroll
            for (int i = 0; i <= NEED_HALO; i++) S += NEED_P(i);
            uint32_t cm = 0;
#pragma unroll
            for (int k = 0; k < NEED_SPAN + NEED_HALO; k++) {
                const int i = NEED_HALO + k;
                // the look-back stops at the first word of the record: drop the entries k .. seg-1
                const uint64_t below = bnd & ((2ull << i) - 1);
                const int seg = below ? 63 - __clzll((long long)below) : 0;
                int Sk = S;
                if (seg > k) {
                    const int drop = seg - k;              // 1..16 leading bytes of the 17-byte window
                    // bytes k..k+7 and k+8..k+15 of the string
                    const int q = k >> 3, r = 8 * (k & 7);
                    uint64_t a = pc[q] >> r, b2 = pc[q + 1] >> r;
                    if (r) { a |= pc[q + 1] << (64 - r); b2 |= pc[q + 2 < NEED_N / 8 ? q + 2 : q + 1] << (64 - r); }
                    const uint64_t ma = drop >= 8 ? ~0ull : (1ull << (8 * drop)) - 1;
                    const uint64_t mb = drop > 8 ? (drop >= 16 ? ~0ull : (1ull << (8 * (drop - 8))) - 1) : 0ull;
                    Sk -= (int)((((a & ma) * 0x0101010101010101ull) >> 56) + (((b2 & mb) * 0x0101010101010101ull) >> 56));
                }
                cm |= (uint32_t)(((vld >> i) & 1ull) && Sk >= min_sample_hits) << k;
                if (k + 1 < NEED_SPAN + NEED_HALO) S += NEED_P(NEED_HALO + k + 1) - NEED_P(k);
            }
#pragma unroll
            for (int j = 0; j < NEED_SPAN; j++) {
                const int i = NEED_HALO + j;
                const uint64_t above = bnd >> (i + 1);     // boundaries after entry i
                const int t = above ? __ffsll((long long)above) - 1 : 64;     // entries i .. i+t share the record
                const int reach = t < NEED_HALO ? t : NEED_HALO;
                const uint32_t any = (cm >> j) & ((2u << reach) - 1);
                bits |= (uint32_t)(((vld >> i) & 1ull) && any != 0) << j;
            }
        }
        if (w0 + NEED_SPAN <= nwords) {
            uint2 nb;
            nb.x = (bits & 1u) | ((bits & 2u) << 7) | ((bits & 4u) << 14) | ((bits & 8u) << 21);
            nb.y = ((bits >> 4) & 1u) | ((bits & 32u) << 3) | ((bits & 64u) << 10) | ((bits & 128u) << 17);
            *reinterpret_cast<uint2*>(need + w0) = nb;
            const uint4 z = make_uint4(0, 0, 0, 0);
            reinterpret_cast<uint4*>(single + w0)[0] = z; reinterpret_cast<uint4*>(single + w0)[1] = z;
            reinterpret_cast<uint4*>(trio + w0)[0] = z;   reinterpret_cast<uint4*>(trio + w0)[1] = z;
            reinterpret_cast<uint4*>(good + w0)[0] = z;   reinterpret_cast<uint4*>(good + w0)[1] = z;
        } else {
            for (int j = 0; w0 + j < nwords; j++) {
                need[w0 + j] = (uint8_t)((bits >> j) & 1u);
                single[w0 + j] = 0; trio[w0 + j] = 0; good[w0 + j] = 0;
            }
        }
    }
    // append the needed words to the list: exclusive prefix of the counts inside the warp, one atomic per warp
    const int lane = threadIdx.x & 31;
    const uint32_t cnt = __popc(bits);
    uint32_t inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
    if (total == 0) return;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(need_count, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 0);
    EREF_ASSERT(base + total <= nwords);                  // the list holds at most every word once
    uint32_t* dst = list + base + (inc - cnt);
    while (bits) {
        const int j = __ffs(bits) - 1;
        bits &= bits - 1;
        *dst++ = (uint32_t)(w0 + j);
        // records without a listed word keep good == 0 everywhere: k_intervals answers {0, 0} for them without reading it
        const uint32_t g = word_genome[w0 + j];
        if (g != 0xFFFFFFFFu) genome_listed[g] = 1u;
    }
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

// The same probes for the words k_need listed.  LPW lanes share a word (32/LPW positions each) so that a short
// list still spreads over the whole GPU; the loop is warp-uniform (32/LPW list entries per warp and round), the
// partial bit-words are OR-ed with shuffles.
template <int LPW>
__global__ void __launch_bounds__(256)
k_probe_list(Planes pl, CoderMasks m, const uint32_t* __restrict__ bm3, const uint32_t* __restrict__ list,
             const unsigned long long* __restrict__ n_list, uint32_t* __restrict__ single, uint32_t* __restrict__ trio)
{
    constexpr int PER = 32 / LPW;                          // positions per lane
    constexpr int EPW = 32 / LPW;                          // list entries per warp and round
    constexpr int BATCH = PER < 8 ? PER : 8;
    const unsigned long long n = *n_list;
    const int lane = threadIdx.x & 31, q = lane & (LPW - 1);
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long base = warp * EPW; base < n; base += nwarps * EPW) {
        const unsigned long long e = base + (lane / LPW);
        uint32_t one = 0, three = 0, w = 0;
        if (e < n) {
            w = list[e];
            const uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]);
            const uint32_t mine = PER == 32 ? vm : (vm >> (32 - PER - PER * q)) & ((1u << (PER & 31)) - 1u);
            if (mine) {
                const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
#pragma unroll
                for (int s0 = 0; s0 < PER; s0 += BATCH) {
                    uint32_t code[BATCH * 3];
#pragma unroll
                    for (int j = 0; j < BATCH; j++) {
                        const int s = PER * q + s0 + j;
                        WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
                        const bool v = (vm >> (31 - s)) & 1u;
#pragma unroll
                        for (int i = 0; i < 3; i++) code[j * 3 + i] = v ? stream_code(m, b, i) : 0u;
                    }
#pragma unroll
                    for (int j = 0; j < BATCH; j++) {
                        const uint32_t h0 = probe(bm3, code[j * 3]), h1 = probe(bm3, code[j * 3 + 1]),
                                       h2 = probe(bm3, code[j * 3 + 2]);
                        one |= (h0 | h1 | h2) << (31 - PER * q - s0 - j);
                        three |= (h0 & h1 & h2) << (31 - PER * q - s0 - j);
                    }
                }
            }
        }
#pragma unroll
        for (int d = 1; d < LPW; d <<= 1) {
            one |= __shfl_xor_sync(0xffffffffu, one, d);
            three |= __shfl_xor_sync(0xffffffffu, three, d);
        }
        if (e < n && q == 0) { single[w] = one; trio[w] = three; }
    }
}

// ---- P3': the same probes with the table walked one part at a time ----------------------------------
// A B200 answers random 4-byte probes at ~58 G/s when the bitmap is 512 MiB (DRAM, 64-128 B moved per probe), ~123
// G/s at 128 MiB and ~288 G/s at <= 64 MiB (L2) — tools/probe_bench.cu.  So the listed words are probed in `parts`
// rounds: in round r every listed word computes its codes again (cheap) and probes only those whose top bits equal
// r, which keeps the working set of a round at 512/parts MiB.  Measured on the bench workload (2.5 % of the words
// listed), one thread per word: 0.65 ms in one round, 0.55 in two, 0.43 in four and in eight, 0.73 in sixteen; on a
// dense table (BASELINE configs[3], 90 % of the words listed): 9.0 / 4.75 / 4.3 / 8.2 ms for 1 / 4 / 8 / 16 rounds.
// Four lanes per word (8 positions each) cost 0.50 and 6.1 ms in four rounds: the per-word overhead repeats in
// every round.  Streaming each part into L2 first (k_table_prefetch, EREF_PROBE_PREFETCH=1) does not pay.
// Hits are collected per stream in h0/h1/h2 (round 0 stores, later rounds OR; a word belongs to one thread or one
// lane group, so no atomics), the last round turns them into single = h0|h1|h2 and trio = h0&h1&h2.
__global__ void __launch_bounds__(256)
k_table_prefetch(const uint32_t* __restrict__ part, uint32_t lines)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < lines; i += gridDim.x * blockDim.x)
        asm volatile("prefetch.global.L2 [%0];" :: "l"(part + (size_t)i * 32));
}

// Which part does a code fall into?  The top bits of min(fwd, rc) are min(top bits of fwd, top bits of rc) (the smaller
// number has the smaller-or-equal prefix), the top bit z of fwd is the alphabet bit of the window's base z, and the
// top bit z of rc that of its base 31-z (complemented) — so for all 32 window starts of a word at once these are the
// alphabet planes shifted by z and by 31-z.  part_mask returns the starts (bit 31-s) of stream i whose code's top LOG2P
// bits equal `part`, at ~10 logic operations per bit instead of a full code per start: a round then forms only the
// ~1/2^LOG2P of the codes it probes.
template <int LOG2P>
__device__ __forceinline__ uint32_t part_mask(const CoderMasks& m, int i, uint32_t hi0, uint32_t hi1, uint32_t lo0, uint32_t lo1,
                                              uint32_t part)
{
    uint32_t f[LOG2P], r[LOG2P];
#pragma unroll
    for (int z = 0; z < LOG2P; z++) {
        const uint32_t s0 = 0u - ((m.sel0[i] >> (31 - z)) & 1u), s2 = 0u - ((m.sel2[i] >> (31 - z)) & 1u);
        const uint32_t H = window32(hi0, hi1, z), L = window32(lo0, lo1, z);
        f[z] = bitsel(s2, ~L, bitsel(s0, ~(H ^ L), ~H));
        const uint32_t Hr = window32(hi0, hi1, 31 - z), Lr = window32(lo0, lo1, 31 - z);
        r[z] = bitsel(s2, Lr, bitsel(s0, ~(Hr ^ Lr), Hr));
    }
    // lt: fwd prefix < rc prefix, decided at the first differing bit from the top
    uint32_t lt = 0, eq = 0xFFFFFFFFu;
#pragma unroll
    for (int z = 0; z < LOG2P; z++) {
        lt |= eq & ~f[z] & r[z];
        eq &= ~(f[z] ^ r[z]);
    }
    uint32_t in = 0xFFFFFFFFu;
#pragma unroll
    for (int z = 0; z < LOG2P; z++) {
        const uint32_t mz = bitsel(lt, f[z], r[z]);
        const uint32_t want = 0u - ((part >> (LOG2P - 1 - z)) & 1u);
        in &= ~(mz ^ want);
    }
    return in;
}

// list == nullptr: every word of the planes (the unpruned scan), else the words k_need listed
template <int LOG2P>
__global__ void __launch_bounds__(256)
k_probe_part(Planes pl, CoderMasks m, const uint32_t* __restrict__ bm3, const uint32_t* __restrict__ list,
             const unsigned long long* __restrict__ n_list, uint32_t part, bool first, bool last,
             uint32_t* __restrict__ h0w, uint32_t* __restrict__ h1w, uint32_t* __restrict__ h2w,
             uint32_t* __restrict__ single, uint32_t* __restrict__ trio)
{
    const unsigned long long n = list ? *n_list : pl.nwords;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint64_t w = list ? list[e] : e;
        EREF_ASSERT(w < pl.nwords);
        uint32_t acc[3] = {0, 0, 0};
        const uint32_t vm = valid_windows(pl.ok[w], pl.ok[w + 1]);
        if (vm) {
            const uint32_t hi0 = pl.hi[w], hi1 = pl.hi[w + 1], lo0 = pl.lo[w], lo1 = pl.lo[w + 1];
#pragma unroll
            for (int i = 0; i < 3; i++) {
                uint32_t todo = vm & part_mask<LOG2P>(m, i, hi0, hi1, lo0, lo1, part);
                while (todo) {                             // four probes in flight per pass
                    uint32_t code[4];
                    int pos[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        code[q] = 0u; pos[q] = 0;
                        if (todo) {
                            const int s = __clz(todo);
                            todo &= ~(0x80000000u >> s);
                            const WindowBits b = window_bits(window32(hi0, hi1, s), window32(lo0, lo1, s));
                            code[q] = stream_code(m, b, i);
                            pos[q] = s;
                        }
                    }
                    uint32_t hit[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) hit[q] = probe(bm3, code[q]);
#pragma unroll
                    for (int q = 0; q < 4; q++) acc[i] |= hit[q] << (31 - pos[q]);
                }
            }
        }
        uint32_t a0 = acc[0], a1 = acc[1], a2 = acc[2];
        if (!first) { a0 |= h0w[w]; a1 |= h1w[w]; a2 |= h2w[w]; }
        if (last) { single[w] = a0 | a1 | a2; trio[w] = a0 & a1 & a2; }
        else { h0w[w] = a0; h1w[w] = a1; h2w[w] = a2; }
    }
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

__device__ __forceinline__ uint32_t window_word(const uint32_t* __restrict__ single, const uint32_t* __restrict__ trio,
                                                const uint32_t* __restrict__ word_genome,
                                                const uint32_t* __restrict__ g_word, const uint32_t* __restrict__ g_len,
                                                uint64_t w, int one_min, int three_min)
{
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
    uint32_t out = 0;
#pragma unroll
    for (int s = 0; s < 32; s++) {
        if (c1 >= one_min && c3 >= three_min && p0 + s < len) out |= 0x80000000u >> s;
        if (s < 31) {
            c1 += (int)((s_cur >> (30 - s)) & 1u) - (int)((s_leave >> (50 - s)) & 1ull);
            c3 += (int)((t_cur >> (30 - s)) & 1u) - (int)((t_leave >> (50 - s)) & 1ull);
        }
    }
    return out;
}

// every word (unpruned scan)
__global__ void __launch_bounds__(256)
k_window(const uint32_t* __restrict__ single, const uint32_t* __restrict__ trio,
         const uint32_t* __restrict__ word_genome, const uint32_t* __restrict__ g_word,
         const uint32_t* __restrict__ g_len, uint64_t nwords, int one_min, int three_min,
         uint32_t* __restrict__ good)
{
    uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= nwords) return;
    good[w] = window_word(single, trio, word_genome, g_word, g_len, w, one_min, three_min);
}

// the listed words only (good is zero elsewhere, written by k_need)
__global__ void __launch_bounds__(256)
k_window_list(const uint32_t* __restrict__ single, const uint32_t* __restrict__ trio,
              const uint32_t* __restrict__ list, const unsigned long long* __restrict__ n_list,
              const uint32_t* __restrict__ word_genome, const uint32_t* __restrict__ g_word,
              const uint32_t* __restrict__ g_len, int one_min, int three_min, uint32_t* __restrict__ good)
{
    const unsigned long long n = *n_list;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        const uint32_t w = list[e];
        good[w] = window_word(single, trio, word_genome, g_word, g_len, w, one_min, three_min);
    }
}

// ---- P5: good runs -> padded, merged intervals -> {ordinal, frag, el, ref_len} (:568-617) ------------
// One warp per record walks its good words in order; the (rare) rising/falling edges are replayed by
// every lane in lock step, so the state machine stays warp-uniform.
constexpr int IV_UNROLL = 32;                         // 128-byte rows of a record in flight per warp: the kernel lasts as long as the
                                                      // longest listed record, i.e. (its words / (32 * IV_UNROLL)) dependent round trips

__global__ void __launch_bounds__(128)
k_intervals(const uint32_t* __restrict__ good, const uint32_t* __restrict__ g_word,
            const uint32_t* __restrict__ g_len, uint32_t n_genomes, const uint32_t* __restrict__ genome_listed,
            uint32_t* __restrict__ results)
{
    const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= n_genomes) return;
    const uint64_t gs = g_word[g];
    const int len = (int)g_len[g];
    if (genome_listed && !genome_listed[g]) {             // candidate filter on: no word of this record was listed
        if (lane == 0) reinterpret_cast<uint4*>(results)[g] = make_uint4(g + 1, 0u, 0u, (uint32_t)len);
        return;
    }
    const uint32_t nw = ((uint32_t)len + 31u) >> 5;
    int start = 0, last_end = 0, frag = 0, el = 0;
    bool open = false;
    uint32_t carry = 0;                               // good bit of the previous position
    for (uint32_t base0 = 0; base0 < nw; base0 += 32 * IV_UNROLL) {
        uint32_t xs[IV_UNROLL];
#pragma unroll
        for (int u = 0; u < IV_UNROLL; u++) {
            const uint32_t idx = base0 + u * 32 + lane;
            xs[u] = idx < nw ? good[gs + idx] : 0u;
        }
        uint32_t any = 0;
#pragma unroll
        for (int u = 0; u < IV_UNROLL; u++) any |= xs[u];
        if (!__any_sync(0xffffffffu, any != 0) && !carry) continue;      // all clear and no run in progress
#pragma unroll
        for (int u = 0; u < IV_UNROLL; u++) {
            const uint32_t base = base0 + u * 32;
            if (base >= nw) break;
            const uint32_t idx = base + lane;
            uint32_t x = xs[u];
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

// records of one database piece appended to the run's list: ordinals continue in file order (:887, :1175)
__global__ void k_add_ordinal(uint32_t* __restrict__ results, uint32_t n, uint32_t base)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) results[(size_t)i * 4] += base;
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

cudaError_t launch_need(const uint8_t* sample_hits, const uint32_t* word_genome, uint64_t nwords, int min_sample_hits,
                        uint8_t* need, uint32_t* list, unsigned long long* need_count, uint32_t* single,
                        uint32_t* trio, uint32_t* good, uint32_t* genome_listed, cudaStream_t st)
{
    if (!nwords) return cudaSuccess;
    k_need<<<grid_for((nwords + NEED_SPAN - 1) / NEED_SPAN, 256), 256, 0, st>>>(sample_hits, word_genome, nwords,
                                                                              min_sample_hits, need, list, need_count,
                                                                              single, trio, good, genome_listed);
    return cudaGetLastError();
}

cudaError_t launch_probe_full(Planes pl, CoderMasks m, const uint32_t* bm3, const uint8_t* need,
                              uint32_t* single, uint32_t* trio, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_probe_full<<<grid_for(pl.nwords, 256), 256, 0, st>>>(pl, m, bm3, need, single, trio);
    return cudaGetLastError();
}

static unsigned list_grid(int per_sm)
{
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    return (unsigned)(sms * per_sm);
}

cudaError_t launch_probe_list(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* list,
                              const unsigned long long* n_list, uint32_t* single, uint32_t* trio, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    static int lpw = 0, per_sm = 0;
    if (!lpw) {
        const char* v = getenv("EREF_PROBE_LPW");
        lpw = v ? atoi(v) : 4;
        v = getenv("EREF_PROBE_CTAS");
        per_sm = v ? atoi(v) : 16;                      // measured: 0.70 ms at 8, 0.65 ms at 16 (bench workload)
        if (per_sm < 1 || per_sm > 64) per_sm = 16;
    }
    const unsigned grid = list_grid(per_sm);
    switch (lpw) {
    case 1: k_probe_list<1><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, single, trio); break;
    case 2: k_probe_list<2><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, single, trio); break;
    case 8: k_probe_list<8><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, single, trio); break;
    default: k_probe_list<4><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, single, trio); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_probe_parts(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* list,
                               const unsigned long long* n_list, int parts, uint32_t* h0w, uint32_t* h1w, uint32_t* h2w,
                               uint32_t* single, uint32_t* trio, cudaStream_t st, int* launches)
{
    if (!pl.nwords) return cudaSuccess;
    static int per_sm = 0, prefetch = 0;          // measured: the prefetch does not pay (0.52 ms with, 0.50 without)
    if (!per_sm) {
        const char* v = getenv("EREF_PROBE_CTAS");
        per_sm = v ? atoi(v) : 16;
        if (per_sm < 1 || per_sm > 64) per_sm = 16;
        v = getenv("EREF_PROBE_PREFETCH");
        if (v) prefetch = atoi(v);
    }
    int log2p = 0;
    while ((1 << log2p) < parts) log2p++;
    const uint32_t part_words = TABLE_WORDS >> log2p;
    // a listed scan is sized by the SMs (the list length lives on the device), an unpruned one by the words
    const unsigned grid = list ? list_grid(per_sm) : (unsigned)std::min<uint64_t>((pl.nwords + 255) / 256, (uint64_t)list_grid(64));
    for (int r = 0; r < parts; r++) {
        if (prefetch) k_table_prefetch<<<list_grid(8), 256, 0, st>>>(bm3 + (size_t)r * part_words, part_words / 32);
        const uint32_t part = (uint32_t)r;
        const bool first = r == 0, last = r == parts - 1;
        switch (log2p) {
        case 1: k_probe_part<1><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, part, first, last, h0w, h1w, h2w, single, trio); break;
        case 2: k_probe_part<2><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, part, first, last, h0w, h1w, h2w, single, trio); break;
        case 3: k_probe_part<3><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, part, first, last, h0w, h1w, h2w, single, trio); break;
        case 4: k_probe_part<4><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, part, first, last, h0w, h1w, h2w, single, trio); break;
        case 5: k_probe_part<5><<<grid, 256, 0, st>>>(pl, m, bm3, list, n_list, part, first, last, h0w, h1w, h2w, single, trio); break;
        default: return cudaErrorInvalidValue;
        }
        if (launches) *launches += 1 + (prefetch ? 1 : 0);
    }
    return cudaGetLastError();
}

cudaError_t launch_window(const uint32_t* single, const uint32_t* trio, const uint32_t* word_genome,
                          const uint32_t* g_word, const uint32_t* g_len, uint64_t nwords, int one_min, int three_min,
                          uint32_t* good, cudaStream_t st)
{
    if (!nwords) return cudaSuccess;
    k_window<<<grid_for(nwords, 256), 256, 0, st>>>(single, trio, word_genome, g_word, g_len, nwords, one_min, three_min,
                                                    good);
    return cudaGetLastError();
}

cudaError_t launch_window_list(const uint32_t* single, const uint32_t* trio, const uint32_t* list,
                               const unsigned long long* n_list, const uint32_t* word_genome, const uint32_t* g_word,
                               const uint32_t* g_len, uint64_t nwords, int one_min, int three_min, uint32_t* good,
                               cudaStream_t st)
{
    if (!nwords) return cudaSuccess;
    k_window_list<<<list_grid(8), 256, 0, st>>>(single, trio, list, n_list, word_genome, g_word, g_len, one_min, three_min,
                                               good);
    return cudaGetLastError();
}

cudaError_t launch_intervals(const uint32_t* good, const uint32_t* g_word, const uint32_t* g_len,
                             uint32_t n_genomes, const uint32_t* genome_listed, uint32_t* results, cudaStream_t st)
{
    if (!n_genomes) return cudaSuccess;
    k_intervals<<<grid_for((uint64_t)n_genomes * 32, 128), 128, 0, st>>>(good, g_word, g_len, n_genomes, genome_listed, results);
    return cudaGetLastError();
}

cudaError_t launch_add_ordinal(uint32_t* results, uint32_t n, uint32_t base, cudaStream_t st)
{
    if (!n || !base) return cudaSuccess;
    k_add_ordinal<<<grid_for(n, 256), 256, 0, st>>>(results, n, base);
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
