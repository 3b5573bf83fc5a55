// frame.cu — text framing and packing kernels (north_star part 1): FASTQ/FASTA bytes in HBM ->
// three bit-planes (hi, lo, ok).  Replaces the getline loops of bin/extract_ref.cpp:686-756 (FASTA
// records) and :940-947 (FASTQ sequence lines).
//
// Every kernel here is text-driven: a CTA of 256 threads owns an 8 KiB tile, a thread owns 32 bytes
// (two LDG.128) and reduces them to 32-bit masks with SIMD-within-a-register arithmetic (four bytes per
// integer op, bits gathered by one multiply), so the cost is a few integer instructions per byte and the
// text is streamed at HBM speed.  Line/record state that crosses thread and tile boundaries is carried
// by prefix scans over an associative "chunk summary" (FastaSum / TileSum).
#include "eref_internal.cuh"

namespace {

// ---- 32 bytes -> masks (byte k of the chunk <-> bit 31-k) ------------------------------------------

// 32 bytes of text starting at off (off is a multiple of 32; text is 16-byte aligned).
// Bytes at or beyond n read as 0 (an invalid base that is not a newline).
__device__ __forceinline__ void load32(const uint8_t* __restrict__ text, uint64_t n, uint64_t off,
                                       uint32_t w[8])
{
    if (off + 32 <= n) {
        const uint4* p = reinterpret_cast<const uint4*>(text + off);
        uint4 a = __ldg(p), b = __ldg(p + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
        w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) w[i] = 0;
#pragma unroll
        for (int k = 0; k < 32; k++)
            if (off + k < n) w[k >> 2] |= (uint32_t)text[off + k] << ((k & 3) * 8);
    }
}

// Per-byte flags live in bit 7 of their byte.  A byte-equality test under a 7-bit mask is one LOP3 + one add:
// ((w ^ pattern) & mask7) + 0x7F sets bit 7 iff the masked byte differs (no carry leaves the byte); OR-ing w in
// also rejects bytes >= 0x80.
constexpr uint32_t H7 = 0x80808080u;

__device__ __forceinline__ uint32_t eq7(uint32_t w, uint32_t pattern, uint32_t mask7)
{
    const uint32_t t = ((w ^ pattern) & mask7) + 0x7F7F7F7Fu;
    return ~(t | w) & H7;
}

// Two sets of per-byte flags (a in bits 7,15,23,31; b in bits 3,11,19,27) -> one multiply gathers the a flags
// into bits 28..31 and the b flags into bits 24..27, FIRST byte of the word in the LOWEST bit (all partial
// products land on distinct bits: no carries).  The chunk masks are therefore assembled byte k <-> bit k and
// bit-reversed once at the end.
__device__ __forceinline__ uint32_t gather2(uint32_t a7, uint32_t b3) { return (a7 | b3) * 0x00204081u; }

template <int J> __device__ __forceinline__ uint32_t place_a(uint32_t prod)    // bits 31..28 -> bits 4J+3..4J
{
    return (prod >> (28 - 4 * J)) & (0xFu << (4 * J));
}
template <int J> __device__ __forceinline__ uint32_t place_b(uint32_t prod)    // bits 27..24 -> bits 4J+3..4J
{
    return (J <= 6 ? prod >> (J <= 6 ? 24 - 4 * J : 0) : prod << (J <= 6 ? 0 : 4 * J - 24)) & (0xFu << (4 * J));
}

struct Masks { uint32_t nl, hi, lo, ok; };

// nl: byte == '\n'.  ok: byte is one of ACGTacgt (:978).  hi/lo: 2-bit base code A=00 C=01 G=10 T=11 from
// ASCII bits 2 and 1 (hi = bit2, lo = bit1 ^ bit2), zero where ok is zero.
template <int J>
__device__ __forceinline__ void masks_word(uint32_t w, Masks& m)
{
    const uint32_t nl = eq7(w, 0x0A0A0A0Au, 0x7F7F7F7Fu);
    const uint32_t aceg = eq7(w, 0x41414141u, 0x59595959u);          // 0x41 0x43 0x45 0x47, either case
    const uint32_t tt = eq7(w, 0x54545454u, 0x5F5F5F5Fu);            // 0x54, either case
    const uint32_t h2 = w << 5, h1 = w << 6;                         // ASCII bit 2 / bit 1 in bit 7 of the byte
    const uint32_t ok = (aceg & ~(h2 & ~h1)) | tt;                   // drop 0x45/0x65 ('E'): bit 2 set, bit 1 clear
    const uint32_t p0 = gather2(h2 & ok, ((h1 ^ h2) & ok) >> 4);
    const uint32_t p1 = gather2(ok, nl >> 4);
    m.hi |= place_a<J>(p0); m.lo |= place_b<J>(p0);
    m.ok |= place_a<J>(p1); m.nl |= place_b<J>(p1);
}

__device__ __forceinline__ Masks chunk_masks(const uint32_t w[8])
{
    Masks m{0, 0, 0, 0};
    masks_word<0>(w[0], m); masks_word<1>(w[1], m); masks_word<2>(w[2], m); masks_word<3>(w[3], m);
    masks_word<4>(w[4], m); masks_word<5>(w[5], m); masks_word<6>(w[6], m); masks_word<7>(w[7], m);
    m.nl = __brev(m.nl); m.hi = __brev(m.hi); m.lo = __brev(m.lo); m.ok = __brev(m.ok);
    return m;
}

template <int J>
__device__ __forceinline__ uint32_t nl_word(uint32_t w)
{
    return place_a<J>(eq7(w, 0x0A0A0A0Au, 0x7F7F7F7Fu) * 0x00204081u);
}

__device__ __forceinline__ uint32_t chunk_newlines(const uint32_t w[8])
{
    return __brev(nl_word<0>(w[0]) | nl_word<1>(w[1]) | nl_word<2>(w[2]) | nl_word<3>(w[3]) |
                  nl_word<4>(w[4]) | nl_word<5>(w[5]) | nl_word<6>(w[6]) | nl_word<7>(w[7]));
}

// positions [0, k) of a chunk as a mask (k in 0..32)
// (one clamped funnel shift: the low word of (0xFFFFFFFF:0) >> min(k, 32))
__device__ __forceinline__ uint32_t before(int k) { return __funnelshift_rc(0u, 0xFFFFFFFFu, k); }

__device__ __forceinline__ uint32_t count_nl(const uint32_t w[8])
{
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) c += __popc(eq7(w[i], 0x0A0A0A0Au, 0x7F7F7F7Fu));
    return c;
}

// exclusive prefix sum over the 256 threads of the CTA; *total receives the CTA sum
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total)
{
    __shared__ uint32_t warp_sums[TILE_THREADS / 32];
    __shared__ uint32_t block_total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t s = lane < TILE_THREADS / 32 ? warp_sums[lane] : 0;
        uint32_t si = s;
#pragma unroll
        for (int d = 1; d < TILE_THREADS / 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, si, d);
            if (lane >= d) si += t;
        }
        if (lane < TILE_THREADS / 32) warp_sums[lane] = si - s;
        if (lane == TILE_THREADS / 32 - 1) block_total = si;
    }
    __syncthreads();
    uint32_t r = warp_sums[warp] + inc - v;
    if (total) *total = block_total;
    __syncthreads();
    return r;
}

// ---- FASTQ ---------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(TILE_THREADS)
k_count_newlines(const uint8_t* __restrict__ text, uint64_t n, uint32_t* __restrict__ tile_counts)
{
    uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    uint32_t w[8];
    uint32_t c = 0;
    if (off < n) { load32(text, n, off, w); c = count_nl(w); }
    uint32_t total;
    block_exclusive_scan(c, &total);
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

// FASTQ: a byte is a base of the query iff it lies on a line with index % 4 == 1 (:947) and is one
// of ACGTacgt (:978).  Only the sequence lines go into the planes, each followed by ONE invalid position
// (its newline), which breaks every 32-mer that would span two reads — headers, '+' lines and
// qualities are dropped, so every plane word is dense with window starts.

// Is text line `line` a sequence line that is counted?  keep (may be null = all): one bit per sequence line, in
// file order — the reference's rand() % 100 < ratio draw for that read when it down-samples (:955-960).
__device__ __forceinline__ bool seq_line(uint32_t line, const uint32_t* __restrict__ keep)
{
    if ((line & 3u) != 1u) return false;
    const uint32_t i = line >> 2;
    return !keep || ((__ldg(keep + (i >> 5)) >> (i & 31)) & 1u);
}

// bytes of the chunk that lie on a (kept) sequence line; with_nl: including the newline that ends such a line
__device__ __forceinline__ uint32_t seq_line_bytes(uint32_t nl, uint32_t line, bool with_nl, const uint32_t* __restrict__ keep)
{
    uint32_t seq = 0;
    if (!nl) return seq_line(line, keep) ? 0xFFFFFFFFu : 0u;
    uint32_t rest = nl;
    int start = 0;
    while (true) {
        const int k = rest ? __clz(rest) : 32;
        if (seq_line(line, keep)) seq |= before(with_nl ? min(k + 1, 32) : k) & ~before(start);
        if (k == 32) break;
        rest &= ~(0x80000000u >> k);
        start = k + 1; line++;
    }
    return seq;
}

// per tile: plane positions it produces (bytes of the kept sequence lines and their newlines); also
// cal_sam_ratio's sum (:1138-1140): all bytes of ALL sequence lines, valid or not
__global__ void __launch_bounds__(TILE_THREADS)
k_fastq_kept(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ tile_prefix,
             const uint32_t* __restrict__ keep, unsigned long long* __restrict__ tile_kept,
             unsigned long long* __restrict__ seq_bytes)
{
    __shared__ uint32_t s_part[TILE_THREADS / 32], s_seq[TILE_THREADS / 32];
    const uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    uint32_t w[8];
    load32(text, n, off, w);
    const uint32_t nl = chunk_newlines(w);
    const uint32_t line = tile_prefix[blockIdx.x] + block_exclusive_scan(__popc(nl), nullptr);
    uint32_t kept = 0, nseq = 0;
    if (off < n) {
        const uint32_t inr = before((int)min((uint64_t)32, n - off));
        if (keep) {
            kept = __popc(seq_line_bytes(nl, line, true, keep) & inr);
            nseq = __popc(seq_line_bytes(nl, line, false, nullptr) & inr);
        } else {                                          // every sequence line is kept: one pass gives both
            const uint32_t with_nl = seq_line_bytes(nl, line, true, nullptr) & inr;
            kept = __popc(with_nl);
            nseq = __popc(with_nl & ~nl);
        }
    }
    kept = __reduce_add_sync(0xffffffffu, kept);
    nseq = __reduce_add_sync(0xffffffffu, nseq);
    if ((threadIdx.x & 31) == 0) { s_part[threadIdx.x >> 5] = kept; s_seq[threadIdx.x >> 5] = nseq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0, q = 0;
#pragma unroll
        for (int k = 0; k < TILE_THREADS / 32; k++) { t += s_part[k]; q += s_seq[k]; }
        tile_kept[blockIdx.x] = t;
        if (q) atomicAdd(seq_bytes, (unsigned long long)q);       // one atomic per tile: they all hit one address
    }
}

constexpr int RING = 512;     // plane words a tile can touch (see k_pack_fasta), power of two

// Keep the bits of the masks selected by sel, packed towards bit 31 in order; returns how many.  sel has few runs
// (a newline or a header splits a chunk), so the loop runs once or twice.
__device__ __forceinline__ int compress_masks(Masks& m, uint32_t sel)
{
    if (sel == 0xFFFFFFFFu) return 32;
    uint32_t hi = 0, lo = 0, ok = 0, rest = sel;
    int n = 0;
    while (rest) {
        const int k0 = __clz(rest);
        const uint32_t inv = ~(rest << k0);
        const int len = inv ? __clz(inv) : 32;
        const uint32_t keep = before(len);
        hi |= ((m.hi << k0) & keep) >> n; lo |= ((m.lo << k0) & keep) >> n; ok |= ((m.ok << k0) & keep) >> n;
        n += len;
        rest &= ~(~before(k0) & before(k0 + len));
    }
    m.hi = hi; m.lo = lo; m.ok = ok;
    return n;
}

// A run of `len` consecutive chunk bytes starting at byte k0 goes to plane bits dest .. dest+len-1: OR it into
// the CTA's shared-memory piece of the planes (word index modulo RING) and track the words touched.
__device__ __forceinline__ void ring_emit(const Masks& m, int k0, int len, uint64_t dest, uint32_t* s_hi, uint32_t* s_lo,
                                          uint32_t* s_ok, uint32_t& wmin, uint32_t& wmax, uint32_t& i0, uint32_t& i1, bool& two)
{
    const uint32_t wi = (uint32_t)(dest >> 5);
    const int sh = (int)(dest & 31);
    const uint32_t keep = before(len);
    const uint32_t xh = (m.hi << k0) & keep, xl = (m.lo << k0) & keep, xo = (m.ok << k0) & keep;
    i0 = wi & (RING - 1); i1 = (wi + 1) & (RING - 1);
    if (xh >> sh) atomicOr(&s_hi[i0], xh >> sh);
    if (xl >> sh) atomicOr(&s_lo[i0], xl >> sh);
    if (xo >> sh) atomicOr(&s_ok[i0], xo >> sh);
    wmin = min(wmin, wi); wmax = max(wmax, wi);
    two = sh + len > 32;
    if (two) {
        const int up = 32 - sh;
        if (xh << up) atomicOr(&s_hi[i1], xh << up);
        if (xl << up) atomicOr(&s_lo[i1], xl << up);
        if (xo << up) atomicOr(&s_ok[i1], xo << up);
        wmax = max(wmax, wi + 1);
    }
}

// block-wide min / max of the touched word range; false when the CTA touched nothing
__device__ __forceinline__ bool ring_range(uint32_t& wmin, uint32_t& wmax)
{
    __shared__ uint32_t s_min[TILE_THREADS / 32], s_max[TILE_THREADS / 32];
    wmin = __reduce_min_sync(0xffffffffu, wmin);
    wmax = __reduce_max_sync(0xffffffffu, wmax);
    if ((threadIdx.x & 31) == 0) { s_min[threadIdx.x >> 5] = wmin; s_max[threadIdx.x >> 5] = wmax; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TILE_THREADS / 32; k++) { wmin = min(wmin, s_min[k]); wmax = max(wmax, s_max[k]); }
    EREF_ASSERT(wmin == 0xFFFFFFFFu || wmax - wmin < (uint32_t)RING);      // a tile's plane words fit the ring
    return wmin != 0xFFFFFFFFu;
}

__global__ void __launch_bounds__(TILE_THREADS)
k_pack_fastq(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ tile_prefix,
             const uint32_t* __restrict__ keep, const unsigned long long* __restrict__ tile_kept_pre, Planes pl)
{
    __shared__ __align__(16) uint32_t s_hi[RING], s_lo[RING], s_ok[RING];
    {
        const uint2 z = make_uint2(0u, 0u);
        reinterpret_cast<uint2*>(s_hi)[threadIdx.x] = z; reinterpret_cast<uint2*>(s_lo)[threadIdx.x] = z;
        reinterpret_cast<uint2*>(s_ok)[threadIdx.x] = z;
    }
    const uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    uint32_t w[8];
    load32(text, n, off, w);
    Masks m = chunk_masks(w);
    const uint32_t line = tile_prefix[blockIdx.x] + block_exclusive_scan(__popc(m.nl), nullptr);
    uint32_t kept = 0;
    if (off < n) {
        kept = seq_line_bytes(m.nl, line, true, keep) & before((int)min((uint64_t)32, n - off));
        m.ok &= kept; m.hi &= m.ok; m.lo &= m.ok;         // (a newline is never ACGT: it stays an invalid position)
    }
    const uint64_t dest0 = tile_kept_pre[blockIdx.x] + block_exclusive_scan(__popc(kept), nullptr);
    uint32_t wmin = 0xFFFFFFFFu, wmax = 0;
    uint32_t rest = kept;
    while (rest) {
        const int k0 = __clz(rest);
        const uint32_t inv = ~(rest << k0);
        const int len = inv ? __clz(inv) : 32;
        rest &= ~(~before(k0) & before(k0 + len));
        uint32_t i0, i1; bool two;
        ring_emit(m, k0, len, dest0 + __popc(kept & before(k0)), s_hi, s_lo, s_ok, wmin, wmax, i0, i1, two);
    }
    if (!ring_range(wmin, wmax)) return;
    for (uint32_t wd = wmin + threadIdx.x; wd <= wmax; wd += TILE_THREADS) {
        EREF_ASSERT(wd < pl.cap);
        const uint32_t i = wd & (RING - 1);
        const uint32_t vh = s_hi[i], vl = s_lo[i], vo = s_ok[i];
        if (wd == wmin || wd == wmax) {                   // shared with the neighbouring tiles: OR into the zeroed planes
            if (vh) atomicOr(pl.hi + wd, vh);
            if (vl) atomicOr(pl.lo + wd, vl);
            if (vo) atomicOr(pl.ok + wd, vo);
        } else {
            pl.hi[wd] = vh; pl.lo[wd] = vl; pl.ok[wd] = vo;
        }
    }
}

// ---- FASTA ---------------------------------------------------------------------------------------------
// Lines end at '\n' only; a line is a header iff its first byte is '>' (:687); every other byte of every
// other line is appended verbatim to the current record (:754), so '\r', blanks and N count as (invalid)
// bases.  Record 0 is whatever precedes the first header; record r >= 1 starts at the r-th header.
//
// Per chunk: has = a line starts inside the chunk; state = the last line started inside it is a header;
// hdrs = header lines started; seqA = non-newline bytes before the first line start (they belong to the
// line in progress and count only if that line is not a header); seqB = bytes of non-header lines after
// the first line start.  A tile is summarised the same way (TileSum in eref_internal.cuh).

struct FastaChunk {
    Masks m;
    uint32_t inr;       // bytes below n
    uint32_t hs;        // line starts inside the chunk that are header lines
    uint32_t hm;        // bytes of header lines started inside the chunk
    uint32_t pm;        // bytes before the first line start
    uint32_t has, state, hdrs, seqA, seqB;
};

// FULL = false: only the newline mask is formed (hi/lo/ok stay 0) — enough for the summaries.
template <bool FULL>
__device__ __forceinline__ FastaChunk fasta_chunk(const uint8_t* __restrict__ text, uint64_t n, uint64_t off)
{
    FastaChunk c;
    uint32_t w[8];
    load32(text, n, off, w);
    if (FULL) c.m = chunk_masks(w);
    else { c.m.hi = c.m.lo = c.m.ok = 0; c.m.nl = chunk_newlines(w); }
    const int lim = off < n ? (int)min((uint64_t)32, n - off) : 0;
    c.inr = before(lim);
    // the byte before the chunk decides whether byte 0 starts a line
    uint32_t prev_nl = __shfl_up_sync(0xffffffffu, c.m.nl & 1u, 1);
    if ((threadIdx.x & 31) == 0) prev_nl = off == 0 ? 1u : (off - 1 < n ? (uint32_t)(__ldg(text + off - 1) == 0x0Au) : 0u);
    const uint32_t ls = ((c.m.nl >> 1) | (prev_nl << 31)) & c.inr;
    c.hs = 0; c.hm = 0; c.has = 0; c.state = 0; c.hdrs = 0; c.seqB = 0;
    const uint32_t body = ~c.m.nl & c.inr;               // bytes that are not newlines
    if (!ls) {
        c.pm = 0xFFFFFFFFu;
        c.seqA = __popc(body);
    } else {
        uint32_t rest = ls;
        c.pm = before(__clz(rest));
        c.has = 1;
        while (rest) {
            const int k = __clz(rest);
            rest &= ~(0x80000000u >> k);
            const bool hdr = __ldg(text + off + k) == (uint8_t)'>';
            if (hdr) {
                c.hs |= 0x80000000u >> k;
                c.hm |= ~before(k) & before(rest ? __clz(rest) : 32);
            }
            c.state = hdr;
        }
        c.hdrs = __popc(c.hs);
        c.seqA = __popc(body & c.pm);
        c.seqB = __popc(body & ~c.pm & ~c.hm);
    }
    return c;
}

// Is the line in progress at the start of this thread's chunk a header?  Answered from the chunks before it
// in the tile: *known = false when no line starts between the tile's first byte and the chunk (the tile's
// own incoming state applies).  Also returns the tile's (has, state) pair in *tile_flags (bit 0, bit 1).
__device__ __forceinline__ uint32_t incoming_state(const FastaChunk& c, bool* known, uint32_t* tile_flags)
{
    __shared__ unsigned long long s_warp;                 // one byte per warp: bit 0 has, bit 1 state
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t hasm = __ballot_sync(0xffffffffu, c.has);
    const uint32_t stm = __ballot_sync(0xffffffffu, c.has && c.state);
    if (lane == 0) {
        const uint32_t last = hasm ? (stm >> (31 - __clz(hasm))) & 1u : 0u;
        reinterpret_cast<uint8_t*>(&s_warp)[warp] = (uint8_t)((hasm ? 1u : 0u) | (last << 1));
    }
    __syncthreads();
    const unsigned long long all = s_warp;
    const unsigned long long hb = all & 0x0101010101010101ull;
    if (tile_flags) {
        uint32_t f = 0;
        if (hb) { const int w = (63 - __clzll(hb)) >> 3; f = 1u | ((uint32_t)((all >> (8 * w + 1)) & 1ull) << 1); }
        *tile_flags = f;
    }
    const uint32_t prior = hasm & ((1u << lane) - 1u);
    if (prior) { *known = true; return (stm >> (31 - __clz(prior))) & 1u; }
    const unsigned long long pw = hb & ((1ull << (8 * warp)) - 1ull);
    if (pw) { *known = true; const int w = (63 - __clzll(pw)) >> 3; return (uint32_t)((all >> (8 * w + 1)) & 1ull); }
    *known = false;
    return 0;
}

// K1: one summary per tile
__global__ void __launch_bounds__(TILE_THREADS)
k_fasta_tiles(const uint8_t* __restrict__ text, uint64_t n, TileSum* __restrict__ tiles)
{
    __shared__ uint32_t s_a[TILE_THREADS / 32], s_b[TILE_THREADS / 32];
    const uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    const FastaChunk c = fasta_chunk<false>(text, n, off);
    bool known;
    uint32_t flags;
    const uint32_t in_hdr = incoming_state(c, &known, &flags);
    // bytes before the tile's first line start wait for the tile's incoming state (seqA); the rest is decided here
    uint32_t a = known ? 0u : c.seqA;
    uint32_t b = ((known && !in_hdr) ? c.seqA : 0u) + c.seqB + (c.hdrs << 16);
    a = __reduce_add_sync(0xffffffffu, a);
    b = __reduce_add_sync(0xffffffffu, b);
    if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a; s_b[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = 0; b = 0;
#pragma unroll
        for (int k = 0; k < TILE_THREADS / 32; k++) { a += s_a[k]; b += s_b[k]; }
        TileSum t;
        t.flags = flags; t.hdrs = b >> 16; t.seqA = a; t.seqB = b & 0xFFFFu;
        tiles[blockIdx.x] = t;
    }
}

// what a thread knows about the text before its chunk, from the tile prefix and the in-tile scan
struct Incoming { uint32_t in_header, rec, rank; };

__device__ __forceinline__ Incoming incoming(const FastaChunk& c, const TileSum& tp)
{
    bool known;
    uint32_t in_hdr = incoming_state(c, &known, nullptr);
    if (!known) in_hdr = (tp.flags & 1u) ? (tp.flags >> 1) & 1u : 0u;
    const uint32_t mine = (in_hdr ? 0u : c.seqA) + c.seqB + (c.hdrs << 16);      // < 2^29 per tile
    const uint32_t ex = block_exclusive_scan(mine, nullptr);
    Incoming in;
    in.in_header = in_hdr;
    in.rec = tp.hdrs + (ex >> 16);
    in.rank = (uint32_t)(tp.seqA + tp.seqB) + (ex & 0xFFFFu);
    return in;
}

// K3: first sequence rank and header offset of every record that starts in a tile with headers
__global__ void __launch_bounds__(TILE_THREADS)
k_fasta_records(const uint8_t* __restrict__ text, uint64_t n, const TileSum* __restrict__ tiles,
                const TileSum* __restrict__ tile_pre, uint32_t* __restrict__ rec_first,
                uint32_t* __restrict__ rec_hdr_off)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) { rec_first[0] = 0; rec_hdr_off[0] = 0xFFFFFFFFu; }
    if (tiles[blockIdx.x].hdrs == 0) return;
    const uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    const FastaChunk c = fasta_chunk<false>(text, n, off);
    const Incoming in = incoming(c, tile_pre[blockIdx.x]);
    if (!c.hs) return;
    const uint32_t seq = ~c.m.nl & c.inr & ~c.hm & (in.in_header ? ~c.pm : 0xFFFFFFFFu);
    uint32_t rest = c.hs;
    uint32_t r = in.rec;
    while (rest) {
        const int k = __clz(rest);
        rest &= ~(0x80000000u >> k);
        r++;
        rec_first[r] = in.rank + __popc(seq & before(k));
        rec_hdr_off[r] = (uint32_t)(off + k);
    }
}

// record length, keep flag (len > k, :697/:761) and packed size in words: (len>>5)+1 words always
// leaves at least one invalid pad position after the record, so no 32-mer spans two records.
__global__ void k_record_info(const uint32_t* __restrict__ rec_first, uint32_t n_recs, uint64_t total_seq,
                              uint32_t* __restrict__ rec_len, uint32_t* __restrict__ rec_kept,
                              uint32_t* __restrict__ rec_words, unsigned long long* __restrict__ kept_bases)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long mine = 0;
    if (r < n_recs) {
        uint32_t next = r + 1 < n_recs ? rec_first[r + 1] : (uint32_t)total_seq;
        uint32_t len = next - rec_first[r];
        bool kept = len > (uint32_t)EREF_K;
        rec_len[r] = len;
        rec_kept[r] = kept;
        rec_words[r] = kept ? (len >> 5) + 1 : 0;
        mine = kept ? len : 0;
    } else if (r == n_recs) {
        rec_len[r] = 0; rec_kept[r] = 0; rec_words[r] = 0;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(kept_bases, mine);
}

// kept records -> genome table; also names the owner of each record's last word, which holds only pad
// bits when len % 32 == 0 and is then never touched by k_pack_fasta
__global__ void k_genome_table(const uint32_t* __restrict__ rec_len, const uint32_t* __restrict__ rec_kept,
                               const uint32_t* __restrict__ rec_kept_ex, const uint32_t* __restrict__ rec_words_ex,
                               const uint32_t* __restrict__ rec_first, uint32_t n_recs,
                               uint32_t* __restrict__ g_rec, uint32_t* __restrict__ g_len,
                               uint32_t* __restrict__ g_word, uint32_t* __restrict__ g_first,
                               uint32_t* __restrict__ word_genome, unsigned long long* __restrict__ totals)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_recs) return;
    if (r == n_recs) {                                    // end sentinel + the counts the host reads back
        g_word[rec_kept_ex[r]] = rec_words_ex[r];
        totals[0] = rec_kept_ex[r]; totals[1] = rec_words_ex[r];
        return;
    }
    if (!rec_kept[r]) return;
    uint32_t g = rec_kept_ex[r];
    g_rec[g] = r; g_len[g] = rec_len[r]; g_word[g] = rec_words_ex[r]; g_first[g] = rec_first[r];
    word_genome[rec_words_ex[r] + (rec_len[r] >> 5)] = g;
}

// K5: text -> planes.  Within a chunk the sequence bytes of one line are one run of consecutive bytes
// and land on consecutive plane bits (newlines and headers simply vanish), so a run is placed with one
// shift: bit position = 32 * first word of the record + rank inside the record.  The CTA assembles its
// (contiguous, < 512 words) piece of the planes in shared memory and writes whole words; the first and
// last word may be shared with the neighbouring tiles and are OR-ed into the pre-zeroed planes.
__global__ void __launch_bounds__(TILE_THREADS)
k_pack_fasta(const uint8_t* __restrict__ text, uint64_t n, const TileSum* __restrict__ tile_pre,
             const uint32_t* __restrict__ rec_kept, const uint32_t* __restrict__ rec_kept_ex,
             const uint32_t* __restrict__ rec_words_ex, const uint32_t* __restrict__ rec_first,
             Planes pl, uint32_t* __restrict__ word_genome)
{
    __shared__ __align__(16) uint32_t s_hi[RING], s_lo[RING], s_ok[RING], s_g[RING];
    static_assert(RING == 2 * TILE_THREADS, "two words of each ring array per thread");
    {   // clear the ring with 8-byte stores: one per thread and array
        const uint2 z = make_uint2(0u, 0u), f = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
        reinterpret_cast<uint2*>(s_hi)[threadIdx.x] = z; reinterpret_cast<uint2*>(s_lo)[threadIdx.x] = z;
        reinterpret_cast<uint2*>(s_ok)[threadIdx.x] = z; reinterpret_cast<uint2*>(s_g)[threadIdx.x] = f;
    }
    const uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    const FastaChunk c = fasta_chunk<true>(text, n, off);
    const Incoming in = incoming(c, tile_pre[blockIdx.x]);         // (syncs: the ring is zeroed)
    const uint32_t seq = ~c.m.nl & c.inr & ~c.hm & (in.in_header ? ~c.pm : 0xFFFFFFFFu);
    uint32_t wmin = 0xFFFFFFFFu, wmax = 0;
    if (!c.hs) {
        // no record starts inside the chunk: its sequence bytes are consecutive in the planes (newlines vanish)
        if (seq && rec_kept[in.rec]) {
            Masks m = c.m;
            const int len = compress_masks(m, seq);
            const uint64_t dest = (uint64_t)rec_words_ex[in.rec] * 32 + (in.rank - rec_first[in.rec]);
            uint32_t i0, i1; bool two;
            ring_emit(m, 0, len, dest, s_hi, s_lo, s_ok, wmin, wmax, i0, i1, two);
            const uint32_t g = rec_kept_ex[in.rec];
            s_g[i0] = g;
            if (two) s_g[i1] = g;
        }
    } else {
        uint32_t rest = seq;
        while (rest) {                                    // one run per line piece; a header in between changes the record
            const int k0 = __clz(rest);
            const uint32_t inv = ~(rest << k0);
            const int len = inv ? __clz(inv) : 32;
            rest &= ~(~before(k0) & before(k0 + len));
            const uint32_t r = in.rec + __popc(c.hs & before(k0));
            if (!rec_kept[r]) continue;
            const uint32_t rank = in.rank + __popc(seq & before(k0));
            const uint64_t dest = (uint64_t)rec_words_ex[r] * 32 + (rank - rec_first[r]);
            uint32_t i0, i1; bool two;
            ring_emit(c.m, k0, len, dest, s_hi, s_lo, s_ok, wmin, wmax, i0, i1, two);
            const uint32_t g = rec_kept_ex[r];
            s_g[i0] = g;
            if (two) s_g[i1] = g;
        }
    }
    if (!ring_range(wmin, wmax)) return;                  // no sequence byte of a kept record in this tile
    for (uint32_t w = wmin + threadIdx.x; w <= wmax; w += TILE_THREADS) {
        EREF_ASSERT(w < pl.cap);
        const uint32_t i = w & (RING - 1);
        const uint32_t vh = s_hi[i], vl = s_lo[i], vo = s_ok[i], g = s_g[i];
        if (w == wmin || w == wmax) {
            if (vh) atomicOr(pl.hi + w, vh);
            if (vl) atomicOr(pl.lo + w, vl);
            if (vo) atomicOr(pl.ok + w, vo);
        } else {
            pl.hi[w] = vh; pl.lo[w] = vl; pl.ok[w] = vo;
        }
        if (g != 0xFFFFFFFFu) word_genome[w] = g;
    }
}

} // namespace

static inline unsigned grid_for(uint64_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

cudaError_t launch_count_newlines(const uint8_t* text, uint64_t n, uint32_t* tile_counts, uint64_t ntiles,
                                  cudaStream_t st)
{
    if (!ntiles) return cudaSuccess;
    k_count_newlines<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_counts);
    return cudaGetLastError();
}

cudaError_t launch_fastq_kept(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix, const uint32_t* keep,
                              unsigned long long* tile_kept, unsigned long long* seq_bytes, uint64_t ntiles, cudaStream_t st)
{
    if (!ntiles) return cudaSuccess;
    k_fastq_kept<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_prefix, keep, tile_kept, seq_bytes);
    return cudaGetLastError();
}

cudaError_t launch_pack_fastq(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix, const uint32_t* keep,
                              const unsigned long long* tile_kept_pre, Planes pl, cudaStream_t st)
{
    uint64_t ntiles = (n + TILE_BYTES - 1) / TILE_BYTES;
    if (!ntiles) return cudaSuccess;
    k_pack_fastq<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_prefix, keep, tile_kept_pre, pl);
    return cudaGetLastError();
}

cudaError_t launch_fasta_tiles(const uint8_t* text, uint64_t n, TileSum* tiles, uint64_t ntiles, cudaStream_t st)
{
    if (!ntiles) return cudaSuccess;
    k_fasta_tiles<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tiles);
    return cudaGetLastError();
}

cudaError_t launch_fasta_records(const uint8_t* text, uint64_t n, const TileSum* tiles, const TileSum* tile_pre,
                                 uint64_t ntiles, uint32_t* rec_first, uint32_t* rec_hdr_off, cudaStream_t st)
{
    if (!ntiles) {                                        // empty text: record 0 only
        cudaError_t e = cudaMemsetAsync(rec_first, 0, 4, st);
        if (e != cudaSuccess) return e;
        return cudaMemsetAsync(rec_hdr_off, 0xFF, 4, st);
    }
    k_fasta_records<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tiles, tile_pre, rec_first, rec_hdr_off);
    return cudaGetLastError();
}

cudaError_t launch_record_info(const uint32_t* rec_first, uint32_t n_recs, uint64_t total_seq,
                               uint32_t* rec_len, uint32_t* rec_kept, uint32_t* rec_words,
                               unsigned long long* kept_bases, cudaStream_t st)
{
    k_record_info<<<grid_for((uint64_t)n_recs + 1, 256), 256, 0, st>>>(rec_first, n_recs, total_seq, rec_len,
                                                                     rec_kept, rec_words, kept_bases);
    return cudaGetLastError();
}

cudaError_t launch_genome_table(const uint32_t* rec_len, const uint32_t* rec_kept, const uint32_t* rec_kept_ex,
                                const uint32_t* rec_words_ex, const uint32_t* rec_first, uint32_t n_recs,
                                uint32_t* g_rec, uint32_t* g_len, uint32_t* g_word, uint32_t* g_first,
                                uint32_t* word_genome, unsigned long long* totals, cudaStream_t st)
{
    k_genome_table<<<grid_for((uint64_t)n_recs + 1, 256), 256, 0, st>>>(rec_len, rec_kept, rec_kept_ex,
                                                                      rec_words_ex, rec_first, n_recs, g_rec,
                                                                      g_len, g_word, g_first, word_genome, totals);
    return cudaGetLastError();
}

cudaError_t launch_pack_fasta(const uint8_t* text, uint64_t n, const TileSum* tile_pre, uint64_t ntiles,
                              const uint32_t* rec_kept, const uint32_t* rec_kept_ex, const uint32_t* rec_words_ex,
                              const uint32_t* rec_first, Planes pl, uint32_t* word_genome, cudaStream_t st)
{
    if (!ntiles) return cudaSuccess;
    k_pack_fasta<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_pre, rec_kept, rec_kept_ex, rec_words_ex,
                                                           rec_first, pl, word_genome);
    return cudaGetLastError();
}
