// frame.cu — text framing and packing kernels (north_star part 1): FASTQ/FASTA bytes in HBM ->
// three bit-planes (hi, lo, ok).  Replaces the getline loops of bin/extract_ref.cpp:686-756 (FASTA
// records) and :940-947 (FASTQ sequence lines).
#include "eref_internal.cuh"

namespace {

constexpr uint64_t SEQ_MASK = (1ull << 40) - 1;      // line_val/line_cum: low 40 bits = sequence bytes

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
        for (uint64_t k = off; k < n; k++) {
            uint32_t j = (uint32_t)(k - off);
            w[j >> 2] |= (uint32_t)text[k] << ((j & 3) * 8);
        }
    }
}

__device__ __forceinline__ uint32_t count_nl(const uint32_t w[8])
{
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) c += __popc(__vcmpeq4(w[i], 0x0A0A0A0Au) & 0x01010101u);
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
// of ACGTacgt (:978).  Everything else (headers, '+', qualities, '\n', '\r', N) clears the ok bit,
// which breaks every 32-mer that would span it.
__global__ void __launch_bounds__(TILE_THREADS)
k_pack_fastq(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ tile_prefix,
             Planes pl, unsigned long long* __restrict__ seq_bytes)
{
    uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    uint32_t w[8];
    uint32_t c = 0;
    const bool live = off < n;
    if (live) { load32(text, n, off, w); c = count_nl(w); }
    uint32_t line = tile_prefix[blockIdx.x] + block_exclusive_scan(c, nullptr);
    uint32_t hi = 0, lo = 0, ok = 0, nseq = 0;
    if (live) {
        const uint32_t lim = (uint32_t)min((uint64_t)32, n - off);
#pragma unroll
        for (int k = 0; k < 32; k++) {
            uint32_t ch = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            bool seq_line = (line & 3u) == 1u;
            if (ch == 0x0Au) line++;
            else if (seq_line && (uint32_t)k < lim) nseq++;
            uint32_t b2 = (ch >> 2) & 1u, b1 = (ch >> 1) & 1u;
            hi |= b2 << (31 - k);
            lo |= (b1 ^ b2) << (31 - k);
            ok |= (uint32_t)(seq_line && is_acgt(ch)) << (31 - k);
        }
        uint64_t word = off >> 5;
        pl.hi[word] = hi; pl.lo[word] = lo; pl.ok[word] = ok;
    }
    // cal_sam_ratio's sum (:1138-1140): all bytes of sequence lines, valid or not
#pragma unroll
    for (int d = 16; d; d >>= 1) nseq += __shfl_xor_sync(0xffffffffu, nseq, d);
    if ((threadIdx.x & 31) == 0 && nseq) atomicAdd(seq_bytes, (unsigned long long)nseq);
}

__global__ void __launch_bounds__(TILE_THREADS)
k_write_newlines(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ tile_prefix,
                 uint32_t* __restrict__ nl_pos)
{
    uint64_t off = (uint64_t)blockIdx.x * TILE_BYTES + (uint64_t)threadIdx.x * 32;
    uint32_t w[8];
    uint32_t c = 0;
    const bool live = off < n;
    if (live) { load32(text, n, off, w); c = count_nl(w); }
    uint32_t idx = tile_prefix[blockIdx.x] + block_exclusive_scan(c, nullptr);
    if (live && c) {
#pragma unroll
        for (int k = 0; k < 32; k++) {
            uint32_t ch = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            if (ch == 0x0Au) nl_pos[idx++] = (uint32_t)(off + k);
        }
    }
}

// One value per text line: (is_header << 40) | sequence_bytes.  A line is a header iff its first
// byte is '>' (:687); every other line is appended verbatim to the current record (:754).
__global__ void k_line_info(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ nl_pos,
                            uint32_t n_nl, uint32_t n_lines, uint64_t* __restrict__ line_val)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_lines) return;
    if (i == n_lines) { line_val[i] = 0; return; }       // sentinel so the scan yields the totals
    uint64_t start = i ? (uint64_t)nl_pos[i - 1] + 1 : 0;
    uint64_t end = i < n_nl ? (uint64_t)nl_pos[i] : n;
    bool hdr = end > start && text[start] == '>';
    line_val[i] = hdr ? (1ull << 40) : (end - start);
}

// record r >= 1 starts at the r-th header line; record 0 is whatever precedes the first header.
__global__ void k_record_starts(const uint64_t* __restrict__ line_val, const uint64_t* __restrict__ line_cum,
                                uint32_t n_lines, uint32_t* __restrict__ rec_hdr_line,
                                uint32_t* __restrict__ rec_first)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { rec_hdr_line[0] = 0xFFFFFFFFu; rec_first[0] = 0; }
    if (i >= n_lines) return;
    if (line_val[i] >> 40) {
        uint32_t r = (uint32_t)(line_cum[i] >> 40) + 1;
        rec_hdr_line[r] = i;
        rec_first[r] = (uint32_t)(line_cum[i] & SEQ_MASK);
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

__global__ void k_genome_table(const uint32_t* __restrict__ rec_len, const uint32_t* __restrict__ rec_kept,
                               const uint32_t* __restrict__ rec_kept_ex, const uint32_t* __restrict__ rec_words_ex,
                               const uint32_t* __restrict__ rec_first, uint32_t n_recs,
                               uint32_t* __restrict__ g_rec, uint32_t* __restrict__ g_len,
                               uint32_t* __restrict__ g_word, uint32_t* __restrict__ g_first)
{
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_recs) return;
    if (r == n_recs) { g_word[rec_kept_ex[r]] = rec_words_ex[r]; return; }   // end sentinel
    if (!rec_kept[r]) return;
    uint32_t g = rec_kept_ex[r];
    g_rec[g] = r; g_len[g] = rec_len[r]; g_word[g] = rec_words_ex[r]; g_first[g] = rec_first[r];
}

// One thread per packed word: locate the record and the text line holding the word's first base,
// then walk the text skipping newlines (the only bytes between two sequence lines of a record).
__global__ void __launch_bounds__(256)
k_pack_fasta(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ nl_pos, uint32_t n_nl,
             const uint64_t* __restrict__ line_cum, uint32_t n_lines, const uint32_t* __restrict__ rec_hdr_line,
             uint32_t n_recs, const uint32_t* __restrict__ g_rec, const uint32_t* __restrict__ g_len,
             const uint32_t* __restrict__ g_word, const uint32_t* __restrict__ g_first, uint32_t n_genomes,
             Planes pl, uint32_t* __restrict__ word_genome)
{
    uint64_t W = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (W >= pl.nwords) return;
    // genome = last g with g_word[g] <= W
    uint32_t lo = 0, hi = n_genomes;                 // invariant: g_word[lo] <= W < g_word[hi]
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if ((uint64_t)g_word[mid] <= W) lo = mid; else hi = mid;
    }
    const uint32_t g = lo;
    word_genome[W] = g;
    const uint32_t len = g_len[g];
    const uint32_t p0 = (uint32_t)(W - g_word[g]) << 5;
    uint32_t vhi = 0, vlo = 0, vok = 0;
    if (p0 < len) {
        const uint32_t r = g_rec[g];
        const uint64_t rank = (uint64_t)g_first[g] + p0;
        uint32_t a = r ? rec_hdr_line[r] + 1 : 0;                     // first line of the record
        uint32_t b = r + 1 < n_recs ? rec_hdr_line[r + 1] : n_lines;  // one past its last line
        // last line i in [a, b) with seqcum(i) <= rank: the line that holds byte `rank`
        while (b - a > 1) {
            uint32_t mid = (a + b) >> 1;
            if ((line_cum[mid] & SEQ_MASK) <= rank) a = mid; else b = mid;
        }
        uint64_t pos = (a ? (uint64_t)nl_pos[a - 1] + 1 : 0) + (rank - (line_cum[a] & SEQ_MASK));
        const uint32_t cnt = min(32u, len - p0);
        uint32_t k = 0;
        while (k < cnt && pos < n) {
            uint32_t ch = text[pos++];
            if (ch == 0x0Au) continue;
            uint32_t b2 = (ch >> 2) & 1u, b1 = (ch >> 1) & 1u;
            vhi |= b2 << (31 - k);
            vlo |= (b1 ^ b2) << (31 - k);
            vok |= (uint32_t)is_acgt(ch) << (31 - k);
            k++;
        }
    }
    pl.hi[W] = vhi; pl.lo[W] = vlo; pl.ok[W] = vok;
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

cudaError_t launch_pack_fastq(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix, Planes pl,
                              unsigned long long* seq_bytes, cudaStream_t st)
{
    uint64_t ntiles = (n + TILE_BYTES - 1) / TILE_BYTES;
    if (!ntiles) return cudaSuccess;
    k_pack_fastq<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_prefix, pl, seq_bytes);
    return cudaGetLastError();
}

cudaError_t launch_write_newlines(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix,
                                  uint32_t* nl_pos, uint64_t ntiles, cudaStream_t st)
{
    if (!ntiles) return cudaSuccess;
    k_write_newlines<<<(unsigned)ntiles, TILE_THREADS, 0, st>>>(text, n, tile_prefix, nl_pos);
    return cudaGetLastError();
}

cudaError_t launch_line_info(const uint8_t* text, uint64_t n, const uint32_t* nl_pos, uint32_t n_nl,
                             uint32_t n_lines, uint64_t* line_val, cudaStream_t st)
{
    k_line_info<<<grid_for((uint64_t)n_lines + 1, 256), 256, 0, st>>>(text, n, nl_pos, n_nl, n_lines, line_val);
    return cudaGetLastError();
}

cudaError_t launch_record_starts(const uint64_t* line_val, const uint64_t* line_cum, uint32_t n_lines,
                                 uint32_t* rec_hdr_line, uint32_t* rec_first, cudaStream_t st)
{
    k_record_starts<<<grid_for((uint64_t)n_lines + 1, 256), 256, 0, st>>>(line_val, line_cum, n_lines,
                                                                        rec_hdr_line, rec_first);
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
                                const uint32_t* rec_words_ex, const uint32_t* rec_first,
                                const uint32_t* /*rec_hdr_line*/, uint32_t n_recs, uint32_t* g_rec,
                                uint32_t* g_len, uint32_t* g_word, uint32_t* g_first, cudaStream_t st)
{
    k_genome_table<<<grid_for((uint64_t)n_recs + 1, 256), 256, 0, st>>>(rec_len, rec_kept, rec_kept_ex,
                                                                      rec_words_ex, rec_first, n_recs, g_rec,
                                                                      g_len, g_word, g_first);
    return cudaGetLastError();
}

cudaError_t launch_pack_fasta(const uint8_t* text, uint64_t n, const uint32_t* nl_pos, uint32_t n_nl,
                              const uint64_t* line_cum, uint32_t n_lines, const uint32_t* rec_hdr_line,
                              uint32_t n_recs, const uint32_t* g_rec, const uint32_t* g_len,
                              const uint32_t* g_word, const uint32_t* g_first, uint32_t n_genomes, Planes pl,
                              uint32_t* word_genome, cudaStream_t st)
{
    if (!pl.nwords) return cudaSuccess;
    k_pack_fasta<<<grid_for(pl.nwords, 256), 256, 0, st>>>(text, n, nl_pos, n_nl, line_cum, n_lines,
                                                           rec_hdr_line, n_recs, g_rec, g_len, g_word, g_first,
                                                           n_genomes, pl, word_genome);
    return cudaGetLastError();
}
