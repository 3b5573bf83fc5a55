// eref_internal.cuh — shared declarations of liberef_gpu.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/eref_gpu.h"

// ---- bounds-checked build (make -C palace_b200/csrc bounds -> ../liberef_gpu_bounds.so) ----------------------------------
// compute-sanitizer is not available on every pool, so the kernels carry their own checks: every index whose bound is
// a capacity (ring words per tile, staging slots, code-list slots, list entries, plane words) is asserted under
// -DEREF_BOUNDS; a failed assert aborts the kernel and the next API call reports cudaErrorAssert.  The checks compile
// to nothing in the product build.
#ifdef EREF_BOUNDS
#include <cassert>
#define EREF_ASSERT(x) assert(x)
#else
#define EREF_ASSERT(x) ((void)0)
#endif

// ---- layout constants ---------------------------------------------------------------------------
// Text tiles: one CTA of 256 threads frames 8 KiB of text, 32 bytes per thread, so one thread owns
// exactly one 32-position word of the bit-planes.
constexpr int      TILE_THREADS = 256;
constexpr int      TILE_BYTES   = TILE_THREADS * 32;
constexpr uint32_t TABLE_WORDS  = 1u << 27;          // 2^32 one-bit slots per level (bin/extract_ref.cpp:25)
constexpr int      WINDOW       = 500;               // :510
constexpr int      WIN_WORDS    = 16;                // a 500-position window reaches back 16 words
constexpr uint32_t SUMMARY_WORDS = 1u << 23;         // level-3 summary: 1 bit per 16 slots = 32 MiB
constexpr int      MAX_BUCKETS = 256;                // coarse / fine partition fan-out of the table build (table.cu)
#ifndef EREF_SLICE_BITS
#define EREF_SLICE_BITS 19
#endif
constexpr int      SLICE_BITS  = EREF_SLICE_BITS;    // a table slice = 2^SLICE_BITS slots: at 19 its three levels (192 KiB) fill one CTA's shared memory; 18 = two CTAs per SM
constexpr int      NSLICES     = 1 << (32 - SLICE_BITS);
constexpr int      SLICE_WORDS = 1 << (SLICE_BITS - 5);
constexpr uint32_t SLICE_MASK  = (1u << SLICE_BITS) - 1u;
constexpr int      FINE_CHUNK  = 8192;               // codes a CTA of the fine partition sorts in shared memory
constexpr int      MAX_LISTS   = 16;                 // query files counted by one table build
constexpr int      SAMPLE_STRIDE_MIN = 4, SAMPLE_STRIDE_MAX = 8;   // candidate filter probes every s-th position
constexpr int      SAMPLE_MIN_HITS = 6;              // smallest hit threshold worth a sparser sampling stride

// Per-stream select masks of the coder assignment: bit (31-z) of sel0[i] (sel2[i]) is set iff code
// stream i uses 1-bit alphabet 0 (2) at window position z; alphabet 1 everywhere else (:972).
struct CoderMasks {
    uint32_t sel0[3];
    uint32_t sel2[3];
};

// Three bit-planes over a position space, MSB-first: position p lives in word p>>5, bit 31-(p&31).
//   hi/lo : 2-bit base code A=00 C=01 G=10 T=11;  ok : 1 iff the byte is one of ACGTacgt (and, for
//   FASTQ, lies on a sequence line).  One zero pad word follows the last word.
struct Planes {
    uint32_t* hi;
    uint32_t* lo;
    uint32_t* ok;
    uint64_t  nwords;        // without the pad word
    uint64_t  cap;           // words allocated per plane (bounds-checked build)
};

// FASTA framing state of a text tile (frame.cu), combined left to right by TileOp: flags bit 0 = a line
// starts inside the tile, bit 1 = the last line started inside it is a header; hdrs = header lines started;
// seqA = non-newline bytes before the first line start, seqB = bytes of non-header lines after it.
struct TileSum {
    uint32_t flags, hdrs;
    unsigned long long seqA, seqB;
};

struct TileOp {
    __host__ __device__ __forceinline__ TileSum operator()(const TileSum& x, const TileSum& y) const
    {
        TileSum r;
        r.flags = (y.flags & 1u) ? y.flags : (x.flags | (y.flags & 1u));
        r.hdrs = x.hdrs + y.hdrs;
        if (x.flags & 1u) { r.seqA = x.seqA; r.seqB = x.seqB + y.seqB + ((x.flags & 2u) ? 0ull : y.seqA); }
        else              { r.seqA = x.seqA + y.seqA; r.seqB = y.seqB; }
        return r;
    }
};

__device__ __forceinline__ uint32_t bitsel(uint32_t mask, uint32_t a, uint32_t b)
{
    return (a & mask) | (b & ~mask);                 // one LOP3
}

// Valid-window mask for the 32 window starts of one word: bit 31-s set iff ok bits s..s+31 of the
// 64-bit string (ok0:ok1) are all ones, i.e. the 32-mer starting s positions into the word has no
// invalid base (:978-982).
__device__ __forceinline__ uint32_t valid_windows(uint32_t ok0, uint32_t ok1)
{
    uint64_t t = ((uint64_t)ok0 << 32) | ok1;
    t &= t << 1; t &= t << 2; t &= t << 4; t &= t << 8; t &= t << 16;
    return (uint32_t)(t >> 32);
}

// 32-position window starting s positions into (w0:w1), first position in the MSB.
__device__ __forceinline__ uint32_t window32(uint32_t w0, uint32_t w1, int s)
{
    return __funnelshift_l(w1, w0, s);
}

// The canonical code of stream i for the window (hi, lo) — restates the z-loop of
// bin/extract_ref.cpp:971-994 through bit-planes:
//   alphabet0 = ~(hi^lo) (A,T->1), alphabet1 = ~hi (A,C->1), alphabet2 = ~lo (A,G->1);
//   complementing a base keeps alphabet0 and flips alphabets 1 and 2, so the reverse-complement
//   code is the same select over bit-reversed planes rh = brev(hi), rl = brev(lo):
//   alphabet0' = ~(rh^rl), alphabet1' = rh, alphabet2' = rl.
struct WindowBits {
    uint32_t a0, a1, a2;     // forward alphabets
    uint32_t r0, r1, r2;     // reverse-complement alphabets
};

__device__ __forceinline__ WindowBits window_bits(uint32_t hi, uint32_t lo)
{
    WindowBits b;
    b.a0 = ~(hi ^ lo); b.a1 = ~hi; b.a2 = ~lo;
    uint32_t rh = __brev(hi), rl = __brev(lo);
    b.r0 = ~(rh ^ rl); b.r1 = rh; b.r2 = rl;
    return b;
}

__device__ __forceinline__ uint32_t stream_code(const CoderMasks& m, const WindowBits& b, int i)
{
    uint32_t fwd = bitsel(m.sel2[i], b.a2, bitsel(m.sel0[i], b.a0, b.a1));
    uint32_t rc  = bitsel(m.sel2[i], b.r2, bitsel(m.sel0[i], b.r0, b.r1));
    return min(fwd, rc);                             // :989-994, unsigned compare
}

// byte -> (2-bit code, validity).  A=0x41 C=0x43 G=0x47 T=0x54 (and +0x20): bits 2..1 give
// A=00 C=01 G=11 T=10; hi = bit2, lo = bit1 ^ bit2.
__device__ __forceinline__ bool is_acgt(uint32_t c)
{
    uint32_t u = c & 0xDFu;
    return u == 0x41u || u == 0x43u || u == 0x47u || u == 0x54u;
}

#define CUDA_TRY(expr)                                                                     \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess) return eref_fail_cuda(e_, #expr, __FILE__, __LINE__);       \
    } while (0)

int eref_fail_cuda(cudaError_t e, const char* what, const char* file, int line);
int eref_fail(int code, const char* fmt, ...);

// ---- kernel launchers (each returns a cudaError_t from the launch) ------------------------------
// frame.cu
cudaError_t launch_count_newlines(const uint8_t* text, uint64_t n, uint32_t* tile_counts,
                                  uint64_t ntiles, cudaStream_t st);
cudaError_t launch_fastq_kept(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix, const uint32_t* keep,
                              unsigned long long* tile_kept, unsigned long long* seq_bytes, uint64_t ntiles, cudaStream_t st);
cudaError_t launch_pack_fastq(const uint8_t* text, uint64_t n, const uint32_t* tile_prefix, const uint32_t* keep,
                              const unsigned long long* tile_kept_pre, Planes pl, cudaStream_t st);
cudaError_t launch_fasta_tiles(const uint8_t* text, uint64_t n, TileSum* tiles, uint64_t ntiles, cudaStream_t st);
cudaError_t launch_fasta_records(const uint8_t* text, uint64_t n, const TileSum* tiles, const TileSum* tile_pre,
                                 uint64_t ntiles, uint32_t* rec_first, uint32_t* rec_hdr_off, cudaStream_t st);
cudaError_t launch_record_info(const uint32_t* rec_first, uint32_t n_recs, uint64_t total_seq,
                               uint32_t* rec_len, uint32_t* rec_kept, uint32_t* rec_words,
                               unsigned long long* kept_bases, cudaStream_t st);
cudaError_t launch_genome_table(const uint32_t* rec_len, const uint32_t* rec_kept, const uint32_t* rec_kept_ex,
                                const uint32_t* rec_words_ex, const uint32_t* rec_first, uint32_t n_recs,
                                uint32_t* g_rec, uint32_t* g_len, uint32_t* g_word, uint32_t* g_first,
                                uint32_t* word_genome, unsigned long long* totals, cudaStream_t st);
cudaError_t launch_pack_fasta(const uint8_t* text, uint64_t n, const TileSum* tile_pre, uint64_t ntiles,
                              const uint32_t* rec_kept, const uint32_t* rec_kept_ex, const uint32_t* rec_words_ex,
                              const uint32_t* rec_first, Planes pl, uint32_t* word_genome, cudaStream_t st);
// table.cu
cudaError_t launch_count_kmers(Planes pl, CoderMasks m, uint32_t* bm1, uint32_t* bm2, uint32_t* bm3,
                               cudaStream_t st);
// s0/ns: the table slices whose codes are kept (all of them unless the build is sharded over GPUs)
cudaError_t launch_slice_hist(Planes pl, const unsigned long long* plane_bits, CoderMasks m, unsigned long long* g_hist,
                              uint32_t s0, uint32_t ns, cudaStream_t st);
cudaError_t launch_coarse_scatter(Planes pl, const unsigned long long* plane_bits, const uint32_t* skip, CoderMasks m, int cbits,
                                  const unsigned long long* slice_off, unsigned long long* g_tail, uint32_t* codes_out,
                                  unsigned long long out_cap, uint32_t s0, uint32_t ns, cudaStream_t st);
cudaError_t launch_fine_partition(const uint32_t* in, const unsigned long long* slice_off, const unsigned long long* coarse_n,
                                  const uint32_t* chunk_start, uint32_t nchunks, int cbits, unsigned long long* g_tail,
                                  uint32_t* out, unsigned long long out_cap, cudaStream_t st);
cudaError_t launch_slice_offsets(const unsigned long long* hist, int cbits, unsigned long long* off, unsigned long long* cnt,
                                 unsigned long long* coarse_n, uint32_t* chunk_start, unsigned long long* totals,
                                 unsigned long long cap_codes, uint32_t* overflow, cudaStream_t st);
// slice-ordered code lists of the files one table build counts: list l holds cnt[l][s] codes of slice s at
// codes[l] + off[l][s] (16-byte aligned)
struct CountLists {
    const uint32_t* codes[MAX_LISTS];
    const unsigned long long* off[MAX_LISTS];
    const unsigned long long* cnt[MAX_LISTS];
    unsigned long long cap[MAX_LISTS];      // codes allocated per list (bounds-checked build)
    int n;
};
// Sharded table build: one GPU ("part") counts the slices [s0, s0 + ns) and stores them into the table of every GPU
// of the build through peer pointers (index 0 = its own table; n = 1: a local build).  flags[p] points at GPU p's
// flag words (one per part): part `part` of build `epoch` has landed there once flags[p][part] >= epoch.
constexpr int MAX_PARTS = 16;
struct TableShare {
    uint32_t* bm3[MAX_PARTS];
    uint32_t* summary[MAX_PARTS];
    uint32_t* flags[MAX_PARTS];
    unsigned int* done;              // CTAs of the current build that have finished (this GPU's own counter)
    int n, part;
    uint32_t s0, ns, epoch;
};
cudaError_t launch_count_slices(const CountLists& lists, const TableShare& share, unsigned long long* level_pop,
                                cudaStream_t st);
cudaError_t launch_table_wait(const uint32_t* flags, int n, uint32_t epoch, uint32_t* err, cudaStream_t st);
// shared reads: peer lists (arena layout) whose slices [s0, s0 + ns) are pulled into local staging, one slot per list
struct PullLists {
    const uint8_t* src[MAX_LISTS];
    uint8_t* dst;
    size_t list_stride;
    unsigned long long cap_codes;
    uint32_t s0, ns;
    int n;
};
cudaError_t preload_share_kernels();
cudaError_t launch_lists_signal(const TableShare& share, uint32_t flag0, uint32_t epoch, cudaStream_t st);
cudaError_t launch_pull_lists(const PullLists& pl, cudaStream_t st);
cudaError_t launch_build_summary(const uint32_t* bm3, uint32_t* summary, uint32_t summary_words, cudaStream_t st);
cudaError_t launch_popcount(const uint32_t* bm, uint64_t nwords, unsigned long long* out, cudaStream_t st);
// scan.cu
cudaError_t launch_probe_sampled(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* summary, int stride,
                                 uint8_t* sample_hits, cudaStream_t st);
cudaError_t launch_need(const uint8_t* sample_hits, const uint32_t* word_genome, uint64_t nwords, int min_sample_hits,
                        uint8_t* need, uint32_t* list, unsigned long long* need_count, uint32_t* single,
                        uint32_t* trio, uint32_t* good, uint32_t* genome_listed, cudaStream_t st);
cudaError_t launch_probe_full(Planes pl, CoderMasks m, const uint32_t* bm3, const uint8_t* need,
                              uint32_t* single, uint32_t* trio, cudaStream_t st);
cudaError_t launch_probe_list(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* list,
                              const unsigned long long* n_list, uint32_t* single, uint32_t* trio, cudaStream_t st);
cudaError_t launch_probe_parts(Planes pl, CoderMasks m, const uint32_t* bm3, const uint32_t* list,
                               const unsigned long long* n_list, int parts, uint32_t* h0w, uint32_t* h1w, uint32_t* h2w,
                               uint32_t* single, uint32_t* trio, cudaStream_t st, int* launches);
cudaError_t launch_window(const uint32_t* single, const uint32_t* trio, const uint32_t* word_genome,
                          const uint32_t* g_word, const uint32_t* g_len, uint64_t nwords, int one_min, int three_min,
                          uint32_t* good, cudaStream_t st);
cudaError_t launch_window_list(const uint32_t* single, const uint32_t* trio, const uint32_t* list,
                               const unsigned long long* n_list, const uint32_t* word_genome, const uint32_t* g_word,
                               const uint32_t* g_len, uint64_t nwords, int one_min, int three_min, uint32_t* good,
                               cudaStream_t st);
// genome_listed: null, or one word per record, non-zero iff k_need listed a word of it (others answer {0, 0} unread)
cudaError_t launch_intervals(const uint32_t* good, const uint32_t* g_word, const uint32_t* g_len,
                             uint32_t n_genomes, const uint32_t* genome_listed, uint32_t* results, cudaStream_t st);
cudaError_t launch_add_ordinal(uint32_t* results, uint32_t n, uint32_t base, cudaStream_t st);
cudaError_t launch_record_codes(Planes pl, CoderMasks m, uint64_t first_word, uint32_t len,
                                uint32_t* codes, cudaStream_t st);
cudaError_t launch_index_body(Planes pl, CoderMasks m, const uint32_t* g_word, const uint32_t* g_len,
                              const unsigned long long* cell_pre, uint32_t g0, uint32_t g1, unsigned long long ncells,
                              uint32_t* out, cudaStream_t st);
