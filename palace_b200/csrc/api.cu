// api.cu — the C ABI of liberef_gpu.so (include/eref_gpu.h): context, workspaces, phase ordering.
// Prefix sums over tile/line/record tables use cub::DeviceScan (CUDA toolkit); every kernel that
// touches sequence data is hand-written in frame.cu / table.cu / scan.cu.
#include "eref_internal.cuh"
#include <cub/device/device_scan.cuh>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <chrono>
#include <vector>
#include <algorithm>
#include <new>

namespace {
thread_local char g_err[512] = "";
}

int eref_fail(int code, const char* fmt, ...)
{
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

int eref_fail_cuda(cudaError_t e, const char* what, const char* file, int line)
{
    snprintf(g_err, sizeof g_err, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return EREF_ERR_CUDA;
}

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool view = false;                         // a window into memory someone else owns (fixed size, never freed here)
    void attach(void* q, size_t bytes) { release(); p = q; cap = bytes; view = true; }
    cudaError_t ensure(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (view) return cudaErrorMemoryAllocation;
        if (p) { cudaError_t e = cudaFree(p); p = nullptr; cap = 0; if (e != cudaSuccess) return e; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    // grow keeping the first `keep` bytes (stream-ordered copy; the old block is freed after it)
    cudaError_t grow(size_t bytes, size_t keep, cudaStream_t st)
    {
        if (bytes <= cap) return cudaSuccess;
        void* q = nullptr;
        const size_t want = bytes + bytes / 2 + 256;
        cudaError_t e = cudaMalloc(&q, want);
        if (e != cudaSuccess) return e;
        if (p && keep) e = cudaMemcpyAsync(q, p, keep, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess && p) { e = cudaStreamSynchronize(st); if (e == cudaSuccess) e = cudaFree(p); }
        if (e != cudaSuccess) { cudaFree(q); return e; }
        p = q; cap = want;
        return cudaSuccess;
    }
    void release() { if (p && !view) cudaFree(p); p = nullptr; cap = 0; view = false; }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

enum Ev { E0 = 0, E1, E2, E3, E4, E5, E6, E7, E8, E9, E10, E11,
          R1, R2, R3, R4,                          // eref_ref_set: text landed, framed (census stream), packed; R4 = packing starts (E0 = before the upload)
          S0, S1, S2, S3, S4, S5, S6, S7,          // scan: start, summary, sampled, need, probed, windows, records, copied out
          EV_COUNT };

} // namespace

struct eref_ctx {
    int device = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr, copy_stream = nullptr;
    unsigned long long* h_pin = nullptr;       // pinned scratch for the small device->host reads
    bool host_tables_valid = false;            // h_g_* mirrors fetched lazily (test hooks only)
    std::vector<cudaEvent_t> copy_ev;          // eref_run_host: one event per staged H2D copy
    DevBuf q_text2, run_results;
    size_t run_total = 0;                      // records of the last eref_run_host, in run_results
    // down-sampling masks handed in by the caller for the next eref_run_host (eref_run_host_keep)
    std::vector<uint32_t> pend_keep1, pend_keep2;
    uint64_t pend_lines1 = 0, pend_lines2 = 0;
    bool pend_keep = false;
    CoderMasks masks{};
    float hit_ratio = 0, perfect_hit_ratio = 0;
    int one_min = 0, three_min = 0, sample_stride = 4;
    uint64_t down_sampling_size = 2000000000ull;       // :1230 (EREF_DOWN_SAMPLING_SIZE: test hook)
    unsigned flags = 0;
    cudaEvent_t ev[EV_COUNT]{};
    // query table: level L bitmap = "code seen >= L times"
    uint32_t* bm[3] = {nullptr, nullptr, nullptr};
    // shared scratch
    DevBuf tile_counts, tile_prefix, cub_tmp, counters;      // counters: 4 x u64
    // query side
    DevBuf q_text, q_hi, q_lo, q_ok, tile_kept, tile_kept_pre, q_keep;
    // Table build (table.cu): every file added since the last reset is kept as a slice-ordered code list (codes
    // radix-partitioned by their top 13 bits); a build counts all lists slice by slice in shared memory and
    // writes level 3 and its summary.  Levels 1/2 exist in HBM only for the direct build (EREF_FLAG_DIRECT_COUNT).
    struct CodeList { DevBuf codes, off, cnt; unsigned long long total = 0; bool total_pending = false; };
    bool retry_pending = false;
    bool force_exact = false;                  // size the code buffers from the real count (one host wait per file)
    uint32_t* guard = nullptr;                 // device word: a file's codes did not fit buffers sized before they were counted
    std::vector<CodeList> lists;
    size_t n_lists = 0;
    bool table_dirty = true;                   // files added / table reset since the last build
    bool want_pop = false, have_pop = false;   // level populations are counted only once eref_query_popcount has asked
    DevBuf part_ctr;                           // u64: hist[NSLICES] | slice tails[NSLICES] | coarse tails[256] | coarse_n[256]; u32 chunk_start[257]
    DevBuf codes_coarse;                       // coarse-partitioned codes of the file being added
    DevBuf level_pop;                          // 3 x u64: codes at level >= 1, 2, 3 after the last build
    cudaEvent_t evf[2]{}, table_ready = nullptr;
    cudaStream_t table_stream = nullptr;
    bool table_inflight = false, count_time_pending = false;
    cudaEvent_t aux_ev[2]{};
    uint32_t* summary = nullptr;               // 1 bit per 16 level-3 slots
    bool summary_stale = true;
    // Sharded build (eref_table_set_part): this GPU counts the slices [slice0, slice0 + nslices) and stores them into
    // every peer's table; share_words = 16 flag words (one per part) | done counter | error word
    int parts = 1, part = 0;
    double part_w = 0.0;                       // balance of the parts: weight of code mass against width
    bool build_sharded = true;                 // this query's build is shared out (false: small query, every GPU counts all)
    uint32_t slice0 = 0, nslices = NSLICES;
    struct Peer { uint32_t* bm3 = nullptr; uint32_t* summary = nullptr; uint32_t* flags = nullptr; bool ipc = false; };
    Peer peers[MAX_PARTS];
    uint32_t* share_words = nullptr;
    uint32_t table_epoch = 0;
    bool join_pending = false;                 // a sharded build was queued and the main stream has not joined it yet
    eref_table_hook hook = nullptr;
    void* hook_arg = nullptr;
    // Shared reads (eref_query_set_share): this GPU partitions only ITS share of every query file, into lists that live
    // in one exported arena; a build pulls the codes of this GPU's table slices out of every peer's lists.
    // arena layout per list: off[NSLICES + 1] u64 | cnt[NSLICES] u64 | codes[arena_cap_codes] u32, list_stride bytes apart
    DevBuf list_arena, pulled;
    size_t list_stride = 0;
    unsigned long long arena_cap_codes = 0;
    int arena_lists = 0;
    bool share_reads = false;
    struct ListPeer { uint8_t* arena = nullptr; bool ipc = false; };
    ListPeer list_peers[MAX_PARTS];
    uint32_t lists_epoch = 0;
    int probe_parts = 8;                       // EREF_PROBE_PARTS: rounds of the listed-word probe, 1 = whole table at once
    int bucket_bits = 32 - SLICE_BITS - 8;     // EREF_BUCKET_BITS: 2^bits coarse buckets, 256 slices per bucket by default
    // reference side
    DevBuf r_text, tile_sum, tile_pre, rec_hdr_off, rec_first, rec_len, rec_kept, rec_words,
        rec_kept_ex, rec_words_ex, g_rec, g_len, g_word, g_first, r_hi, r_lo, r_ok, word_genome,
        sample_hits, need, need_list, genome_listed, hits0, hits1, hits2, single, trio, good, results, codes_tmp;
    const uint8_t* ref_text_dev = nullptr;    // where the current reference text lives
    uint64_t ref_text_n = 0;
    bool have_ref = false;
    uint32_t n_recs = 0, n_genomes = 0;
    uint64_t ref_words = 0, ref_bases = 0;
    std::vector<uint32_t> h_g_len, h_g_word, h_g_rec, h_rec_hdr_off;
    std::vector<unsigned long long> h_cell_pre;
    DevBuf cell_pre;
    bool cell_pre_valid = false;
    eref_phase_times t{};
    uint64_t launches = 0;
    // Phase times are collected lazily (collect_times): the calls below do not wait for their own kernels any more
    // than they must, so the events of a call are read only after a later synchronisation covers them.
    std::vector<cudaEvent_t> qev;              // 4 events per query file since the last reset: start, packed, partitioned, end
    bool q_pending = false, ref_pending = false, scan_pending = false, scan_pending_d2h = false;
    cudaEvent_t totals_ev = nullptr;           // (unused since the framing census has a stream of its own)
    // framing census of the reference text: own stream, ordered behind the text and behind the previous scan
    cudaStream_t frame_stream = nullptr;
    cudaEvent_t ev_scan_done = nullptr, ev_frame_done = nullptr, ev_text_free = nullptr;
    bool scan_done_valid = false, text_free_valid = false;
    bool serial = false;                       // eref_set_serial: no overlap between the streams (per-kernel timing)
    bool text_is_resident = false;             // EREF_FLAG_RESIDENT_TEXT: device texts are complete when the call is made
    DevBuf frame_cub_tmp;
};

namespace {

#define LAUNCH(ctx, expr)                                                                  \
    do {                                                                                   \
        (ctx)->launches++;                                                                 \
        CUDA_TRY(expr);                                                                    \
    } while (0)

int start_table_update(eref_ctx* c);
int check_guard(eref_ctx* c);
void rebalance_parts(eref_ctx* c, uint64_t first_file_bytes);
int wait_table(eref_ctx* c, cudaStream_t st);
int join_table(eref_ctx* c, cudaStream_t st);
// layout of part_ctr after the u64 arrays: u32 chunk_start[260], then u64 totals[3]
inline unsigned long long* chunk_start_end(uint32_t* chunk_start) { return reinterpret_cast<unsigned long long*>(chunk_start + 260); }

// (a scan owns its temporary storage while it runs: scans on different streams need different buffers)
template <class T>
int exclusive_scan(eref_ctx* c, const T* in, T* out, size_t n, cudaStream_t st = nullptr, DevBuf* tmp = nullptr)
{
    if (!st) st = c->stream;
    if (!tmp) tmp = &c->cub_tmp;
    size_t bytes = 0;
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, st));
    CUDA_TRY(tmp->ensure(bytes));
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp->p, bytes, in, out, (int)n, st));
    return 0;
}

float ev_ms(eref_ctx* c, int a, int b)
{
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[a], c->ev[b]);
    return ms;
}

cudaEvent_t* list_events(eref_ctx* c, size_t k)
{
    while (c->qev.size() < 4 * (k + 1)) {
        cudaEvent_t e = nullptr;
        if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
        c->qev.push_back(e);
    }
    return c->qev.data() + 4 * k;
}

// share_words: [0, 16) table-part flags | 16 done counter | 17 error word | 18 capacity guard | [32, 48) lists-ready flags
constexpr int SHARE_WORDS = 64, LISTS_FLAG0 = 32;
constexpr int HPIN_SEQ = 32;                   // h_pin[HPIN_SEQ + k]: sequence-line bytes of query file k
constexpr int HPIN_CODES = 48;                 // h_pin[HPIN_CODES + k]: codes of query file k
constexpr int HPIN_GUARD = 14;                 // h_pin[HPIN_GUARD]: the capacity guard word, read back with the scan

void table_done(eref_ctx* c);

// Read the events of the calls made since the last collection.  Only valid after a host synchronisation that
// covers the main stream (and through it the table stream).
void collect_times(eref_ctx* c)
{
    auto ms = [](cudaEvent_t a, cudaEvent_t b) { float v = 0; if (cudaEventElapsedTime(&v, a, b) != cudaSuccess) { cudaGetLastError(); v = 0; } return v; };
    if (c->q_pending) {
        float pack = 0, bucket = 0;
        unsigned long long seq = 0, codes = 0;
        for (size_t k = 0; k < c->n_lists; k++) {
            cudaEvent_t* e = c->qev.data() + 4 * k;
            pack += ms(e[0], e[1]); bucket += ms(e[1], e[2]);
            if (c->lists[k].total_pending) { c->lists[k].total = c->h_pin[HPIN_CODES + k]; c->lists[k].total_pending = false; }
            seq += c->h_pin[HPIN_SEQ + k]; codes += c->lists[k].total;
        }
        // sums over the files added since the last reset
        c->t.query_pack_ms = pack; c->t.query_bucket_ms = bucket;
        c->t.query_seq_bytes = seq; c->t.query_codes = codes;
        c->q_pending = false;
    }
    if (c->ref_pending) {
        c->t.ref_h2d_ms = ms(c->ev[E0], c->ev[R1]);
        c->t.ref_frame_ms = ms(c->ev[R1], c->ev[R2]);     // on the census stream, beside the query work
        c->t.ref_pack_ms = ms(c->ev[R4], c->ev[R3]);
        c->ref_pending = false;
    }
    if (c->scan_pending) {
        if (c->h_pin[HPIN_GUARD] & 0xFFFFFFFFull) c->retry_pending = true;
        table_done(c);
        c->t.scan_summary_ms = ms(c->ev[S0], c->ev[S1]);
        c->t.scan_sample_ms = ms(c->ev[S1], c->ev[S2]);
        c->t.scan_cand_ms = ms(c->ev[S2], c->ev[S3]);
        c->t.scan_probe_ms = ms(c->ev[S3], c->ev[S4]);
        c->t.scan_window_ms = ms(c->ev[S4], c->ev[S5]);
        c->t.scan_interval_ms = ms(c->ev[S5], c->ev[S6]);
        c->t.scan_d2h_ms = c->scan_pending_d2h ? ms(c->ev[S6], c->ev[S7]) : 0.f;
        const bool pruned = !(c->flags & EREF_FLAG_NO_PRUNE) && c->three_min - (WINDOW - WINDOW / c->sample_stride) > 0;
        c->t.ref_need_words = pruned ? c->h_pin[6] : c->ref_words;
        c->scan_pending = false;
    }
}

// after collect_times: did a query file outgrow buffers that were sized before its codes were counted?
int check_guard(eref_ctx* c)
{
    if (!c->retry_pending) return 0;
    c->retry_pending = false;
    cudaMemsetAsync(c->guard, 0, 4, c->stream);
    c->h_pin[HPIN_GUARD] = 0;
    c->force_exact = true;
    return eref_fail(EREF_ERR_RETRY, "a query file holds more sequence than quality bytes: its k-mers did not fit the "
                                     "buffers sized from the text length; eref_query_reset and add the files again "
                                     "(exact sizes are used from now on)");
}

// newline census of a text: tile_prefix[t] = newlines before tile t, tile_prefix[ntiles] = total
int newline_census(eref_ctx* c, const uint8_t* text, uint64_t n, uint64_t* ntiles_out)
{
    uint64_t ntiles = (n + TILE_BYTES - 1) / TILE_BYTES;
    if (ntiles >= (1ull << 31)) return eref_fail(EREF_ERR_LIMIT, "text of %llu bytes has too many tiles", (unsigned long long)n);
    CUDA_TRY(c->tile_counts.ensure((ntiles + 1) * 4));
    CUDA_TRY(c->tile_prefix.ensure((ntiles + 1) * 4));
    CUDA_TRY(cudaMemsetAsync(c->tile_counts.as<uint32_t>() + ntiles, 0, 4, c->stream));
    LAUNCH(c, launch_count_newlines(text, n, c->tile_counts.as<uint32_t>(), ntiles, c->stream));
    int rc = exclusive_scan<uint32_t>(c, c->tile_counts.as<uint32_t>(), c->tile_prefix.as<uint32_t>(), ntiles + 1);
    if (rc) return rc;
    *ntiles_out = ntiles;
    return 0;
}

int check_text_ptr(const void* p, size_t n)
{
    if (n && !p) return eref_fail(EREF_ERR_ARG, "null text pointer with n=%zu", n);
    if (((uintptr_t)p & 15) != 0) return eref_fail(EREF_ERR_ARG, "device text must be 16-byte aligned");
    return 0;
}

// cal_sam_ratio's pass over one FASTQ text (:1124-1148): bytes on its sequence lines and how many such lines it has
int census_impl(eref_ctx* c, const uint8_t* text, uint64_t n, uint64_t* seq_bytes_out, uint64_t* seq_lines_out)
{
    cudaStream_t st = c->stream;
    uint64_t ntiles = 0;
    int rc = newline_census(c, text, n, &ntiles);
    if (rc) return rc;
    CUDA_TRY(c->tile_kept.ensure((ntiles + 1) * 8));
    unsigned long long* seq_bytes = c->counters.as<unsigned long long>();
    CUDA_TRY(cudaMemsetAsync(seq_bytes, 0, 8, st));
    LAUNCH(c, launch_fastq_kept(text, n, c->tile_prefix.as<uint32_t>(), nullptr, c->tile_kept.as<unsigned long long>(), seq_bytes,
                                ntiles, st));
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 5, seq_bytes, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 12, c->tile_prefix.as<uint32_t>() + ntiles, 4, cudaMemcpyDeviceToHost, st));
    uint8_t last = '\n';
    if (n) CUDA_TRY(cudaMemcpyAsync(c->h_pin + 13, text + n - 1, 1, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (n) last = (uint8_t)(c->h_pin[13] & 0xFF);
    // getline yields one line per newline plus an unterminated last one; lines 1, 5, 9, ... are sequence lines (:947)
    const uint64_t lines = (c->h_pin[12] & 0xFFFFFFFFull) + (n && last != '\n' ? 1 : 0);
    if (seq_bytes_out) *seq_bytes_out = c->h_pin[5];
    if (seq_lines_out) *seq_lines_out = (lines + 2) / 4;
    return 0;
}

int query_add_impl(eref_ctx* c, const uint8_t* text, uint64_t n, const uint32_t* keep_host = nullptr, uint64_t n_keep = 0)
{
    cudaStream_t st = c->stream;
    int wrc = wait_table(c, st);              // an update still in flight reads the code lists this call rewrites
    if (wrc) return wrc;
    const bool direct = (c->flags & EREF_FLAG_DIRECT_COUNT) != 0;
    // (shared reads: the shares differ in size from GPU to GPU, the cut points must not — take the reserved file size)
    if (c->parts > 1 && c->n_lists == 0)
        rebalance_parts(c, c->share_reads ? (uint64_t)((c->arena_cap_codes - 4ull * NSLICES) / 3 * 2) * (uint64_t)c->parts : n);
    if (!direct && c->n_lists >= (size_t)MAX_LISTS)
        return eref_fail(EREF_ERR_LIMIT, "more than %d query files since the last eref_query_reset", MAX_LISTS);
    // events of this file: E1..E4 for the direct build (read at the end of this call), a set of its own otherwise
    cudaEvent_t* qe = direct ? &c->ev[E1] : list_events(c, c->n_lists);
    if (!qe) return eref_fail(EREF_ERR_CUDA, "cudaEventCreate failed");
    CUDA_TRY(cudaEventRecord(qe[0], st));
    uint64_t ntiles = 0;
    int rc = newline_census(c, text, n, &ntiles);
    if (rc) return rc;
    // plane positions per tile (sequence-line bytes + their newlines) and their prefix: where each tile packs to
    CUDA_TRY(c->tile_kept.ensure((ntiles + 1) * 8)); CUDA_TRY(c->tile_kept_pre.ensure((ntiles + 1) * 8));
    unsigned long long* tile_kept = c->tile_kept.as<unsigned long long>();
    unsigned long long* tile_kept_pre = c->tile_kept_pre.as<unsigned long long>();
    CUDA_TRY(cudaMemsetAsync(tile_kept + ntiles, 0, 8, st));
    const uint32_t* keep = nullptr;                     // down-sampling: one bit per sequence line (:955-960)
    if (keep_host) {
        const size_t words = (size_t)((n / 4 + 2 + 31) / 32) + 1;       // the kernels index the mask by line number: cover
                                                                        // every line a text of n bytes can have, zero = drop
        const size_t have = (size_t)((n_keep + 31) / 32);
        CUDA_TRY(c->q_keep.ensure(words * 4));
        CUDA_TRY(cudaMemsetAsync(c->q_keep.p, 0, words * 4, st));
        if (have) CUDA_TRY(cudaMemcpyAsync(c->q_keep.p, keep_host, (have < words ? have : words) * 4, cudaMemcpyHostToDevice, st));
        keep = c->q_keep.as<uint32_t>();
    }
    unsigned long long* seq_bytes = c->counters.as<unsigned long long>();
    CUDA_TRY(cudaMemsetAsync(seq_bytes, 0, 8, st));
    LAUNCH(c, launch_fastq_kept(text, n, c->tile_prefix.as<uint32_t>(), keep, tile_kept, seq_bytes, ntiles, st));
    rc = exclusive_scan<unsigned long long>(c, tile_kept, tile_kept_pre, ntiles + 1);
    if (rc) return rc;
    const unsigned long long* plane_bits = tile_kept_pre + ntiles;     // total, known to the device only for now
    Planes pl;
    const uint64_t words_ub = (n + 31) / 32 + 2;
    CUDA_TRY(c->q_hi.ensure(words_ub * 4)); CUDA_TRY(c->q_lo.ensure(words_ub * 4)); CUDA_TRY(c->q_ok.ensure(words_ub * 4));
    pl.hi = c->q_hi.as<uint32_t>(); pl.lo = c->q_lo.as<uint32_t>(); pl.ok = c->q_ok.as<uint32_t>();
    pl.nwords = 0; pl.cap = words_ub;
    CUDA_TRY(cudaMemsetAsync(pl.hi, 0, words_ub * 4, st));           // tiles OR their boundary words in; pad words stay 0
    CUDA_TRY(cudaMemsetAsync(pl.lo, 0, words_ub * 4, st));
    CUDA_TRY(cudaMemsetAsync(pl.ok, 0, words_ub * 4, st));
    LAUNCH(c, launch_pack_fastq(text, n, c->tile_prefix.as<uint32_t>(), keep, tile_kept_pre, pl, st));
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 11, plane_bits, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(qe[1], st));
    c->summary_stale = true;
    unsigned long long total_codes = 0;
    if (direct) {
        CUDA_TRY(cudaStreamSynchronize(st));
        pl.nwords = (c->h_pin[11] + 31) >> 5;
        CUDA_TRY(cudaEventRecord(qe[2], st));
        LAUNCH(c, launch_count_kmers(pl, c->masks, c->bm[0], c->bm[1], c->bm[2], st));
    } else {
        // codes -> slice-ordered list: histogram over the 8192 table slices, exact offsets, coarse scatter from the
        // planes, fine partition through shared memory; counted by the next table build
        const int cbits = c->bucket_bits;
        // shared reads: every GPU keeps ALL slices of its share of the reads; else only the slices of its table part
        const uint32_t f_s0 = c->share_reads ? 0u : c->slice0, f_ns = c->share_reads ? (uint32_t)NSLICES : c->nslices;
        CUDA_TRY(c->part_ctr.ensure(((size_t)NSLICES * 2 + 512 + 4) * 8 + 260 * 4));
        unsigned long long* hist = c->part_ctr.as<unsigned long long>();
        unsigned long long* slice_tail = hist + NSLICES;
        unsigned long long* coarse_tail = slice_tail + NSLICES;
        unsigned long long* coarse_n = coarse_tail + 256;
        uint32_t* chunk_start = reinterpret_cast<uint32_t*>(coarse_n + 256);
        CUDA_TRY(cudaMemsetAsync(hist, 0, ((size_t)NSLICES * 2 + 256) * 8, st));
        LAUNCH(c, launch_slice_hist(pl, plane_bits, c->masks, hist, f_s0, f_ns, st));
        unsigned long long* totals = chunk_start_end(chunk_start);     // {codes, padded size, chunks}
        if (c->lists.size() <= c->n_lists) c->lists.resize(c->n_lists + 1);
        eref_ctx::CodeList& L = c->lists[c->n_lists];
        if (c->share_reads) {
            if ((int)c->n_lists >= c->arena_lists)
                return eref_fail(EREF_ERR_LIMIT, "shared reads: eref_query_reserve made room for %d files per build", c->arena_lists);
            uint8_t* base = c->list_arena.as<uint8_t>() + c->n_lists * c->list_stride;     // the exported arena: fixed addresses
            L.off.attach(base, ((size_t)NSLICES + 1) * 8);
            L.cnt.attach(base + ((size_t)NSLICES + 1) * 8, (size_t)NSLICES * 8);
            L.codes.attach(base + ((size_t)2 * NSLICES + 2) * 8, (size_t)c->arena_cap_codes * 4);
        } else if (L.codes.view) { L.codes.release(); L.off.release(); L.cnt.release(); }
        CUDA_TRY(L.off.ensure(((size_t)NSLICES + 1) * 8)); CUDA_TRY(L.cnt.ensure((size_t)NSLICES * 8));

        // Sizes.  The number of codes is known to the device only.  A FASTQ record spends as many bytes on its quality
        // line as on its sequence line, so a file of n bytes holds fewer than n/2 sequence bytes and 3n/2 codes: when the
        // buffers already hold that much (they only grow), everything is queued without asking — k_slice_offsets
        // checks the real size against the capacity, and a file that breaks the rule (quality lines shorter than
        // sequence lines) is reported at the next synchronisation and re-added with exact sizes.  Otherwise the host
        // waits for the count once and sizes the buffers for the estimate, so the next file of this size goes through.
        // (a sharded build keeps the codes of its slice range only; canonical codes thin out linearly towards the top
        // of the code space: the range [x0, x1) of it holds (1 - x0)^2 - (1 - x1)^2 of all codes)
        const double x0 = (double)c->slice0 / NSLICES, x1 = (double)(c->slice0 + c->nslices) / NSLICES;
        const double part_frac = c->parts > 1 && !c->share_reads ? std::min(1.0, 1.15 * ((1 - x0) * (1 - x0) - (1 - x1) * (1 - x1))) : 1.0;
        const unsigned long long est_codes = (unsigned long long)(3.0 * (double)(n / 2 + 1) * part_frac) + 4ull * NSLICES + 64;
        const unsigned long long cap_codes = std::min(L.codes.cap, c->codes_coarse.cap) / 4 > 4
                                                 ? std::min(L.codes.cap, c->codes_coarse.cap) / 4 - 4 : 0;
        const bool optimistic = !c->force_exact && cap_codes >= est_codes;
        LAUNCH(c, launch_slice_offsets(hist, cbits, L.off.as<unsigned long long>(), L.cnt.as<unsigned long long>(), coarse_n,
                                       chunk_start, totals, optimistic ? cap_codes : 0ull, optimistic ? c->guard : nullptr, st));
        CUDA_TRY(cudaMemcpyAsync(c->h_pin + HPIN_CODES + c->n_lists, totals, 8, cudaMemcpyDeviceToHost, st));
        uint32_t nchunks;
        Planes pl_scatter = pl;
        if (optimistic) {
            pl_scatter.nwords = words_ub;                  // grid bound; the kernel reads the packed size itself
            nchunks = (uint32_t)std::min<unsigned long long>(cap_codes / FINE_CHUNK + (1u << cbits) + 1, 0x7FFFFFFFull);
            L.total = 0; L.total_pending = true;
        } else {
            CUDA_TRY(cudaMemcpyAsync(c->h_pin + 8, totals, 24, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
            total_codes = c->h_pin[8];
            const unsigned long long padded = c->h_pin[9];
            nchunks = (uint32_t)c->h_pin[10];
            pl.nwords = (c->h_pin[11] + 31) >> 5;
            pl_scatter = pl;
            L.total = total_codes; L.total_pending = false;
            const size_t want = (size_t)std::max<unsigned long long>(padded, c->force_exact || L.codes.view ? 0ull : est_codes) * 4 + 16;
            if (L.codes.view && want > L.codes.cap)
                return eref_fail(EREF_ERR_LIMIT, "shared reads: a file of %llu bytes needs %zu bytes of code list, eref_query_reserve "
                                                 "made room for %zu", (unsigned long long)n, want, L.codes.cap);
            CUDA_TRY(L.codes.ensure(want));
            CUDA_TRY(c->codes_coarse.ensure(want));
        }
        LAUNCH(c, launch_coarse_scatter(pl_scatter, optimistic ? plane_bits : nullptr, optimistic ? c->guard : nullptr, c->masks,
                                        cbits, L.off.as<unsigned long long>(), coarse_tail, c->codes_coarse.as<uint32_t>(),
                                        c->codes_coarse.cap / 4, f_s0, f_ns, st));
        LAUNCH(c, launch_fine_partition(c->codes_coarse.as<uint32_t>(), L.off.as<unsigned long long>(), coarse_n, chunk_start,
                                        nchunks, cbits, slice_tail, L.codes.as<uint32_t>(), L.codes.cap / 4, st));
        // the partition kernels are queued; nothing below waits for them (times and byte counts: collect_times)
        CUDA_TRY(cudaEventRecord(qe[2], st));
        CUDA_TRY(cudaMemcpyAsync(c->h_pin + HPIN_SEQ + c->n_lists, seq_bytes, 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaEventRecord(qe[3], st));
        c->n_lists++;
        c->table_dirty = true;
        c->q_pending = true;
        return 0;
    }
    CUDA_TRY(cudaEventRecord(qe[3], st));
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 5, seq_bytes, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    c->t.query_seq_bytes = c->h_pin[5];
    c->t.query_codes = total_codes;
    c->t.query_h2d_ms = ev_ms(c, E0, E1);
    c->t.query_pack_ms = ev_ms(c, E1, E2);
    c->t.query_bucket_ms = ev_ms(c, E2, E3);
    c->t.query_count_ms = ev_ms(c, E3, E4);
    return 0;
}

// sequence-line bytes of the file added last (cal_sam_ratio's sum): waits for that file's kernels
int last_file_seq_bytes(eref_ctx* c, unsigned long long* out)
{
    if ((c->flags & EREF_FLAG_DIRECT_COUNT) || !c->n_lists) { *out = c->t.query_seq_bytes; return 0; }
    CUDA_TRY(cudaEventSynchronize(c->qev[4 * (c->n_lists - 1) + 3]));
    *out = c->h_pin[HPIN_SEQ + c->n_lists - 1];
    return 0;
}


// Bring the table up to date, asynchronously: one CTA per 2^19-slot slice counts that slice's codes from every
// file added since the last reset in shared memory and writes the level-3 slice and its summary
// (k_count_slices).  The build is queued on the table stream behind everything already queued on the main
// stream, so the reference framing/packing kernels the caller queues next can run beside it; wait_table()
// makes a stream wait for the result.
int start_table_update(eref_ctx* c)
{
    if (c->flags & EREF_FLAG_DIRECT_COUNT) return 0;
    if (!c->table_dirty) return 0;
    cudaStream_t st = c->stream, ts = c->table_stream;
    CUDA_TRY(cudaEventRecord(c->aux_ev[0], st));
    CUDA_TRY(cudaStreamWaitEvent(ts, c->aux_ev[0], 0));
    CUDA_TRY(cudaEventRecord(c->evf[0], ts));
    CountLists cl{};
    cl.n = (int)c->n_lists;
    for (size_t k = 0; k < c->n_lists; k++) {
        cl.codes[k] = c->lists[k].codes.as<uint32_t>();
        cl.off[k] = c->lists[k].off.as<unsigned long long>();
        cl.cnt[k] = c->lists[k].cnt.as<unsigned long long>();
        cl.cap[k] = c->lists[k].codes.cap / 4;
    }
    if (c->share_reads && c->parts > 1) {
        // Shared reads: this GPU's lists hold all slices of ITS share of the reads.  Tell every peer they are complete,
        // wait for theirs, pull the codes of this GPU's table slices out of each peer list (one contiguous range per list)
        // into local staging, and count own lists + staged copies.
        const int files = (int)c->n_lists;
        if (files * c->parts > MAX_LISTS)
            return eref_fail(EREF_ERR_LIMIT, "shared reads: %d files x %d GPUs exceed %d lists per build", files, c->parts, MAX_LISTS);
        TableShare sig{};
        sig.part = c->part;
        sig.flags[0] = c->share_words; sig.n = 1;
        for (int p = 0; p < c->parts; p++)
            if (p != c->part) {
                if (!c->peers[p].flags || !c->list_peers[p].arena)
                    return eref_fail(EREF_ERR_STATE, "shared reads: peer %d is not wired (eref_table_set_peer, eref_lists_set_peer)", p);
                sig.flags[sig.n++] = c->peers[p].flags;
            }
        c->lists_epoch++;
        LAUNCH(c, launch_lists_signal(sig, LISTS_FLAG0, c->lists_epoch, st));      // behind this GPU's partition kernels
        CUDA_TRY(cudaEventRecord(c->aux_ev[0], st));
        CUDA_TRY(cudaStreamWaitEvent(ts, c->aux_ev[0], 0));
        LAUNCH(c, launch_table_wait(c->share_words + LISTS_FLAG0, c->parts, c->lists_epoch, c->share_words + MAX_PARTS + 1, ts));
        PullLists pull{};
        CUDA_TRY(c->pulled.ensure((size_t)(c->parts - 1) * files * c->list_stride));
        pull.dst = c->pulled.as<uint8_t>(); pull.list_stride = c->list_stride; pull.cap_codes = c->arena_cap_codes;
        pull.s0 = c->slice0; pull.ns = c->nslices;
        for (int p = 0; p < c->parts; p++)
            if (p != c->part)
                for (int f = 0; f < files; f++) {
                    pull.src[pull.n] = c->list_peers[p].arena + (size_t)f * c->list_stride;
                    uint8_t* d = pull.dst + (size_t)pull.n * c->list_stride;
                    cl.off[cl.n] = reinterpret_cast<const unsigned long long*>(d);
                    cl.cnt[cl.n] = cl.off[cl.n] + NSLICES + 1;
                    cl.codes[cl.n] = reinterpret_cast<const uint32_t*>(d + ((size_t)2 * NSLICES + 2) * 8);
                    cl.cap[cl.n] = c->arena_cap_codes;
                    cl.n++; pull.n++;
                }
        LAUNCH(c, launch_pull_lists(pull, ts));
    }
    if (c->want_pop) CUDA_TRY(cudaMemsetAsync(c->level_pop.p, 0, 24, ts));
    // destinations of the finished slices: this GPU's table, and in a sharded build with peers every other GPU's
    TableShare sh{};
    sh.bm3[0] = c->bm[2]; sh.summary[0] = c->summary; sh.flags[0] = c->share_words;
    sh.n = 1; sh.part = c->part; sh.s0 = c->slice0; sh.ns = c->nslices;
    sh.done = c->share_words + MAX_PARTS;
    const bool sharded = c->parts > 1 && c->build_sharded;
    bool all_peers = sharded;
    for (int p = 0; p < c->parts && all_peers; p++) if (p != c->part && !c->peers[p].bm3) all_peers = false;
    if (sharded) {
        if (!all_peers && !c->hook)
            return eref_fail(EREF_ERR_STATE, "sharded table build: neither every peer (eref_table_set_peer) nor a completion hook "
                                             "(eref_table_set_hook) is set");
        c->table_epoch++;
        sh.epoch = c->table_epoch;
        if (all_peers)
            for (int p = 0; p < c->parts; p++)
                if (p != c->part) { sh.bm3[sh.n] = c->peers[p].bm3; sh.summary[sh.n] = c->peers[p].summary; sh.flags[sh.n] = c->peers[p].flags; sh.n++; }
        c->join_pending = true;
    }
    LAUNCH(c, launch_count_slices(cl, sh, c->want_pop ? c->level_pop.as<unsigned long long>() : nullptr, ts));
    c->have_pop = c->want_pop;
    CUDA_TRY(cudaEventRecord(c->evf[1], ts));
    CUDA_TRY(cudaEventRecord(c->table_ready, ts));
    c->table_dirty = false;
    c->summary_stale = false;
    c->table_inflight = true;
    c->count_time_pending = true;
    if (c->serial) CUDA_TRY(cudaStreamWaitEvent(st, c->table_ready, 0));   // nothing runs beside the count
    return 0;
}

// make `st` wait for the table update queued last (no-op when none is outstanding)
int wait_table(eref_ctx* c, cudaStream_t st)
{
    if (c->table_inflight) CUDA_TRY(cudaStreamWaitEvent(st, c->table_ready, 0));
    return 0;
}

// Make `st` wait for the table to be complete.  Local build: the update queued last.  Sharded build: also every other
// GPU's part — with peer pointers the builders stored their slices straight into this table and raised a flag each
// (k_table_wait acquires them); without, the caller's hook completes the table on `st` (e.g. an in-place all-gather).
int join_table(eref_ctx* c, cudaStream_t st)
{
    int rc = wait_table(c, st);
    if (rc || !c->join_pending) return rc;
    c->join_pending = false;
    bool all_peers = true;
    for (int p = 0; p < c->parts; p++) if (p != c->part && !c->peers[p].bm3) all_peers = false;
    if (all_peers) {
        LAUNCH(c, launch_table_wait(c->share_words, c->parts, c->table_epoch, c->share_words + MAX_PARTS + 1, st));
        return 0;
    }
    if (!c->hook) return eref_fail(EREF_ERR_STATE, "sharded table build without peers or a completion hook");
    rc = c->hook(c->hook_arg, (void*)st);
    if (rc) return eref_fail(EREF_ERR_STATE, "the table completion hook failed (%d)", rc);
    return 0;
}

// First table slice of part p of `parts`.  Canonical codes are min(fwd, rc): their density falls linearly, 2(1 - x), over
// the code space, so equal slice ranges would give the first GPU of two 75 % of the codes; but a part also costs per
// slice — clearing and writing out 2^19 slots and, above all, storing them into N - 1 peers over NVLink (egress is
// balanced only for equal widths).  The cut points equalise w * mass + (1 - w) * width.
uint32_t part_first_slice(int p, int parts, double w)
{
    if (p <= 0) return 0;
    if (p >= parts) return NSLICES;
    const double t = (double)p / parts;
    const double x = w > 1e-6 ? ((1 + w) - sqrt((1 + w) * (1 + w) - 4 * w * t)) / (2 * w) : t;
    long long sl = llround(x * NSLICES);
    if (sl < p) sl = p;
    if (sl > NSLICES - (parts - p)) sl = NSLICES - (parts - p);
    return (uint32_t)sl;
}

// The weight for a query of `query_bytes` of FASTQ text on `parts` GPUs, from the measured costs (B200):
// partitioning + counting ~0.5 ms per 100 MB of FASTQ text, spread by code mass; per unit of width 0.3 ms of local
// clearing/copy-out plus 0.6 ms for every peer the 544 MiB table goes to.  Small queries on many GPUs end up with
// nearly equal widths (the exchange dominates), large queries with nearly equal code mass.  EREF_PART_WEIGHT overrides.
double part_weight(int parts, uint64_t query_bytes)
{
    static const double forced = [] { const char* v = getenv("EREF_PART_WEIGHT"); return v ? atof(v) : -1.0; }();
    if (forced >= 0) return forced > 0.95 ? 0.95 : forced;
    const double a = 0.5 * (double)query_bytes / 1.04e8, b = 0.3 + 0.6 * (parts - 1);
    const double w = a / (a + b);
    return w > 0.95 ? 0.95 : w;
}

void set_part_range(eref_ctx* c)
{
    c->slice0 = part_first_slice(c->part, c->parts, c->part_w);
    c->nslices = part_first_slice(c->part + 1, c->parts, c->part_w) - c->slice0;
}

// Sharded build, first file after a reset: the parts are cut for the size of this query (every GPU sees the same file
// sizes in the same order, so every GPU cuts alike; a sample is two files of about this size)
// Sharing the build out costs an exchange of the 544 MiB table (measured on 8 B200s: 0.74 ms per build, NVLink-bound
// at ~640 GB/s of peer stores per GPU; 0.35 ms on 2) and saves (1 - 1/N) of partitioning and counting (~0.8 ms per
// 100 MB of FASTQ text): below EREF_SHARD_MIN_BYTES of query text (default 160 MB; 0 = always share) every GPU counts
// the whole table for itself, as the reference's threads share one table built once (:1257-1291).
void rebalance_parts(eref_ctx* c, uint64_t first_file_bytes)
{
    const char* mv = getenv("EREF_SHARD_MIN_BYTES");
    const double min_bytes = mv ? atof(mv) : 1.6e8;
    c->build_sharded = c->share_reads || 2.0 * (double)first_file_bytes >= min_bytes;   // (shared reads divide ALL query work)
    if (!c->build_sharded) { c->slice0 = 0; c->nslices = NSLICES; return; }
    c->part_w = part_weight(c->parts, 2 * first_file_bytes);
    set_part_range(c);
}

// after a host synchronisation that covers the table stream: collect the update's duration
void table_done(eref_ctx* c)
{
    c->table_inflight = false;
    if (c->count_time_pending) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, c->evf[0], c->evf[1]) == cudaSuccess) c->t.query_count_ms = ms;
        else cudaGetLastError();
        c->count_time_pending = false;
    }
}

int finalize_table(eref_ctx* c)
{
    int rc = start_table_update(c);
    if (rc) return rc;
    if (c->table_inflight) {
        CUDA_TRY(cudaStreamSynchronize(c->table_stream));
        table_done(c);
    }
    return 0;
}

constexpr uint64_t REF_TEXT_LIMIT = 0xFFFFFF00ull;     // text offsets and plane positions of one shard are 32-bit

// Framing and packing of one reference text.
//
// The framing census (per-tile summaries, their prefix, the record table) needs nothing from the query side, and the
// host must read its totals to size the planes and the record arrays.  It therefore runs on a stream of its own
// (frame_stream), ordered only behind (a) the text being there — `text_ready`, an event of whoever produced it, or
// nothing for a text that is already complete in device memory — and (b) the previous scan, which still reads the
// record tables this call overwrites.  While the host waits for the totals the main stream keeps working through
// whatever was queued before (in a screening step: both FASTQ files), so the two host round trips cost no idle GPU
// time; the main stream joins the census just before the packing kernel.
int ref_set_impl(eref_ctx* c, const uint8_t* text, uint64_t n, cudaEvent_t text_ready, const void* host_text = nullptr)
{
    cudaStream_t st = c->stream, fs = c->frame_stream;
    c->have_ref = false;
    if (n >= REF_TEXT_LIMIT)
        return eref_fail(EREF_ERR_LIMIT, "reference text of %llu bytes: one shard must stay below 4 GiB",
                         (unsigned long long)n);
    int trc = start_table_update(c);          // pending query files are counted while this text is framed and packed
    if (trc) return trc;
    if (c->scan_done_valid) CUDA_TRY(cudaStreamWaitEvent(fs, c->ev_scan_done, 0));
    if (text_ready) CUDA_TRY(cudaStreamWaitEvent(fs, text_ready, 0));
    if (c->serial) {                          // the census starts only when the main stream has drained
        CUDA_TRY(cudaEventRecord(c->aux_ev[1], st));
        CUDA_TRY(cudaStreamWaitEvent(fs, c->aux_ev[1], 0));
    }
    CUDA_TRY(cudaEventRecord(c->ev[E0], fs));
    if (host_text) {
        // eref_ref_set: the upload rides on the census stream too (behind the last reader of the text buffer)
        if (c->text_free_valid) CUDA_TRY(cudaStreamWaitEvent(fs, c->ev_text_free, 0));
        if (n) CUDA_TRY(cudaMemcpyAsync(const_cast<uint8_t*>(text), host_text, n, cudaMemcpyHostToDevice, fs));
    }
    CUDA_TRY(cudaEventRecord(c->ev[R1], fs));
    // 1. per-tile framing summaries and their prefix: line/header state, records and sequence bytes before each tile
    const uint64_t ntiles = (n + TILE_BYTES - 1) / TILE_BYTES;
    CUDA_TRY(c->tile_sum.ensure((ntiles + 1) * sizeof(TileSum)));
    CUDA_TRY(c->tile_pre.ensure((ntiles + 1) * sizeof(TileSum)));
    TileSum* tiles = c->tile_sum.as<TileSum>();
    TileSum* tile_pre = c->tile_pre.as<TileSum>();
    CUDA_TRY(cudaMemsetAsync(tiles + ntiles, 0, sizeof(TileSum), fs));          // sentinel: its prefix is the total
    LAUNCH(c, launch_fasta_tiles(text, n, tiles, ntiles, fs));
    {
        size_t bytes = 0;
        const TileSum zero{0, 0, 0, 0};
        CUDA_TRY(cub::DeviceScan::ExclusiveScan(nullptr, bytes, tiles, tile_pre, TileOp(), zero, (int)(ntiles + 1), fs));
        CUDA_TRY(c->frame_cub_tmp.ensure(bytes));
        CUDA_TRY(cub::DeviceScan::ExclusiveScan(c->frame_cub_tmp.p, bytes, tiles, tile_pre, TileOp(), zero, (int)(ntiles + 1), fs));
        c->launches += 2;
    }
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 16, tile_pre + ntiles, sizeof(TileSum), cudaMemcpyDeviceToHost, fs));
    CUDA_TRY(cudaStreamSynchronize(fs));
    TileSum total;
    memcpy(&total, c->h_pin + 16, sizeof total);
    const uint32_t n_recs = total.hdrs + 1;               // record 0 = whatever precedes the first header
    const uint64_t total_seq = total.seqA + total.seqB;
    // 2. records: first sequence byte, header offset, length, keep flag, packed size
    const size_t rn = (size_t)n_recs + 1;
    CUDA_TRY(c->rec_hdr_off.ensure(rn * 4)); CUDA_TRY(c->rec_first.ensure(rn * 4));
    CUDA_TRY(c->rec_len.ensure(rn * 4)); CUDA_TRY(c->rec_kept.ensure(rn * 4));
    CUDA_TRY(c->rec_words.ensure(rn * 4)); CUDA_TRY(c->rec_kept_ex.ensure(rn * 4));
    CUDA_TRY(c->rec_words_ex.ensure(rn * 4));
    CUDA_TRY(c->g_rec.ensure(rn * 4)); CUDA_TRY(c->g_len.ensure(rn * 4));
    CUDA_TRY(c->g_word.ensure(rn * 4)); CUDA_TRY(c->g_first.ensure(rn * 4));
    const size_t words_ub = (size_t)(total_seq / 32) + n_recs + 2;   // every kept record: (len>>5)+1 words
    CUDA_TRY(c->r_hi.ensure(words_ub * 4)); CUDA_TRY(c->r_lo.ensure(words_ub * 4));
    CUDA_TRY(c->r_ok.ensure(words_ub * 4)); CUDA_TRY(c->word_genome.ensure(words_ub * 4));
    Planes pl;
    pl.hi = c->r_hi.as<uint32_t>(); pl.lo = c->r_lo.as<uint32_t>(); pl.ok = c->r_ok.as<uint32_t>();
    pl.nwords = 0; pl.cap = words_ub;
    // the planes are zeroed here, beside the query work (tiles OR their border words into them; pad words stay 0)
    CUDA_TRY(cudaMemsetAsync(pl.hi, 0, words_ub * 4, fs));
    CUDA_TRY(cudaMemsetAsync(pl.lo, 0, words_ub * 4, fs));
    CUDA_TRY(cudaMemsetAsync(pl.ok, 0, words_ub * 4, fs));
    LAUNCH(c, launch_fasta_records(text, n, tiles, tile_pre, ntiles, c->rec_first.as<uint32_t>(),
                                   c->rec_hdr_off.as<uint32_t>(), fs));
    unsigned long long* totals = c->counters.as<unsigned long long>() + 3;      // bases | genomes | words
    CUDA_TRY(cudaMemsetAsync(totals, 0, 24, fs));
    LAUNCH(c, launch_record_info(c->rec_first.as<uint32_t>(), n_recs, total_seq, c->rec_len.as<uint32_t>(),
                                 c->rec_kept.as<uint32_t>(), c->rec_words.as<uint32_t>(), totals, fs));
    int rc = exclusive_scan<uint32_t>(c, c->rec_kept.as<uint32_t>(), c->rec_kept_ex.as<uint32_t>(), rn, fs, &c->frame_cub_tmp);
    if (rc) return rc;
    rc = exclusive_scan<uint32_t>(c, c->rec_words.as<uint32_t>(), c->rec_words_ex.as<uint32_t>(), rn, fs, &c->frame_cub_tmp);
    if (rc) return rc;
    LAUNCH(c, launch_genome_table(c->rec_len.as<uint32_t>(), c->rec_kept.as<uint32_t>(), c->rec_kept_ex.as<uint32_t>(),
                                  c->rec_words_ex.as<uint32_t>(), c->rec_first.as<uint32_t>(), n_recs,
                                  c->g_rec.as<uint32_t>(), c->g_len.as<uint32_t>(), c->g_word.as<uint32_t>(),
                                  c->g_first.as<uint32_t>(), c->word_genome.as<uint32_t>(), totals + 1, fs));
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 2, totals, 24, cudaMemcpyDeviceToHost, fs));
    CUDA_TRY(cudaEventRecord(c->ev[R2], fs));
    CUDA_TRY(cudaEventRecord(c->ev_frame_done, fs));
    // 3. text -> planes, on the main stream once the census is there
    CUDA_TRY(cudaStreamWaitEvent(st, c->ev_frame_done, 0));
    CUDA_TRY(cudaEventRecord(c->ev[R4], st));
    LAUNCH(c, launch_pack_fasta(text, n, tile_pre, ntiles, c->rec_kept.as<uint32_t>(), c->rec_kept_ex.as<uint32_t>(),
                                c->rec_words_ex.as<uint32_t>(), c->rec_first.as<uint32_t>(), pl,
                                c->word_genome.as<uint32_t>(), st));
    CUDA_TRY(cudaEventRecord(c->ev[R3], st));
    CUDA_TRY(cudaEventRecord(c->ev_text_free, st));
    c->text_free_valid = true;
    CUDA_TRY(cudaStreamSynchronize(fs));      // the sizes are here; the main stream is still busy
    const uint64_t bases = c->h_pin[2];
    const uint32_t n_genomes = (uint32_t)c->h_pin[3], n_words = (uint32_t)c->h_pin[4];
    c->host_tables_valid = false;             // ordinal -> length / packed offset / header offset: fetched on demand
    c->cell_pre_valid = false;
    c->ref_text_dev = text; c->ref_text_n = n;
    c->n_recs = n_recs; c->n_genomes = n_genomes;
    c->ref_words = n_words; c->ref_bases = bases;
    c->have_ref = true;
    c->ref_pending = true;
    c->t.ref_bases = bases; c->t.ref_words = n_words;
    return 0;
}

int scan_impl(eref_ctx* c)
{
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "eref_scan before eref_ref_set");
    int frc = start_table_update(c);
    if (frc) return frc;
    cudaStream_t st = c->stream;
    frc = join_table(c, st);
    if (frc) return frc;
    const uint64_t nw = c->ref_words;
    const uint32_t G = c->n_genomes;
    Planes pl;
    pl.hi = c->r_hi.as<uint32_t>(); pl.lo = c->r_lo.as<uint32_t>(); pl.ok = c->r_ok.as<uint32_t>();
    pl.nwords = nw; pl.cap = c->r_hi.cap / 4;
    CUDA_TRY(c->single.ensure((nw + 64) * 4)); CUDA_TRY(c->trio.ensure((nw + 64) * 4));
    CUDA_TRY(c->good.ensure((nw + 64) * 4));
    CUDA_TRY(c->results.ensure(((size_t)G + 1) * 16));
    const int min_sample_hits = c->three_min - (WINDOW - WINDOW / c->sample_stride);
    const bool prune = !(c->flags & EREF_FLAG_NO_PRUNE) && min_sample_hits > 0;
    unsigned long long* need_count = c->counters.as<unsigned long long>() + 1;
    CUDA_TRY(cudaEventRecord(c->ev[S0], st));
    if (prune && c->summary_stale) {
        LAUNCH(c, launch_build_summary(c->bm[2], c->summary, SUMMARY_WORDS, st));
        c->summary_stale = false;
    }
    CUDA_TRY(cudaEventRecord(c->ev[S1], st));
    if (prune) {
        // the sampled probe, then the exact candidate filter: k_need lists the words whose windows may still be good
        // and zeroes single/trio/good everywhere; the full probe and the window rule run on the listed words only
        CUDA_TRY(c->sample_hits.ensure(nw + 64)); CUDA_TRY(c->need.ensure(nw + 64));
        CUDA_TRY(c->need_list.ensure((nw + 64) * 4));
        CUDA_TRY(cudaMemsetAsync(need_count, 0, 8, st));
        CUDA_TRY(c->genome_listed.ensure(((size_t)G + 1) * 4));
        CUDA_TRY(cudaMemsetAsync(c->genome_listed.p, 0, ((size_t)G + 1) * 4, st));
        LAUNCH(c, launch_probe_sampled(pl, c->masks, c->bm[2], c->summary, c->sample_stride, c->sample_hits.as<uint8_t>(), st));
        CUDA_TRY(cudaEventRecord(c->ev[S2], st));
        LAUNCH(c, launch_need(c->sample_hits.as<uint8_t>(), c->word_genome.as<uint32_t>(), nw, min_sample_hits,
                              c->need.as<uint8_t>(), c->need_list.as<uint32_t>(), need_count, c->single.as<uint32_t>(),
                              c->trio.as<uint32_t>(), c->good.as<uint32_t>(), c->genome_listed.as<uint32_t>(), st));
        CUDA_TRY(cudaEventRecord(c->ev[S3], st));
        if (c->probe_parts > 1) {
            // the table is walked one L2-sized part at a time (scan.cu, P3')
            CUDA_TRY(c->hits0.ensure((nw + 64) * 4)); CUDA_TRY(c->hits1.ensure((nw + 64) * 4));
            CUDA_TRY(c->hits2.ensure((nw + 64) * 4));
            int extra = 0;
            cudaError_t pe = launch_probe_parts(pl, c->masks, c->bm[2], c->need_list.as<uint32_t>(), need_count,
                                                c->probe_parts, c->hits0.as<uint32_t>(), c->hits1.as<uint32_t>(),
                                                c->hits2.as<uint32_t>(), c->single.as<uint32_t>(), c->trio.as<uint32_t>(),
                                                st, &extra);
            if (pe != cudaSuccess) return eref_fail(EREF_ERR_CUDA, "probe parts: %s", cudaGetErrorString(pe));
            c->launches += extra;
        } else {
            LAUNCH(c, launch_probe_list(pl, c->masks, c->bm[2], c->need_list.as<uint32_t>(), need_count,
                                        c->single.as<uint32_t>(), c->trio.as<uint32_t>(), st));
        }
        CUDA_TRY(cudaEventRecord(c->ev[S4], st));
        LAUNCH(c, launch_window_list(c->single.as<uint32_t>(), c->trio.as<uint32_t>(), c->need_list.as<uint32_t>(), need_count,
                                     c->word_genome.as<uint32_t>(), c->g_word.as<uint32_t>(), c->g_len.as<uint32_t>(), nw,
                                     c->one_min, c->three_min, c->good.as<uint32_t>(), st));
    } else {
        CUDA_TRY(cudaEventRecord(c->ev[S2], st));
        CUDA_TRY(cudaEventRecord(c->ev[S3], st));
        if (c->probe_parts > 1) {
            // every word, the table walked one L2-sized part at a time like the listed scan
            CUDA_TRY(c->hits0.ensure((nw + 64) * 4)); CUDA_TRY(c->hits1.ensure((nw + 64) * 4));
            CUDA_TRY(c->hits2.ensure((nw + 64) * 4));
            int extra = 0;
            cudaError_t pe = launch_probe_parts(pl, c->masks, c->bm[2], nullptr, nullptr, c->probe_parts, c->hits0.as<uint32_t>(),
                                                c->hits1.as<uint32_t>(), c->hits2.as<uint32_t>(), c->single.as<uint32_t>(),
                                                c->trio.as<uint32_t>(), st, &extra);
            if (pe != cudaSuccess) return eref_fail(EREF_ERR_CUDA, "probe parts: %s", cudaGetErrorString(pe));
            c->launches += extra;
        } else
        LAUNCH(c, launch_probe_full(pl, c->masks, c->bm[2], nullptr, c->single.as<uint32_t>(), c->trio.as<uint32_t>(), st));
        CUDA_TRY(cudaEventRecord(c->ev[S4], st));
        LAUNCH(c, launch_window(c->single.as<uint32_t>(), c->trio.as<uint32_t>(), c->word_genome.as<uint32_t>(),
                                c->g_word.as<uint32_t>(), c->g_len.as<uint32_t>(), nw, c->one_min, c->three_min,
                                c->good.as<uint32_t>(), st));
    }
    CUDA_TRY(cudaEventRecord(c->ev[S5], st));
    LAUNCH(c, launch_intervals(c->good.as<uint32_t>(), c->g_word.as<uint32_t>(), c->g_len.as<uint32_t>(), G,
                               prune ? c->genome_listed.as<uint32_t>() : nullptr, c->results.as<uint32_t>(), st));
    CUDA_TRY(cudaEventRecord(c->ev[S6], st));
    CUDA_TRY(cudaEventRecord(c->ev_scan_done, st));
    c->scan_done_valid = true;
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + HPIN_GUARD, c->guard, 4, cudaMemcpyDeviceToHost, st));
    return 0;
}

int fetch_host_tables(eref_ctx* c)
{
    if (c->host_tables_valid) return 0;
    const size_t gn = (size_t)c->n_genomes + 1, rn = (size_t)c->n_recs + 1;
    c->h_g_len.assign(gn, 0); c->h_g_word.assign(gn, 0); c->h_g_rec.assign(gn, 0); c->h_rec_hdr_off.assign(rn, 0);
    if (c->n_genomes) {
        CUDA_TRY(cudaMemcpy(c->h_g_len.data(), c->g_len.p, (size_t)c->n_genomes * 4, cudaMemcpyDeviceToHost));
        CUDA_TRY(cudaMemcpy(c->h_g_word.data(), c->g_word.p, gn * 4, cudaMemcpyDeviceToHost));
        CUDA_TRY(cudaMemcpy(c->h_g_rec.data(), c->g_rec.p, (size_t)c->n_genomes * 4, cudaMemcpyDeviceToHost));
    }
    CUDA_TRY(cudaMemcpy(c->h_rec_hdr_off.data(), c->rec_hdr_off.p, (size_t)c->n_recs * 4, cudaMemcpyDeviceToHost));
    c->host_tables_valid = true;
    return 0;
}

} // namespace

extern "C" {

const char* eref_last_error(void) { return g_err; }

int eref_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int eref_create(eref_ctx** out, int device, const int16_t coder[96], float hit_ratio, float perfect_hit_ratio,
                unsigned flags)
{
    if (!out || !coder) return eref_fail(EREF_ERR_ARG, "eref_create: null argument");
    *out = nullptr;
    CoderMasks m{};
    for (int z = 0; z < EREF_K; z++)
        for (int i = 0; i < 3; i++) {
            int a = coder[z * 3 + i];
            if (a < 0 || a > 2)
                return eref_fail(EREF_ERR_ARG, "coder[%d] = %d is not one of the three 1-bit alphabets", z * 3 + i, a);
            if (a == 0) m.sel0[i] |= 1u << (EREF_K - 1 - z);
            if (a == 2) m.sel2[i] |= 1u << (EREF_K - 1 - z);
        }
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return eref_fail(EREF_ERR_ARG, "device %d of %d", device, ndev);
    CUDA_TRY(cudaSetDevice(device));
    eref_ctx* c = new (std::nothrow) eref_ctx();
    if (!c) return eref_fail(EREF_ERR_ARG, "out of host memory");
    c->device = device; c->masks = m; c->flags = flags;
    c->hit_ratio = hit_ratio; c->perfect_hit_ratio = perfect_hit_ratio;
    c->one_min = (int)(WINDOW * hit_ratio);              // :513 — int * float, truncated
    c->three_min = (int)(WINDOW * perfect_hit_ratio);    // :514
    // sampling stride of the candidate filter: any 500 consecutive positions hold >= floor(500/s) sampled ones, so a good
    // window (>= three_min positions hit by all three streams) has >= three_min - (500 - floor(500/s)) sampled hits.
    // Take the sparsest stride that still leaves a threshold of a few hits (EREF_SAMPLE_STRIDE overrides).
    c->sample_stride = SAMPLE_STRIDE_MIN;
    for (int s = SAMPLE_STRIDE_MAX; s > SAMPLE_STRIDE_MIN; s--)
        if (c->three_min - (WINDOW - WINDOW / s) >= SAMPLE_MIN_HITS) { c->sample_stride = s; break; }
    if (const char* v = getenv("EREF_SAMPLE_STRIDE")) { int k = atoi(v); if (k >= SAMPLE_STRIDE_MIN && k <= SAMPLE_STRIDE_MAX) c->sample_stride = k; }
    cudaError_t e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < EV_COUNT && e == cudaSuccess; i++) e = cudaEventCreate(&c->ev[i]);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaEventCreate(&c->evf[i]);
    for (int i = 0; i < 2 && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&c->aux_ev[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->table_ready, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->totals_ev, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->table_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->frame_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_scan_done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_frame_done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_text_free, cudaEventDisableTiming);
    if (const char* v = getenv("EREF_DOWN_SAMPLING_SIZE")) { long long k = atoll(v); if (k > 0) c->down_sampling_size = (uint64_t)k; }
    if (const char* v = getenv("EREF_PROBE_PARTS")) { int k = atoi(v); if (k == 1 || k == 2 || k == 4 || k == 8 || k == 16 || k == 32) c->probe_parts = k; }
    if (const char* v = getenv("EREF_BUCKET_BITS")) { int k = atoi(v); if (k >= 32 - SLICE_BITS - 8 && k <= 8) c->bucket_bits = k; }
    for (int l = (flags & EREF_FLAG_DIRECT_COUNT) ? 0 : 2; l < 3 && e == cudaSuccess; l++)
        e = cudaMalloc(&c->bm[l], (size_t)TABLE_WORDS * 4);
    if (e == cudaSuccess) e = c->level_pop.ensure(64);

    if (e == cudaSuccess) e = c->counters.ensure(64);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->h_pin, (8 + MAX_BUCKETS) * 8, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaMalloc(&c->summary, (size_t)SUMMARY_WORDS * 4);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->share_words, SHARE_WORDS * 4);
    if (e == cudaSuccess) e = cudaMemset(c->share_words, 0, SHARE_WORDS * 4);
    if (e == cudaSuccess) c->guard = c->share_words + MAX_PARTS + 2;
    if (getenv("EREF_EXACT_SIZES")) c->force_exact = true;
    if (flags & EREF_FLAG_LEVEL_POP) c->want_pop = true;
    if (flags & EREF_FLAG_RESIDENT_TEXT) c->text_is_resident = true;
    if (e != cudaSuccess) { eref_destroy(c); return eref_fail_cuda(e, "context allocation", __FILE__, __LINE__); }
    c->stream = c->own_stream;
    *out = c;
    int rc = eref_query_reset(c);
    if (rc) { eref_destroy(c); *out = nullptr; }
    return rc;
}

void eref_destroy(eref_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->table_stream) cudaStreamSynchronize(c->table_stream);
    if (c->frame_stream) { cudaStreamSynchronize(c->frame_stream); cudaStreamDestroy(c->frame_stream); }
    for (cudaEvent_t e : {c->ev_scan_done, c->ev_frame_done, c->ev_text_free}) if (e) cudaEventDestroy(e);
    c->frame_cub_tmp.release();
    DevBuf* bufs[] = {&c->tile_counts, &c->tile_prefix, &c->cub_tmp, &c->counters, &c->tile_kept, &c->tile_kept_pre, &c->q_keep, &c->q_text, &c->q_hi, &c->q_lo,
                      &c->q_ok, &c->r_text, &c->tile_sum, &c->tile_pre, &c->rec_hdr_off, &c->rec_first,
                      &c->rec_len, &c->rec_kept, &c->rec_words, &c->rec_kept_ex, &c->rec_words_ex, &c->g_rec,
                      &c->g_len, &c->g_word, &c->g_first, &c->r_hi, &c->r_lo, &c->r_ok, &c->word_genome,
                      &c->sample_hits, &c->need, &c->need_list, &c->hits0, &c->hits1, &c->hits2, &c->single, &c->trio, &c->good, &c->results,
                      &c->codes_tmp, &c->cell_pre};
    for (DevBuf* b : bufs) b->release();
    for (int p = 0; p < MAX_PARTS; p++) if (c->list_peers[p].ipc) cudaIpcCloseMemHandle(c->list_peers[p].arena);
    for (auto& L : c->lists) { L.codes.release(); L.off.release(); L.cnt.release(); }
    c->list_arena.release(); c->pulled.release();
    for (int p = 0; p < MAX_PARTS; p++)
        if (c->peers[p].ipc) {
            cudaIpcCloseMemHandle(c->peers[p].bm3); cudaIpcCloseMemHandle(c->peers[p].summary);
            cudaIpcCloseMemHandle(c->peers[p].flags);
        }
    for (int l = 0; l < 3; l++) if (c->bm[l]) cudaFree(c->bm[l]);
    if (c->summary) cudaFree(c->summary);
    if (c->share_words) cudaFree(c->share_words);
    if (c->h_pin) cudaFreeHost(c->h_pin);
    c->part_ctr.release(); c->codes_coarse.release(); c->level_pop.release();
    for (auto& L : c->lists) { L.codes.release(); L.off.release(); L.cnt.release(); }
    for (int i = 0; i < 2; i++) if (c->evf[i]) cudaEventDestroy(c->evf[i]);
    for (int i = 0; i < 2; i++) if (c->aux_ev[i]) cudaEventDestroy(c->aux_ev[i]);
    if (c->table_ready) cudaEventDestroy(c->table_ready);
    if (c->totals_ev) cudaEventDestroy(c->totals_ev);
    for (cudaEvent_t e : c->qev) cudaEventDestroy(e);
    if (c->table_stream) cudaStreamDestroy(c->table_stream);
    for (int i = 0; i < EV_COUNT; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (cudaEvent_t e : c->copy_ev) cudaEventDestroy(e);
    c->q_text2.release(); c->run_results.release();
    delete c;
}

int eref_set_stream(eref_ctx* c, void* cuda_stream)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return 0;
}

int eref_query_reset(eref_ctx* c)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    CUDA_TRY(cudaSetDevice(c->device));
    c->summary_stale = true;
    c->n_lists = 0;
    c->q_pending = false;
    c->table_dirty = true;                     // rebuilt (to the empty table if nothing is added) before the next use
    if (c->flags & EREF_FLAG_DIRECT_COUNT) {
        int wrc = wait_table(c, c->stream);
        if (wrc) return wrc;
        for (int l = 0; l < 3; l++) CUDA_TRY(cudaMemsetAsync(c->bm[l], 0, (size_t)TABLE_WORDS * 4, c->stream));
    }
    return 0;
}

int eref_query_add(eref_ctx* c, const char* fastq_host, size_t n)
{
    if (!c || (n && !fastq_host)) return eref_fail(EREF_ERR_ARG, "eref_query_add: null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(c->q_text.ensure(n + 64));
    CUDA_TRY(cudaEventRecord(c->ev[E0], c->stream));
    if (n) CUDA_TRY(cudaMemcpyAsync(c->q_text.p, fastq_host, n, cudaMemcpyHostToDevice, c->stream));
    return query_add_impl(c, c->q_text.as<uint8_t>(), n);
}

int eref_query_add_device(eref_ctx* c, const void* fastq_dev, size_t n)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    int rc = check_text_ptr(fastq_dev, n);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaEventRecord(c->ev[E0], c->stream));
    return query_add_impl(c, static_cast<const uint8_t*>(fastq_dev), n);
}

int eref_query_add_masked(eref_ctx* c, const char* fastq_host, size_t n, const uint32_t* keep_bits, uint64_t n_keep)
{
    if (!c || (n && !fastq_host)) return eref_fail(EREF_ERR_ARG, "eref_query_add_masked: null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(c->q_text.ensure(n + 64));
    CUDA_TRY(cudaEventRecord(c->ev[E0], c->stream));
    if (n) CUDA_TRY(cudaMemcpyAsync(c->q_text.p, fastq_host, n, cudaMemcpyHostToDevice, c->stream));
    return query_add_impl(c, c->q_text.as<uint8_t>(), n, keep_bits, n_keep);
}

int eref_query_add_masked_device(eref_ctx* c, const void* fastq_dev, size_t n, const uint32_t* keep_bits, uint64_t n_keep)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    int rc = check_text_ptr(fastq_dev, n);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaEventRecord(c->ev[E0], c->stream));
    return query_add_impl(c, static_cast<const uint8_t*>(fastq_dev), n, keep_bits, n_keep);
}

int eref_query_census(eref_ctx* c, const char* fastq_host, size_t n, uint64_t* seq_bytes, uint64_t* seq_lines)
{
    if (!c || (n && !fastq_host)) return eref_fail(EREF_ERR_ARG, "eref_query_census: null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    int wrc = wait_table(c, c->stream);
    if (wrc) return wrc;
    CUDA_TRY(c->q_text.ensure(n + 64));
    if (n) CUDA_TRY(cudaMemcpyAsync(c->q_text.p, fastq_host, n, cudaMemcpyHostToDevice, c->stream));
    return census_impl(c, c->q_text.as<uint8_t>(), n, seq_bytes, seq_lines);
}

int eref_query_census_device(eref_ctx* c, const void* fastq_dev, size_t n, uint64_t* seq_bytes, uint64_t* seq_lines)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    int rc = check_text_ptr(fastq_dev, n);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(c->device));
    return census_impl(c, static_cast<const uint8_t*>(fastq_dev), n, seq_bytes, seq_lines);
}

// read_fastq's byte budget (:942-945): the getline loop stops in front of the first line whose PRECEDING lines add up
// to >= end characters, newlines not counted.  main hands file_size(fq1) to both files (:1265-1291), so fq2 is cut short
// when it is the longer one.  A line that starts at byte c behind nl(c) newlines is read iff c - nl(c) < end; returns the
// first line start that is not (the length of the prefix that is read), or n.  Host code, no CUDA.
// Cost: nothing when n <= end; else newlines are counted block-wise from the start (no line start below `end` can be
// cut) and the walk stops as soon as n - end of them were seen — behind that many newlines no line start can reach
// the budget — which for two mates of similar size is after a small part of the text.
size_t eref_fastq_limit(const char* text, size_t n, uint64_t end)
{
    if (!text || n <= end) return n;
    if (end == 0) return 0;                              // the very first line already meets an empty budget
    const uint64_t d = n - end;
    size_t pos = 0;
    uint64_t cnt = 0;
    while (pos < end) {
        const size_t len = std::min<size_t>((size_t)1 << 16, (size_t)end - pos);
        const char* p = text + pos;
        uint32_t k = 0;
        for (size_t i = 0; i < len; i++) k += p[i] == '\n';
        cnt += k;
        pos += len;
        if (cnt >= d) return n;
    }
    while (pos < n) {
        const char* q = (const char*)memchr(text + pos, '\n', n - pos);
        if (!q) return n;
        cnt++;
        const size_t c = (size_t)(q - text) + 1;
        if (c < n && c - cnt >= end) return c;
        if (cnt >= d) return n;
        pos = c;
    }
    return n;
}

int eref_sample_ratio(uint64_t seq_bytes_fq1, uint64_t down_sampling_size, int* ratio)
{
    if (!ratio) return eref_fail(EREF_ERR_ARG, "null argument");
    if (seq_bytes_fq1 == 0)
        return eref_fail(EREF_ERR_UNSUPPORTED, "fq1 holds no sequence lines (the reference divides by zero here, :1145)");
    const long sample_size = (long)seq_bytes_fq1 * 2;                       // :1144 "pair"
    *ratio = (int)(100L * (long)down_sampling_size / sample_size);          // :1145, long division narrowed to int
    return 0;
}

void eref_draw_keep(int ratio, uint64_t n_lines, uint32_t* keep_bits)
{
    for (uint64_t i = 0; i < (n_lines + 31) / 32; i++) keep_bits[i] = 0;
    for (uint64_t i = 0; i < n_lines; i++)
        if (rand() % 100 < ratio) keep_bits[i >> 5] |= 1u << (i & 31);      // :955-957, one draw per sequence line
}

int eref_query_seq_bytes(eref_ctx* c, uint64_t* out)
{
    if (!c || !out) return eref_fail(EREF_ERR_ARG, "null argument");
    unsigned long long v = 0;
    int rc = last_file_seq_bytes(c, &v);
    if (rc) return rc;
    *out = v;
    return 0;
}

int eref_ref_set(eref_ctx* c, const char* fasta_host, size_t n)
{
    if (!c || (n && !fasta_host)) return eref_fail(EREF_ERR_ARG, "eref_ref_set: null argument");
    if (n >= REF_TEXT_LIMIT)
        return eref_fail(EREF_ERR_LIMIT, "reference text of %llu bytes: one shard must stay below 4 GiB",
                         (unsigned long long)n);
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(c->r_text.ensure(n + 64));
    return ref_set_impl(c, c->r_text.as<uint8_t>(), n, nullptr, fasta_host);
}

int eref_ref_set_device(eref_ctx* c, const void* fasta_dev, size_t n)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    int rc = check_text_ptr(fasta_dev, n);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(c->device));
    // whatever was queued on the context's stream before this call may have produced the text
    CUDA_TRY(cudaEventRecord(c->aux_ev[1], c->stream));
    return ref_set_impl(c, static_cast<const uint8_t*>(fasta_dev), n, c->text_is_resident ? nullptr : c->aux_ev[1]);
}

int eref_ref_count(eref_ctx* c, size_t* n_kept, size_t* n_records, uint64_t* kept_bases)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "no reference set");
    if (n_kept) *n_kept = c->n_genomes;
    if (n_records) *n_records = c->n_recs;
    if (kept_bases) *kept_bases = c->ref_bases;
    return 0;
}

// tail of a synchronous scan: the filter's word count and the sharded build's error word come back with the records,
// the stream is waited for and the phase times are read
static int finish_scan_sync(eref_ctx* c)
{
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 6, c->counters.as<unsigned long long>() + 1, 8, cudaMemcpyDeviceToHost, c->stream));
    if (c->parts > 1) CUDA_TRY(cudaMemcpyAsync(c->h_pin + 7, c->share_words + MAX_PARTS, 8, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaEventRecord(c->ev[S7], c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->scan_pending = true; c->scan_pending_d2h = true;
    collect_times(c);
    if (int grc = check_guard(c)) return grc;
    if (c->parts > 1 && (c->h_pin[7] >> 32)) {
        const unsigned who = (unsigned)(c->h_pin[7] >> 32) - 1;
        cudaMemsetAsync(c->share_words + MAX_PARTS + 1, 0, 4, c->stream);
        return eref_fail(EREF_ERR_STATE, "sharded table build: part %u never arrived in this GPU's table (is every GPU "
                                         "of the build running the same sequence of calls?)", who);
    }
    return 0;
}

int eref_scan(eref_ctx* c, eref_result* out, size_t cap, size_t* n_out)
{
    if (!c || (cap && !out)) return eref_fail(EREF_ERR_ARG, "eref_scan: null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = scan_impl(c);
    if (rc) return rc;
    size_t n = c->n_genomes < cap ? c->n_genomes : cap;
    if (n) CUDA_TRY(cudaMemcpyAsync(out, c->results.p, n * sizeof(eref_result), cudaMemcpyDeviceToHost, c->stream));
    if (n_out) *n_out = c->n_genomes;
    return finish_scan_sync(c);
}

// eref_run_host: scan the current reference (one database piece) and append its records, ordinals continuing in file
// order, to the run's list in device memory; waits for the scan (the next piece reuses the buffers)
static int scan_append(eref_ctx* c, size_t* total)
{
    int rc = scan_impl(c);
    if (rc) return rc;
    const size_t G = c->n_genomes;
    CUDA_TRY(c->run_results.grow((*total + G + 1) * sizeof(eref_result), *total * sizeof(eref_result), c->stream));
    if (G) {
        uint32_t* dst = c->run_results.as<uint32_t>() + *total * 4;
        CUDA_TRY(cudaMemcpyAsync(dst, c->results.p, G * sizeof(eref_result), cudaMemcpyDeviceToDevice, c->stream));
        LAUNCH(c, launch_add_ordinal(dst, (uint32_t)G, (uint32_t)*total, c->stream));
    }
    *total += G;
    return finish_scan_sync(c);
}

/* The whole path in one call, from HOST buffers, with the H2D copies overlapped with the kernels: both
 * FASTQ files and the FASTA text (cut into `chunks` pieces at record boundaries) are queued on a copy
 * stream; the compute stream counts each file as it lands, brings the table up to date while the database
 * is still in flight, and frames+scans every database piece as soon as its copy completes.  Pieces are
 * independent shards (a record's result depends only on its own bases, :846-896), so their records are
 * concatenated with ordinals continuing in file order.  Host buffers should be pinned for the overlap. */
static int run_host_impl(eref_ctx* c, const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
                         int chunks, eref_result* out, size_t cap, size_t* n_out);

int eref_run_host(eref_ctx* c, const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
                  int chunks, eref_result* out, size_t cap, size_t* n_out)
{
    int rc = run_host_impl(c, fq1, n1, fq2, n2, fasta, nf, chunks, out, cap, n_out);
    // a query file outgrew buffers sized from its text length: once more with exact sizes (a sharded build cannot
    // repeat on its own, the other GPUs would have to follow)
    if (rc == EREF_ERR_RETRY && c->parts == 1) rc = run_host_impl(c, fq1, n1, fq2, n2, fasta, nf, chunks, out, cap, n_out);
    return rc;
}

static int run_host_impl(eref_ctx* c, const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
                         int chunks, eref_result* out, size_t cap, size_t* n_out)
{
    if (!c || (n1 && !fq1) || (n2 && !fq2) || (nf && !fasta) || (cap && !out))
        return eref_fail(EREF_ERR_ARG, "eref_run_host: null argument");
    if (c->share_reads)
        return eref_fail(EREF_ERR_STATE, "eref_run_host takes whole query files (cal_sam_ratio needs all of fq1): turn shared "
                                         "reads off (eref_query_set_share) or use the phase calls");
    CUDA_TRY(cudaSetDevice(c->device));
    // fasta == NULL: screen this sample against the reference that is already set (one database, many samples)
    const bool resident = !fasta && nf == 0;
    if (resident && !c->have_ref)
        return eref_fail(EREF_ERR_STATE, "eref_run_host without a database needs one set by eref_ref_set");
    if (chunks < 1) chunks = 1;
    if ((uint64_t)chunks < (nf >> 31) + 1) chunks = (int)std::min<uint64_t>((nf >> 31) + 1, 64);   // pieces of <= 2 GiB
    if (chunks > 64) chunks = 64;
    // piece boundaries: offsets that start a header line ('>' right after a newline)
    // Equal pieces.  Measured alternative that lost: geometrically shrinking pieces (a small last piece should leave
    // a short tail after the last byte lands) — every piece costs ~0.5 ms of launches and host round trips, so the
    // small pieces queue up behind each other and the tail grows (0.65 ms with 4 equal pieces, 0.76-1.1 ms shrinking).
    std::vector<size_t> cut(1, 0);
    for (int k = 1; k < chunks; k++) {
        size_t pos = nf / chunks * k;
        if (pos < cut.back()) pos = cut.back();
        size_t found = nf;
        while (pos < nf) {
            const char* nl = (const char*)memchr(fasta + pos, '\n', nf - pos);
            if (!nl) break;
            size_t q = (size_t)(nl - fasta) + 1;
            if (q < nf && fasta[q] == '>') { found = q; break; }
            pos = q;
        }
        if (found > cut.back() && found < nf) cut.push_back(found);
    }
    cut.push_back(nf);
    const size_t npieces = cut.size() - 1;
    while (c->copy_ev.size() < npieces + 2) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->copy_ev.push_back(e);
    }
    CUDA_TRY(c->q_text.ensure(n1 + 64));
    CUDA_TRY(c->q_text2.ensure(n2 + 64));
    if (!resident) CUDA_TRY(c->r_text.ensure(nf + 256 * (npieces + 1)));
    cudaStream_t cs = c->copy_stream, st = c->stream;
    // the copy stream must not overwrite buffers the compute stream may still be reading
    CUDA_TRY(cudaEventRecord(c->ev[E9], st));
    CUDA_TRY(cudaStreamWaitEvent(cs, c->ev[E9], 0));
    static const bool trace = getenv("EREF_TRACE_E2E") != nullptr;
    const auto h0 = std::chrono::steady_clock::now();
    // whatever way this call ends, no copy from the caller's buffers is still in flight when it returns
    struct CopyGuard { cudaStream_t s; ~CopyGuard() { cudaStreamSynchronize(s); } } copy_guard{cs};
    CUDA_TRY(cudaEventRecord(c->ev[E10], cs));
    // cudaMemcpyDefault: the texts may also be device pointers (e.g. a query all-gathered over NVLink by the caller)
    if (n1) CUDA_TRY(cudaMemcpyAsync(c->q_text.p, fq1, n1, cudaMemcpyDefault, cs));
    CUDA_TRY(cudaEventRecord(c->copy_ev[0], cs));
    if (n2) CUDA_TRY(cudaMemcpyAsync(c->q_text2.p, fq2, n2, cudaMemcpyDefault, cs));
    CUDA_TRY(cudaEventRecord(c->copy_ev[1], cs));
    std::vector<uint8_t*> piece_dev(npieces);
    size_t doff = 0;
    for (size_t k = 0; k < npieces; k++) {
        piece_dev[k] = c->r_text.as<uint8_t>() + doff;
        const size_t len = cut[k + 1] - cut[k];
        if (len) CUDA_TRY(cudaMemcpyAsync(piece_dev[k], fasta + cut[k], len, cudaMemcpyDefault, cs));
        CUDA_TRY(cudaEventRecord(c->copy_ev[2 + k], cs));
        doff += (len + 255) & ~(size_t)255;
    }
    CUDA_TRY(cudaEventRecord(c->ev[E11], cs));
    // main reads fq2 only up to fq1's size (:1265-1291, eref_fastq_limit): whatever lies behind that was uploaded but
    // is not counted.  Looked up here, while the first copy is still in flight.  A query in DEVICE memory cannot be
    // walked by the host: its caller passes the limited size (the executable does).
    {
        cudaPointerAttributes at{};
        const bool on_device = n2 && cudaPointerGetAttributes(&at, fq2) == cudaSuccess && at.type == cudaMemoryTypeDevice;
        cudaGetLastError();
        if (n2 && !on_device) n2 = eref_fastq_limit(fq2, n2, n1);
    }
    const auto h1 = std::chrono::steady_clock::now();
    int rc = eref_query_reset(c);
    if (rc) return rc;
    CUDA_TRY(cudaStreamWaitEvent(st, c->copy_ev[0], 0));
    CUDA_TRY(cudaEventRecord(c->ev[E0], st));
    rc = query_add_impl(c, c->q_text.as<uint8_t>(), n1);
    if (rc) return rc;
    unsigned long long seq1 = 0;
    rc = last_file_seq_bytes(c, &seq1);       // cal_sam_ratio needs fq1's byte count before fq2 is counted
    if (rc) return rc;
    eref_phase_times sum = c->t;
    // cal_sam_ratio (:1124-1148): 100 * 2e9 / (2 * sequence bytes of fq1)
    int ratio = 0;
    rc = eref_sample_ratio(seq1, c->down_sampling_size, &ratio);
    if (rc) return rc;
    std::vector<uint32_t> keep1, keep2;
    uint64_t lines1 = 0, lines2 = 0;
    if (ratio < 100) {
        // Down-sampling (:955-960): every sequence line of fq1, then of fq2, draws rand() % 100 and is counted iff the
        // draw is below the ratio.  Replay the draws on the host in the reference's order (its main() seeds rand with
        // 1, :1236-1238; EREF_FLAG_KEEP_RAND_STATE: the caller has positioned rand() already, like a cold reference run
        // whose coder draw re-seeded it), then count both files again under the masks.
        if (c->pend_keep) {
            // the caller drew the masks (several GPUs screen the same sample: the draws happen once, :955-957)
            keep1.swap(c->pend_keep1); keep2.swap(c->pend_keep2);
            lines1 = c->pend_lines1; lines2 = c->pend_lines2;
            c->pend_keep = false;
        } else {
            rc = census_impl(c, c->q_text.as<uint8_t>(), n1, nullptr, &lines1);
            if (rc) return rc;
            CUDA_TRY(cudaStreamWaitEvent(st, c->copy_ev[1], 0));
            rc = census_impl(c, c->q_text2.as<uint8_t>(), n2, nullptr, &lines2);
            if (rc) return rc;
            if (!(c->flags & EREF_FLAG_KEEP_RAND_STATE)) srand(1);
            keep1.resize((lines1 + 31) / 32 + 1); keep2.resize((lines2 + 31) / 32 + 1);
            eref_draw_keep(ratio, lines1, keep1.data());
            eref_draw_keep(ratio, lines2, keep2.data());
        }
        rc = eref_query_reset(c);
        if (rc) return rc;
        CUDA_TRY(cudaEventRecord(c->ev[E0], st));
        rc = query_add_impl(c, c->q_text.as<uint8_t>(), n1, keep1.data(), lines1);
        if (rc) return rc;
    }
    CUDA_TRY(cudaStreamWaitEvent(st, c->copy_ev[1], 0));
    CUDA_TRY(cudaEventRecord(c->ev[E0], st));
    rc = ratio < 100 ? query_add_impl(c, c->q_text2.as<uint8_t>(), n2, keep2.data(), lines2)
                     : query_add_impl(c, c->q_text2.as<uint8_t>(), n2);
    if (rc) return rc;
    rc = start_table_update(c);               // counted while the first database piece is framed and packed
    if (rc) return rc;
    sum.ref_frame_ms = sum.ref_pack_ms = 0;
    sum.scan_summary_ms = sum.scan_sample_ms = sum.scan_cand_ms = sum.scan_probe_ms = sum.scan_window_ms = 0;
    sum.scan_interval_ms = sum.scan_d2h_ms = 0;
    sum.ref_bases = sum.ref_words = sum.ref_need_words = 0;
    size_t total = 0;
    if (resident) {
        rc = scan_append(c, &total);
        if (rc) return rc;
        sum.scan_summary_ms = c->t.scan_summary_ms; sum.scan_sample_ms = c->t.scan_sample_ms;
        sum.scan_cand_ms = c->t.scan_cand_ms; sum.scan_probe_ms = c->t.scan_probe_ms;
        sum.scan_window_ms = c->t.scan_window_ms; sum.scan_interval_ms = c->t.scan_interval_ms;
        sum.scan_d2h_ms = c->t.scan_d2h_ms; sum.query_count_ms = c->t.query_count_ms;
        sum.ref_bases = c->t.ref_bases; sum.ref_words = c->t.ref_words; sum.ref_need_words = c->t.ref_need_words;
    }
    for (size_t k = 0; k < npieces && !resident; k++) {
        rc = ref_set_impl(c, piece_dev[k], cut[k + 1] - cut[k], c->copy_ev[2 + k]);
        if (rc) return rc;
        rc = scan_append(c, &total);
        if (rc) return rc;
        sum.ref_frame_ms += c->t.ref_frame_ms; sum.ref_pack_ms += c->t.ref_pack_ms;
        sum.scan_summary_ms += c->t.scan_summary_ms; sum.scan_sample_ms += c->t.scan_sample_ms;
        sum.scan_cand_ms += c->t.scan_cand_ms; sum.scan_probe_ms += c->t.scan_probe_ms;
        sum.scan_window_ms += c->t.scan_window_ms; sum.scan_interval_ms += c->t.scan_interval_ms;
        sum.scan_d2h_ms += c->t.scan_d2h_ms;
        sum.query_count_ms = c->t.query_count_ms;
        sum.ref_bases += c->t.ref_bases; sum.ref_words += c->t.ref_words; sum.ref_need_words += c->t.ref_need_words;
    }
    c->run_total = total;
    // the records of all pieces leave the device in one copy (they also stay there: eref_run_results_device)
    const size_t wrote = total < cap ? total : cap;
    if (wrote) {
        CUDA_TRY(cudaMemcpyAsync(out, c->run_results.p, wrote * sizeof(eref_result), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
    }
    if (trace) {
        const auto h2 = std::chrono::steady_clock::now();
        float copy_ms = 0, tail_ms = 0;
        cudaEventElapsedTime(&copy_ms, c->ev[E10], c->ev[E11]);
        cudaEventElapsedTime(&tail_ms, c->ev[E11], c->ev[S7]);   // (S7: end of the last piece's scan)
        fprintf(stderr, "eref_run_host: queued copies after %.3f ms, returned after %.3f ms; copies took %.3f ms on the "
                        "device, last copy -> last result %.3f ms\n",
                std::chrono::duration<double, std::milli>(h1 - h0).count(),
                std::chrono::duration<double, std::milli>(h2 - h0).count(), copy_ms, tail_ms);
    }
    // the query times are sums over both files (collected by the first eref_scan's synchronisation)
    sum.query_pack_ms = c->t.query_pack_ms; sum.query_bucket_ms = c->t.query_bucket_ms;
    sum.query_seq_bytes = c->t.query_seq_bytes; sum.query_codes = c->t.query_codes;
    if (c->flags & EREF_FLAG_DIRECT_COUNT) sum.query_count_ms = c->t.query_count_ms;
    c->t = sum;
    if (!resident) c->have_ref = npieces == 1;   // per-record lookups only make sense for a single piece
    if (n_out) *n_out = total;
    return 0;
}

int eref_run_host_keep(eref_ctx* c, const uint32_t* keep1, uint64_t lines1, const uint32_t* keep2, uint64_t lines2)
{
    if (!c || (lines1 && !keep1) || (lines2 && !keep2)) return eref_fail(EREF_ERR_ARG, "eref_run_host_keep: null argument");
    c->pend_keep1.assign(keep1, keep1 + (lines1 + 31) / 32); c->pend_keep1.push_back(0);
    c->pend_keep2.assign(keep2, keep2 + (lines2 + 31) / 32); c->pend_keep2.push_back(0);
    c->pend_lines1 = lines1; c->pend_lines2 = lines2;
    c->pend_keep = true;
    return 0;
}

int eref_run_results_device(eref_ctx* c, const uint32_t** results_dev, size_t* n_out)
{
    if (!c || !results_dev) return eref_fail(EREF_ERR_ARG, "eref_run_results_device: null argument");
    *results_dev = c->run_results.as<uint32_t>();
    if (n_out) *n_out = c->run_total;
    return 0;
}

int eref_scan_device_async(eref_ctx* c, const uint32_t** results_dev, size_t* n_out)
{
    if (!c || !results_dev) return eref_fail(EREF_ERR_ARG, "eref_scan_device: null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = scan_impl(c);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->h_pin + 6, c->counters.as<unsigned long long>() + 1, 8, cudaMemcpyDeviceToHost, c->stream));
    c->scan_pending = true; c->scan_pending_d2h = false;
    *results_dev = c->results.as<uint32_t>();
    if (n_out) *n_out = c->n_genomes;
    return 0;
}

int eref_scan_device(eref_ctx* c, const uint32_t** results_dev, size_t* n_out)
{
    int rc = eref_scan_device_async(c, results_dev, n_out);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    collect_times(c);
    return check_guard(c);
}

int eref_get_codes(eref_ctx* c, uint32_t ordinal, uint32_t* codes_host, size_t cap, size_t* n_out)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "no reference set");
    if (ordinal < 1 || ordinal > c->n_genomes) return eref_fail(EREF_ERR_ARG, "ordinal %u of %u", ordinal, c->n_genomes);
    CUDA_TRY(cudaSetDevice(c->device));
    int hrc = fetch_host_tables(c);
    if (hrc) return hrc;
    const uint32_t len = c->h_g_len[ordinal - 1];
    const size_t n = (size_t)(len - EREF_K + 1) * 3;
    if (n_out) *n_out = n;
    if (!codes_host || cap < n) return 0;
    CUDA_TRY(c->codes_tmp.ensure(n * 4));
    Planes pl;
    pl.hi = c->r_hi.as<uint32_t>(); pl.lo = c->r_lo.as<uint32_t>(); pl.ok = c->r_ok.as<uint32_t>();
    pl.nwords = c->ref_words; pl.cap = c->r_hi.cap / 4;
    LAUNCH(c, launch_record_codes(pl, c->masks, c->h_g_word[ordinal - 1], len, c->codes_tmp.as<uint32_t>(), c->stream));
    CUDA_TRY(cudaMemcpyAsync(codes_host, c->codes_tmp.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

int eref_ref_record(eref_ctx* c, uint32_t ordinal, uint64_t* hdr_off, uint32_t* hdr_len, uint32_t* len)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "no reference set");
    if (ordinal < 1 || ordinal > c->n_genomes) return eref_fail(EREF_ERR_ARG, "ordinal %u of %u", ordinal, c->n_genomes);
    CUDA_TRY(cudaSetDevice(c->device));
    int hrc = fetch_host_tables(c);
    if (hrc) return hrc;
    const uint32_t r = c->h_g_rec[ordinal - 1];
    if (len) *len = c->h_g_len[ordinal - 1];
    uint64_t off = 0; uint32_t hl = 0;
    if (r > 0) {                                      // record 0 has no header line
        off = c->h_rec_hdr_off[r];
        // the header runs to the next newline (or the end of the text): look for it 256 bytes at a time
        char buf[256];
        uint64_t pos = off;
        bool found = false;
        while (pos < c->ref_text_n && !found) {
            const size_t k = (size_t)std::min<uint64_t>(sizeof buf, c->ref_text_n - pos);
            CUDA_TRY(cudaMemcpy(buf, c->ref_text_dev + pos, k, cudaMemcpyDeviceToHost));
            const void* nl = memchr(buf, '\n', k);
            if (nl) { pos += (uint64_t)((const char*)nl - buf); found = true; }
            else pos += k;
        }
        hl = (uint32_t)(pos - off);
    }
    if (hdr_off) *hdr_off = off;
    if (hdr_len) *hdr_len = hl;
    return 0;
}

int eref_ref_records(eref_ctx* c, uint32_t* len_out, uint32_t* hdr_off_out, size_t cap, size_t* n_out)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "no reference set");
    CUDA_TRY(cudaSetDevice(c->device));
    const size_t n = c->n_recs < cap ? c->n_recs : cap;
    if (n && len_out) CUDA_TRY(cudaMemcpy(len_out, c->rec_len.p, n * 4, cudaMemcpyDeviceToHost));
    if (n && hdr_off_out) CUDA_TRY(cudaMemcpy(hdr_off_out, c->rec_hdr_off.p, n * 4, cudaMemcpyDeviceToHost));
    if (n_out) *n_out = c->n_recs;
    return 0;
}

int eref_index_body(eref_ctx* c, uint32_t first, uint32_t* out_host, size_t cap_words, size_t* n_words, uint32_t* next)
{
    if (!c || !out_host || !n_words || !next) return eref_fail(EREF_ERR_ARG, "eref_index_body: null argument");
    if (!c->have_ref) return eref_fail(EREF_ERR_STATE, "no reference set");
    if (first < 1 || first > c->n_genomes + 1) return eref_fail(EREF_ERR_ARG, "ordinal %u of %u", first, c->n_genomes);
    CUDA_TRY(cudaSetDevice(c->device));
    int hrc = fetch_host_tables(c);
    if (hrc) return hrc;
    const uint32_t G = c->n_genomes;
    if (!c->cell_pre_valid) {                             // cells (length word + window starts) before each kept record
        c->h_cell_pre.assign((size_t)G + 1, 0);
        for (uint32_t g = 0; g < G; g++) c->h_cell_pre[g + 1] = c->h_cell_pre[g] + (c->h_g_len[g] - (EREF_K - 2));
        CUDA_TRY(c->cell_pre.ensure(((size_t)G + 1) * 8));
        CUDA_TRY(cudaMemcpy(c->cell_pre.p, c->h_cell_pre.data(), ((size_t)G + 1) * 8, cudaMemcpyHostToDevice));
        c->cell_pre_valid = true;
    }
    uint32_t g0 = first - 1, g1 = g0;
    size_t words = 0;
    while (g1 < G) {
        const size_t w = 1 + 3 * (size_t)(c->h_g_len[g1] - (EREF_K - 1));
        if (words + w > cap_words) break;
        words += w; g1++;
    }
    if (g1 == g0 && g0 < G)
        return eref_fail(EREF_ERR_LIMIT, "record %u needs %zu words, the buffer holds %zu", first,
                         1 + 3 * (size_t)(c->h_g_len[g0] - (EREF_K - 1)), cap_words);
    if (words) {
        CUDA_TRY(c->codes_tmp.ensure(words * 4));
        Planes pl;
        pl.hi = c->r_hi.as<uint32_t>(); pl.lo = c->r_lo.as<uint32_t>(); pl.ok = c->r_ok.as<uint32_t>();
        pl.nwords = c->ref_words; pl.cap = c->r_hi.cap / 4;
        LAUNCH(c, launch_index_body(pl, c->masks, c->g_word.as<uint32_t>(), c->g_len.as<uint32_t>(),
                                    c->cell_pre.as<unsigned long long>(), g0, g1, c->h_cell_pre[g1] - c->h_cell_pre[g0],
                                    c->codes_tmp.as<uint32_t>(), c->stream));
        CUDA_TRY(cudaMemcpyAsync(out_host, c->codes_tmp.p, words * 4, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    *n_words = words;
    *next = g1 + 1;
    return 0;
}

int eref_query_popcount(eref_ctx* c, int level, uint64_t* out)
{
    if (!c || !out || level < 1 || level > 3) return eref_fail(EREF_ERR_ARG, "eref_query_popcount: bad argument");
    CUDA_TRY(cudaSetDevice(c->device));
    if (c->parts > 1 && !c->want_pop)
        return eref_fail(EREF_ERR_STATE, "level populations of a sharded build need EREF_FLAG_LEVEL_POP at eref_create "
                                         "(a rebuild for the count would write into the peers' tables)");
    if (!(c->flags & EREF_FLAG_DIRECT_COUNT) && !c->want_pop) { c->want_pop = true; c->table_dirty = true; }
    if (!(c->flags & EREF_FLAG_DIRECT_COUNT) && !c->have_pop && c->parts == 1) c->table_dirty = true;
    int frc = finalize_table(c);
    if (frc) return frc;
    unsigned long long h = 0;
    if (!(c->flags & EREF_FLAG_DIRECT_COUNT)) {          // the build counted the levels while it had them in shared memory
        CUDA_TRY(cudaMemcpy(&h, c->level_pop.as<unsigned long long>() + (level - 1), 8, cudaMemcpyDeviceToHost));
        *out = h;
        return 0;
    }
    unsigned long long* ctr = c->counters.as<unsigned long long>() + 2;
    CUDA_TRY(cudaMemsetAsync(ctr, 0, 8, c->stream));
    LAUNCH(c, launch_popcount(c->bm[level - 1], TABLE_WORDS, ctr, c->stream));
    CUDA_TRY(cudaMemcpyAsync(&h, ctr, 8, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    *out = h;
    return 0;
}

int eref_set_serial(eref_ctx* c, int on)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    c->serial = on != 0;
    return 0;
}

int eref_sync(eref_ctx* c)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (c->table_inflight) { CUDA_TRY(cudaStreamSynchronize(c->table_stream)); table_done(c); }
    collect_times(c);
    return check_guard(c);
}

int eref_table_export(eref_ctx* c, void* ipc_handles, void** bm3_dev, void** summary_dev, void** flags_dev)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (c->flags & EREF_FLAG_DIRECT_COUNT) return eref_fail(EREF_ERR_UNSUPPORTED, "the direct build cannot be sharded");
    CUDA_TRY(cudaSetDevice(c->device));
    if (ipc_handles) {
        static_assert(3 * sizeof(cudaIpcMemHandle_t) == EREF_TABLE_HANDLE_BYTES, "handle blob size");
        cudaIpcMemHandle_t h[3];
        CUDA_TRY(cudaIpcGetMemHandle(&h[0], c->bm[2]));
        CUDA_TRY(cudaIpcGetMemHandle(&h[1], c->summary));
        CUDA_TRY(cudaIpcGetMemHandle(&h[2], c->share_words));
        memcpy(ipc_handles, h, sizeof h);
    }
    if (bm3_dev) *bm3_dev = c->bm[2];
    if (summary_dev) *summary_dev = c->summary;
    if (flags_dev) *flags_dev = c->share_words;
    return 0;
}

int eref_table_set_part(eref_ctx* c, int part, int parts)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (parts < 1 || parts > MAX_PARTS || part < 0 || part >= parts)
        return eref_fail(EREF_ERR_ARG, "eref_table_set_part: part %d of %d (at most %d parts)", part, parts, MAX_PARTS);
    if (parts > 1 && (c->flags & EREF_FLAG_DIRECT_COUNT))
        return eref_fail(EREF_ERR_UNSUPPORTED, "the direct build cannot be sharded");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = eref_sync(c);
    if (rc) return rc;
    c->parts = parts; c->part = part;
    if (parts > 1) CUDA_TRY(preload_share_kernels());
    c->part_w = part_weight(parts, 0);
    c->build_sharded = true;
    set_part_range(c);
    for (int p = 0; p < MAX_PARTS; p++) {
        if (c->list_peers[p].ipc) cudaIpcCloseMemHandle(c->list_peers[p].arena);
        c->list_peers[p] = eref_ctx::ListPeer();
    }
    c->share_reads = false;
    for (int p = 0; p < MAX_PARTS; p++) {           // peers are set after the part (a new layout drops the old wiring)
        if (c->peers[p].ipc) { cudaIpcCloseMemHandle(c->peers[p].bm3); cudaIpcCloseMemHandle(c->peers[p].summary); cudaIpcCloseMemHandle(c->peers[p].flags); }
        c->peers[p] = eref_ctx::Peer();
    }
    c->table_dirty = true; c->join_pending = false;
    c->n_lists = 0;                             // code lists hold only the codes of the part they were made for
    return 0;
}

int eref_table_set_peer(eref_ctx* c, int peer_part, int peer_device, const void* ipc_handles, void* bm3_dev,
                        void* summary_dev, void* flags_dev)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (peer_part < 0 || peer_part >= c->parts || peer_part == c->part)
        return eref_fail(EREF_ERR_ARG, "eref_table_set_peer: part %d (this context is part %d of %d)", peer_part, c->part, c->parts);
    CUDA_TRY(cudaSetDevice(c->device));
    eref_ctx::Peer& P = c->peers[peer_part];
    if (P.ipc) { cudaIpcCloseMemHandle(P.bm3); cudaIpcCloseMemHandle(P.summary); cudaIpcCloseMemHandle(P.flags); }
    P = eref_ctx::Peer();
    if (ipc_handles) {
        cudaIpcMemHandle_t h[3];
        memcpy(h, ipc_handles, sizeof h);
        void *a = nullptr, *b = nullptr, *f = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&a, h[0], cudaIpcMemLazyEnablePeerAccess));
        CUDA_TRY(cudaIpcOpenMemHandle(&b, h[1], cudaIpcMemLazyEnablePeerAccess));
        CUDA_TRY(cudaIpcOpenMemHandle(&f, h[2], cudaIpcMemLazyEnablePeerAccess));
        P.bm3 = (uint32_t*)a; P.summary = (uint32_t*)b; P.flags = (uint32_t*)f; P.ipc = true;
        return 0;
    }
    if (!bm3_dev || !summary_dev || !flags_dev) return eref_fail(EREF_ERR_ARG, "eref_table_set_peer: null table pointer");
    if (peer_device >= 0 && peer_device != c->device) {
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, c->device, peer_device));
        if (!can) return eref_fail(EREF_ERR_UNSUPPORTED, "GPU %d cannot access GPU %d's memory", c->device, peer_device);
        cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) return eref_fail_cuda(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
    }
    P.bm3 = (uint32_t*)bm3_dev; P.summary = (uint32_t*)summary_dev; P.flags = (uint32_t*)flags_dev;
    return 0;
}

int eref_query_reserve(eref_ctx* c, size_t max_file_bytes, int max_files)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (max_files < 1 || max_files > MAX_LISTS) return eref_fail(EREF_ERR_ARG, "eref_query_reserve: %d files (1..%d)", max_files, MAX_LISTS);
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = eref_sync(c);
    if (rc) return rc;
    // room for the codes of a well-formed FASTQ file of that size (3 per sequence byte, fewer than half the bytes are
    // sequence), every slice padded to 4 codes
    const unsigned long long cap = ((3ull * (max_file_bytes / 2 + 1) + 4ull * NSLICES + 64 + 3) & ~3ull) + 4;
    const size_t stride = ((((size_t)2 * NSLICES + 2) * 8 + (size_t)cap * 4) + 255) & ~(size_t)255;
    for (auto& L : c->lists) if (L.codes.view) { L.codes.release(); L.off.release(); L.cnt.release(); }
    c->n_lists = 0; c->table_dirty = true;
    c->list_arena.release();                   // (a new arena has a new address: peers must be wired again)
    CUDA_TRY(c->list_arena.ensure(stride * max_files));
    c->list_stride = stride; c->arena_cap_codes = cap; c->arena_lists = max_files;
    CUDA_TRY(c->codes_coarse.ensure((size_t)cap * 4 + 16));
    return 0;
}

int eref_lists_export(eref_ctx* c, void* ipc_handle, void** arena_dev)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (!c->list_arena.p) return eref_fail(EREF_ERR_STATE, "eref_lists_export before eref_query_reserve");
    CUDA_TRY(cudaSetDevice(c->device));
    if (ipc_handle) {
        static_assert(sizeof(cudaIpcMemHandle_t) == EREF_LISTS_HANDLE_BYTES, "handle blob size");
        cudaIpcMemHandle_t h;
        CUDA_TRY(cudaIpcGetMemHandle(&h, c->list_arena.p));
        memcpy(ipc_handle, &h, sizeof h);
    }
    if (arena_dev) *arena_dev = c->list_arena.p;
    return 0;
}

int eref_lists_set_peer(eref_ctx* c, int peer_part, const void* ipc_handle, void* arena_dev)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (peer_part < 0 || peer_part >= c->parts || peer_part == c->part)
        return eref_fail(EREF_ERR_ARG, "eref_lists_set_peer: part %d (this context is part %d of %d)", peer_part, c->part, c->parts);
    CUDA_TRY(cudaSetDevice(c->device));
    eref_ctx::ListPeer& P = c->list_peers[peer_part];
    if (P.ipc) cudaIpcCloseMemHandle(P.arena);
    P = eref_ctx::ListPeer();
    if (ipc_handle) {
        cudaIpcMemHandle_t h;
        memcpy(&h, ipc_handle, sizeof h);
        void* a = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&a, h, cudaIpcMemLazyEnablePeerAccess));
        P.arena = (uint8_t*)a; P.ipc = true;
        return 0;
    }
    if (!arena_dev) return eref_fail(EREF_ERR_ARG, "eref_lists_set_peer: null arena pointer");
    P.arena = (uint8_t*)arena_dev;             // (peer access was enabled by eref_table_set_peer)
    return 0;
}

int eref_query_set_share(eref_ctx* c, int on)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (on && (c->parts < 2 || !c->list_arena.p))
        return eref_fail(EREF_ERR_STATE, "shared reads need eref_table_set_part (parts > 1) and eref_query_reserve first");
    if (on && (c->flags & EREF_FLAG_DIRECT_COUNT)) return eref_fail(EREF_ERR_UNSUPPORTED, "the direct build cannot share reads");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = eref_sync(c);
    if (rc) return rc;
    c->share_reads = on != 0;
    c->n_lists = 0; c->table_dirty = true;
    return 0;
}

int eref_table_set_hook(eref_ctx* c, eref_table_hook fn, void* arg)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    c->hook = fn; c->hook_arg = arg;
    return 0;
}

int eref_table_part_range(eref_ctx* c, int part, size_t* table_off, size_t* table_bytes, size_t* summary_off, size_t* summary_bytes)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    if (part >= c->parts) return eref_fail(EREF_ERR_ARG, "part %d of %d", part, c->parts);
    const uint32_t s0 = part < 0 ? c->slice0 : part_first_slice(part, c->parts, c->part_w);
    const uint32_t ns = part < 0 ? c->nslices : part_first_slice(part + 1, c->parts, c->part_w) - s0;
    if (table_off) *table_off = (size_t)s0 * SLICE_WORDS * 4;
    if (table_bytes) *table_bytes = (size_t)ns * SLICE_WORDS * 4;
    if (summary_off) *summary_off = (size_t)s0 * (SLICE_WORDS / 16) * 4;
    if (summary_bytes) *summary_bytes = (size_t)ns * (SLICE_WORDS / 16) * 4;
    return 0;
}

int eref_table_join(eref_ctx* c)
{
    if (!c) return eref_fail(EREF_ERR_ARG, "null context");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = start_table_update(c);
    if (rc) return rc;
    return join_table(c, c->stream);
}

int eref_timings(eref_ctx* c, eref_phase_times* out)
{
    if (!c || !out) return eref_fail(EREF_ERR_ARG, "null argument");
    if (c->q_pending || c->ref_pending || c->scan_pending) {
        CUDA_TRY(cudaSetDevice(c->device));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        collect_times(c);
    }
    *out = c->t;
    out->kernel_launches = c->launches;
    return check_guard(c);
}

int eref_format_line(const eref_result* r, char* buf, size_t cap)
{
    if (!r || !buf) return 0;
    float ratio = (float)r->el / (float)r->ref_len;           // :615
    if (!(r->el > 0 && ratio > 0.75)) return 0;               // :616
    // :617 — ostream << float at default precision is printf %g
    return snprintf(buf, cap, "ref_index\t%u\t%u\t%u\t%u\t%g\n", r->ordinal, r->frag, r->el, r->ref_len, (double)ratio);
}

} // extern "C"
