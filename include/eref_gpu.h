/*
 * eref_gpu.h — C ABI of liberef_gpu.so: the B200 (sm_100a) implementation of PALACE's `eref`
 * k-mer reference-identification step.
 *
 * The reference has no library/FFI boundary for this path: `eref` is a process
 * (bin/extract_ref.cpp:1210-1334, called by palace:475-477).  The drop-in is therefore the `eref`
 * executable built from palace_b200/csrc/eref_main.cpp, and THIS header is the one thin layer
 * between that host code (or any other host: ctypes, cgo, JNI ...) and the CUDA kernels.
 * Each entry point names the reference code it replaces.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * error code (message via eref_last_error()); no C++ types or exceptions cross the boundary; the
 * caller owns host buffers, the context owns device memory; one context = one GPU; calls on one
 * context come from one thread at a time.  There is NO CPU fallback: without a CUDA device every
 * compute entry point fails with EREF_ERR_CUDA.
 */
#ifndef EREF_GPU_H
#define EREF_GPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define EREF_K 32                      /* bin/extract_ref.cpp:24 */
#define EREF_OK 0
#define EREF_ERR_ARG   (-1)
#define EREF_ERR_CUDA  (-2)
#define EREF_ERR_STATE (-3)
#define EREF_ERR_LIMIT (-4)            /* input beyond an implementation limit (e.g. FASTA shard >= 4 GiB) */
#define EREF_ERR_UNSUPPORTED (-5)      /* reference behaviour not implemented yet: fails loudly */
#define EREF_ERR_RETRY (-6)            /* eref_scan*: a query file outgrew buffers sized from its text length (its sequence
                                          lines are longer than its quality lines); eref_query_reset, add the files again
                                          — exact sizes are used from then on.  eref_run_host retries by itself. */

/* eref_create flags */
#define EREF_FLAG_NO_PRUNE 1u          /* probe all three streams at every position (no candidate filter) */
#define EREF_FLAG_DIRECT_COUNT 2u      /* count k-mers straight into the table (no code-range partitioning) */
#define EREF_FLAG_KEEP_RAND_STATE 4u   /* eref_run_host: do not srand(1) before the down-sampling draws (:1236-1238) */
#define EREF_FLAG_LEVEL_POP 8u         /* count the codes at each level during every table build (eref_query_popcount) */
#define EREF_FLAG_RESIDENT_TEXT 16u    /* texts handed to eref_ref_set_device are complete in device memory when the call is
                                          made (not merely ordered on the context's stream): their framing may then start at
                                          once, beside the query work still queued on the stream */

typedef struct eref_ctx eref_ctx;

/* One record per kept reference (len > 32), ordinal order — the numbers of the stdout line at :617 */
typedef struct {
    uint32_t ordinal;                  /* 1-based, only records with len > 32 count (:887, :1175) */
    uint32_t frag;                     /* frag_index (:509) */
    uint32_t el;                       /* covered length (:611-614) */
    uint32_t ref_len;                  /* record length (:694) */
} eref_result;

/* CUDA-event times of the last call of each phase, in milliseconds, and launch counts */
typedef struct {
    float query_h2d_ms, query_pack_ms, query_bucket_ms, query_count_ms;   /* pack/bucket: sums over the files added since the last reset */
    float ref_h2d_ms, ref_frame_ms, ref_pack_ms;                          /* last eref_ref_set*   */
    float scan_summary_ms, scan_sample_ms, scan_cand_ms, scan_probe_ms, scan_window_ms, scan_interval_ms,
          scan_d2h_ms;                                                    /* last eref_scan*      */
    uint64_t kernel_launches;                               /* since eref_create */
    uint64_t ref_bases, ref_words, ref_need_words;          /* last scan */
    uint64_t query_seq_bytes;                               /* last eref_query_add* (cal_sam_ratio's sum) */
    uint64_t query_codes;                                   /* last eref_query_add*: codes counted (3 per valid 32-mer) */
} eref_phase_times;

const char* eref_last_error(void);
int  eref_device_count(void);

/* replaces generate_coder/generate_base/generate_complement + saved_random_coder state
 * (:1216-1218, :1258) and the argv ratios (:1228-1229).  coder = 96 entries in {0,1,2}:
 * coder[z*3+i] is the 1-bit alphabet used at window position z by code stream i (:972). */
int  eref_create(eref_ctx** out, int device, const int16_t coder[96], float hit_ratio,
                 float perfect_hit_ratio, unsigned flags);
void eref_destroy(eref_ctx* ctx);
/* run all work of this context on an existing CUDA stream (cudaStream_t); NULL = context's own */
int  eref_set_stream(eref_ctx* ctx, void* cuda_stream);

/* memset(kmer_count_table) at :1257 */
int  eref_query_reset(eref_ctx* ctx);
/* read_fastq at one thread (:905-1008) over the text it is given: frames lines (index%4==1),
 * forms the three canonical codes of every valid 32-mer and counts them, saturating at 3.
 * (main gives read_fastq a byte budget, :1265-1291: for the second file pass eref_fastq_limit's size, below.) */
int  eref_query_add(eref_ctx* ctx, const char* fastq_host, size_t n);
int  eref_query_add_device(eref_ctx* ctx, const void* fastq_dev, size_t n);
/* Down-sampling (:955-960, :1124-1148, :1239-1242).  When fq1 holds more than
 * down_sampling_size / 2 bases the reference keeps a read only if rand() % 100 < ratio, one draw per sequence line,
 * fq1 first then fq2, rand seeded with 1 by main().  The host replays it:
 *   eref_query_census*   cal_sam_ratio's pass: bytes on sequence lines (index % 4 == 1) and the number of such lines
 *   eref_sample_ratio    100 * down_sampling_size / (2 * bytes), as the reference's long division (2e9 there)
 *   eref_draw_keep       n_lines draws of libc rand() % 100 < ratio -> one bit per sequence line (LSB first)
 *   eref_query_add_masked*  read_fastq counting only the sequence lines whose bit is set (NULL = all)
 * eref_run_host does all of this itself. */
int  eref_query_census(eref_ctx* ctx, const char* fastq_host, size_t n, uint64_t* seq_bytes, uint64_t* seq_lines);
int  eref_query_census_device(eref_ctx* ctx, const void* fastq_dev, size_t n, uint64_t* seq_bytes, uint64_t* seq_lines);
int  eref_sample_ratio(uint64_t seq_bytes_fq1, uint64_t down_sampling_size, int* ratio);
void eref_draw_keep(int ratio, uint64_t n_lines, uint32_t* keep_bits);
int  eref_query_add_masked(eref_ctx* ctx, const char* fastq_host, size_t n, const uint32_t* keep_bits, uint64_t n_keep);
int  eref_query_add_masked_device(eref_ctx* ctx, const void* fastq_dev, size_t n, const uint32_t* keep_bits, uint64_t n_keep);
/* read_fastq's byte budget (:942-945 `if (add_size >= end) break; add_size += reads_seq.length();`): the prefix of a
 * FASTQ text the reference reads when main hands it the range [0, end) — every line in front of which fewer than `end`
 * characters (newlines not counted) were read.  main passes end = file_size(fq1) for fq1 AND fq2 (:1265-1291), so fq2
 * loses its tail when it is the longer file by more than its number of lines.  Returns the prefix length (n when
 * nothing is cut).  Host arithmetic, no GPU.  eref_run_host applies it to a host fq2 itself; callers of the phase
 * functions (eref_query_add* = read_fastq over the whole text) and callers that hand eref_run_host a DEVICE fq2 pass
 * eref_fastq_limit(fq2, n2, n1) as fq2's size. */
size_t eref_fastq_limit(const char* fastq_host, size_t n, uint64_t end);
/* cal_sam_ratio's sum (:1133-1142) for the file of the last eref_query_add* call */
int  eref_query_seq_bytes(eref_ctx* ctx, uint64_t* out);

/* read_ref's framing (:686-805) over one FASTA text (whole DB or one shard starting at a '>' line):
 * records, lengths, drop len <= 32, ordinals; packs the kept records into HBM. */
int  eref_ref_set(eref_ctx* ctx, const char* fasta_host, size_t n);
int  eref_ref_set_device(eref_ctx* ctx, const void* fasta_dev, size_t n);
int  eref_ref_count(eref_ctx* ctx, size_t* n_kept, size_t* n_records, uint64_t* kept_bases);

/* read_index + slide_window (:813-903, :504-624) for every kept record of the current reference
 * against the current table.  Writes min(cap, n_kept) records; *n_out = n_kept. */
int  eref_scan(eref_ctx* ctx, eref_result* out, size_t cap, size_t* n_out);
/* same, results left on the device as n_kept x 4 uint32 (for an NCCL allgather by the caller) */
int  eref_scan_device(eref_ctx* ctx, const uint32_t** results_dev, size_t* n_out);
/* same without waiting: the records are valid in stream order on the context's stream (eref_set_stream), e.g. for
 * a collective the caller queues on that stream */
int  eref_scan_device_async(eref_ctx* ctx, const uint32_t** results_dev, size_t* n_out);

/* main() of the reference from the FASTQ passes on (:1242-1329) in one call, from HOST buffers (pinned for
 * overlap): H2D copies run on a copy stream while the kernels work; the FASTA text is cut into `chunks`
 * pieces at record boundaries and each piece is framed and scanned as soon as it has landed.  Writes
 * min(cap, total) records in file order with global ordinals; *n_out = total.  Includes the table reset
 * the cal_sam_ratio pass and, when it yields a ratio below 100, the rand()-replayed down-sampling
 * (EREF_ERR_UNSUPPORTED when the reference would divide by 0: fq1 without sequence lines).
 * fasta == NULL (nf == 0): the sample is screened against the reference already set with eref_ref_set* — one
 * database kept in HBM, many samples — and only the two FASTQ uploads are pipelined. */
int  eref_run_host(eref_ctx* ctx, const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta,
                   size_t nf, int chunks, eref_result* out, size_t cap, size_t* n_out);

/* the records of the last eref_run_host, still in device memory (n x 4 uint32, file-order ordinals): what a multi-GPU
 * caller hands to its all-gather.  out/cap of eref_run_host may be NULL/0 when only these are wanted. */
int  eref_run_results_device(eref_ctx* ctx, const uint32_t** results_dev, size_t* n_out);
/* Down-sampling masks for the NEXT eref_run_host (one bit per sequence line, LSB first, as eref_draw_keep writes them):
 * when several GPUs screen the same sample the rand() draws of :955-957 must happen once, not once per GPU.  Used only
 * if that call's ratio is below 100. */
int  eref_run_host_keep(eref_ctx* ctx, const uint32_t* keep1, uint64_t lines1, const uint32_t* keep2, uint64_t lines2);

/* the per-record body of <db>.k32.index.dat (:711-738): (len-31)*3 codes, 0 = invalid window.
 * ordinal is 1-based within the current reference text.  Test hook and index writer (f2). */
int  eref_get_codes(eref_ctx* ctx, uint32_t ordinal, uint32_t* codes_host, size_t cap, size_t* n_out);
/* name offset/length of a kept record's header line inside the text given to eref_ref_set* */
int  eref_ref_record(eref_ctx* ctx, uint32_t ordinal, uint64_t* hdr_off, uint32_t* hdr_len, uint32_t* len);

/* every record of the current reference text in file order (kept or not; record 0 = whatever precedes the first
 * header): its length (:694) and the offset of its header line (0xFFFFFFFF for record 0).  Writes min(cap, n)
 * entries; *n_out = n.  With these a host writes <db>.genome.len.txt (:697, :800). */
int  eref_ref_records(eref_ctx* ctx, uint32_t* len_out, uint32_t* hdr_off_out, size_t cap, size_t* n_out);
/* the body of <db>.k32.index.dat (:698, :711-738) for as many whole kept records as fit into cap_words,
 * starting at ordinal first (1-based): per record one word with its length, then 3 codes per window start
 * (0 = invalid window).  *n_words = words written, *next = ordinal to continue with (n_kept + 1 at the end).
 * A host streams the chunks to disk behind the 400-byte coder header to share a DB directory with the CPU
 * eref. */
int  eref_index_body(eref_ctx* ctx, uint32_t first, uint32_t* out_host, size_t cap_words, size_t* n_words, uint32_t* next);

/* Sharded table build.  The reference shares ONE table among its scan threads (:1257-1291 build it, :1309-1329 read
 * it); with the database sharded over GPUs every GPU needs the whole table, but the saturating count of a code does
 * not depend on any other code (:995-997), so GPU `part` of `parts` counts only the table slices
 * of its part (a contiguous range; the cut points balance code mass against the per-slice cost, eref_table_part_range)
 * and the finished level-3 slices are exchanged instead of every GPU counting everything.  Two ways to complete the tables:
 *   peers  eref_table_export on every context, eref_table_set_peer for every other part: the counting kernel stores
 *          each finished slice into every GPU's table through peer pointers (NVLink) and raises a flag there; the
 *          scan waits for all flags on the device.  ipc_handles (EREF_TABLE_HANDLE_BYTES, from the peer's
 *          eref_table_export) for a context in another process, or its raw device pointers + device id in the same one.
 *   hook   eref_table_set_hook: called on the context's stream once the local part is complete and before the first
 *          probe; the callee completes the table there (e.g. one NCCL broadcast per part over eref_table_part_range).
 * Every GPU of a build must make the same sequence of calls, and a collective or barrier ordered on the streams must
 * separate a scan from the NEXT table build (the all-gather of the result records is one): a builder writes into
 * its peers' tables.  eref_query_popcount then reports this part's codes only. */
#define EREF_TABLE_HANDLE_BYTES 192
typedef int (*eref_table_hook)(void* arg, void* cuda_stream);
int  eref_table_export(eref_ctx* ctx, void* ipc_handles, void** bm3_dev, void** summary_dev, void** flags_dev);
int  eref_table_set_part(eref_ctx* ctx, int part, int parts);
int  eref_table_set_peer(eref_ctx* ctx, int peer_part, int peer_device, const void* ipc_handles, void* bm3_dev,
                         void* summary_dev, void* flags_dev);
int  eref_table_set_hook(eref_ctx* ctx, eref_table_hook fn, void* arg);
/* byte ranges of part `part` (-1: this context's own) inside the level-3 table and its summary.  The parts are NOT of
 * equal size and are re-cut for every query (at the first file added after a reset, from its size): canonical codes
 * thin out towards the top of the code space, so the cut points balance code mass against the per-slice cost of the
 * exchange.  A completion hook must therefore ask for the ranges when it runs, not when it is installed. */
int  eref_table_part_range(eref_ctx* ctx, int part, size_t* table_off, size_t* table_bytes, size_t* summary_off, size_t* summary_bytes);
/* Shared reads — the whole query side divided by N.  With the table build shared out every GPU still packs the whole
 * query and forms all its codes, keeping those of its part.  With shared reads the CALLER gives every GPU 1/N of each
 * query file (any cut at a 4-line record boundary is a valid FASTQ text; every GPU adds the same number of files in the
 * same order); a GPU partitions the codes of its share over ALL table slices into lists that live in one exported arena,
 * and a build first waits (device flags) until every GPU's lists are complete, pulls the codes of ITS table slices out
 * of each peer's lists — one contiguous range per list, over NVLink — and counts them with its own.  Everything else is
 * the sharded build above (peer stores of the finished slices, flags, the barrier contract).
 *   eref_query_reserve   fixed-address arena for max_files lists of files up to max_file_bytes (same arguments on every GPU)
 *   eref_lists_export / eref_lists_set_peer   like eref_table_export / eref_table_set_peer (which must come first)
 *   eref_query_set_share on/off; while on, eref_query_add* take this GPU's share, the byte count of cal_sam_ratio is this
 *                        share's only (down-sampling decisions are the caller's), and eref_run_host is refused. */
#define EREF_LISTS_HANDLE_BYTES 64
int  eref_query_reserve(eref_ctx* ctx, size_t max_file_bytes, int max_files);
int  eref_lists_export(eref_ctx* ctx, void* ipc_handle, void** arena_dev);
int  eref_lists_set_peer(eref_ctx* ctx, int peer_part, const void* ipc_handle, void* arena_dev);
int  eref_query_set_share(eref_ctx* ctx, int on);
/* queue the build of the files added so far and make the context's stream wait for the complete table */
int  eref_table_join(eref_ctx* ctx);
/* on: the table build and the framing census no longer run beside the main stream's kernels, so the phase times of
 * eref_timings are per-kernel durations (profiling); off (default): they overlap and a step is shorter than their sum */
int  eref_set_serial(eref_ctx* ctx, int on);
/* wait for everything queued on the context */
int  eref_sync(eref_ctx* ctx);

/* table inspection (tests): number of codes with count >= level (1..3) */
int  eref_query_popcount(eref_ctx* ctx, int level, uint64_t* out);

/* Phase times are read from CUDA events lazily: the call waits for the context's stream when events of earlier
 * calls are still outstanding. */
int  eref_timings(eref_ctx* ctx, eref_phase_times* out);

/* the stdout line of :615-617; returns bytes written, 0 if the record is not selected */
int  eref_format_line(const eref_result* r, char* buf, size_t cap);

/* eref_query_census_device without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Needs no context: it works on the CURRENT device, queues everything on cuda_stream (0 = default stream)
 * and returns without waiting.  fastq_dev: n bytes of FASTQ text in device memory, 16-byte aligned;
 * out_dev: two uint64 in device memory, [0] = bytes on sequence lines, [1] = sequence lines — the two numbers
 * eref_query_census[_device] return.  The per-tile scratch of the pass lives in one file-scope __device__ array of
 * 512 KiB (csrc/noalloc.cu) that every GPU using the library holds until the process exits, which bounds the text:
 * n > EREF_NOALLOC_MAX_BYTES is refused with EREF_ERR_LIMIT.  Every call shares that array: calls ordered on one stream
 * are safe, calls that overlap on the GPU (different streams or threads) race and give wrong counts.
 * Calls no cudaMalloc / cudaFree / allocating library routine. */
#define EREF_NOALLOC_MAX_BYTES (256ull << 20)
int  eref_census_noalloc(const void* fastq_dev, size_t n, uint64_t* out_dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
