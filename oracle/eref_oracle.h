/*
 * eref_oracle.h — CPU restatement of PALACE's `eref` algorithm (bin/extract_ref.cpp).
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (palace_b200/, include/) may include,
 * link or execute this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker / the timed CPU baseline.
 *
 * Parity pin: this restatement is checked (tests/test_oracle.py) against the UNMODIFIED
 * reference compiled from /root/reference/bin/extract_ref.cpp into oracle/_ref/eref_ref:
 *   - the body of the reference's own <db>.k32.index.dat (every code of every position),
 *   - the reference's stdout on seeded inputs (committed under tests/golden/).
 * The reference ships no tests or golden vectors of its own.
 *
 * All citations are to /root/reference/bin/extract_ref.cpp.
 */
#ifndef EREF_ORACLE_H
#define EREF_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define EO_K 32

typedef struct {
    uint32_t ordinal;   /* 1-based, counts only records with len > 32  (:887, :1175) */
    uint32_t frag;      /* frag_index                                    (:509)       */
    uint32_t el;        /* covered length                                (:611-614)   */
    uint32_t ref_len;   /* record length incl. invalid bytes             (:694)       */
} eo_result;

/* 32 x rand()%6 draws after srand(seed), permutation table of :1084  (random_coder :1082-1102) */
void eo_random_coder(unsigned seed, int16_t coder[96]);
/* low 16 bits of the first 96 u32 words of an existing index (saved_random_coder :1104-1122) */
int eo_read_coder(const char* index_path, int16_t coder[96]);

/* literal restatement of the inner z-loop (:971-994 / :717-735): returns 1 if the window is valid */
int eo_code3_naive(const int16_t coder[96], const unsigned char* s, uint32_t out[3]);
/* the same through bit-planes (SURVEY.md §3.3 identities); used for bulk work */
int eo_code3_fast(const int16_t coder[96], const unsigned char* s, uint32_t out[3]);

/* 2^32 saturating byte counters (:25-26). */
uint8_t* eo_table_new(void);
void eo_table_free(uint8_t* t);
/* read_fastq at one thread (:940-1004): every line with index%4==1 is a sequence line. */
void eo_count_fastq(uint8_t* table, const int16_t coder[96], const char* text, size_t n);
/* the same under down-sampling (:955-960): down_sam_ratio < 0 = count every read, no rand() draws */
void eo_count_fastq_sampled(uint8_t* table, const int16_t coder[96], const char* text, size_t n, int down_sam_ratio);
/* the count without draws on `threads` threads, each owning a range of the code space (same table, byte for byte) */
void eo_count_fastq_mt(uint8_t* table, const int16_t coder[96], const char* text, size_t n, int threads);
/* diagnostic: what the last 31 positions of every record count as (0 = canonical "no hit"; 3 = hit by all coders):
 * the reference reads them uninitialised (:856 vs :527-531) */
void eo_set_tail_hits(int v);
/* the prefix of a FASTQ text read_fastq processes for the byte range [0, end) (:942-945); main passes
 * end = file_size(fq1) for both files (:1265-1291), eo_run* apply it */
size_t eo_fastq_limit(const char* text, size_t n, long end);
/* cal_sam_ratio's sum (:1133-1142): bytes on lines with index%4==1 */
long eo_fastq_seq_bytes(const char* text, size_t n);

/* FASTA records the way read_ref frames them (:686-756, :757-805).  Returns a malloc'ed
 * concatenation of the kept (len > 32) records' raw sequence bytes; offsets has n_kept+1 entries. */
int eo_parse_fasta(const char* text, size_t n, unsigned char** seq_out, uint64_t** offsets_out,
                   size_t* n_kept, size_t* n_records);

/* codes of one record as the index stores them (:711-738): (len-31)*3 u32, 0 = invalid window */
void eo_record_codes(const int16_t coder[96], const unsigned char* seq, uint32_t len, uint32_t* codes);

/* read_index + slide_window (:846-896, :504-624) for one record; tail positions are "no hit" */
void eo_scan_record(const uint8_t* table, const int16_t coder[96], const unsigned char* seq,
                    uint32_t len, float hit_ratio, float perfect_hit_ratio, eo_result* out);

/* whole path: fq1, fq2, fasta -> results for EVERY kept record (ordinal order). threads>1 only
 * parallelises the per-record scan (results identical by construction).  malloc'ed *out. */
int eo_run(const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
           const int16_t coder[96], float hit_ratio, float perfect_hit_ratio, int threads,
           eo_result** out, size_t* n_out);

/* the same with down_sampling_size (:1230, 2e9 in the reference) as a parameter, as a WARM reference run does it:
 * srand(1) (:1236-1238), ratio = 100 * down_sampling_size / (2 * sequence bytes of fq1) (:1145), and below 100 one
 * rand() % 100 draw per sequence line of fq1 then fq2 (:955-960).  Returns -3 where the reference divides by 0. */
int eo_run_sampled(const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
                   const int16_t coder[96], float hit_ratio, float perfect_hit_ratio, int threads,
                   long down_sampling_size, eo_result** out, size_t* n_out);

/* the stdout line of :617 (or 0 bytes if the record is not selected); returns bytes written */
int eo_format_line(const eo_result* r, char* buf, size_t cap);

/* <db>.k32.index.dat in the reference's layout (:680-682, :710-737) */
int eo_write_index(const char* path, const int16_t coder[96], const char* fasta, size_t nf);

#ifdef __cplusplus
}
#endif
#endif
