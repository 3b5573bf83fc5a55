/*
 * eref_oracle.c — CPU restatement of PALACE `eref` (see eref_oracle.h for the rules).
 * TEST INFRASTRUCTURE ONLY: the checker and the timed CPU baseline, never the product.
 * Citations: /root/reference/bin/extract_ref.cpp.
 */
#define _GNU_SOURCE
#include "eref_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#include <math.h>
#endif

/* ---- coders ------------------------------------------------------------------------------ */

/* generate_coder (:1010-1054): three 1-bit alphabets; anything outside ACGTacgt is invalid. */
static int base_bit(int which, unsigned char x)
{
    switch (x) {
    case 'A': case 'a': return 1;
    case 'T': case 't': return which == 0;
    case 'C': case 'c': return which == 1;
    case 'G': case 'g': return which == 2;
    default: return -1;
    }
}

/* generate_complement (:1065-1080): upper-case complement, 0 for anything else. */
static unsigned char comp_base(unsigned char x)
{
    switch (x) {
    case 'A': case 'a': return 'T';
    case 'T': case 't': return 'A';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    default: return 0;
    }
}

void eo_random_coder(unsigned seed, int16_t coder[96])
{
    /* :1084 — the six orderings of {0,1,2}, in the reference's order */
    static const int16_t perm6[6][3] = {{0,1,2},{0,2,1},{1,2,0},{1,0,2},{2,0,1},{2,1,0}};
    srand(seed);                                   /* :1088-1089 (seed = time(0) there) */
    for (int z = 0; z < EO_K; z++) {
        int r = rand() % 6;                        /* :1091 */
        for (int i = 0; i < 3; i++) coder[z * 3 + i] = perm6[r][i];
    }
}

int eo_read_coder(const char* index_path, int16_t coder[96])
{
    FILE* f = fopen(index_path, "rb");
    if (!f) return -1;
    uint32_t w[96];
    size_t got = fread(w, 4, 96, f);
    fclose(f);
    if (got != 96) return -2;
    for (int j = 0; j < 96; j++) coder[j] = (int16_t)(w[j] & 0xffffu);   /* :1113 short = u32 */
    return 0;
}

int eo_code3_naive(const int16_t coder[96], const unsigned char* s, uint32_t out[3])
{
    int ok = 1;
    for (int i = 0; i < 3; i++) {
        uint32_t fwd = 0, rc = 0;
        int valid = 1;
        for (int z = 0; z < EO_K; z++) {
            int m = base_bit(coder[z * 3 + i], s[z]);
            if (m < 0) { valid = 0; break; }                            /* :978-982 */
            int n = base_bit(coder[(EO_K - 1 - z) * 3 + i], comp_base(s[z]));
            fwd += (uint32_t)m << (EO_K - 1 - z);                       /* base[z] = 2^(31-z), :1060 */
            rc  += (uint32_t)n << z;                                    /* base[31-z] = 2^z          */
        }
        out[i] = valid ? (fwd > rc ? rc : fwd) : 0;                      /* :989-994, :733-735 */
        ok &= valid;
    }
    return ok;
}

/* bit-plane form: 2-bit alphabet A=00 C=01 G=10 T=11 (hi,lo); then
 * coder0 = ~(hi^lo), coder1 = ~hi, coder2 = ~lo; complement flips coder1/coder2 only. */
typedef struct { uint32_t m[3][3]; } eo_masks;

static uint32_t brev32(uint32_t x)
{
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
    return (x >> 16) | (x << 16);
}

static void make_masks(const int16_t coder[96], eo_masks* mk)
{
    memset(mk, 0, sizeof *mk);
    for (int z = 0; z < EO_K; z++)
        for (int i = 0; i < 3; i++) {
            int c = coder[z * 3 + i];
            if (c >= 0 && c < 3) mk->m[i][c] |= 1u << (EO_K - 1 - z);
        }
}

static inline void codes_from_windows(const eo_masks* mk, uint32_t hi, uint32_t lo, uint32_t out[3])
{
    uint32_t b0 = ~(hi ^ lo), b1 = ~hi, b2 = ~lo;
    uint32_t r0 = brev32(b0), r1 = ~brev32(b1), r2 = ~brev32(b2);
    for (int i = 0; i < 3; i++) {
        uint32_t fwd = (b0 & mk->m[i][0]) | (b1 & mk->m[i][1]) | (b2 & mk->m[i][2]);
        uint32_t rc  = (r0 & mk->m[i][0]) | (r1 & mk->m[i][1]) | (r2 & mk->m[i][2]);
        out[i] = fwd > rc ? rc : fwd;
    }
}

static inline int base2(unsigned char x)
{
    switch (x) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return -1;
    }
}

int eo_code3_fast(const int16_t coder[96], const unsigned char* s, uint32_t out[3])
{
    eo_masks mk; make_masks(coder, &mk);
    uint32_t hi = 0, lo = 0;
    for (int z = 0; z < EO_K; z++) {
        int b = base2(s[z]);
        if (b < 0) { out[0] = out[1] = out[2] = 0; return 0; }
        hi = (hi << 1) | (uint32_t)(b >> 1);
        lo = (lo << 1) | (uint32_t)(b & 1);
    }
    codes_from_windows(&mk, hi, lo, out);
    return 1;
}

/* rolling walk over one sequence; calls fn(j, codes, valid) for every window start j */
#define ROLL_BEGIN(seq, len)                                                        \
    uint32_t hi_ = 0, lo_ = 0; int run_ = 0;                                        \
    for (uint32_t p_ = 0; p_ < (len); p_++) {                                       \
        int b_ = base2((seq)[p_]);                                                  \
        if (b_ < 0) { run_ = 0; b_ = 0; } else run_++;                              \
        hi_ = (hi_ << 1) | (uint32_t)(b_ >> 1); lo_ = (lo_ << 1) | (uint32_t)(b_ & 1); \
        if (p_ + 1 < EO_K) continue;                                                \
        uint32_t j_ = p_ + 1 - EO_K; int valid_ = run_ >= EO_K;
#define ROLL_END }

/* ---- query table -------------------------------------------------------------------------- */

uint8_t* eo_table_new(void) { return (uint8_t*)calloc((size_t)1 << 32, 1); }
void eo_table_free(uint8_t* t) { free(t); }

static void count_read(uint8_t* table, const eo_masks* mk, const unsigned char* s, uint32_t len)
{
    uint32_t c[3];
    ROLL_BEGIN(s, len)
        (void)j_;
        if (valid_) {
            codes_from_windows(mk, hi_, lo_, c);
            for (int i = 0; i < 3; i++)
                if (table[c[i]] < 3) table[c[i]]++;                     /* :995-997 */
        }
    ROLL_END
}

void eo_count_fastq(uint8_t* table, const int16_t coder[96], const char* text, size_t n)
{
    eo_count_fastq_sampled(table, coder, text, n, -1);
}

/* down_sam_ratio < 0: no draws at all (every read counted; the reference's result whenever ratio >= 100).
 * Otherwise one r = rand() % 100 per sequence line and the read is counted iff r < down_sam_ratio (:955-960). */
void eo_count_fastq_sampled(uint8_t* table, const int16_t coder[96], const char* text, size_t n, int down_sam_ratio)
{
    eo_masks mk; make_masks(coder, &mk);
    size_t pos = 0, line = 0;
    while (pos < n) {                                   /* getline loop, :940 */
        const char* nl = memchr(text + pos, '\n', n - pos);
        size_t end = nl ? (size_t)(nl - text) : n;
        if (line % 4 == 1 && (down_sam_ratio < 0 || rand() % 100 < down_sam_ratio))     /* :947, :955-957 */
            count_read(table, &mk, (const unsigned char*)text + pos, (uint32_t)(end - pos));
        line++;
        pos = end + 1;
    }
}

/* The same count with T threads, for the case without draws: the saturating count of a slot depends only on how
 * often ITS code occurs (:995-997), so thread t walks every sequence line and applies exactly the codes whose top
 * byte falls in its range — every table byte has one writer and ends up as in the serial loop. */
static void count_read_range(uint8_t* table, const eo_masks* mk, const unsigned char* s, uint32_t len, uint32_t lo8, uint32_t hi8)
{
    uint32_t c[3];
    ROLL_BEGIN(s, len)
        (void)j_;
        if (valid_) {
            codes_from_windows(mk, hi_, lo_, c);
            for (int i = 0; i < 3; i++) {
                const uint32_t top = c[i] >> 24;
                if (top >= lo8 && top < hi8 && table[c[i]] < 3) table[c[i]]++;
            }
        }
    ROLL_END
}

void eo_count_fastq_mt(uint8_t* table, const int16_t coder[96], const char* text, size_t n, int threads)
{
    if (threads <= 1) { eo_count_fastq_sampled(table, coder, text, n, -1); return; }
    if (threads > 64) threads = 64;
    eo_masks mk; make_masks(coder, &mk);
    /* canonical codes are min(fwd, rc): their density falls linearly with the code, so equal-width ranges would
     * leave the low threads with most of the work; cut at equal areas under 2(1-x) instead */
    uint32_t cut[65];
    for (int t = 0; t <= threads; t++) {
        double x = 1.0 - sqrt(1.0 - (double)t / threads);
        cut[t] = t == threads ? 256u : (uint32_t)(x * 256.0);
    }
    #pragma omp parallel num_threads(threads)
    {
        const int t = omp_get_thread_num();
        const uint32_t lo8 = cut[t], hi8 = cut[t + 1];
        size_t pos = 0, line = 0;
        while (pos < n && hi8 > lo8) {
            const char* nl = memchr(text + pos, '\n', n - pos);
            size_t end = nl ? (size_t)(nl - text) : n;
            if (line % 4 == 1) count_read_range(table, &mk, (const unsigned char*)text + pos, (uint32_t)(end - pos), lo8, hi8);
            line++;
            pos = end + 1;
        }
    }
}

/* The part of a FASTQ text that read_fastq processes when it is handed the byte range [0, end) (:940-946): the
 * getline loop stops in front of the first line whose PRECEDING lines add up to >= end characters, newlines not
 * counted (`if (add_size >= end) break; add_size += reads_seq.length();`).  main computes `size = file_size(fq1)` once
 * (:1265) and passes it as `end` for fq1 AND for fq2 (:1269-1291), so at one thread fq1 is always read whole and fq2
 * loses its tail whenever it is the longer file by more than its number of lines. */
size_t eo_fastq_limit(const char* text, size_t n, long end)
{
    size_t pos = 0; long add_size = 0;
    while (pos < n) {
        if (add_size >= end) return pos;                        /* :942-944 */
        const char* nl = memchr(text + pos, '\n', n - pos);
        size_t e = nl ? (size_t)(nl - text) : n;
        add_size += (long)(e - pos);                            /* :945 */
        pos = e + 1;
    }
    return n;
}

long eo_fastq_seq_bytes(const char* text, size_t n)
{
    size_t pos = 0, line = 0; long sum = 0;
    while (pos < n) {
        const char* nl = memchr(text + pos, '\n', n - pos);
        size_t end = nl ? (size_t)(nl - text) : n;
        if (line % 4 == 1) sum += (long)(end - pos);
        line++;
        pos = end + 1;
    }
    return sum;
}

/* ---- FASTA framing -------------------------------------------------------------------------- */

int eo_parse_fasta(const char* text, size_t n, unsigned char** seq_out, uint64_t** offsets_out,
                   size_t* n_kept, size_t* n_records)
{
    unsigned char* seq = (unsigned char*)malloc(n + 1);
    size_t cap = 1024, kept = 0, recs = 0;
    uint64_t* off = (uint64_t*)malloc((cap + 1) * sizeof *off);
    if (!seq || !off) return -1;
    size_t w = 0, rec_start = 0, pos = 0;
    off[0] = 0;
    /* A record ends at the next header line or at EOF (:687-752, :757-805).  The text before the
     * first header is a record too (normally empty); only records with len > 32 are kept (:697). */
    for (;;) {
        int at_end = pos >= n;
        int is_hdr = !at_end && text[pos] == '>';
        if (at_end || is_hdr) {
            recs++;
            if (w - rec_start > EO_K) {
                if (kept == cap) { cap *= 2; off = (uint64_t*)realloc(off, (cap + 1) * sizeof *off); }
                off[++kept] = w;
                rec_start = w;
            } else {
                w = rec_start;                      /* dropped: does not consume an ordinal */
            }
            if (at_end) break;
        }
        const char* nl = memchr(text + pos, '\n', n - pos);
        size_t end = nl ? (size_t)(nl - text) : n;
        if (!is_hdr) { memcpy(seq + w, text + pos, end - pos); w += end - pos; }   /* :754 */
        pos = end + 1;
    }
    *seq_out = seq; *offsets_out = off; *n_kept = kept;
    if (n_records) *n_records = recs;
    return 0;
}

void eo_record_codes(const int16_t coder[96], const unsigned char* seq, uint32_t len, uint32_t* codes)
{
    eo_masks mk; make_masks(coder, &mk);
    ROLL_BEGIN(seq, len)
        if (valid_) codes_from_windows(&mk, hi_, lo_, codes + (size_t)j_ * 3);
        else codes[(size_t)j_ * 3] = codes[(size_t)j_ * 3 + 1] = codes[(size_t)j_ * 3 + 2] = 0;
    ROLL_END
}

/* ---- probe + sliding window ------------------------------------------------------------------ */

static void window_rule(const uint8_t* nhit /* per position: number of coders that hit, 0..3 */,
                        uint32_t len, float hit_ratio, float perfect_hit_ratio, eo_result* out)
{
    const int window = 500;
    const int one_min = (int)(window * hit_ratio);              /* :513, float multiply */
    const int three_min = (int)(window * perfect_hit_ratio);    /* :514 */
    int one = 0, three = 0, open = 0, good = 0;
    int start = 0, last_end = 0, frag = 0;
    long el = 0;
    int cur_start = 0;                                           /* start of the stored interval */
    for (int j = 0; j < (int)len; j++) {
        one += nhit[j] > 0; three += nhit[j] == 3;
        if (j >= window) { one -= nhit[j - window] > 0; three -= nhit[j - window] == 3; }  /* :548-559 */
        good = one >= one_min && three >= three_min;            /* :561 */
        if (!open && good) {                                     /* :568-574 */
            start = j - 2 * window; if (start < 1) start = 1;
            open = 1;
        } else if (open && !good) {                              /* :575-589 */
            int end = j + 2 * window; if (end > (int)len) end = (int)len;
            if (frag > 0 && start - last_end < window) { el += end - last_end; }
            else { frag++; cur_start = start; el += end - start; }
            last_end = end; open = 0;
        }
    }
    if (open && good) {                                          /* :599-609 */
        int end = (int)len;
        if (frag > 0 && start - last_end < window) { el += end - last_end; }
        else { frag++; cur_start = start; el += end - start; }
        last_end = end;
    }
    (void)cur_start;
    out->frag = (uint32_t)frag; out->el = (uint32_t)el; out->ref_len = len;
}

/* The reference never writes the hit flags of a record's last 31 positions (:846-870 fills (ref_len-31)*3 of the
 * ref_len*3 bytes it allocates, :856) but slide_window reads them (:527-531): whatever the heap held counts.  The
 * canonical value here is "no hit" (0).  eo_set_tail_hits(3) makes them all count as hit by three coders instead — the
 * other end of what the reference can print — so that a fuzzer can tell a difference that this undefined read explains
 * (the reference's lines lie between the two) from a real one. */
static int g_tail_hits = 0;
void eo_set_tail_hits(int v) { g_tail_hits = v; }

void eo_scan_record(const uint8_t* table, const int16_t coder[96], const unsigned char* seq,
                    uint32_t len, float hit_ratio, float perfect_hit_ratio, eo_result* out)
{
    eo_masks mk; make_masks(coder, &mk);
    uint8_t* nhit = (uint8_t*)calloc(len, 1);                    /* tail positions stay 0 */
    if (g_tail_hits) for (uint32_t j = len >= 31 ? len - 31 : 0; j < len; j++) nhit[j] = (uint8_t)g_tail_hits;
    uint32_t c[3];
    ROLL_BEGIN(seq, len)
        if (valid_) {
            codes_from_windows(&mk, hi_, lo_, c);
            int h = 0;
            for (int i = 0; i < 3; i++)
                h += (c[i] != 0 && table[c[i]] == 3);            /* :861-866, :531 */
            nhit[j_] = (uint8_t)h;
        }
    ROLL_END
    window_rule(nhit, len, hit_ratio, perfect_hit_ratio, out);
    free(nhit);
}

int eo_run(const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
           const int16_t coder[96], float hit_ratio, float perfect_hit_ratio, int threads,
           eo_result** out, size_t* n_out)
{
    return eo_run_sampled(fq1, n1, fq2, n2, fasta, nf, coder, hit_ratio, perfect_hit_ratio, threads, 2000000000L, out, n_out);
}

/* main() of a WARM reference run (index on disk) with down_sampling_size (:1230) as a parameter: srand(1) (:1236-1238),
 * cal_sam_ratio (:1124-1148), then both FASTQ files under that ratio. */
int eo_run_sampled(const char* fq1, size_t n1, const char* fq2, size_t n2, const char* fasta, size_t nf,
                   const int16_t coder[96], float hit_ratio, float perfect_hit_ratio, int threads,
                   long down_sampling_size, eo_result** out, size_t* n_out)
{
    uint8_t* table = eo_table_new();
    if (!table) return -1;
    long sample_size = eo_fastq_seq_bytes(fq1, n1) * 2;          /* :1144 */
    if (sample_size == 0) { eo_table_free(table); return -3; }   /* the reference divides by zero here (:1145) */
    int ratio = (int)(100 * down_sampling_size / sample_size);   /* :1145 */
    srand(1);                                                    /* :1236-1238 */
    const long size = (long)n1;                                  /* :1265 file_size(fq1), the `end` of BOTH files */
    n1 = eo_fastq_limit(fq1, n1, size);                          /* (never shortens fq1) */
    n2 = eo_fastq_limit(fq2, n2, size);                          /* :1281-1288: fq2 read up to fq1's size */
    if (ratio < 100 || threads <= 1) {
        eo_count_fastq_sampled(table, coder, fq1, n1, ratio < 100 ? ratio : -1);   /* :1269-1279 */
        eo_count_fastq_sampled(table, coder, fq2, n2, ratio < 100 ? ratio : -1);   /* :1281-1291 */
    } else {                                                     /* no draws: the count parallelises by code range */
        eo_count_fastq_mt(table, coder, fq1, n1, threads);
        eo_count_fastq_mt(table, coder, fq2, n2, threads);
    }
    unsigned char* seq; uint64_t* off; size_t kept;
    if (eo_parse_fasta(fasta, nf, &seq, &off, &kept, NULL)) { eo_table_free(table); return -2; }
    eo_result* res = (eo_result*)calloc(kept ? kept : 1, sizeof *res);
    if (threads < 1) threads = 1;
    #pragma omp parallel for schedule(dynamic, 4) num_threads(threads)
    for (long g = 0; g < (long)kept; g++) {
        eo_scan_record(table, coder, seq + off[g], (uint32_t)(off[g + 1] - off[g]),
                       hit_ratio, perfect_hit_ratio, &res[g]);
        res[g].ordinal = (uint32_t)g + 1;                        /* :1175, :887 */
    }
    free(seq); free(off); eo_table_free(table);
    *out = res; *n_out = kept;
    return 0;
}

int eo_format_line(const eo_result* r, char* buf, size_t cap)
{
    float ratio = (float)r->el / (float)r->ref_len;              /* :615 */
    if (!(r->el > 0 && ratio > 0.75)) return 0;                  /* :616 (float vs double 0.75) */
    /* ostream << float at default precision == %g */
    return snprintf(buf, cap, "ref_index\t%u\t%u\t%u\t%u\t%g\n", r->ordinal, r->frag, r->el,
                    r->ref_len, (double)ratio);
}

int eo_write_index(const char* path, const int16_t coder[96], const char* fasta, size_t nf)
{
    unsigned char* seq; uint64_t* off; size_t kept;
    if (eo_parse_fasta(fasta, nf, &seq, &off, &kept, NULL)) return -2;
    FILE* f = fopen(path, "wb");
    if (!f) { free(seq); free(off); return -1; }
    /* :680-682 writes 4 bytes starting at each short: low half = entry j, high half = entry j+1 */
    int16_t cc[101]; memset(cc, 0, sizeof cc); memcpy(cc, coder, 96 * sizeof(int16_t));
    for (int j = 0; j < 100; j++) {
        uint32_t w = (uint16_t)cc[j] | ((uint32_t)(uint16_t)cc[j + 1] << 16);
        fwrite(&w, 4, 1, f);
    }
    for (size_t g = 0; g < kept; g++) {
        uint32_t len = (uint32_t)(off[g + 1] - off[g]);
        size_t nc = (size_t)(len - EO_K + 1) * 3;
        uint32_t* codes = (uint32_t*)malloc(nc * 4);
        eo_record_codes(coder, seq + off[g], len, codes);
        fwrite(&len, 4, 1, f);                                   /* :710 */
        fwrite(codes, 4, nc, f);                                 /* :736 */
        free(codes);
    }
    fclose(f); free(seq); free(off);
    return 0;
}
