// eref_main.cpp — the drop-in `eref` executable for the PALACE pipeline (palace:475-477):
//
//   eref <fq1> <fq2> <phagedb.fasta> <interval_out> <hit_ratio> <perfect_hit_ratio> <threads>  > ref_names.txt
//
// Same seven positional arguments, same stdout lines and side-effect files as
// bin/extract_ref.cpp:1210-1334; all sequence work runs on B200s through the C ABI of
// liberef_gpu.so (include/eref_gpu.h).  There is no CPU path: without a GPU it exits non-zero.
//
// How a run goes: the three input files are read (and, if gzip/BGZF, inflated) by threads of their own WHILE the
// CUDA contexts are created; every GPU then runs the pipelined entry point eref_run_host on its shard of the
// database (uploads overlapped with kernels, pieces of <= 2 GiB).  With more than one GPU each uploads 1/N of the
// query text and an NCCL all-gather replicates it over NVLink, the table build is sharded (every GPU counts 1/N of
// the table slices and stores them into all tables through peer pointers), and ONE NCCL all-gather merges the
// per-record results.
//
// Environment (the palace call line stays unchanged):
//   EREF_GPUS=N|all     GPUs to shard the database over (default 1; more only when a shard would exceed 8 GiB of FASTA:
//                       the GPU work of a 1.5 Gbp database is tens of milliseconds, every extra context costs more)
//   EREF_CODER_SEED=S   seed of the coder draw when <db>.k32.index.dat does not exist (default time(0), :1088)
//   EREF_NO_PRUNE=1     probe all three streams everywhere (disables the exact candidate filter)
//   EREF_DIRECT_COUNT=1 count k-mers straight into the table (disables the code-range bucketing; one GPU only)
//   EREF_REPLICATED_TABLE=1  N > 1: every GPU counts the whole table instead of 1/N of it
//   EREF_TIMING=1       wall-clock breakdown and per-phase CUDA-event times on stderr
//   EREF_PIN=0|1        page-lock the input buffers before the upload (default 1)
//   EREF_WRITE_INDEX=1  also write <db>.k32.index.dat + <db>.genome.len.txt in the reference's format (when absent),
//                       so that a CPU eref can share the DB directory (12 bytes per base: off by default)
//   EREF_REFS_FASTA=path, EREF_REFS_PERCENT=path   also write what the next pipeline step makes of the result lines
//                       (share/palace/scripts/get_ref_by_index.py, palace:482-498): the selected records as single-line
//                       FASTA and `name<TAB>percentage`
#include "eref_gpu.h"
#include <cuda_runtime.h>
#include <nccl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#include <algorithm>
#include <atomic>
#include <cctype>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

using Clock = std::chrono::steady_clock;
double ms_between(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); }

struct HostFile {                       // whole file in page-aligned host memory, page-locked later (pin())
    char* p = nullptr;
    size_t n = 0, cap = 0;
    bool pinned = false, was_gzip = false;
    std::string err;
    // (never unregistered or freed: the buffers live until the process exits, and unpinning gigabytes costs time)
    bool reserve(size_t want)
    {
        if (want <= cap) return true;
        const size_t ncap = ((p ? want + want / 2 : want) + 4095) & ~(size_t)4095;
        void* q = nullptr;
        if (posix_memalign(&q, 4096, ncap) != 0) return false;
        if (p) { memcpy(q, p, n); free(p); }
        p = (char*)q; cap = ncap;
        return true;
    }
    void pin()
    {
        if (!p || !n || pinned) return;
        if (cudaHostRegister(p, std::min(cap, (n + 4095) & ~(size_t)4095), cudaHostRegisterDefault) == cudaSuccess) pinned = true;
        else cudaGetLastError();                 // pageable memory still works, only slower
    }
};

// ---- gzip input: the reference reads plain text only; users' files are .gz -------------
// BGZF (bgzip) files carry every block's compressed size in a gzip extra field ("BC"), so the blocks are found
// without inflating and inflated by several threads; any other gzip stream (one or more members) is inflated
// sequentially by the file's own reader thread — the three input files are read in parallel anyway.

struct BgzfBlock { size_t in_off, in_len, out_off, out_len; };

bool bgzf_blocks(const unsigned char* z, size_t zn, std::vector<BgzfBlock>& blocks, size_t& total)
{
    size_t pos = 0;
    total = 0;
    while (pos < zn) {
        if (zn - pos < 28 || z[pos] != 0x1f || z[pos + 1] != 0x8b || z[pos + 2] != 8 || !(z[pos + 3] & 4)) return false;
        const size_t xlen = z[pos + 10] | (z[pos + 11] << 8);
        size_t x = pos + 12, xend = x + xlen, bsize = 0;
        if (xend > zn) return false;
        while (x + 4 <= xend) {
            const size_t slen = z[x + 2] | (z[x + 3] << 8);
            if (z[x] == 'B' && z[x + 1] == 'C' && slen == 2 && x + 6 <= xend) bsize = (size_t)(z[x + 4] | (z[x + 5] << 8)) + 1;
            x += 4 + slen;
        }
        if (!bsize || pos + bsize > zn || bsize < xlen + 20) return false;
        const unsigned char* tail = z + pos + bsize - 4;
        const size_t isize = tail[0] | (tail[1] << 8) | (tail[2] << 16) | ((size_t)tail[3] << 24);
        blocks.push_back({pos + 12 + xlen, bsize - xlen - 20, total, isize});
        total += isize;
        pos += bsize;
    }
    return !blocks.empty();
}

bool inflate_bgzf(const unsigned char* z, const std::vector<BgzfBlock>& blocks, char* out, int threads, std::string& err)
{
    std::atomic<size_t> next{0};
    std::atomic<int> bad{0};
    auto work = [&] {
        z_stream zs;
        for (size_t i; (i = next.fetch_add(1)) < blocks.size() && !bad.load();) {
            const BgzfBlock& b = blocks[i];
            memset(&zs, 0, sizeof zs);
            if (inflateInit2(&zs, -15) != Z_OK) { bad.store(1); return; }
            zs.next_in = const_cast<unsigned char*>(z + b.in_off); zs.avail_in = (uInt)b.in_len;
            zs.next_out = (unsigned char*)out + b.out_off; zs.avail_out = (uInt)b.out_len;
            const int r = inflate(&zs, Z_FINISH);
            if (r != Z_STREAM_END || zs.total_out != b.out_len) bad.store(1);
            inflateEnd(&zs);
            // raw deflate checks nothing: compare the block's CRC-32 (the four bytes behind the deflate stream) like
            // gzip -d does for the stream it inflates
            const unsigned char* c = z + b.in_off + b.in_len;
            const uLong stored = (uLong)c[0] | ((uLong)c[1] << 8) | ((uLong)c[2] << 16) | ((uLong)c[3] << 24);
            if (!bad.load() && crc32(crc32(0L, Z_NULL, 0), (const unsigned char*)out + b.out_off, (uInt)b.out_len) != stored) bad.store(2);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (bad.load()) { err = bad.load() == 2 ? "corrupt BGZF block (CRC mismatch)" : "corrupt BGZF block"; return false; }
    return true;
}

bool inflate_stream(const unsigned char* z, size_t zn, HostFile& f, std::string& err)
{
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, 15 + 32) != Z_OK) { err = "zlib init failed"; return false; }
    // a single-member file states its size (mod 2^32) in the last four bytes: a good first reservation
    size_t hint = zn >= 4 ? (size_t)(z[zn - 4] | (z[zn - 3] << 8) | (z[zn - 2] << 16) | ((size_t)z[zn - 1] << 24)) : 0;
    if (hint < zn) hint = zn * 4;
    if (!f.reserve(hint + 64)) { inflateEnd(&zs); err = "out of memory"; return false; }
    size_t in = 0;
    f.n = 0;
    for (;;) {
        if (f.cap - f.n < (1u << 20) && !f.reserve(f.cap + f.cap / 2)) { inflateEnd(&zs); err = "out of memory"; return false; }
        const size_t in_now = std::min<size_t>(zn - in, 1u << 30), out_now = std::min<size_t>(f.cap - f.n, 1u << 30);
        zs.next_in = const_cast<unsigned char*>(z + in); zs.avail_in = (uInt)in_now;
        zs.next_out = (unsigned char*)f.p + f.n; zs.avail_out = (uInt)out_now;
        const int r = inflate(&zs, Z_NO_FLUSH);
        in += in_now - zs.avail_in; f.n += out_now - zs.avail_out;
        if (r == Z_STREAM_END) {
            if (in >= zn) break;
            if (inflateReset(&zs) != Z_OK) { inflateEnd(&zs); err = "zlib reset failed"; return false; }   // next member
            continue;
        }
        if (r != Z_OK && r != Z_BUF_ERROR) { inflateEnd(&zs); err = std::string("inflate: ") + (zs.msg ? zs.msg : "error"); return false; }
        if (in >= zn && zs.avail_out) { inflateEnd(&zs); err = "truncated gzip stream"; return false; }
    }
    inflateEnd(&zs);
    return true;
}

bool read_file(const std::string& path, HostFile& f, int gz_threads)
{
    FILE* fp = fopen(path.c_str(), "rb");
    if (!fp) { f.err = "cannot open " + path + ": " + strerror(errno); return false; }
    struct stat st;
    if (fstat(fileno(fp), &st) != 0) { fclose(fp); f.err = "cannot size " + path; return false; }
    HostFile raw;
    HostFile& dst = f;
    const size_t sz = (size_t)st.st_size;
    unsigned char magic[2] = {0, 0};
    const size_t mg = fread(magic, 1, 2, fp);
    const bool gz = mg == 2 && magic[0] == 0x1f && magic[1] == 0x8b;
    HostFile& into = gz ? raw : dst;
    if (!into.reserve(sz + 64)) { fclose(fp); f.err = "out of memory reading " + path; return false; }
    memcpy(into.p, magic, mg);
    size_t got = mg;
    while (got < sz) {
        const size_t r = fread(into.p + got, 1, sz - got, fp);
        if (!r) break;
        got += r;
    }
    fclose(fp);
    if (got != sz) { f.err = "short read on " + path; return false; }
    into.n = sz;
    if (!gz) return true;
    f.was_gzip = true;
    std::vector<BgzfBlock> blocks;
    size_t total = 0;
    const unsigned char* z = (const unsigned char*)raw.p;
    if (bgzf_blocks(z, sz, blocks, total)) {
        if (!f.reserve(total + 64)) { f.err = "out of memory inflating " + path; return false; }
        f.n = total;
        if (!inflate_bgzf(z, blocks, f.p, gz_threads, f.err)) { f.err = path + ": " + f.err; return false; }
        return true;
    }
    if (!inflate_stream(z, sz, f, f.err)) { f.err = path + ": " + f.err; return false; }
    return true;
}

// random_coder (:1082-1102): srand(seed); 32 draws of rand()%6 pick one ordering of {0,1,2} each
void draw_coder(unsigned seed, int16_t coder[96])
{
    static const int16_t orderings[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 2, 0}, {1, 0, 2}, {2, 0, 1}, {2, 1, 0}};
    srand(seed);
    for (int z = 0; z < EREF_K; z++) {
        int r = rand() % 6;
        for (int i = 0; i < 3; i++) coder[z * 3 + i] = orderings[r][i];
    }
}

// saved_random_coder (:1104-1122): low 16 bits of the first 96 words of an existing index
bool read_index_header(const std::string& index, int16_t coder[96])
{
    FILE* fp = fopen(index.c_str(), "rb");
    if (!fp) return false;
    uint32_t w[96];
    size_t got = fread(w, 4, 96, fp);
    fclose(fp);
    if (got != 96) return false;
    for (int j = 0; j < 96; j++) coder[j] = (int16_t)(w[j] & 0xffffu);
    return true;
}

bool read_sidecar(const std::string& path, int16_t coder[96])
{
    std::ifstream in(path);
    if (!in) return false;
    for (int j = 0; j < 96; j++) { int v; if (!(in >> v)) return false; coder[j] = (int16_t)v; }
    return true;
}

// shard boundaries: byte offsets that start a header line ('>' right after a newline)
std::vector<size_t> shard_offsets(const char* fa, size_t n, int shards)
{
    std::vector<size_t> cut(1, 0);
    for (int s = 1; s < shards; s++) {
        size_t pos = std::max(cut.back(), n / shards * s);
        size_t found = n;
        while (pos < n) {
            const char* nl = (const char*)memchr(fa + pos, '\n', n - pos);
            if (!nl) break;
            size_t q = (size_t)(nl - fa) + 1;
            if (q < n && fa[q] == '>') { found = q; break; }
            pos = q;
        }
        cut.push_back(found);
    }
    cut.push_back(n);
    return cut;
}

// get_read_ID(line).substr(1) (:246-254, :689): the header up to the first '/', ' ' or tab, without '>'
std::string record_name(const char* p, size_t n)
{
    size_t e = n;
    for (char d : {'/', ' ', '\t'}) {
        const void* q = memchr(p, d, e);
        if (q) e = (size_t)((const char*)q - p);
    }
    return e > 1 ? std::string(p + 1, e - 1) : std::string();
}

// <db>.k32.index.dat (400-byte coder header :680-682, then per kept record its length and
// 3 codes per window start :698, :711-738) and <db>.genome.len.txt (:697, :800), produced on GPU 0 shard by shard.
// Both files are written under temporary names and renamed together, so a failed run leaves neither.
bool write_index_files(const std::string& index_name, const std::string& len_name, const char* fa, size_t n,
                       const int16_t coder[96], std::string& err)
{
    const std::string tmp = index_name + ".part", ltmp = len_name + ".part";
    FILE* fi = fopen(tmp.c_str(), "wb");
    FILE* fl = fopen(ltmp.c_str(), "w");
    if (!fi || !fl) {
        err = "cannot create " + (fi ? ltmp : tmp);
        if (fi) { fclose(fi); remove(tmp.c_str()); }
        if (fl) { fclose(fl); remove(ltmp.c_str()); }
        return false;
    }
    uint32_t hdr[100] = {0};
    for (int j = 0; j < 96; j++) hdr[j] = (uint32_t)(uint16_t)coder[j] | (j < 95 ? (uint32_t)(uint16_t)coder[j + 1] << 16 : 0u);
    bool ok = fwrite(hdr, 4, 100, fi) == 100;
    eref_ctx* ctx = nullptr;
    if (ok && eref_create(&ctx, 0, coder, 0.9f, 0.85f, 0)) { err = std::string("eref_create: ") + eref_last_error(); ok = false; }
    const int shards = (int)(n >> 30) + 1;               // every piece stays far below the 4 GiB shard limit
    std::vector<size_t> cut = shard_offsets(fa, n, shards);
    const size_t cap = (size_t)1 << 26;                  // 256 MiB of body per round trip
    std::vector<uint32_t> buf(cap), rlen, roff;
    unsigned long long run = 0, hdrs_before = 0;
    for (int s = 0; ok && s < shards; s++) {
        const char* text = fa + cut[s];
        const size_t len = cut[s + 1] - cut[s];
        if (eref_ref_set(ctx, text, len)) { err = std::string("eref_ref_set: ") + eref_last_error(); ok = false; break; }
        size_t kept = 0, nrec = 0;
        eref_ref_count(ctx, &kept, &nrec, nullptr);
        uint32_t first = 1;
        while (ok && first <= kept) {
            size_t nw = 0; uint32_t next = first;
            int rc = eref_index_body(ctx, first, buf.data(), buf.size(), &nw, &next);
            if (rc == EREF_ERR_LIMIT && buf.size() < ((size_t)1 << 31)) { buf.resize(buf.size() * 4); continue; }
            if (rc) { err = std::string("eref_index_body: ") + eref_last_error(); ok = false; break; }
            ok = fwrite(buf.data(), 4, nw, fi) == nw;
            if (!ok) err = "short write on " + tmp;
            first = next;
        }
        rlen.resize(nrec); roff.resize(nrec);
        if (ok && eref_ref_records(ctx, rlen.data(), roff.data(), nrec, &nrec)) { err = eref_last_error(); ok = false; }
        for (size_t r = 0; ok && r < nrec; r++) {
            run += rlen[r];
            if (rlen[r] <= (uint32_t)EREF_K) continue;
            std::string name = "start";
            if (r > 0) {
                const char* h = text + roff[r];
                const void* nl = memchr(h, '\n', len - roff[r]);
                name = record_name(h, nl ? (size_t)((const char*)nl - h) : len - roff[r]);
            }
            fprintf(fl, "%s\t%llu\t%u\t%llu\n", name.c_str(), hdrs_before + r, rlen[r], run);
        }
        hdrs_before += nrec ? nrec - 1 : 0;
    }
    if (ctx) eref_destroy(ctx);
    ok = (fclose(fi) == 0) && ok;
    ok = (fclose(fl) == 0) && ok;
    if (ok && rename(ltmp.c_str(), len_name.c_str()) != 0) { err = "cannot rename " + ltmp; ok = false; }
    if (ok && rename(tmp.c_str(), index_name.c_str()) != 0) { err = "cannot rename " + tmp; ok = false; }
    if (!ok) { remove(tmp.c_str()); remove(ltmp.c_str()); }
    return ok;
}

// ---- the next pipeline step, fused (share/palace/scripts/get_ref_by_index.py) -------------
// The script maps a printed index to the index-th LINE of <db>.fai (:40-49) — samtools faidx writes one line per FASTA
// record, in file order, named by the header up to the first white space — then looks that name up among the records
// (:78-82) and writes ">name\n<sequence on one line>\n" and "name\tpercentage\n" in ascending index order (:74).  Records
// of <= 32 bases take a .fai line but no eref ordinal (:697), so the index drifts from the record eref scored whenever
// such records precede it: reference behaviour, reproduced here by construction (the index-th header line of the text).

// Python's repr(float(s)) for the %g text eref printed (:617 -> get_ref_by_index.py:27-33, :86)
std::string py_float_repr(const char* g_text)
{
    const double v = strtod(g_text, nullptr);
    char buf[64];
    int prec = 1;
    for (; prec < 17; prec++) { snprintf(buf, sizeof buf, "%.*e", prec - 1, v); if (strtod(buf, nullptr) == v) break; }
    snprintf(buf, sizeof buf, "%.*e", prec - 1, v);
    // digits and decimal exponent of the shortest round-trip form
    std::string digits;
    for (const char* q = buf; *q && *q != 'e'; q++) if (*q >= '0' && *q <= '9') digits.push_back(*q);
    const int exp10 = atoi(strchr(buf, 'e') + 1);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = v < 0 ? "-" : "";
    if (v == 0) return "0.0";
    if (exp10 < -4 || exp10 >= 16) {                     // repr switches to exponent form here
        out += digits.substr(0, 1);
        if (digits.size() > 1) out += "." + digits.substr(1);
        char e[16]; snprintf(e, sizeof e, "e%c%02d", exp10 < 0 ? '-' : '+', exp10 < 0 ? -exp10 : exp10);
        return out + e;
    }
    if (exp10 < 0) return out + "0." + std::string((size_t)(-exp10 - 1), '0') + digits;
    if ((int)digits.size() <= exp10 + 1) return out + digits + std::string((size_t)(exp10 + 1 - (int)digits.size()), '0') + ".0";
    return out + digits.substr(0, (size_t)exp10 + 1) + "." + digits.substr((size_t)exp10 + 1);
}

// offsets of the header lines number want[0] < want[1] < ... (1-based, counted over every line that starts with '>'),
// found by `threads` threads counting per chunk first
std::vector<size_t> header_offsets(const char* fa, size_t n, const std::vector<uint32_t>& want, int threads)
{
    std::vector<size_t> out(want.size(), (size_t)-1);
    if (want.empty() || !n) return out;
    threads = std::max(1, std::min(threads, (int)(n >> 24) + 1));
    std::vector<size_t> lo(threads + 1), cnt(threads, 0);
    for (int t = 0; t <= threads; t++) lo[t] = n / threads * t;
    lo[threads] = n;
    auto each = [&](int t, auto&& fn) {
        size_t pos = lo[t];
        while (pos < lo[t + 1]) {
            const char* q = (const char*)memchr(fa + pos, '>', lo[t + 1] - pos);
            if (!q) break;
            const size_t at = (size_t)(q - fa);
            if (at == 0 || fa[at - 1] == '\n') fn(at);
            pos = at + 1;
        }
    };
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back([&, t] { each(t, [&](size_t) { cnt[t]++; }); });
    for (auto& th : pool) th.join();
    pool.clear();
    std::vector<size_t> before(threads + 1, 0);
    for (int t = 0; t < threads; t++) before[t + 1] = before[t] + cnt[t];
    for (int t = 0; t < threads; t++)
        pool.emplace_back([&, t] {
            size_t k = before[t];
            size_t w = (size_t)(std::lower_bound(want.begin(), want.end(), (uint32_t)(k + 1)) - want.begin());
            each(t, [&](size_t at) { k++; if (w < want.size() && want[w] == k) out[w++] = at; });
        });
    for (auto& th : pool) th.join();
    return out;
}

bool write_refs(const std::string& fasta_out, const std::string& percent_out, const char* fa, size_t n,
                const std::vector<eref_result>& merged, int threads, std::string& err)
{
    std::map<uint32_t, std::string> pct;                 // index -> percentage text, ascending (:74); later lines win (:35)
    char line[256];
    for (const eref_result& r : merged) {
        if (eref_format_line(&r, line, sizeof line) <= 0) continue;
        // the last token that parses as a float (:27-33) is the %g ratio at the end of the line
        char* tab = strrchr(line, '\t');
        pct[r.ordinal] = py_float_repr(tab + 1);
    }
    std::vector<uint32_t> want;
    for (auto& kv : pct) want.push_back(kv.first);
    const std::vector<size_t> off = header_offsets(fa, n, want, threads);
    const std::string ftmp = fasta_out + ".part", ptmp = percent_out + ".part";
    FILE* ff = fasta_out.empty() ? nullptr : fopen(ftmp.c_str(), "w");
    FILE* fp = percent_out.empty() ? nullptr : fopen(ptmp.c_str(), "w");
    if ((!fasta_out.empty() && !ff) || (!percent_out.empty() && !fp)) {
        err = "cannot create " + (ff || fasta_out.empty() ? ptmp : ftmp);
        if (ff) fclose(ff);
        if (fp) fclose(fp);
        return false;
    }
    std::string seq;
    for (size_t w = 0; w < want.size(); w++) {
        if (off[w] == (size_t)-1) {                      // fewer .fai lines than the index (:91)
            fprintf(stderr, "Warning: Index %u not found in FAI file\n", want[w]);
            continue;
        }
        const char* h = fa + off[w];
        const char* nl = (const char*)memchr(h, '\n', n - off[w]);
        const size_t hl = nl ? (size_t)(nl - h) : n - off[w];
        size_t e = 1;                                    // name = header up to the first white space
        while (e < hl && !isspace((unsigned char)h[e])) e++;
        const std::string name(h + 1, e - 1);
        // the record's lines up to the next header, joined; trailing white space of a line, blanks and '\r' dropped
        seq.clear();
        size_t pos = off[w] + hl + (nl ? 1 : 0);
        while (pos < n && fa[pos] != '>') {
            const char* q = (const char*)memchr(fa + pos, '\n', n - pos);
            size_t le = q ? (size_t)(q - fa) : n, te = le;
            while (te > pos && isspace((unsigned char)fa[te - 1])) te--;
            for (size_t k = pos; k < te; k++) if (fa[k] != ' ' && fa[k] != '\r') seq.push_back(fa[k]);
            pos = le + 1;
        }
        if (ff) { fprintf(ff, ">%s\n", name.c_str()); fwrite(seq.data(), 1, seq.size(), ff); fputc('\n', ff); }
        if (fp) fprintf(fp, "%s\t%s\n", name.c_str(), pct[want[w]].c_str());
    }
    bool ok = true;
    if (ff) ok = (fclose(ff) == 0) && ok;
    if (fp) ok = (fclose(fp) == 0) && ok;
    if (ok && ff && rename(ftmp.c_str(), fasta_out.c_str()) != 0) ok = false;
    if (ok && fp && rename(ptmp.c_str(), percent_out.c_str()) != 0) ok = false;
    if (!ok) err = "cannot write " + fasta_out + " / " + percent_out;
    return ok;
}

struct Barrier {
    std::mutex m; std::condition_variable cv; int count, waiting = 0, gen = 0;
    explicit Barrier(int n) : count(n) {}
    void wait()
    {
        std::unique_lock<std::mutex> lk(m);
        int g = gen;
        if (++waiting == count) { gen++; waiting = 0; cv.notify_all(); }
        else cv.wait(lk, [&] { return g != gen; });
    }
};

} // namespace

int main(int argc, char** argv)
{
    if (argc < 8) {
        fprintf(stderr, "usage: eref <fq1> <fq2> <phagedb.fasta> <interval_out> <hit_ratio> <perfect_hit_ratio> <threads>\n");
        return 2;
    }
    const std::string fq1 = argv[1], fq2 = argv[2], fasta = argv[3], interval_name = argv[4];
    float hit_ratio, perfect_hit_ratio;
    int host_threads = 1;
    try {
        hit_ratio = (float)std::stod(argv[5]);               // :1228
        perfect_hit_ratio = (float)std::stod(argv[6]);       // :1229
        host_threads = (int)std::stod(argv[7]);              // :1231 threads: a hint; here: host helper threads (inflate, f1)
    } catch (const std::exception& e) {
        fprintf(stderr, "eref: bad numeric argument: %s\n", e.what());
        return 2;
    }
    host_threads = std::max(1, std::min(host_threads, 64));
    // stdout is the result channel (palace redirects it into ref_names.txt): keep libraries (e.g. NCCL's
    // version banner) off it by parking the real stdout and pointing fd 1 at stderr until the results are ready
    fflush(stdout);
    const int result_fd = dup(1);
    if (result_fd < 0 || dup2(2, 1) < 0) { fprintf(stderr, "eref: cannot set up the result channel\n"); return 3; }
    const bool timing = getenv("EREF_TIMING") != nullptr;
    const auto t_start = Clock::now();

    // ---- inputs: one reader thread per file, started before anything touches CUDA ----
    HostFile f1, f2, fa;
    const int gz_threads = std::max(1, host_threads / 2);
    std::thread r1([&] { read_file(fq1, f1, gz_threads); });
    std::thread r2([&] { read_file(fq2, f2, gz_threads); });
    std::thread r3([&] { read_file(fasta, fa, std::max(gz_threads, host_threads - 2)); });
    auto join_readers = [&] { if (r1.joinable()) r1.join(); if (r2.joinable()) r2.join(); if (r3.joinable()) r3.join(); };

    int ndev = eref_device_count();
    if (ndev <= 0) { join_readers(); fprintf(stderr, "eref: no CUDA device visible (this build has no CPU path)\n"); return 3; }

    { std::ofstream trunc(interval_name, std::ios::out | std::ios::trunc); }   // :825 created, left empty

    // ---- coder assignment (:1245-1258) ----
    int16_t coder[96];
    const std::string index_name = fasta + ".k" + std::to_string(EREF_K) + ".index.dat";
    const std::string sidecar = fasta + ".k" + std::to_string(EREF_K) + ".coder";
    bool coder_drawn_here = false;                   // then rand() continues from the coder draw, like a cold reference run
    if (!read_index_header(index_name, coder) && !read_sidecar(sidecar, coder)) {
        coder_drawn_here = true;
        const char* s = getenv("EREF_CODER_SEED");
        draw_coder(s ? (unsigned)strtoul(s, nullptr, 10) : (unsigned)time(nullptr), coder);
        std::ofstream out(sidecar);
        for (int j = 0; j < 96; j++) out << coder[j] << (j % 3 == 2 ? '\n' : ' ');
        if (!out) fprintf(stderr, "eref: warning: could not persist the coder assignment to %s\n", sidecar.c_str());
    }

    // ---- how many GPUs: one, unless asked or a shard would not be comfortable on one ----
    int shards = 1;
    {
        struct stat st;
        if (stat(fasta.c_str(), &st) == 0) shards = (int)std::min<long long>(ndev, ((long long)st.st_size >> 33) + 1);
    }
    if (const char* g = getenv("EREF_GPUS")) {
        shards = strcmp(g, "all") == 0 ? ndev : atoi(g);
        if (shards < 1 || shards > ndev) { join_readers(); fprintf(stderr, "eref: EREF_GPUS=%s but %d device(s) visible\n", g, ndev); return 3; }
    }
    const bool shard_table = shards > 1 && !getenv("EREF_REPLICATED_TABLE") && !getenv("EREF_DIRECT_COUNT");
    const unsigned flags = (getenv("EREF_NO_PRUNE") ? EREF_FLAG_NO_PRUNE : 0u) |
                           (getenv("EREF_DIRECT_COUNT") ? EREF_FLAG_DIRECT_COUNT : 0u) |
                           (coder_drawn_here ? EREF_FLAG_KEEP_RAND_STATE : 0u);
    uint64_t down_sampling_size = 2000000000ull;         // :1230
    if (const char* v = getenv("EREF_DOWN_SAMPLING_SIZE")) { long long k = atoll(v); if (k > 0) down_sampling_size = (uint64_t)k; }

    // ---- contexts (and the NCCL communicators) come up while the files are still being read ----
    std::vector<eref_ctx*> ctx(shards, nullptr);
    std::vector<std::string> errors(shards);
    std::atomic<int> failed{0};
    std::vector<ncclComm_t> comms(shards);
    bool comms_ok = false;
    {
        std::vector<std::thread> pool;
        for (int i = 0; i < shards; i++)
            pool.emplace_back([&, i] {
                if (eref_create(&ctx[i], i, coder, hit_ratio, perfect_hit_ratio, flags)) {
                    errors[i] = std::string("eref_create: ") + eref_last_error(); failed.store(1);
                }
            });
        for (auto& t : pool) t.join();
        if (shards > 1 && !failed.load()) {
            std::vector<int> devs(shards);
            for (int i = 0; i < shards; i++) devs[i] = i;
            ncclResult_t r = ncclCommInitAll(comms.data(), shards, devs.data());
            if (r != ncclSuccess) { errors[0] = std::string("ncclCommInitAll: ") + ncclGetErrorString(r); failed.store(1); }
            else comms_ok = true;
        }
    }
    const auto t_ctx = Clock::now();
    join_readers();
    const auto t_read = Clock::now();
    for (HostFile* f : {&f1, &f2, &fa})
        if (!f->err.empty()) { fprintf(stderr, "eref: %s\n", f->err.c_str()); failed.store(1); }
    auto cleanup = [&] {
        for (eref_ctx* c : ctx) if (c) eref_destroy(c);
        if (comms_ok) for (auto& c : comms) ncclCommDestroy(c);
    };
    if (failed.load()) {
        for (int i = 0; i < shards; i++) if (!errors[i].empty()) fprintf(stderr, "eref: gpu %d: %s\n", i, errors[i].c_str());
        cleanup();
        return 3;
    }
    // main hands file_size(fq1) to read_fastq as the byte budget of BOTH files (:1265-1291, :942-945): whatever of fq2
    // lies behind it is never read.  Cut it here, once, so that every path below (upload slices, census, masks) agrees.
    f2.n = eref_fastq_limit(f2.p, f2.n, f1.n);
    if (!getenv("EREF_PIN") || atoi(getenv("EREF_PIN")) != 0) {
        cudaSetDevice(0);
        std::thread p1([&] { cudaSetDevice(0); f1.pin(); }), p2([&] { cudaSetDevice(0); f2.pin(); });
        fa.pin();
        p1.join(); p2.join();
    }
    const auto t_pin = Clock::now();

    // ---- sharded table build: every context learns where the other tables live (same process: raw pointers) ----
    if (shard_table) {
        std::vector<void*> bm3(shards), summ(shards), flg(shards);
        for (int i = 0; i < shards && !failed.load(); i++)
            if (eref_table_export(ctx[i], nullptr, &bm3[i], &summ[i], &flg[i]) || eref_table_set_part(ctx[i], i, shards)) {
                errors[i] = std::string("table export: ") + eref_last_error(); failed.store(1);
            }
        for (int i = 0; i < shards && !failed.load(); i++)
            for (int j = 0; j < shards && !failed.load(); j++)
                if (i != j && eref_table_set_peer(ctx[i], j, j, nullptr, bm3[j], summ[j], flg[j])) {
                    errors[i] = std::string("eref_table_set_peer: ") + eref_last_error(); failed.store(1);
                }
    }

    std::vector<size_t> cut = shard_offsets(fa.p, fa.n, shards);
    std::vector<size_t> kept(shards, 0);
    std::vector<eref_phase_times> times(shards);
    std::vector<eref_result> merged;                 // filled by shard 0
    Barrier bar(shards);
    size_t max_kept = 0;
    bool go = false;
    std::vector<uint32_t> keep1, keep2;              // down-sampling masks, drawn once by shard 0
    uint64_t lines1 = 0, lines2 = 0;
    int ratio = 100;

    auto worker = [&](int i) {
        eref_ctx* c = ctx[i];
        auto fail = [&](const char* what) { errors[i] = std::string(what) + ": " + eref_last_error(); failed.store(1); };
        bool ok = failed.load() == 0;
        if (shards == 1) {
            // one GPU: the pipelined entry point straight from the host buffers; the records stay on the device
            size_t n = 0;
            if (ok && eref_run_host(c, f1.p, f1.n, f2.p, f2.n, fa.p, fa.n, 4, nullptr, 0, &n)) { fail("eref_run_host"); ok = false; }
            const uint32_t* dres = nullptr;
            if (ok && eref_run_results_device(c, &dres, &n)) { fail("eref_run_results_device"); ok = false; }
            if (ok) {
                merged.resize(n);
                if (n && cudaMemcpy(merged.data(), dres, n * sizeof(eref_result), cudaMemcpyDeviceToHost) != cudaSuccess) {
                    errors[i] = "copying the records back failed"; failed.store(1);
                }
            }
            eref_timings(c, &times[i]);
            return;
        }
        // ---- N GPUs.  The query: every GPU uploads 1/N of each FASTQ text, one NCCL all-gather per file replicates it
        // over NVLink (the texts land in place: slot i of GPU i's buffer is its own upload).
        cudaSetDevice(i);
        cudaStream_t st = nullptr;
        char *q1 = nullptr, *q2 = nullptr;
        void *send = nullptr, *recv = nullptr;
        const size_t slot1 = ((f1.n + shards - 1) / shards + 15) & ~(size_t)15, slot2 = ((f2.n + shards - 1) / shards + 15) & ~(size_t)15;
        ok = ok && cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess &&
             cudaMalloc((void**)&q1, slot1 * shards + 64) == cudaSuccess && cudaMalloc((void**)&q2, slot2 * shards + 64) == cudaSuccess;
        if (!ok && errors[i].empty()) { errors[i] = "query buffers: out of device memory"; failed.store(1); }
        auto upload_slice = [&](char* q, const HostFile& f, size_t slot) {
            const size_t lo = std::min(f.n, slot * i), hi = std::min(f.n, slot * (i + 1));
            return hi == lo || cudaMemcpyAsync(q + slot * i, f.p + lo, hi - lo, cudaMemcpyHostToDevice, st) == cudaSuccess;
        };
        if (ok && !(upload_slice(q1, f1, slot1) && upload_slice(q2, f2, slot2))) { errors[i] = "query upload failed"; failed.store(1); ok = false; }
        bar.wait();
        const bool all_up = failed.load() == 0;       // one decision for every shard: all enter the collectives or none
        if (all_up) {
            ncclGroupStart();
            ncclResult_t ra = ncclAllGather(q1 + slot1 * i, q1, slot1, ncclChar, comms[i], st);
            ncclResult_t rb = ncclAllGather(q2 + slot2 * i, q2, slot2, ncclChar, comms[i], st);
            ncclGroupEnd();
            if (ra != ncclSuccess || rb != ncclSuccess) { errors[i] = "ncclAllGather(query) failed"; failed.store(1); }
            if (cudaStreamSynchronize(st) != cudaSuccess) { errors[i] = "query all-gather failed"; failed.store(1); }
        }
        ok = all_up && failed.load() == 0;
        // cal_sam_ratio + the down-sampling draws (:1124-1148, :955-960): once, by shard 0, shared by every shard
        if (ok && i == 0) {
            uint64_t seq = 0, l1 = 0, l2 = 0;
            if (eref_query_census_device(c, q1, f1.n, &seq, &l1) || eref_sample_ratio(seq, down_sampling_size, &ratio)) { fail("cal_sam_ratio"); }
            else if (ratio < 100) {
                if (eref_query_census_device(c, q2, f2.n, nullptr, &l2)) fail("eref_query_census");
                else {
                    if (!coder_drawn_here) srand(1);               // :1236-1238
                    keep1.assign((l1 + 31) / 32 + 1, 0); keep2.assign((l2 + 31) / 32 + 1, 0);
                    eref_draw_keep(ratio, l1, keep1.data()); eref_draw_keep(ratio, l2, keep2.data());
                    lines1 = l1; lines2 = l2;
                }
            }
        }
        bar.wait();
        ok = failed.load() == 0;
        if (ok && ratio < 100 && eref_run_host_keep(c, keep1.data(), lines1, keep2.data(), lines2)) { fail("eref_run_host_keep"); ok = false; }
        // the pipelined call on this GPU's shard (pieces of <= 2 GiB); with the sharded table every GPU builds its part
        // inside this call and waits on the device for the others' parts, so all shards make the call or none
        bar.wait();
        const bool all_go = failed.load() == 0;
        size_t n = 0;
        if (all_go) {
            if (eref_run_host(c, q1, f1.n, q2, f2.n, fa.p + cut[i], cut[i + 1] - cut[i], 4, nullptr, 0, &n)) { fail("eref_run_host"); }
            else kept[i] = n;
        }
        bar.wait();                                   // every shard's record count is known
        if (i == 0) { max_kept = *std::max_element(kept.begin(), kept.end()); go = failed.load() == 0; }
        bar.wait();
        if (go) {
            // the one collective on the results: allgather of fixed-size records, shards padded to max_kept
            const uint32_t* dres = nullptr;
            const size_t slot = std::max<size_t>(max_kept, 1) * sizeof(eref_result);
            bool good = eref_run_results_device(c, &dres, &n) == 0;
            good = good && cudaMalloc(&send, slot) == cudaSuccess && cudaMalloc(&recv, slot * shards) == cudaSuccess &&
                   cudaMemsetAsync(send, 0, slot, st) == cudaSuccess &&
                   (n == 0 || cudaMemcpyAsync(send, dres, n * sizeof(eref_result), cudaMemcpyDeviceToDevice, st) == cudaSuccess);
            if (!good && errors[i].empty()) { errors[i] = "allgather staging failed"; failed.store(1); }
            // every rank must enter the collective, even after a local failure, or the others hang
            ncclResult_t r = ncclAllGather(send, recv, slot, ncclChar, comms[i], st);
            if (r != ncclSuccess) { errors[i] = std::string("ncclAllGather: ") + ncclGetErrorString(r); failed.store(1); }
            cudaStreamSynchronize(st);
            if (i == 0 && failed.load() == 0) {
                std::vector<eref_result> all(std::max<size_t>(max_kept, 1) * shards);
                cudaMemcpy(all.data(), recv, slot * shards, cudaMemcpyDeviceToHost);
                uint32_t base = 0;
                for (int s = 0; s < shards; s++) {
                    for (size_t k = 0; k < kept[s]; k++) {
                        eref_result rr = all[(size_t)s * std::max<size_t>(max_kept, 1) + k];
                        rr.ordinal += base;           // ordinals continue across shards (file order)
                        merged.push_back(rr);
                    }
                    base += (uint32_t)kept[s];
                }
            }
        }
        eref_timings(c, &times[i]);
        if (send) cudaFree(send);
        if (recv) cudaFree(recv);
        if (q1) cudaFree(q1);
        if (q2) cudaFree(q2);
        if (st) cudaStreamDestroy(st);
    };

    std::vector<std::thread> pool;
    for (int i = 0; i < shards; i++) pool.emplace_back(worker, i);
    for (auto& t : pool) t.join();
    const auto t_gpu = Clock::now();

    if (failed.load()) {
        for (int i = 0; i < shards; i++)
            if (!errors[i].empty()) fprintf(stderr, "eref: gpu %d: %s\n", i, errors[i].c_str());
        cleanup();
        return 3;                                     // nothing was written to stdout
    }

    std::string out;
    char line[256];
    for (const eref_result& r : merged) {
        int k = eref_format_line(&r, line, sizeof line);    // :615-617
        if (k > 0) out.append(line, (size_t)k);
    }
    fflush(stdout);
    for (size_t done = 0; done < out.size();) {
        ssize_t w = write(result_fd, out.data() + done, out.size() - done);
        if (w <= 0) { fprintf(stderr, "eref: write to stdout failed\n"); cleanup(); return 3; }
        done += (size_t)w;
    }
    close(result_fd);

    int rc_extra = 0;
    const char* refs_fa = getenv("EREF_REFS_FASTA");
    const char* refs_pc = getenv("EREF_REFS_PERCENT");
    if (refs_fa || refs_pc) {
        std::string werr;
        if (!write_refs(refs_fa ? refs_fa : "", refs_pc ? refs_pc : "", fa.p, fa.n, merged, host_threads, werr)) {
            fprintf(stderr, "eref: %s\n", werr.c_str());
            rc_extra = 3;                             // asked-for outputs are missing: the pipeline must not go on
        }
    }

    if (getenv("EREF_WRITE_INDEX") && access(index_name.c_str(), F_OK) != 0) {
        std::string werr;
        for (eref_ctx*& c : ctx) { if (c) eref_destroy(c); c = nullptr; }     // hand the memory back first
        if (!write_index_files(index_name, fasta + ".genome.len.txt", fa.p, fa.n, coder, werr))
            fprintf(stderr, "eref: warning: index not written: %s\n", werr.c_str());   // the results above stand
    }

    if (timing) {
        const auto t_end = Clock::now();
        fprintf(stderr, "eref timing: contexts%s %.1f ms | files read%s after %.1f ms | page-locking %.1f ms | gpu+merge %.1f ms | "
                        "output %.1f ms | total %.1f ms, shards %d%s\n",
                shards > 1 ? " + NCCL" : "", ms_between(t_start, t_ctx),
                (f1.was_gzip || f2.was_gzip || fa.was_gzip) ? " + inflated" : "", ms_between(t_start, t_read),
                ms_between(t_read, t_pin), ms_between(t_pin, t_gpu), ms_between(t_gpu, t_end), ms_between(t_start, t_end), shards,
                shard_table ? ", table build sharded" : "");
        for (int i = 0; i < shards; i++) {
            const eref_phase_times& t = times[i];
            fprintf(stderr,
                    "  gpu %d: fq2 h2d %.3f pack %.3f bucket %.3f count %.3f | ref h2d %.3f frame %.3f pack %.3f | scan summary %.3f sample %.3f cand %.3f "
                    "probe %.3f window %.3f intervals %.3f d2h %.3f ms | bases %llu need_words %llu/%llu launches %llu\n",
                    i, t.query_h2d_ms, t.query_pack_ms, t.query_bucket_ms, t.query_count_ms, t.ref_h2d_ms, t.ref_frame_ms, t.ref_pack_ms,
                    t.scan_summary_ms, t.scan_sample_ms, t.scan_cand_ms, t.scan_probe_ms, t.scan_window_ms, t.scan_interval_ms, t.scan_d2h_ms,
                    (unsigned long long)t.ref_bases, (unsigned long long)t.ref_need_words, (unsigned long long)t.ref_words,
                    (unsigned long long)t.kernel_launches);
        }
    }
    cleanup();
    return rc_extra;
}
