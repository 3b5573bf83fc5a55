// eref_main.cpp — the drop-in `eref` executable for the PALACE pipeline (palace:475-477):
//
//   eref <fq1> <fq2> <phagedb.fasta> <interval_out> <hit_ratio> <perfect_hit_ratio> <threads>  > ref_names.txt
//
// Same seven positional arguments, same stdout lines and side-effect files as
// bin/extract_ref.cpp:1210-1334; all sequence work runs on B200s through the C ABI of
// liberef_gpu.so (include/eref_gpu.h).  There is no CPU path: without a GPU it exits non-zero.
//
// Environment (the palace call line stays unchanged):
//   EREF_GPUS=N|all     GPUs to shard the database over (default: one per 128 MiB of FASTA, <= visible)
//   EREF_CODER_SEED=S   seed of the coder draw when <db>.k32.index.dat does not exist (default time(0), :1088)
//   EREF_NO_PRUNE=1     probe all three streams everywhere (disables the exact candidate filter)
//   EREF_DIRECT_COUNT=1 count k-mers straight into the table (disables the code-range bucketing)
//   EREF_TIMING=1       per-phase CUDA-event times on stderr
//   EREF_WRITE_INDEX=1  also write <db>.k32.index.dat + <db>.genome.len.txt in the reference's format (when absent),
//                       so that a CPU eref can share the DB directory (12 bytes per base: off by default)
#include "eref_gpu.h"
#include <cuda_runtime.h>
#include <nccl.h>
#include <unistd.h>
#include <zlib.h>
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <string>
#include <thread>
#include <vector>
#include <condition_variable>
#include <mutex>

namespace {

struct HostFile {                       // whole file in pinned host memory
    char* p = nullptr;
    size_t n = 0;
    bool pinned = false;
    ~HostFile() { if (p) { if (pinned) cudaFreeHost(p); else free(p); } }
};

// a gzip (or bgzip, or multi-member) file is inflated on the way in; the reference itself only
// reads plain text, so for plain files nothing changes.  The text ends up in pinned memory like any other input.
bool read_gzip(const std::string& path, HostFile& f, std::string& err)
{
    gzFile gz = gzopen(path.c_str(), "rb");
    if (!gz) { err = "cannot open " + path + ": " + strerror(errno); return false; }
    gzbuffer(gz, 1 << 20);
    std::vector<char> buf;
    size_t got = 0;
    for (;;) {
        if (buf.size() - got < (1u << 24)) buf.resize(buf.size() + buf.size() / 2 + (1u << 26));
        const int r = gzread(gz, buf.data() + got, (unsigned)std::min<size_t>(buf.size() - got, 1u << 30));
        if (r < 0) { int e = 0; err = path + ": " + gzerror(gz, &e); gzclose(gz); return false; }
        if (r == 0) break;
        got += (size_t)r;
    }
    gzclose(gz);
    void* q = nullptr;
    if (cudaHostAlloc(&q, got + 64, cudaHostAllocDefault) == cudaSuccess) f.pinned = true;
    else { cudaGetLastError(); q = malloc(got + 64); }
    if (!q) { err = "out of memory reading " + path; return false; }
    memcpy(q, buf.data(), got);
    f.p = (char*)q; f.n = got;
    return true;
}

bool read_file(const std::string& path, HostFile& f, std::string& err)
{
    FILE* fp = fopen(path.c_str(), "rb");
    if (!fp) { err = "cannot open " + path + ": " + strerror(errno); return false; }
    unsigned char magic[2] = {0, 0};
    const size_t mg = fread(magic, 1, 2, fp);
    if (mg == 2 && magic[0] == 0x1f && magic[1] == 0x8b) { fclose(fp); return read_gzip(path, f, err); }
    fseek(fp, 0, SEEK_END);
    long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    if (sz < 0) { fclose(fp); err = "cannot size " + path; return false; }
    f.n = (size_t)sz;
    void* q = nullptr;
    if (cudaHostAlloc(&q, f.n + 64, cudaHostAllocDefault) == cudaSuccess) { f.pinned = true; }
    else { cudaGetLastError(); q = malloc(f.n + 64); }
    if (!q) { fclose(fp); err = "out of memory reading " + path; return false; }
    f.p = (char*)q;
    size_t got = 0;
    while (got < f.n) {
        size_t r = fread(f.p + got, 1, f.n - got, fp);
        if (!r) break;
        got += r;
    }
    fclose(fp);
    if (got != f.n) { err = "short read on " + path; return false; }
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
bool write_index_files(const std::string& index_name, const std::string& len_name, const char* fa, size_t n,
                       const int16_t coder[96], std::string& err)
{
    const std::string tmp = index_name + ".part";
    FILE* fi = fopen(tmp.c_str(), "wb");
    FILE* fl = fopen(len_name.c_str(), "w");
    if (!fi || !fl) { err = "cannot create " + (fi ? len_name : tmp); if (fi) fclose(fi); if (fl) fclose(fl); return false; }
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
    if (ok && rename(tmp.c_str(), index_name.c_str()) != 0) { err = "cannot rename " + tmp; ok = false; }
    if (!ok) remove(tmp.c_str());
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
    try {
        hit_ratio = (float)std::stod(argv[5]);               // :1228
        perfect_hit_ratio = (float)std::stod(argv[6]);       // :1229
        (void)std::stod(argv[7]);                            // :1231 threads: a hint, unused on the GPU
    } catch (const std::exception& e) {
        fprintf(stderr, "eref: bad numeric argument: %s\n", e.what());
        return 2;
    }
    // stdout is the result channel (palace redirects it into ref_names.txt): keep libraries (e.g. NCCL's
    // version banner) off it by parking the real stdout and pointing fd 1 at stderr until the results are ready
    fflush(stdout);
    const int result_fd = dup(1);
    if (result_fd < 0 || dup2(2, 1) < 0) { fprintf(stderr, "eref: cannot set up the result channel\n"); return 3; }
    const bool timing = getenv("EREF_TIMING") != nullptr;
    auto t_start = std::chrono::steady_clock::now();

    int ndev = eref_device_count();
    if (ndev <= 0) { fprintf(stderr, "eref: no CUDA device visible (this build has no CPU path)\n"); return 3; }

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

    // ---- inputs ----
    HostFile f1, f2, fa;
    std::string err;
    if (!read_file(fq1, f1, err) || !read_file(fq2, f2, err) || !read_file(fasta, fa, err)) {
        fprintf(stderr, "eref: %s\n", err.c_str());
        return 3;
    }
    auto t_read = std::chrono::steady_clock::now();

    int shards = (int)std::min<size_t>((size_t)ndev, std::max<size_t>(1, fa.n >> 27));
    if (const char* g = getenv("EREF_GPUS")) {
        shards = strcmp(g, "all") == 0 ? ndev : atoi(g);
        if (shards < 1 || shards > ndev) { fprintf(stderr, "eref: EREF_GPUS=%s but %d device(s) visible\n", g, ndev); return 3; }
    }
    std::vector<size_t> cut = shard_offsets(fa.p, fa.n, shards);
    const unsigned flags = (getenv("EREF_NO_PRUNE") ? EREF_FLAG_NO_PRUNE : 0u) |
                           (getenv("EREF_DIRECT_COUNT") ? EREF_FLAG_DIRECT_COUNT : 0u) |
                           (coder_drawn_here ? EREF_FLAG_KEEP_RAND_STATE : 0u);
    uint64_t down_sampling_size = 2000000000ull;         // :1230
    if (const char* v = getenv("EREF_DOWN_SAMPLING_SIZE")) { long long k = atoll(v); if (k > 0) down_sampling_size = (uint64_t)k; }
    std::vector<uint32_t> keep1, keep2;                  // down-sampling masks, drawn once by shard 0
    uint64_t lines1 = 0, lines2 = 0;

    std::vector<ncclComm_t> comms(shards);
    if (shards > 1) {
        std::vector<int> devs(shards);
        for (int i = 0; i < shards; i++) devs[i] = i;
        ncclResult_t r = ncclCommInitAll(comms.data(), shards, devs.data());
        if (r != ncclSuccess) { fprintf(stderr, "eref: ncclCommInitAll: %s\n", ncclGetErrorString(r)); return 3; }
    }

    std::vector<size_t> kept(shards, 0);
    std::vector<std::string> errors(shards);
    std::vector<eref_phase_times> times(shards);
    std::vector<eref_result> merged;                 // filled by shard 0
    std::atomic<int> failed{0};
    Barrier bar(shards);
    size_t max_kept = 0;
    bool go = false, single_done = false;

    auto worker = [&](int i) {
        eref_ctx* ctx = nullptr;
        auto fail = [&](const char* what) {
            errors[i] = std::string(what) + ": " + eref_last_error();
            failed.store(1);
        };
        bool ok = true;
        if (eref_create(&ctx, i, coder, hit_ratio, perfect_hit_ratio, flags)) { fail("eref_create"); ok = false; }
        if (ok && shards == 1) {
            // one GPU: the pipelined entry point (copies overlapped with kernels, DB in 4 pieces)
            size_t cap = 1;
            for (size_t k = 0; k < fa.n; k++) cap += fa.p[k] == '>';
            merged.resize(cap);
            size_t n = 0;
            if (eref_run_host(ctx, f1.p, f1.n, f2.p, f2.n, fa.p, fa.n, 4, merged.data(), cap, &n)) { fail("eref_run_host"); ok = false; }
            merged.resize(ok ? std::min(n, cap) : 0);
            single_done = ok;
        }
        if (ok && shards > 1 && eref_query_add(ctx, f1.p, f1.n)) { fail("eref_query_add(fq1)"); ok = false; }
        int ratio = 100;
        if (ok && shards > 1) {
            // cal_sam_ratio (:1124-1148): 100 * 2e9 / (2 * sequence bytes of fq1), integer division
            uint64_t seq = 0;
            eref_query_seq_bytes(ctx, &seq);
            if (eref_sample_ratio(seq, down_sampling_size, &ratio)) { fail("cal_sam_ratio"); ok = false; }
        }
        if (shards > 1) {
            // down-sampling (:955-960): shard 0 replays the reference's rand() draws once, every shard counts under them
            if (ok && ratio < 100 && i == 0) {
                uint64_t l1 = 0, l2 = 0;
                if (eref_query_census(ctx, f1.p, f1.n, nullptr, &l1) || eref_query_census(ctx, f2.p, f2.n, nullptr, &l2)) {
                    fail("eref_query_census"); ok = false;
                } else {
                    if (!coder_drawn_here) srand(1);               // :1236-1238
                    keep1.assign((l1 + 31) / 32 + 1, 0); keep2.assign((l2 + 31) / 32 + 1, 0);
                    eref_draw_keep(ratio, l1, keep1.data()); eref_draw_keep(ratio, l2, keep2.data());
                    lines1 = l1; lines2 = l2;
                }
            }
            bar.wait();                                   // the masks are drawn (or nobody needs them)
            if (ok && ratio < 100) {
                if (failed.load()) ok = false;
                else if (eref_query_reset(ctx) || eref_query_add_masked(ctx, f1.p, f1.n, keep1.data(), lines1) ||
                         eref_query_add_masked(ctx, f2.p, f2.n, keep2.data(), lines2)) { fail("eref_query_add_masked"); ok = false; }
            } else if (ok && eref_query_add(ctx, f2.p, f2.n)) { fail("eref_query_add(fq2)"); ok = false; }
        }
        if (ok && shards > 1 && eref_ref_set(ctx, fa.p + cut[i], cut[i + 1] - cut[i])) { fail("eref_ref_set"); ok = false; }
        if (ok && shards > 1) eref_ref_count(ctx, &kept[i], nullptr, nullptr);
        bar.wait();                                   // every shard's record count is known
        if (i == 0) { max_kept = *std::max_element(kept.begin(), kept.end()); go = failed.load() == 0; }
        bar.wait();
        const bool all_ok = go;                       // one decision for every shard: all enter the collective or none
        if (shards == 1) {
            (void)single_done;                        // already scanned by eref_run_host
        } else if (all_ok) {
            // one collective: allgather of fixed-size records, shards padded to max_kept
            const uint32_t* dres = nullptr;
            size_t n = 0;
            void *send = nullptr, *recv = nullptr;
            cudaStream_t st = nullptr;
            const size_t slot = std::max<size_t>(max_kept, 1) * sizeof(eref_result);
            bool good = eref_scan_device(ctx, &dres, &n) == 0;
            if (!good) fail("eref_scan_device");
            good = good && cudaSetDevice(i) == cudaSuccess && cudaStreamCreate(&st) == cudaSuccess &&
                   cudaMalloc(&send, slot) == cudaSuccess && cudaMalloc(&recv, slot * shards) == cudaSuccess &&
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
            if (send) cudaFree(send);
            if (recv) cudaFree(recv);
            if (st) cudaStreamDestroy(st);
        }
        if (ctx) { eref_timings(ctx, &times[i]); eref_destroy(ctx); }
    };

    std::vector<std::thread> pool;
    for (int i = 0; i < shards; i++) pool.emplace_back(worker, i);
    for (auto& t : pool) t.join();
    if (shards > 1) for (auto& c : comms) ncclCommDestroy(c);

    if (failed.load()) {
        for (int i = 0; i < shards; i++)
            if (!errors[i].empty()) fprintf(stderr, "eref: gpu %d: %s\n", i, errors[i].c_str());
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
        if (w <= 0) { fprintf(stderr, "eref: write to stdout failed\n"); return 3; }
        done += (size_t)w;
    }
    close(result_fd);

    if (getenv("EREF_WRITE_INDEX") && access(index_name.c_str(), F_OK) != 0) {
        std::string werr;
        if (!write_index_files(index_name, fasta + ".genome.len.txt", fa.p, fa.n, coder, werr))
            fprintf(stderr, "eref: warning: index not written: %s\n", werr.c_str());   // the results above stand
    }

    if (timing) {
        auto t_end = std::chrono::steady_clock::now();
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "eref timing: read %.1f ms, gpu+merge %.1f ms, shards %d\n", ms(t_start, t_read), ms(t_read, t_end), shards);
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
    return 0;
}
