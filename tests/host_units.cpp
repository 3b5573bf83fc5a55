// host_units.cpp — TEST harness (CPU only): the host-side helpers of the drop-in executable
// (palace_b200/csrc/eref_main.cpp: input reading + gzip/BGZF inflation, coder header, record sharding, record names,
// the fused next step of SURVEY.md §8(f) row f1) behind a small command line, so that tests/test_host_units.py can
// drive them without a GPU, under AddressSanitizer/UBSan.  The translation unit of the executable is included as it
// is (its main renamed), nothing is re-implemented here; no command below touches CUDA or a compute entry point.
#define main eref_executable_main
#include "../palace_b200/csrc/eref_main.cpp"
#undef main

static int usage()
{
    fprintf(stderr, "usage: host_units inflate <file> <threads> | shards <fasta> <n> | headers <fasta> <threads> <k>... |\n"
                    "       name | pyrepr | coder <seed> | index_header <file> | sidecar <file> |\n"
                    "       refs <fasta> <records.txt> <fasta_out|-> <percent_out|-> <threads>\n");
    return 2;
}

int main(int argc, char** argv)
{
    if (argc < 2) return usage();
    const std::string cmd = argv[1];
    if (cmd == "inflate" && argc == 4) {                 // read_file: plain, gzip (1..n members), BGZF -> stdout
        HostFile f;
        if (!read_file(argv[2], f, atoi(argv[3]))) { fprintf(stderr, "%s\n", f.err.c_str()); return 1; }
        fprintf(stderr, "gzip=%d bytes=%zu\n", (int)f.was_gzip, f.n);
        return fwrite(f.p, 1, f.n, stdout) == f.n ? 0 : 1;
    }
    if (cmd == "shards" && argc == 4) {                  // shard_offsets
        HostFile f;
        if (!read_file(argv[2], f, 1)) { fprintf(stderr, "%s\n", f.err.c_str()); return 1; }
        for (size_t c : shard_offsets(f.p, f.n, atoi(argv[3]))) printf("%zu\n", c);
        return 0;
    }
    if (cmd == "headers" && argc >= 4) {                 // header_offsets of the wanted 1-based header ordinals
        HostFile f;
        if (!read_file(argv[2], f, 1)) { fprintf(stderr, "%s\n", f.err.c_str()); return 1; }
        std::vector<uint32_t> want;
        for (int i = 4; i < argc; i++) want.push_back((uint32_t)strtoul(argv[i], nullptr, 10));
        for (size_t o : header_offsets(f.p, f.n, want, atoi(argv[3]))) printf("%lld\n", o == (size_t)-1 ? -1ll : (long long)o);
        return 0;
    }
    if (cmd == "name") {                                 // record_name of every stdin line
        std::string line;
        while (std::getline(std::cin, line)) printf("%s\n", record_name(line.data(), line.size()).c_str());
        return 0;
    }
    if (cmd == "pyrepr") {                               // py_float_repr of every stdin line
        std::string line;
        while (std::getline(std::cin, line)) printf("%s\n", py_float_repr(line.c_str()).c_str());
        return 0;
    }
    if (cmd == "coder" && argc == 3) {                   // draw_coder
        int16_t coder[96];
        draw_coder((unsigned)strtoul(argv[2], nullptr, 10), coder);
        for (int j = 0; j < 96; j++) printf("%d%c", coder[j], j == 95 ? '\n' : ' ');
        return 0;
    }
    if ((cmd == "index_header" || cmd == "sidecar") && argc == 3) {
        int16_t coder[96];
        if (!(cmd == "sidecar" ? read_sidecar(argv[2], coder) : read_index_header(argv[2], coder))) return 1;
        for (int j = 0; j < 96; j++) printf("%d%c", coder[j], j == 95 ? '\n' : ' ');
        return 0;
    }
    if (cmd == "refs" && argc == 7) {                    // write_refs on records "ordinal frag el ref_len" per line
        HostFile f;
        if (!read_file(argv[2], f, 1)) { fprintf(stderr, "%s\n", f.err.c_str()); return 1; }
        std::vector<eref_result> rec;
        std::ifstream in(argv[3]);
        eref_result r;
        while (in >> r.ordinal >> r.frag >> r.el >> r.ref_len) rec.push_back(r);
        const std::string fo = strcmp(argv[4], "-") ? argv[4] : "", po = strcmp(argv[5], "-") ? argv[5] : "";
        std::string err;
        if (!write_refs(fo, po, f.p, f.n, rec, atoi(argv[6]), err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        // the result lines as the executable prints them (eref_format_line is host code of the library)
        char line[256];
        for (const eref_result& q : rec) { int k = eref_format_line(&q, line, sizeof line); if (k > 0) fwrite(line, 1, (size_t)k, stdout); }
        return 0;
    }
    return usage();
}
