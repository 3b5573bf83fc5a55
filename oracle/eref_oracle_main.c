/*
 * eref_oracle_main.c — command-line front end of the CPU restatement, same seven positional
 * arguments as the reference (bin/extract_ref.cpp:1221-1231).  TEST INFRASTRUCTURE ONLY.
 * Coder: read from <fasta>.k32.index.dat when present, else EREF_CODER_SEED (default 1).
 */
#include "eref_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static char* slurp(const char* path, size_t* n)
{
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "eref_oracle: cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    char* b = (char*)malloc((size_t)sz + 1);
    if (fread(b, 1, (size_t)sz, f) != (size_t)sz) { fprintf(stderr, "short read %s\n", path); exit(2); }
    fclose(f); *n = (size_t)sz; return b;
}

int main(int argc, char** argv)
{
    if (argc < 8) {
        fprintf(stderr, "usage: eref_oracle fq1 fq2 phagedb.fasta interval_out hit_ratio perfect_hit_ratio threads\n");
        return 2;
    }
    size_t n1, n2, nf;
    char* fq1 = slurp(argv[1], &n1); char* fq2 = slurp(argv[2], &n2); char* fa = slurp(argv[3], &nf);
    float hr = (float)strtod(argv[5], NULL), phr = (float)strtod(argv[6], NULL);
    int threads = (int)strtod(argv[7], NULL);
    FILE* t = fopen(argv[4], "w"); if (t) fclose(t);                 /* :825 created, left empty */
    int16_t coder[96];
    char idx[4096]; snprintf(idx, sizeof idx, "%s.k32.index.dat", argv[3]);
    if (eo_read_coder(idx, coder) != 0) {
        const char* s = getenv("EREF_CODER_SEED");
        eo_random_coder(s ? (unsigned)strtoul(s, NULL, 10) : 1u, coder);
        if (getenv("EREF_ORACLE_WRITE_INDEX")) eo_write_index(idx, coder, fa, nf);
    }
    eo_result* res; size_t nres;
    if (eo_run(fq1, n1, fq2, n2, fa, nf, coder, hr, phr, threads, &res, &nres)) return 3;
    char line[256];
    for (size_t i = 0; i < nres; i++)
        if (eo_format_line(&res[i], line, sizeof line) > 0) fputs(line, stdout);
    return 0;
}
