#!/usr/bin/env python
"""Generate the golden fixtures of tests/golden/ by running the UNMODIFIED reference.

Needs oracle/_ref/eref_ref (built by `make -C oracle` from /root/reference/bin/extract_ref.cpp) and
~24 GB of host memory; each reference run costs 15-30 s.  Run from the repo root:

    python tests/golden/make_golden.py

For every case it writes <case>.{fa,r1.fq,r2.fq}.gz (the exact input bytes) and <case>.json:
the coder header the reference drew (time-seeded, bin/extract_ref.cpp:1088), sha256 of the body of
the reference's own <db>.k32.index.dat (every code of every position), and the reference's stdout for
several (hit_ratio, perfect_hit_ratio) pairs at threads=1.
"""
import gzip
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from palace_b200 import synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "eref_ref")
OUT = os.path.dirname(os.path.abspath(__file__))


def edge_case():
    """Hand-made inputs around the framing rules."""
    rng = np.random.default_rng(77)
    acgt = np.frombuffer(b"ACGT", np.uint8)

    def rnd(n):
        return acgt[rng.integers(0, 4, n)].tobytes()

    g_long = rnd(4000)
    g_mid = rnd(1500)
    fa = b""
    fa += rnd(40) + b"\n"                                    # sequence before the first header: record 0
    fa += b">r33 exactly 33\n" + rnd(33) + b"\n"
    fa += b">r32 exactly 32 (dropped, no ordinal)\n" + rnd(32) + b"\n"
    fa += b">empty\n"
    fa += b">long wrapped at 61 with blank lines\n"
    for i in range(0, len(g_long), 61):
        fa += g_long[i:i + 61] + b"\n"
        if i % 610 == 0:
            fa += b"\n"
    fa += b">mid one line, lower case and N and a '>' inside\n"
    m = bytearray(g_mid)
    m[100:160] = bytes(m[100:160]).lower()
    m[700:710] = b"N" * 10
    m[900:901] = b">"
    fa += bytes(m) + b"\n"
    fa += b">crlf record\r\n" + g_long[:700] + b"\r\n" + g_long[700:1400] + b"\r\n"
    fa += b">last without trailing newline\n" + g_long[1000:3000]
    # reads: tile g_long and g_mid densely so they are selected
    recs = []
    rid = 0
    for src in (g_long, bytes(m).upper().replace(b">", b"A")):
        for rep in range(5):
            for s in range(rep * 13, len(src) - 40, 90):
                seq = src[s:s + 150]
                recs.append(seq)
    recs.append(rnd(32)); recs.append(rnd(31)); recs.append(b"ACGTN" * 20); recs.append(b"")
    recs.append(rnd(400)); recs.append(rnd(100).lower())
    out1, out2 = b"", b""
    for k, seq in enumerate(recs):
        rec = b"@e%d\n" % rid + seq + b"\n+\n" + b"I" * len(seq) + b"\n"
        rid += 1
        if k % 2 == 0:
            out1 += rec
        else:
            out2 += rec
    out2 = out2[:-1]                                         # fq2 ends without a newline
    return (np.frombuffer(fa, np.uint8), np.frombuffer(out1, np.uint8), np.frombuffer(out2, np.uint8))


def cases():
    db = synth.make_db(60, 600_000, 1101, min_len=3000, max_len=30000, frac_mutated=0.1, frac_n=0.1,
                       frac_lower=0.05, n_short=3)
    f1, f2, _ = synth.make_reads(db, 450_000, 2101, cov_lo=1.5, cov_hi=6.0)
    yield "basic", db.text, f1, f2, [(0.9, 0.85), (0.5, 0.3), (0.95, 0.95), (0.2, 0.8)]
    db = synth.make_db(30, 300_000, 1102, min_len=3000, max_len=30000, frac_mutated=0.1, n_short=2, width=60)
    f1, f2, _ = synth.make_reads(db, 300_000, 2102, min_len=20, max_len=150, cov_lo=4, cov_hi=12)
    yield "short_reads", db.text, f1, f2, [(0.9, 0.85), (0.6, 0.5)]
    fa, f1, f2 = edge_case()
    yield "edge", fa, f1, f2, [(0.9, 0.85), (0.7, 0.6)]


def main():
    if not os.path.exists(REF):
        sys.exit("build oracle/_ref/eref_ref first: make -C oracle")
    for name, fa, f1, f2, ratios in cases():
        tmp = tempfile.mkdtemp(prefix="golden_")
        try:
            fa.tofile(os.path.join(tmp, "db.fa")); f1.tofile(os.path.join(tmp, "r1.fq")); f2.tofile(os.path.join(tmp, "r2.fq"))
            outs = {}
            for hr, phr in ratios:                       # first run builds the index (cold), the rest reuse it
                r = subprocess.run([REF, "r1.fq", "r2.fq", "db.fa", "tmp.txt", str(hr), str(phr), "1"], cwd=tmp,
                                   capture_output=True, check=True)
                outs[f"{hr},{phr}"] = r.stdout.decode()
                print(name, hr, phr, "->", len(r.stdout.splitlines()), "lines", flush=True)
            idx = np.fromfile(os.path.join(tmp, "db.fa.k32.index.dat"), dtype="<u4")
            meta = {
                "coder": [int(x) for x in (idx[:96] & 0xFFFF)],
                "index_words": int(idx.size),
                "index_body_sha256": hashlib.sha256(idx[100:].tobytes()).hexdigest(),
                "genome_len_txt": open(os.path.join(tmp, "db.fa.genome.len.txt")).read(),
                "stdout": outs,
                "sha256": {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in (("fa", fa), ("r1", f1), ("r2", f2))},
            }
            for k, v in (("fa", fa), ("r1.fq", f1), ("r2.fq", f2)):
                with gzip.GzipFile(os.path.join(OUT, f"{name}.{k}.gz"), "wb", compresslevel=9, mtime=0) as g:
                    g.write(v.tobytes())
            with open(os.path.join(OUT, f"{name}.json"), "w") as f:
                json.dump(meta, f, indent=1)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
