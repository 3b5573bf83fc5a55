#!/usr/bin/env python
"""Test infrastructure: the C restatement against the UNMODIFIED reference binary (oracle/_ref/eref_ref, threads=1) on
seeded inputs whose FORMAT is mangled — the things real files do and a clean generator does not: CRLF line ends, blank
lines, a missing final newline, a truncated last record, sequence in front of the first header, header-only records,
'>' and white space inside sequence lines, lower case, fq2 longer than fq1.  Each case applies a random subset of the
manglings to a clean synthetic sample and compares the index body (every code of every position) and stdout under two
threshold pairs.  Stays clear of the reference's
undefined behaviour: FASTQ lines <= 400 characters, no records of 450..499 bases.
usage: python oracle/fuzz_formats.py [first_seed n_cases]   (each reference run: 15-30 s, ~20 GiB of host memory)"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_util  # noqa: E402
from palace_b200 import eref, synth  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "eref_ref")


def mangle_fasta(text: bytes, rng, log):
    recs = [b">" + r for r in text.split(b"\n>")]
    recs[0] = recs[0][1:]
    recs = [r if r.endswith(b"\n") else r + b"\n" for r in recs]
    n = len(recs)

    def pick():
        return int(rng.integers(0, n))
    if rng.random() < 0.6:                                   # blank lines between the sequence lines of a record
        k = pick(); lines = recs[k].split(b"\n")
        for p in sorted(rng.integers(1, len(lines), 3), reverse=True):
            lines.insert(int(p), b"")
        recs[k] = b"\n".join(lines); log.append(f"blank lines in record {k}")
    if rng.random() < 0.6:                                   # CRLF on every line of a few records
        for k in {pick() for _ in range(3)}:
            recs[k] = recs[k].replace(b"\n", b"\r\n")
        log.append("CRLF records")
    if rng.random() < 0.5:                                   # trailing blanks / tabs on some sequence lines of a record
        k = pick(); lines = recs[k].split(b"\n")
        for p in rng.integers(1, max(len(lines) - 1, 2), 4):
            if 0 < p < len(lines) - 1:
                lines[int(p)] += b" \t"[int(rng.integers(0, 2)):][:1]
        recs[k] = b"\n".join(lines); log.append(f"white space in record {k}")
    if rng.random() < 0.5:                                   # '>' inside a sequence line
        k = pick(); lines = recs[k].split(b"\n")
        if len(lines) > 3:
            p = int(rng.integers(1, len(lines) - 1)); q = int(rng.integers(1, max(len(lines[p]) - 1, 2)))
            lines[p] = lines[p][:q] + b">" + lines[p][q + 1:]
            recs[k] = b"\n".join(lines); log.append(f"'>' inside record {k}")
    if rng.random() < 0.5:                                   # header-only records (no ordinal: length 0)
        for _ in range(2):
            recs.insert(pick(), b">empty_record with words\n")
        log.append("header-only records")
    if rng.random() < 0.5:                                   # sequence in front of the first header (record 0, 'start')
        L = int(rng.choice([20, 33, 200, 1500]))
        recs.insert(0, bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), L)) + b"\n"); log.append(f"{L} bases before the first header")
    out = b"".join(recs)
    if rng.random() < 0.4:
        out = out.rstrip(b"\r\n"); log.append("FASTA without final newline")
    return out


def mangle_fastq(text: bytes, rng, log, which):
    lines = text.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    if rng.random() < 0.4:                                   # lower-case reads
        for k in range(1, len(lines), 4):
            if rng.random() < 0.1:
                lines[k] = lines[k].lower()
        log.append(f"{which}: lower-case reads")
    if rng.random() < 0.3:                                   # a blank line shifts the 4-line phase of everything behind it
        p = int(len(lines) * rng.uniform(0.8, 0.98)); lines.insert(p, b""); log.append(f"{which}: blank line at line {p}")
    if rng.random() < 0.3:                                   # the last record is cut off
        lines = lines[: len(lines) - int(rng.integers(1, 4))]; log.append(f"{which}: truncated last record")
    eol = b"\n"
    if rng.random() < 0.3:
        eol = b"\r\n"; log.append(f"{which}: CRLF")
    out = eol.join(lines) + eol
    if rng.random() < 0.4:
        out = out[: -len(eol)]; log.append(f"{which}: no final newline")
    return out


def main():
    first, n = (int(x) for x in (sys.argv[1:3] + ["900", "8"][len(sys.argv) - 1:]))
    oracle = oracle_util.load()
    bad = lines = 0
    for case in range(n):
        seed = first + case
        rng = np.random.default_rng(seed)
        db = synth.make_db(60, 1_200_000, seed, min_len=3000, max_len=60000, n_short=int(rng.integers(0, 8)))
        f1, f2, _ = synth.make_reads(db, 300_000, seed + 1000, cov_lo=4.0, cov_hi=8.0, max_len=399)
        log = []
        fa = np.frombuffer(mangle_fasta(db.text.tobytes(), rng, log), np.uint8)
        f1 = np.frombuffer(mangle_fastq(f1.tobytes(), rng, log, "fq1"), np.uint8)
        f2 = np.frombuffer(mangle_fastq(f2.tobytes(), rng, log, "fq2"), np.uint8)
        if rng.random() < 0.4:                               # fq2 clearly longer: main's byte budget cuts it (:1265-1291)
            f2 = np.concatenate([f2, f1[: f1.size // 3]]); log.append("fq2 longer than fq1")
        with tempfile.TemporaryDirectory() as d:
            fa.tofile(os.path.join(d, "db.fa")); f1.tofile(os.path.join(d, "r1.fq")); f2.tofile(os.path.join(d, "r2.fq"))
            for hr, phr in (("0.9", "0.85"), ("0.5", "0.3")):
                r = subprocess.run([REF, "r1.fq", "r2.fq", "db.fa", "t.txt", hr, phr, "1"], cwd=d, capture_output=True)
                if r.returncode != 0:
                    print(f"seed {seed}: reference exit {r.returncode} ({'; '.join(log)})", flush=True)
                    bad += 1
                    break
                ref = r.stdout.decode()
                coder = eref.read_coder(os.path.join(d, "db.fa.k32.index.dat"))
                if hr == "0.9":
                    # the reference's own index of the mangled FASTA (:652-811): record lengths and every code of every
                    # position, byte for byte
                    oracle.write_index(os.path.join(d, "mine.dat"), coder, fa)
                    a = np.fromfile(os.path.join(d, "db.fa.k32.index.dat"), "<u4")[100:]
                    b = np.fromfile(os.path.join(d, "mine.dat"), "<u4")[100:]
                    same = a.size == b.size and bool((a == b).all())
                    bad += not same
                    print(f"seed {seed}: index body of {a.size} words {'identical' if same else 'DIFFERENT'}", flush=True)
                mine = oracle.format(oracle.run(f1, f2, fa, coder, hit_ratio=float(hr), perfect_hit_ratio=float(phr)))
                bad += mine != ref
                lines += ref.count("\n")
                print(f"seed {seed} {hr}/{phr}: {ref.count(chr(10))} lines {'identical' if mine == ref else 'DIFFERENT'}  [{'; '.join(log)}]",
                      flush=True)
                if mine != ref:
                    a, b = ref.splitlines(), mine.splitlines()
                    for x in a:
                        if x not in b:
                            print("   reference only:", x)
                    for x in b:
                        if x not in a:
                            print("   oracle only   :", x)
    print(f"{lines} result lines compared in all")
    print("all identical" if not bad else f"{bad} case(s) differ")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
