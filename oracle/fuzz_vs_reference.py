#!/usr/bin/env python
"""Test infrastructure: the C restatement (oracle/liberef_oracle.so) against the UNMODIFIED reference binary
(oracle/_ref/eref_ref, built by `make -C oracle ref` where /root/reference exists) on seeded synthetic inputs of
several shapes — stdout byte for byte, reference at threads=1.  Every case is compared under TWO threshold pairs on the
same index: the pipeline's 0.9/0.85 and a loose one, every third case with fq2 longer than fq1 (main's byte budget
cuts it, :1265-1291), and the reads cover their genomes 4-8x, so that every case prints
lines (a comparison of two empty outputs pins nothing: the run fails if a case prints none under its loose pair).  Each
reference run takes 15-30 s and ~20 GiB of host memory.  usage: python oracle/fuzz_vs_reference.py [first_seed n_cases]"""
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
SHAPES = [  # genomes, DB bases, query bases, short-read-heavy
    (40, 1_000_000, 250_000, False), (150, 3_000_000, 400_000, False), (25, 2_000_000, 600_000, True),
    (300, 2_500_000, 300_000, False)]
LOOSE = [("0.8", "0.5"), ("0.3", "0.2"), ("0.6", "0.55"), ("0.5", "0.3")]


def main():
    first, n = (int(x) for x in (sys.argv[1:3] + ["500", "8"][len(sys.argv) - 1:]))
    oracle = oracle_util.load()
    bad = lines = 0
    for case in range(n):
        seed = first + case
        G, R, Q, short = SHAPES[case % len(SHAPES)]
        db = synth.make_db(G, R, seed)
        kw = {"min_len": 20, "max_len": 150} if short else {}        # BASELINE configs[4]: many reads under 32 bases
        f1, f2, _ = synth.make_reads(db, Q, seed + 1000, cov_lo=4.0, cov_hi=8.0, **kw)
        if case % 3 == 2:
            # fq2 clearly longer than fq1: main's byte budget (file_size(fq1) for both files, :1265-1291) cuts its tail
            nl = np.flatnonzero(f1 == 10)
            cut = int(nl[4 * (nl.size // 4 * 3 // 4) - 1]) + 1
            f1, f2 = f1[:cut].copy(), np.concatenate([f2, f1[cut:]])
        with tempfile.TemporaryDirectory() as d:
            db.text.tofile(os.path.join(d, "db.fa")); f1.tofile(os.path.join(d, "r1.fq")); f2.tofile(os.path.join(d, "r2.fq"))
            for k, (hr, phr) in enumerate([("0.9", "0.85"), LOOSE[case % len(LOOSE)]]):
                ref = subprocess.run([REF, "r1.fq", "r2.fq", "db.fa", "t.txt", hr, phr, "1"], cwd=d, capture_output=True, check=True).stdout.decode()
                coder = eref.read_coder(os.path.join(d, "db.fa.k32.index.dat"))
                mine = oracle.format(oracle.run(f1, f2, db.text, coder, hit_ratio=float(hr), perfect_hit_ratio=float(phr)))
                ok = mine == ref and (k == 0 or ref.count("\n") > 0)
                bad += not ok
                lines += ref.count("\n")
                print(f"seed {seed}: {G} genomes {R} bp, reads {Q} bp{' short' if kw else ''}{' fq2>fq1' if case % 3 == 2 else ''}, ratios {hr}/{phr}: "
                      f"{ref.count(chr(10))} lines {'identical' if mine == ref else 'DIFFERENT'}"
                      f"{'' if ok or mine != ref else ' (but EMPTY: pins nothing)'}", flush=True)
    print(f"{lines} result lines compared in all")
    print("all identical" if not bad else f"{bad} case(s) differ")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
