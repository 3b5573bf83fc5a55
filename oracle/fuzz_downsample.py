#!/usr/bin/env python
"""Test infrastructure: the oracle's down-sampling branch (bin/extract_ref.cpp:955-960, :1124-1148) against DIAGNOSTIC
VARIANTS of the reference — the unmodified source with the single constant down_sampling_size (:1230) lowered, built
from a scratch copy by tests/golden/make_golden_downsample.py: build_variant into the git-ignored oracle/_ref/ — on
seeded inputs: several budgets (ratios from ~30 to ~95), every other case with fq2 the longer file, so that the
rand() draws (one per sequence line READ, fq1 then fq2) meet main's byte budget for fq2 (:1265-1291).  WARM runs only
(index on disk: main seeds rand() with 1; a cold run re-seeds it with time(0) while drawing the coder).
usage: python oracle/fuzz_downsample.py [first_seed n_cases]"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import oracle_util  # noqa: E402
from make_golden_downsample import build_variant  # noqa: E402
from palace_b200 import eref, synth  # noqa: E402

BUDGETS = [250_000, 400_000, 550_000]


def main():
    first, n = (int(x) for x in (sys.argv[1:3] + ["2000", "6"][len(sys.argv) - 1:]))
    oracle = oracle_util.load()
    exes = {b: build_variant(b) for b in BUDGETS}
    bad = lines = 0
    for case in range(n):
        seed = first + case
        budget = BUDGETS[case % len(BUDGETS)]
        db = synth.make_db(40, 1_000_000, seed)
        f1, f2, _ = synth.make_reads(db, 600_000, seed + 1000, cov_lo=8.0, cov_hi=14.0)
        if case % 2 == 1:
            nl = np.flatnonzero(f1 == 10)
            cut = int(nl[4 * (nl.size // 4 * 3 // 4) - 1]) + 1
            f1, f2 = f1[:cut].copy(), np.concatenate([f2, f1[cut:]])
        seq1 = oracle.fastq_seq_bytes(f1)
        ratio = 100 * budget // (2 * seq1)
        with tempfile.TemporaryDirectory() as d:
            db.text.tofile(os.path.join(d, "db.fa")); f1.tofile(os.path.join(d, "r1.fq")); f2.tofile(os.path.join(d, "r2.fq"))

            def run(hr, phr):
                return subprocess.run([exes[budget], "r1.fq", "r2.fq", "db.fa", "t.txt", hr, phr, "1"], cwd=d,
                                      capture_output=True, check=True).stdout.decode()
            run("0.9", "0.85")                               # cold: builds the index, output discarded
            coder = eref.read_coder(os.path.join(d, "db.fa.k32.index.dat"))
            for hr, phr in (("0.9", "0.85"), ("0.5", "0.3")):
                ref = run(hr, phr)
                mine = oracle.format(oracle.run(f1, f2, db.text, coder, hit_ratio=float(hr), perfect_hit_ratio=float(phr),
                                                down_sampling_size=budget))
                full = oracle.format(oracle.run(f1, f2, db.text, coder, hit_ratio=float(hr), perfect_hit_ratio=float(phr)))
                bad += mine != ref
                lines += ref.count("\n")
                print(f"seed {seed}: budget {budget} -> ratio {ratio}{', fq2 > fq1' if case % 2 else ''}, {hr}/{phr}: "
                      f"{ref.count(chr(10))} lines {'identical' if mine == ref else 'DIFFERENT'} "
                      f"({full.count(chr(10))} without down-sampling{'' if full != ref else ': the same, pins little'})", flush=True)
    print(f"{lines} result lines compared in all")
    print("all identical" if not bad else f"{bad} case(s) differ")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
