#!/usr/bin/env python
"""Golden outputs for the reference's byte budget on fq2 (bin/extract_ref.cpp:1265-1291: main computes
file_size(fq1) once and hands it to read_fastq as `end` for fq1 AND fq2; :942-945: the getline loop stops once the
lines read so far add up to `end` characters).  Inputs: derived from the committed "basic" fixture by
tests/conftest.py: fq2_limit_inputs (fq2 twice as long as fq1).  Runs the UNMODIFIED reference
(oracle/_ref/eref_ref, threads=1) and records the coder header it drew and its stdout for several threshold pairs,
plus — for the record — how many bytes of fq2 the budget leaves.  Run from the repo root:

    python tests/golden/make_golden_fq2_limit.py
"""
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
from tests import conftest  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "eref_ref")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fq2_limit.json")
RATIOS = [(0.9, 0.85), (0.5, 0.3), (0.3, 0.2)]


def main():
    if not os.path.exists(REF):
        sys.exit("build oracle/_ref/eref_ref first: make -C oracle")
    if not os.path.exists(OUT):
        with open(OUT, "w") as f:
            json.dump({}, f)
    _, fa, f1, f2 = conftest.fq2_limit_inputs()
    tmp = tempfile.mkdtemp(prefix="golden_fq2_")
    try:
        fa.tofile(os.path.join(tmp, "db.fa")); f1.tofile(os.path.join(tmp, "r1.fq")); f2.tofile(os.path.join(tmp, "r2.fq"))
        outs = {}
        for hr, phr in RATIOS:
            r = subprocess.run([REF, "r1.fq", "r2.fq", "db.fa", "tmp.txt", str(hr), str(phr), "1"], cwd=tmp,
                               capture_output=True, check=True)
            outs[f"{hr},{phr}"] = r.stdout.decode()
            print(hr, phr, "->", len(r.stdout.splitlines()), "lines", flush=True)
        idx = np.fromfile(os.path.join(tmp, "db.fa.k32.index.dat"), dtype="<u4", count=96)
        # lines of fq2 in front of which fewer than size(fq1) characters were read (:942-945), counted here in numpy
        nl = np.flatnonzero(f2 == 10)
        starts = np.concatenate([[0], nl + 1]); starts = starts[starts < f2.size]
        cut = np.flatnonzero(starts - np.arange(starts.size) >= f1.size)
        meta = {
            "derived_from": "basic (tests/conftest.py: fq2_limit_inputs)",
            "coder": [int(x) for x in (idx & 0xFFFF)],
            "stdout": outs,
            "fq1_bytes": int(f1.size), "fq2_bytes": int(f2.size),
            "fq2_bytes_read": int(starts[cut[0]]) if cut.size else int(f2.size),
            "sha256": {k: hashlib.sha256(v.tobytes()).hexdigest() for k, v in (("r1", f1), ("r2", f2))},
        }
        with open(OUT, "w") as f:
            json.dump(meta, f, indent=1)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
