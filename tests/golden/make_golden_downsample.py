#!/usr/bin/env python
"""Golden fixture for the reference's down-sampling branch (bin/extract_ref.cpp:955-960, :1124-1148).

The branch is only reachable when fq1 holds more than 1 Gbase (down_sampling_size = 2e9, :1230), far beyond what
the CPU reference can process here.  This script therefore builds a DIAGNOSTIC VARIANT of the reference — the
unmodified source with the single constant on line 1230 lowered — from a scratch copy under /tmp (nothing of the
reference is copied into the repository; the binary goes to the git-ignored oracle/_ref/), runs it on the inputs
of the committed "basic" fixture and records the stdout of WARM runs (index on disk: rand() is seeded with 1 by
main(), :1236-1238; a cold run re-seeds it with time(0) while drawing the coder and is not reproducible).

    python tests/golden/make_golden_downsample.py
"""
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests.conftest import load_golden  # noqa: E402

SRC = "/root/reference/bin/extract_ref.cpp"
OUT = os.path.dirname(os.path.abspath(__file__))
DOWN_SAMPLING_SIZE = 400_000


def build_variant(DOWN_SAMPLING_SIZE=DOWN_SAMPLING_SIZE):
    text = open(SRC).read()
    pat = r"long down_sampling_size = 2000000000;"
    assert len(re.findall(pat, text)) == 1, "line 1230 not found"
    scratch = tempfile.mkdtemp(prefix="eref_ds_")
    cpp = os.path.join(scratch, "extract_ref_ds.cpp")
    open(cpp, "w").write(text.replace(pat, f"long down_sampling_size = {DOWN_SAMPLING_SIZE};"))
    exe = os.path.join(ROOT, "oracle", "_ref", f"eref_ref_ds{DOWN_SAMPLING_SIZE}")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, cpp, "-lpthread"], check=True)
    shutil.rmtree(scratch)
    return exe


def main():
    exe = build_variant()
    meta, fa, f1, f2 = load_golden("basic")
    tmp = tempfile.mkdtemp(prefix="golden_ds_")
    try:
        fa.tofile(os.path.join(tmp, "db.fa")); f1.tofile(os.path.join(tmp, "r1.fq")); f2.tofile(os.path.join(tmp, "r2.fq"))

        def run(hr, phr):
            return subprocess.run([exe, "r1.fq", "r2.fq", "db.fa", "tmp.txt", str(hr), str(phr), "1"], cwd=tmp,
                                  capture_output=True, check=True).stdout.decode()

        run(0.9, 0.85)                                   # cold: builds the index (time-seeded), output discarded
        outs = {f"{hr},{phr}": run(hr, phr) for hr, phr in [(0.9, 0.85), (0.5, 0.3)]}
        assert outs == {f"{hr},{phr}": run(hr, phr) for hr, phr in [(0.9, 0.85), (0.5, 0.3)]}, "warm runs must repeat"
        idx = np.fromfile(os.path.join(tmp, "db.fa.k32.index.dat"), dtype="<u4", count=100)
        seq1 = sum(len(s) for s in f1.tobytes().split(b"\n")[1::4])
        out = {"inputs": "basic", "down_sampling_size": DOWN_SAMPLING_SIZE, "coder": [int(x) for x in (idx[:96] & 0xFFFF)],
               "ratio": int(100 * DOWN_SAMPLING_SIZE // (2 * seq1)), "stdout": outs}
        for k, v in outs.items():
            print(k, "->", len(v.splitlines()), "lines")
        json.dump(out, open(os.path.join(OUT, "downsample.json"), "w"), indent=1)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
