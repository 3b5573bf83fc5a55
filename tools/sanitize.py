#!/usr/bin/env python
"""Sanitizer runs SURVEY.md §5 asks for, on the GPU box:

  * compute-sanitizer --tool memcheck / racecheck / initcheck over the drop-in executable on the `edge` and
    `short_reads` fixtures (framing borders, ring buffers, shared-memory atomics cascades, list appends), and over the
    sharded table build (two contexts on one GPU through the Python binding);
  * an AddressSanitizer + UBSan build of the host program (eref_main.cpp) on the same fixtures.

Writes gpurun_out/sanitizer_r02.txt (copied to profiles/ afterwards).  Usage: python tools/sanitize.py
"""
import gzip
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
EXE = os.path.join(ROOT, "palace_b200", "bin", "eref")
CS = "/usr/local/cuda/bin/compute-sanitizer"
OUT = os.path.join(ROOT, "gpurun_out", "sanitizer_r02.txt")


def fixture_dir(name):
    d = tempfile.mkdtemp(prefix=f"san_{name}_")
    meta = json.load(open(os.path.join(GOLDEN, name + ".json")))
    for ext, dst in (("fa", "db.fa"), ("r1.fq", "r1.fq"), ("r2.fq", "r2.fq")):
        with gzip.open(os.path.join(GOLDEN, f"{name}.{ext}.gz"), "rb") as g, open(os.path.join(d, dst), "wb") as f:
            f.write(g.read())
    import numpy as np
    hdr = np.zeros(100, "<u4"); hdr[:96] = np.asarray(meta["coder"], np.uint32)
    hdr.tofile(os.path.join(d, "db.fa.k32.index.dat"))
    return d, meta


def run(log, title, cmd, cwd, env=None, expect=None, timeout=900):
    t0 = time.time()
    try:
        r = subprocess.run(cmd, cwd=cwd, capture_output=True, env=env, timeout=timeout)
        rc, so, se = r.returncode, r.stdout.decode(errors="replace"), r.stderr.decode(errors="replace")
    except subprocess.TimeoutExpired:
        rc, so, se = -9, "", "TIMEOUT"
    dt = time.time() - t0
    tail = [l for l in (se + so).splitlines() if "ERROR SUMMARY" in l or "RACECHECK SUMMARY" in l or "ERROR:" in l
            or "runtime error" in l or "AddressSanitizer" in l or "TIMEOUT" in l][:12]
    same = None if expect is None else (so.replace("\r", "") == expect or expect in so)
    log.write(f"## {title}\n$ {' '.join(cmd)}\nrc={rc} wall={dt:.1f}s" + ("" if same is None else f" stdout_identical_to_golden={same}") + "\n")
    for l in tail:
        log.write("  " + l.strip() + "\n")
    if rc != 0 and not tail:
        for l in (se + so).splitlines()[-8:]:
            log.write("  | " + l.rstrip() + "\n")
    log.write("\n")
    log.flush()
    print(title, "rc", rc, f"{dt:.0f}s", tail[-1] if tail else "")


def main():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as log:
        log.write("# sanitizer runs (tools/sanitize.py), " + time.strftime("%Y-%m-%d %H:%M:%S") + "\n\n")
        env = dict(os.environ, EREF_GPUS="1")
        for name in ("edge", "short_reads"):
            d, meta = fixture_dir(name)
            expect = meta["stdout"]["0.9,0.85"]
            args = [EXE, "r1.fq", "r2.fq", "db.fa", "tmp.txt", "0.9", "0.85", "2"]
            for tool in ("memcheck", "racecheck", "initcheck"):
                extra = ["--racecheck-report", "all"] if tool == "racecheck" else []
                run(log, f"compute-sanitizer {tool}: eref on fixture `{name}`", [CS, "--tool", tool, *extra, *args], d, env, expect)
            run(log, f"compute-sanitizer memcheck: eref on fixture `{name}`, unpruned scan + index written",
                [CS, "--tool", "memcheck", *args], d, dict(env, EREF_NO_PRUNE="1", EREF_WRITE_INDEX="0"), expect)
        # the sharded table build (peer stores + flags) under memcheck and racecheck: two contexts on this GPU
        code = ("import sys; sys.path.insert(0, %r)\n"
                "from palace_b200 import eref\nfrom tests.conftest import load_golden\n"
                "meta, fa, f1, f2 = load_golden('edge')\n"
                "t = eref.run_eref(f1, f2, fa, coder=meta['coder'], devices=(0, 0), shard_table=True)\n"
                "assert t == meta['stdout']['0.9,0.85'], t\nprint('sharded ok')\n") % ROOT
        for tool in ("memcheck", "racecheck"):
            run(log, f"compute-sanitizer {tool}: sharded table build, 2 contexts on one GPU (python binding)",
                [CS, "--tool", tool, "--target-processes", "all", sys.executable, "-c", code], ROOT,
                dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS="32"), "sharded ok", timeout=1500)
        # host code: ASan + UBSan build of the executable
        asan = os.path.join(tempfile.mkdtemp(prefix="asan_"), "eref_asan")
        b = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-omit-frame-pointer",
                            "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", "-o", asan,
                            os.path.join(ROOT, "palace_b200", "csrc", "eref_main.cpp"), "-L" + os.path.join(ROOT, "palace_b200"),
                            "-leref_gpu", "-L/usr/local/cuda/lib64", "-lcudart", "-lnccl", "-lz", "-lpthread",
                            "-Wl,-rpath," + os.path.join(ROOT, "palace_b200"), "-Wl,-rpath,/usr/local/cuda/lib64"],
                           capture_output=True)
        log.write(f"## ASan/UBSan build of eref_main.cpp: rc={b.returncode}\n{b.stderr.decode()[-400:]}\n\n")
        if b.returncode == 0:
            aenv = dict(env, ASAN_OPTIONS="protect_shadow_gap=0:detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")
            for name in ("edge", "short_reads", "basic"):
                d, meta = fixture_dir(name)
                run(log, f"ASan+UBSan host build: eref on fixture `{name}` (+ f1 outputs)", [asan, "r1.fq", "r2.fq", "db.fa", "tmp.txt", "0.9", "0.85", "4"],
                    d, dict(aenv, EREF_REFS_FASTA=os.path.join(d, "refs.fa"), EREF_REFS_PERCENT=os.path.join(d, "pct.txt")),
                    meta["stdout"]["0.9,0.85"])
            # gzip path
            d, meta = fixture_dir("edge")
            for f in ("db.fa", "r1.fq", "r2.fq"):
                subprocess.run(["gzip", "-k", f], cwd=d)
            os.rename(os.path.join(d, "db.fa.k32.index.dat"), os.path.join(d, "db.fa.gz.k32.index.dat"))
            run(log, "ASan+UBSan host build: gzip inputs", [asan, "r1.fq.gz", "r2.fq.gz", "db.fa.gz", "tmp.txt", "0.9", "0.85", "4"], d, aenv,
                meta["stdout"]["0.9,0.85"])
    print(open(OUT).read())


if __name__ == "__main__":
    main()
