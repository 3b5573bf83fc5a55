"""CPU suite, part 5: the pure-arithmetic device functions of the CUDA path, compiled for the host.

The kernels cannot run here, but the integer arithmetic they are built on can: tests/device_units.cpp includes
palace_b200/csrc/eref_internal.cuh as it is and the source regions this file cuts out of frame.cu (text -> masks), api.cu (table parts)
and scan.cu (part_mask) — the product's own source text, with host versions of five CUDA integer intrinsics — and
checks them against naive restatements and, for the 32-mer codes, against the oracle's literal z-loop
(oracle/eref_oracle.c: eo_code3_naive, pinned to the reference's own index file by tests/test_oracle.py):

  masks    every byte value at every chunk position -> newline / ACGTacgt / 2-bit code masks (bin/extract_ref.cpp:978)
  codes    window_bits + stream_code on bit-planes == the three canonical codes of :971-994, valid_windows == :978-982
  parts    part_mask<1..5>: the table part of min(fwd, rc) decided from shifted alphabet planes == top bits of the code
  tileop   TileSum/TileOp: folding per-tile FASTA framing summaries in any association == framing the text in one go
  cuts     part_first_slice / part_weight (api.cu, host arithmetic): the table parts of a sharded build partition the
           8192 slices and equalise w * code mass + (1 - w) * width; the numbers DESIGN.md quotes
"""
import os
import re
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT

BUILD = os.path.join(ROOT, "tests", "_build")
EXE = os.path.join(BUILD, "device_units")
SRC = os.path.join(ROOT, "tests", "device_units.cpp")
CSRC = os.path.join(ROOT, "palace_b200", "csrc")


def _cut(path, first, last):
    """the text of `path` from the line that starts with `first` up to (not including) the line that starts with `last`"""
    text = open(path).read()
    a = text.index("\n" + first) + 1
    b = text.index("\n" + last, a) + 1
    return text[a:b]


def snippets():
    frame = _cut(os.path.join(CSRC, "frame.cu"), "constexpr uint32_t H7", "// exclusive prefix sum over the 256 threads")
    scan = _cut(os.path.join(CSRC, "scan.cu"), "template <int LOG2P>\n__device__ __forceinline__ uint32_t part_mask",
                "// list == nullptr: every word of the planes")
    api = _cut(os.path.join(CSRC, "api.cu"), "uint32_t part_first_slice(", "void set_part_range(")
    for name in ("eq7", "gather2", "masks_word", "chunk_masks", "chunk_newlines", "count_nl", "before"):
        assert re.search(r"\b%s\b" % name, frame), name
    assert "__global__" not in frame + scan + api and "threadIdx" not in frame + scan + api
    return ("// cut out of palace_b200/csrc/frame.cu and scan.cu by tests/test_device_units.py — do not edit\n"
            "namespace {\n" + frame + scan + api + "}\n")


@pytest.fixture(scope="session")
def units():
    os.makedirs(BUILD, exist_ok=True)
    inc = os.path.join(BUILD, "device_snippets.inc")
    text = snippets()
    if not os.path.exists(inc) or open(inc).read() != text:
        with open(inc, "w") as f:
            f.write(text)
    deps = [SRC, inc, os.path.join(CSRC, "eref_internal.cuh"), os.path.join(ROOT, "include", "eref_gpu.h")]
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(p) for p in deps):
        r = subprocess.run(["g++", "-O2", "-g", "-fsanitize=undefined", "-fno-sanitize-recover=undefined", "-std=c++17",
                            "-I/usr/local/cuda/include", "-I" + os.path.join(ROOT, "tests"), SRC, "-o", EXE],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-3000:]

    def run(*args, stdin=None):
        r = subprocess.run([EXE, *map(str, args)], input=stdin, capture_output=True)
        assert r.returncode == 0, (r.stdout.decode()[-500:], r.stderr.decode()[-2000:])
        return r.stdout.decode()
    return run


def test_text_to_masks_for_every_byte_value_at_every_position(units):
    assert units("masks", 7, 2).strip() == "masks mismatches 0"


def test_part_masks_equal_the_top_bits_of_the_codes(units):
    assert units("parts", 11, 40).strip() == "parts mismatches 0"


def test_tile_summaries_compose(units):
    assert units("tileop", 5, 3000).strip() == "tileop mismatches 0"


def test_table_parts_partition_the_slices_and_balance_their_cost(units):
    out = units("cuts").splitlines()
    assert out[0] == "cuts mismatches 0" and out[3] == "weight mismatches 0"
    # the cut points DESIGN.md section 6 quotes for a weight of 0.7
    assert out[1].split() == ["0", "2805"]
    assert out[2].split() == ["0", "622", "1288", "2010", "2805", "3700", "4747", "6066"]


def test_bitplane_codes_equal_the_oracles_literal_loop(units, oracle):
    rng = np.random.default_rng(17)
    for seed in (1, 99):
        coder = oracle.random_coder(seed)
        seqs = []
        for k in range(60):
            s = bytearray(rng.choice(np.frombuffer(b"ACGT", np.uint8), 64).tobytes())
            if k % 3 == 1:                                        # lower case counts as the same base (:978)
                for p in rng.integers(0, 64, 10):
                    s[p] = ord(chr(s[p]).lower())
            if k % 3 == 2:                                        # invalid bases knock out every window over them
                for p in rng.integers(0, 64, 2):
                    s[p] = b"N>\nEn"[int(rng.integers(0, 5))]
            if k == 0:
                s = bytearray(b"A" * 64)
            if k == 3:                                            # reverse-complement palindrome: fwd == rc
                half = bytes(s[:16])
                rc = bytes({65: 84, 67: 71, 71: 67, 84: 65}[c] for c in reversed(half.upper()))
                s = bytearray((half.upper() + rc) * 2)
            seqs.append(bytes(s).replace(b"\n", b"x"))            # (a newline would end the harness's input line)
        stdin = (" ".join(str(int(x)) for x in coder) + "\n").encode() + b"\n".join(seqs) + b"\n"
        out = [tuple(int(v) for v in ln.split()) for ln in units("codes", stdin=stdin).splitlines()]
        assert len(out) == 32 * len(seqs)
        checked = 0
        for k, s in enumerate(seqs):
            for st in range(32):
                w = s[st:st + 32]
                valid = all(c in b"ACGTacgt" for c in w)
                got = out[k * 32 + st]
                assert bool(got[0]) == valid, (k, st)
                if valid:
                    ok, codes = oracle.code3(coder, w, naive=True)
                    assert ok and tuple(int(c) for c in codes) == got[1:], (k, st)
                    checked += 1
        assert checked > 800
