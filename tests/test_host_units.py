"""CPU suite, part 4: the C++ host side of the drop-in executable (palace_b200/csrc/eref_main.cpp) without a GPU.

tests/host_units.cpp includes the executable's translation unit as it is and exposes its host helpers — input reading
with gzip / multi-member gzip / BGZF inflation, the coder header (:1082-1122), record sharding
(:1158-1208), record names (:246-254) and the fused next step (row f1, share/palace/scripts/get_ref_by_index.py) — on
a command line; it is built here with AddressSanitizer + UBSan, so every case below is also a memory-safety run of the
product's host code.  Expected values: Python's own gzip/zlib/repr, the Python mirror of the same rules
(palace_b200/eref.py), glibc through ctypes, and tests/golden/f1.json (made by the UNMODIFIED reference script from the
REFERENCE eref's stdout, tests/golden/make_golden_f1.py)."""
import ctypes
import gzip
import hashlib
import json
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from palace_b200 import eref, synth
from tests.conftest import GOLDEN, ROOT, load_golden

BUILD = os.path.join(ROOT, "tests", "_build")
EXE = os.path.join(BUILD, "host_units")
SRC = os.path.join(ROOT, "tests", "host_units.cpp")
MAIN = os.path.join(ROOT, "palace_b200", "csrc", "eref_main.cpp")
LIB = os.path.join(ROOT, "palace_b200", "liberef_gpu.so")
ENV = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:abort_on_error=0", UBSAN_OPTIONS="halt_on_error=1:print_stacktrace=1")


@pytest.fixture(scope="session")
def units():
    if not os.path.exists(LIB):
        pytest.skip("palace_b200/liberef_gpu.so is not built (python -c 'import __graft_entry__ as g; g.build()')")
    newest = max(os.path.getmtime(p) for p in (SRC, MAIN, LIB, os.path.join(ROOT, "include", "eref_gpu.h")))
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < newest:
        os.makedirs(BUILD, exist_ok=True)
        cmd = ["g++", "-O1", "-g", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-std=c++17",
               "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", SRC, "-o", EXE,
               "-L" + os.path.dirname(LIB), "-leref_gpu", "-L/usr/local/cuda/lib64", "-lcudart",
               "-L/usr/lib/x86_64-linux-gnu", "-lnccl", "-lz", "-lpthread",
               "-Wl,-rpath," + os.path.dirname(LIB), "-Wl,-rpath,/usr/local/cuda/lib64"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]

    def run(*args, stdin=None, ok=True):
        r = subprocess.run([EXE, *map(str, args)], input=stdin, capture_output=True, env=ENV)
        if ok:
            assert r.returncode == 0, r.stderr.decode()[-2000:]
            assert b"runtime error" not in r.stderr and b"AddressSanitizer" not in r.stderr, r.stderr.decode()[-2000:]
        return r
    return run


def bgzf(data: bytes, block=30000, eof=True) -> bytes:
    out = []
    offs = list(range(0, len(data), block)) + ([len(data)] if eof else [])
    for off in offs:
        chunk = data[off:off + block] if off < len(data) else b""
        co = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = co.compress(chunk) + co.flush()
        out.append(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", len(body) + 25) + body +
                   struct.pack("<II", zlib.crc32(chunk), len(chunk)))
    return b"".join(out)


# ---- row f3: input reading ---------------------------------------------------------------------------------------

def test_read_file_plain_gzip_members_and_bgzf(units, tmp_path):
    rng = np.random.default_rng(5)
    text = b"".join(b">g%d\n" % i + bytes(rng.choice(np.frombuffer(b"ACGT\n", np.uint8), 70_000)) + b"\n" for i in range(40))
    cases = {
        "plain": text,
        "one.gz": gzip.compress(text),
        "members.gz": gzip.compress(text[:1000]) + gzip.compress(text[1000:900_000]) + gzip.compress(text[900_000:]),
        # the last member is tiny: its ISIZE trailer is a useless size hint, the buffer has to grow while inflating
        "grow.gz": gzip.compress(text) + gzip.compress(b"ACGT\n"),
        "bgzf.gz": bgzf(text),
        "bgzf_big_blocks.gz": bgzf(text, block=65280),
        "bgzf_no_eof.gz": bgzf(text, eof=False),
        "empty": b"",
        "empty.gz": gzip.compress(b""),
        "one_byte": b">",
    }
    want = dict(cases, **{"one.gz": text, "members.gz": text, "grow.gz": text + b"ACGT\n", "bgzf.gz": text,
                          "bgzf_big_blocks.gz": text, "bgzf_no_eof.gz": text, "empty.gz": b""})
    for name, data in cases.items():
        p = tmp_path / name
        p.write_bytes(data)
        for threads in (1, 5):
            r = units("inflate", p, threads)
            assert r.stdout == want[name], name
            assert (b"gzip=1" in r.stderr) == name.endswith(".gz"), name


def test_read_file_refuses_damaged_inputs(units, tmp_path):
    text = b"@r\nACGT\n+\nIIII\n" * 20000
    z = gzip.compress(text)
    good = bgzf(text)
    bad = {
        "truncated.gz": z[: len(z) // 2],
        "garbage.gz": z[:10] + bytes(len(z) - 10),
    }
    first = struct.unpack("<H", good[16:18])[0] + 1      # size of the first block: 18 header bytes, deflate, CRC, ISIZE
    for name, at in (("bgzf_deflate_flip.gz", first + 18 + 9), ("bgzf_crc_flip.gz", first - 7), ("bgzf_isize_flip.gz", first - 3)):
        b = bytearray(good)
        b[at] ^= 0x5A
        bad[name] = bytes(b)
    bad["bgzf_cut.gz"] = good[: len(good) - 40]
    for name, data in bad.items():
        p = tmp_path / name
        p.write_bytes(data)
        r = units("inflate", p, 4, ok=False)
        assert r.returncode == 1 and name.encode() in r.stderr, (name, r.stderr)
        assert b"AddressSanitizer" not in r.stderr and b"runtime error" not in r.stderr
    r = units("inflate", tmp_path / "does_not_exist", 1, ok=False)
    assert r.returncode == 1 and b"cannot open" in r.stderr


# ---- coder header (:1082-1122) ------------------------------------------------------------------------------------

def test_draw_coder_is_the_glibc_draw(units):
    libc = ctypes.CDLL("libc.so.6")
    orderings = [(0, 1, 2), (0, 2, 1), (1, 2, 0), (1, 0, 2), (2, 0, 1), (2, 1, 0)]          # :1084
    for seed in (1, 7, 1793805349):
        libc.srand(ctypes.c_uint(seed))
        want = [v for _ in range(32) for v in orderings[libc.rand() % 6]]
        got = [int(x) for x in units("coder", seed).stdout.split()]
        assert got == want
        assert got == [int(x) for x in eref.random_coder(seed)]


def test_coder_comes_back_from_a_reference_index_header_and_from_the_sidecar(units, tmp_path):
    meta, *_ = load_golden("basic")
    coder = [int(x) for x in meta["coder"]]
    # the reference writes int16 pairs overlapping by one (:680-682 writes 4 bytes from each 2-byte element): the low
    # 16 bits of word j are coder[j], whatever the high half holds
    hdr = np.zeros(100, "<u4")
    hdr[:96] = np.asarray(coder, np.uint32) | (np.arange(96, dtype=np.uint32) * 977 % 65536 << 16)
    hdr.tofile(tmp_path / "db.k32.index.dat")
    assert [int(x) for x in units("index_header", tmp_path / "db.k32.index.dat").stdout.split()] == coder
    (tmp_path / "short.dat").write_bytes(hdr.tobytes()[:200])
    assert units("index_header", tmp_path / "short.dat", ok=False).returncode == 1
    (tmp_path / "db.k32.coder").write_text("".join(f"{a} {b} {c}\n" for a, b, c in zip(*[iter(coder)] * 3)))
    assert [int(x) for x in units("sidecar", tmp_path / "db.k32.coder").stdout.split()] == coder
    (tmp_path / "bad.coder").write_text("0 1 2\n0 1\n")
    assert units("sidecar", tmp_path / "bad.coder", ok=False).returncode == 1


# ---- sharding and names --------------------------------------------------------------------------------------------

def test_shard_offsets_equal_the_python_mirror(units, tmp_path):
    db = synth.make_db(37, 150_000, 9, min_len=40, max_len=30000, n_short=5)
    texts = {"synth": db.text.tobytes(), "edge": load_golden("edge")[1].tobytes(), "one": b">only\nACGT\n",
             "headerless": b"ACGT\nACGT\n", "empty": b""}
    for name, t in texts.items():
        p = tmp_path / (name + ".fa")
        p.write_bytes(t)
        for shards in (1, 2, 3, 8, 64):
            got = [int(x) for x in units("shards", p, shards).stdout.split()]
            assert got == [int(c) for c in eref.shard_offsets(np.frombuffer(t, np.uint8), shards)], (name, shards)


def test_record_name_rule(units):
    """get_read_ID(line).substr(1) (:246-254, :689): up to the first '/', ' ' or tab, without the '>'"""
    lines = [">p1 first phage", ">p2/1\tx", ">p3|x  two", ">", ">/x", "> lead", ">a\tb c/d", ">plain", ">with/slash", ">tab\tonly"]
    import re
    want = [re.split(r"[/ \t]", s)[0][1:] for s in lines]
    got = units("name", stdin="\n".join(lines).encode() + b"\n").stdout.decode().split("\n")[:-1]
    assert got == want


def test_header_offsets_threaded(units, tmp_path):
    """the header-line ordinals the fused next step asks for, found by several threads on a text above 16 MiB"""
    rng = np.random.default_rng(11)
    recs = []
    for i in range(3000):
        body = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), int(rng.integers(1, 14000))))
        if i % 97 == 0:
            body = body[:50] + b">inside a line" + body[50:]     # '>' that does not start a line is not a header
        recs.append(b">r%d\n" % i + body + b"\n")
    text = b"".join(recs)
    assert len(text) > (1 << 24)
    p = tmp_path / "big.fa"
    p.write_bytes(text)
    a = np.frombuffer(text, np.uint8)
    gt = np.flatnonzero(a == ord(">"))
    starts = gt[(gt == 0) | (a[np.maximum(gt, 1) - 1] == ord("\n"))]
    assert starts.size == 3000
    want = [1, 2, 17, 1500, 1501, 2999, 3000, 3001, 9999]
    for threads in (1, 2, 7):
        got = [int(x) for x in units("headers", p, threads, *want).stdout.split()]
        assert got == [int(starts[k - 1]) if k <= starts.size else -1 for k in want], threads


# ---- row f1: the next pipeline step, written by the executable ------------------------------------------------------

def test_py_float_repr_is_pythons_repr(units):
    """get_ref_by_index.py writes repr(float(token)) of the %g text eref printed (:27-33, :86)"""
    rng = np.random.default_rng(3)
    texts = ["1", "0.75", "0.750001", "1e-05", "1.5e-07", "0.0001", "0.00012345", "123456", "1e+06", "1.23457e+06",
             "1e+15", "1e+16", "1.5e+16", "9.99999e+15", "1e+22", "0", "-2.5", "-1e-07", "5e-1", "0.1", "0.3", "100"]
    el = rng.integers(1, 4_000_000, 3000)
    ln = el + rng.integers(0, 1_500_000, 3000)
    texts += ["%g" % float(np.float32(a) / np.float32(b)) for a, b in zip(el, ln)]      # :615-617
    texts += ["%g" % float(x) for x in np.exp(rng.uniform(-30, 50, 1500))]
    got = units("pyrepr", stdin="\n".join(texts).encode() + b"\n").stdout.decode().split("\n")[:-1]
    assert got == [repr(float(t)) for t in texts]


def _records_of(names):
    rec = []
    for line in names.splitlines():
        tok = line.split("\t")
        assert tok[0] == "ref_index" and len(tok) == 6, line
        rec.append(tuple(int(x) for x in tok[1:5]))
    return rec


with open(os.path.join(GOLDEN, "f1.json")) as _f:
    F1 = json.load(_f)


@pytest.mark.parametrize("case", [k for k in F1 if k.startswith("fixture:")] + ["edge", "empty_names"])
def test_write_refs_against_the_reference_script(units, case, tmp_path):
    """write_refs (EREF_REFS_FASTA / EREF_REFS_PERCENT) on the records of the REFERENCE eref's stdout: the two files
    equal what the unmodified get_ref_by_index.py made of that stdout (tests/golden/f1.json), and the records format
    back into the very lines they were parsed from (eref_format_line, :615-617)"""
    c = F1[case]
    if "fixture" in c:
        load_golden(c["fixture"])[1].tofile(tmp_path / "db.fa")
    else:
        (tmp_path / "db.fa").write_text(c["fasta"])
    rec = _records_of(c["names"])
    (tmp_path / "rec.txt").write_text("".join("%d %d %d %d\n" % r for r in rec))
    for threads in (1, 4):
        r = units("refs", tmp_path / "db.fa", tmp_path / "rec.txt", tmp_path / "refs.fa", tmp_path / "pct.txt", threads)
        assert r.stdout.decode() == c["names"]
        assert (tmp_path / "pct.txt").read_text() == c["out_percent"]
        out_fa = (tmp_path / "refs.fa").read_bytes()
        if "out_fasta" in c:
            assert out_fa.decode() == c["out_fasta"]
        else:
            assert out_fa.count(b">") == c["out_fasta_records"]
            assert hashlib.sha256(out_fa).hexdigest() == c["out_fasta_sha256"]
        assert not os.path.exists(tmp_path / "refs.fa.part") and not os.path.exists(tmp_path / "pct.txt.part")
    # an index beyond the last header line: the script's warning (:91), nothing written for it
    (tmp_path / "rec.txt").write_text("".join("%d %d %d %d\n" % r for r in rec) + "99999 1 10 10\n")
    r = units("refs", tmp_path / "db.fa", tmp_path / "rec.txt", "-", tmp_path / "pct2.txt", 2)
    assert b"Warning: Index 99999 not found in FAI file" in r.stderr
    assert (tmp_path / "pct2.txt").read_text() == c["out_percent"]
    # an output directory that does not exist: refused, nothing left behind
    r = units("refs", tmp_path / "db.fa", tmp_path / "rec.txt", tmp_path / "no_dir" / "x.fa", "-", 1, ok=False)
    assert r.returncode == 1 and b"cannot create" in r.stderr


def test_write_refs_equals_the_python_shim_on_random_inputs(units, tmp_path):
    """the stand-alone mirror (palace_b200/get_ref_by_index.py, pinned to the reference script by tests/test_host.py) and
    the executable's write_refs make the same two files of the same result lines, on seeded FASTAs with every framing
    oddity (CRLF, blanks inside lines, blank lines, records of <= 32 bases that shift the index, no final newline)"""
    from palace_b200 import get_ref_by_index as shim
    rng = np.random.default_rng(23)
    for trial in range(12):
        n = int(rng.integers(3, 40))
        parts = []
        for i in range(n):
            ln = int(rng.choice([0, 1, 20, 32, 33, 200, 1500]))
            seq = bytes(rng.choice(np.frombuffer(b"ACGTNacgt", np.uint8), ln))
            width = int(rng.choice([7, 60, 61, 10_000]))
            eol = b"\r\n" if rng.random() < 0.2 else b"\n"
            lines = [seq[k:k + width] for k in range(0, ln, width)]
            if lines and rng.random() < 0.3:
                lines.insert(int(rng.integers(0, len(lines) + 1)), b"")
            if lines and rng.random() < 0.3:
                lines[0] = lines[0][:3] + b" " + lines[0][3:] + b" \t"
            desc = [b"", b" some description", b"\tx/y", b"|z  w"][int(rng.integers(0, 4))]
            parts.append(b">g%d_%d" % (trial, i) + desc + eol + eol.join(lines) + (eol if lines else b""))
        text = b"".join(parts)
        if rng.random() < 0.5:
            text = text.rstrip(b"\r\n")
        (tmp_path / "db.fa").write_bytes(text)
        hdr = [ln.split()[0][1:] if ln.split() else b"" for ln in text.split(b"\n") if ln.startswith(b">")]
        (tmp_path / "db.fa.fai").write_bytes(b"".join(h + b"\t0\t0\t0\t0\n" for h in hdr))
        picks = sorted(set(int(x) for x in rng.integers(1, n + 3, int(rng.integers(0, n + 1)))))
        rec = []
        for k in picks:
            ln = int(rng.integers(40, 5_000_000))
            el = int(ln * rng.uniform(0.751, 1.0)) + 1
            rec.append((k, int(rng.integers(1, 9)), min(el, ln), ln))
        (tmp_path / "rec.txt").write_text("".join("%d %d %d %d\n" % r for r in rec))
        r = units("refs", tmp_path / "db.fa", tmp_path / "rec.txt", tmp_path / "a.fa", tmp_path / "a.txt", 3)
        (tmp_path / "names.txt").write_bytes(r.stdout)
        assert r.stdout.count(b"\n") == len(rec)
        import contextlib
        import io
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            shim.main(["get_ref_by_index.py", str(tmp_path / "db.fa"), str(tmp_path / "db.fa.fai"), str(tmp_path / "names.txt"),
                       str(tmp_path / "b.fa"), str(tmp_path / "b.txt")])
        assert (tmp_path / "a.txt").read_bytes() == (tmp_path / "b.txt").read_bytes(), trial
        assert (tmp_path / "a.fa").read_bytes() == (tmp_path / "b.fa").read_bytes(), trial
