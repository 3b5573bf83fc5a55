"""CPU suite, part 1: the oracle (oracle/eref_oracle.c) against the golden fixtures produced by the
UNMODIFIED reference (tests/golden/make_golden.py), and against its own literal restatement.
No GPU needed."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT


def test_naive_and_bitplane_codes_agree(oracle):
    rng = np.random.default_rng(5)
    alphabet = np.frombuffer(b"ACGTacgtNn\r>", np.uint8)
    for seed in (1, 7, 1234567):
        coder = oracle.random_coder(seed)
        for _ in range(300):
            # mostly valid windows, some with an invalid byte
            w = alphabet[rng.integers(0, 8, 32)].copy()
            if rng.random() < 0.2:
                w[rng.integers(0, 32)] = alphabet[rng.integers(8, alphabet.size)]
            ok_n, c_n = oracle.code3(coder, w.tobytes(), naive=True)
            ok_f, c_f = oracle.code3(coder, w.tobytes(), naive=False)
            assert ok_n == ok_f
            assert (c_n == c_f).all()


def test_code_is_strand_symmetric(oracle):
    """canonical = min(fwd, revcomp) (bin/extract_ref.cpp:989-994): a window and its reverse complement agree"""
    rng = np.random.default_rng(6)
    coder = oracle.random_coder(99)
    comp = bytes.maketrans(b"ACGT", b"TGCA")
    for _ in range(200):
        w = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, 32)].tobytes()
        rc = w.translate(comp)[::-1]
        assert (oracle.code3(coder, w)[1] == oracle.code3(coder, rc)[1]).all()


def test_random_coder_matches_glibc_draw(oracle):
    from palace_b200 import eref
    for seed in (1, 2, 1793805349):
        a = oracle.random_coder(seed)
        b = eref.random_coder(seed)
        assert (a == b).all()
        assert set(a.reshape(32, 3).sum(axis=1)) == {3}          # each row is an ordering of {0,1,2}


def test_oracle_index_body_matches_reference(oracle, golden, tmp_path):
    """every code of every reference position, against the reference's own <db>.k32.index.dat"""
    name, meta, fa, f1, f2 = golden
    p = str(tmp_path / "idx.dat")
    assert oracle.write_index(p, meta["coder"], fa) == 0
    idx = np.fromfile(p, dtype="<u4")
    assert idx.size == meta["index_words"]
    assert (idx[:96] & 0xFFFF).tolist() == meta["coder"]
    assert hashlib.sha256(idx[100:].tobytes()).hexdigest() == meta["index_body_sha256"]


def test_oracle_stdout_matches_reference(oracle, golden):
    name, meta, fa, f1, f2 = golden
    for key, expect in meta["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        res = oracle.run(f1, f2, fa, meta["coder"], hr, phr)
        assert oracle.format(res) == expect, f"{name} {key}"
    if name == "basic":
        assert len(meta["stdout"]["0.9,0.85"].splitlines()) >= 3      # the fixture does select references


def test_oracle_threads_do_not_change_results(oracle, golden):
    name, meta, fa, f1, f2 = golden
    a = oracle.run(f1, f2, fa, meta["coder"], 0.9, 0.85, threads=1)
    b = oracle.run(f1, f2, fa, meta["coder"], 0.9, 0.85, threads=4)
    assert (a == b).all()


def test_fasta_framing_rules(oracle):
    """bin/extract_ref.cpp:686-756: header iff first byte '>', len <= 32 dropped without an ordinal"""
    s40, s33, s32 = b"A" * 40, b"C" * 33, b"G" * 32
    fa = s40[:20] + b"\n" + s40[20:] + b"\n>a\n" + s33 + b"\n>b\n" + s32 + b"\n>c\n\n>d x\nAC>GT" + b"T" * 30 + b"\r\n>e\n" + s40
    kept, recs = oracle.parse_fasta(np.frombuffer(fa, np.uint8))
    assert recs == 6
    assert kept == [s40, s33, b"AC>GT" + b"T" * 30 + b"\r", s40]


def test_fastq_seq_bytes(oracle):
    fq = b"@a\nACGT\n+\nIIII\n@b\nAC\r\n+\nII\n@c\nACG"
    assert oracle.fastq_seq_bytes(np.frombuffer(fq, np.uint8)) == 4 + 3 + 3


@pytest.mark.skipif(not os.environ.get("EREF_TEST_REF"), reason="set EREF_TEST_REF=1 to run the real reference (slow, 24 GB)")
def test_live_reference_matches_oracle(oracle, tmp_path):
    from palace_b200 import synth
    ref = os.path.join(ROOT, "oracle", "_ref", "eref_ref")
    if not os.path.exists(ref):
        pytest.skip("oracle/_ref/eref_ref not built")
    db = synth.make_db(60, 1_500_000, 31)
    f1, f2, _ = synth.make_reads(db, 300_000, 32)
    db.text.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    out = subprocess.run([ref, "r1.fq", "r2.fq", "db.fa", "t.txt", "0.9", "0.85", "1"], cwd=tmp_path,
                         capture_output=True, check=True).stdout.decode()
    from palace_b200 import eref
    coder = eref.read_coder(str(tmp_path / "db.fa.k32.index.dat"))
    assert oracle.format(oracle.run(f1, f2, db.text, coder)) == out


def test_oracle_downsampling_matches_reference(oracle):
    """the rand()-driven down-sampling branch (:955-960) against the diagnostic reference variant with
    down_sampling_size lowered (tests/golden/make_golden_downsample.py)"""
    import json
    from tests.conftest import load_golden, GOLDEN
    ds = json.load(open(os.path.join(GOLDEN, "downsample.json")))
    meta, fa, f1, f2 = load_golden(ds["inputs"])
    for key, expect in ds["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        got = oracle.run(f1, f2, fa, ds["coder"], hr, phr, threads=2, down_sampling_size=ds["down_sampling_size"])
        assert oracle.format(got) == expect, key
    # and it really is a different answer from the unsampled run
    full = oracle.run(f1, f2, fa, ds["coder"], 0.9, 0.85, threads=2)
    assert oracle.format(full) != ds["stdout"]["0.9,0.85"]


def test_oracle_reads_fq2_only_up_to_fq1s_size(oracle):
    """main hands file_size(fq1) to read_fastq as the byte budget of BOTH files (:1265-1291) and the getline loop stops
    once the lines read add up to it (:942-945): with fq2 twice as long as fq1 the reference reads only its head.
    Expected stdout: the unmodified reference on the same inputs (tests/golden/make_golden_fq2_limit.py)."""
    import hashlib
    from tests.conftest import fq2_limit_inputs
    meta, fa, f1, f2 = fq2_limit_inputs()
    assert hashlib.sha256(f1.tobytes()).hexdigest() == meta["sha256"]["r1"]
    assert hashlib.sha256(f2.tobytes()).hexdigest() == meta["sha256"]["r2"]
    assert (f1.size, f2.size) == (meta["fq1_bytes"], meta["fq2_bytes"])
    assert oracle.fastq_limit(f2, f1.size) == meta["fq2_bytes_read"] < f2.size
    assert oracle.fastq_limit(f1, f1.size) == f1.size              # fq1 itself is never cut
    for key, expect in meta["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        assert oracle.format(oracle.run(f1, f2, fa, meta["coder"], hr, phr, threads=2)) == expect, key
    # the budget is what makes the difference: the same reads with the longer file first are all counted
    swapped = oracle.format(oracle.run(f2, f1, fa, meta["coder"], 0.5, 0.3, threads=2))
    assert swapped.count("\n") > meta["stdout"]["0.5,0.3"].count("\n") > 0


def test_tail_hits_diagnostic_only_adds_coverage(oracle):
    """eo_set_tail_hits: what the reference's uninitialised read of a record's last 31 hit flags (:856 vs :527-531) can
    add at most.  Counting them as hits never lowers a record's covered length, leaves records untouched whose last
    window is far from good, and does change short, well-covered records — the ones the extremes fuzz classifies."""
    rng = np.random.default_rng(8)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    fa, reads = [], []
    for i in range(120):
        L = int(rng.integers(260, 420)) if i % 2 else int(rng.integers(2000, 6000))
        seq = acgt[rng.integers(0, 4, L)].tobytes()
        fa.append(b">t%d\n" % i + seq + b"\n")
        if i % 4 < 2:                                              # half of the records tiled 6x by error-free reads
            for _ in range(6 * L // 120):
                st = int(rng.integers(0, max(L - 150, 1)))
                reads.append(seq[st:st + 150])
    fq = np.frombuffer(b"".join(b"@q\n" + r + b"\n+\n" + b"I" * len(r) + b"\n" for r in reads), np.uint8)
    fa = np.frombuffer(b"".join(fa), np.uint8)
    coder = oracle.random_coder(8)
    canonical = oracle.run(fq, fq, fa, coder, 0.5, 0.3, threads=2)
    try:
        oracle.lib.eo_set_tail_hits(3)
        upper = oracle.run(fq, fq, fa, coder, 0.5, 0.3, threads=2)
    finally:
        oracle.lib.eo_set_tail_hits(0)
    assert (upper["el"] >= canonical["el"]).all()
    uncovered = np.arange(120) % 4 >= 2
    assert (upper["el"][uncovered] == 0).all() and (canonical["el"][uncovered] == 0).all()
    assert (upper["el"] > canonical["el"]).any()
    assert (oracle.run(fq, fq, fa, coder, 0.5, 0.3, threads=2) == canonical).all()      # the knob is back at "no hit"
