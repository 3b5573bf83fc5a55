"""GPU suite (-m gpu): the CUDA path through the C ABI against (a) the golden fixtures produced by the
unmodified reference, (b) the CPU oracle on seeded inputs, (c) size-independent properties."""
import os
import subprocess

import numpy as np
import pytest

from palace_b200 import eref, synth
from tests.conftest import GOLDEN_CASES, load_golden

pytestmark = pytest.mark.gpu


# ---- (a) golden fixtures: bit-exact stdout of the real reference ---------------------------------

@pytest.mark.parametrize("prune", [True, False])
def test_golden_stdout(golden, prune):
    name, meta, fa, f1, f2 = golden
    for key, expect in meta["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        got = eref.run_eref(f1, f2, fa, None, hr, phr, 1, coder=meta["coder"], prune=prune)
        assert got == expect, f"{name} {key} prune={prune}"


def test_golden_codes_match_reference_index(golden, oracle):
    """eref_get_codes == the body of the reference's <db>.k32.index.dat (via the pinned oracle), every record"""
    import hashlib
    name, meta, fa, f1, f2 = golden
    seqs, _ = oracle.parse_fasta(fa)
    h = hashlib.sha256()
    with eref.ErefGPU(meta["coder"]) as g:
        g.ref_set(fa)
        kept, recs, bases = g.ref_count()
        assert kept == len(seqs)
        assert bases == sum(len(s) for s in seqs)
        for i, s in enumerate(seqs):
            got = g.get_codes(i + 1)
            assert (got == oracle.record_codes(meta["coder"], s)).all(), f"{name} record {i + 1}"
            h.update(np.uint32(len(s)).tobytes()); h.update(got.tobytes())
    assert h.hexdigest() == meta["index_body_sha256"]


def test_golden_all_records_match_oracle(golden, oracle):
    """not only the selected lines: {ordinal, frag, el, ref_len} of EVERY kept record"""
    name, meta, fa, f1, f2 = golden
    want = oracle.run(f1, f2, fa, meta["coder"], 0.9, 0.85)
    _, got = eref.run_eref(f1, f2, fa, None, 0.9, 0.85, 1, coder=meta["coder"], return_results=True)
    assert got.size == want.size
    assert (got == want).all()


# ---- (b) seeded inputs against the oracle ----------------------------------------------------------

def _case(seed, n_genomes=80, db_bp=2_000_000, q_bp=500_000, **kw):
    db = synth.make_db(n_genomes, db_bp, seed, **kw.pop("db_kw", {}))
    f1, f2, present = synth.make_reads(db, q_bp, seed + 1, **kw)
    return db, f1, f2, present


@pytest.mark.parametrize("seed,ratios", [(11, (0.9, 0.85)), (12, (0.8, 0.5)), (13, (0.3, 0.2)), (14, (0.99, 0.97))])
def test_seeded_parity_all_records(oracle, seed, ratios):
    db, f1, f2, _ = _case(seed, cov_lo=1.0, cov_hi=5.0)
    coder = eref.random_coder(seed)
    want = oracle.run(f1, f2, db.text, coder, *ratios, threads=4)
    for prune in (True, False):
        text, got = eref.run_eref(f1, f2, db.text, None, *ratios, 1, coder=coder, prune=prune, return_results=True)
        assert (got == want).all(), f"seed {seed} prune={prune}"
        assert text == oracle.format(want)
    assert (want["el"] > 0).any()


@pytest.mark.parametrize("stride", [4, 5, 6, 7, 8])
def test_every_sampling_stride_gives_the_same_records(oracle, stride, monkeypatch):
    """the candidate filter is a bound, not an estimate: any stride 4..8, thresholds where the
    filter is sharp (0.99), ordinary (0.85) and vacuous (0.5), same records as the oracle"""
    monkeypatch.setenv("EREF_SAMPLE_STRIDE", str(stride))
    db, f1, f2, _ = _case(15, cov_lo=1.0, cov_hi=6.0)
    coder = eref.random_coder(15)
    for ratios in ((0.9, 0.85), (0.99, 0.99), (0.8, 0.5)):
        want = oracle.run(f1, f2, db.text, coder, *ratios, threads=4)
        _, got = eref.run_eref(f1, f2, db.text, None, *ratios, 1, coder=coder, return_results=True)
        assert (got == want).all(), f"stride {stride} ratios {ratios}"


def test_short_read_heavy_query(oracle):
    """BASELINE configs[4]: reads of 20-150 bases, many shorter than k (no window at all) — the dense FASTQ planes
    hold many one-read words and reads that end inside a word"""
    db, f1, f2, _ = _case(16, q_bp=1_500_000, min_len=20, max_len=150, cov_lo=2.0, cov_hi=8.0)
    coder = eref.random_coder(16)
    for ratios in ((0.9, 0.85), (0.5, 0.3)):
        want = oracle.run(f1, f2, db.text, coder, *ratios, threads=4)
        for pipelined in (False, True):
            _, got = eref.run_eref(f1, f2, db.text, None, *ratios, 1, coder=coder, pipelined=pipelined, return_results=True)
            assert (got == want).all(), f"ratios {ratios} pipelined={pipelined}"
    assert (want["el"] > 0).any()


def test_table_levels_match_oracle_counts(oracle):
    """level-L bitmap popcounts == number of codes the byte table holds at >= L (bin/extract_ref.cpp:995-997)"""
    db, f1, f2, _ = _case(21, n_genomes=20, db_bp=300_000, q_bp=200_000)
    coder = eref.random_coder(5)
    # independent count on the CPU from the oracle's codes: concatenate reads as pseudo records
    import collections
    cnt = collections.Counter()
    for fq in (f1, f2):
        lines = fq.tobytes().split(b"\n")
        for s in lines[1::4]:
            if len(s) >= 32:
                c = oracle.record_codes(coder, s).reshape(-1, 3)
                ok = c.any(axis=1)                  # the oracle writes 0,0,0 for a window with an invalid base
                cnt.update(c[ok].ravel().tolist())
    with eref.ErefGPU(coder) as g:
        g.query_add(f1); g.query_add(f2)
        for level in (1, 2, 3):
            assert g.query_popcount(level) == sum(1 for v in cnt.values() if v >= level)


def test_bucketed_and_direct_builds_give_the_same_table():
    """code-range bucketing only reorders commuting atomics: all three levels must be bit-identical"""
    db, f1, f2, _ = _case(22, n_genomes=40, db_bp=1_000_000, q_bp=800_000)
    coder = eref.random_coder(22)
    pops, recs = [], []
    for bucketed in (True, False):
        with eref.ErefGPU(coder, bucketed=bucketed) as g:
            g.query_add(f1); g.query_add(f2); g.ref_set(db.text)
            pops.append([g.query_popcount(l) for l in (1, 2, 3)])
            recs.append(g.scan().copy())
            assert (g.timings()["query_codes"] > 0) == bucketed
    assert pops[0] == pops[1] and pops[0][2] > 0
    assert (recs[0] == recs[1]).all()


def test_seq_bytes_matches_cal_sam_ratio_sum(oracle, golden):
    name, meta, fa, f1, f2 = golden
    with eref.ErefGPU(meta["coder"]) as g:
        assert g.query_add(f1) == oracle.fastq_seq_bytes(f1)
        assert g.query_add(f2) == oracle.fastq_seq_bytes(f2)


def test_edge_inputs(oracle):
    coder = eref.random_coder(3)
    fq = np.frombuffer(b"@r\n" + b"ACGT" * 10 + b"\n+\n" + b"I" * 40 + b"\n", np.uint8)
    cases = {
        "no records": b"",
        "only short": b">a\nACGT\n>b\n" + b"A" * 32 + b"\n",
        "no trailing newline": b">a\n" + b"ACGT" * 20,
        "blank lines": b"\n\n>a\n\n" + b"ACGT" * 10 + b"\n\n" + b"TTGCA" * 10 + b"\n\n",
        "exactly 33": b">a\n" + b"ACGTA" * 6 + b"CGT\n",
    }
    for name, fa in cases.items():
        fa = np.frombuffer(fa, np.uint8)
        want = oracle.run(fq, fq, fa, coder)
        _, got = eref.run_eref(fq, fq, fa, None, 0.9, 0.85, 1, coder=coder, return_results=True)
        assert got.size == want.size and (got == want).all(), name
    # empty fq1: the reference divides by zero (:1145); the product refuses loudly
    with pytest.raises(eref.ErefError):
        eref.run_eref(np.zeros(0, np.uint8), fq, np.frombuffer(cases["exactly 33"], np.uint8), coder=coder)
    # a coder entry outside {0,1,2} is rejected
    bad = coder.copy(); bad[5] = 7
    with pytest.raises(eref.ErefError):
        eref.ErefGPU(bad)


# ---- (c) properties that hold at any size ----------------------------------------------------------

def test_shard_invariance_and_idempotence(oracle):
    """splitting the DB at record boundaries and merging ordinals gives the same records; a second scan too"""
    db, f1, f2, _ = _case(31, n_genomes=120, db_bp=3_000_000, q_bp=600_000)
    f2 = f2[:eref.fastq_limit(f2, f1.size)]                # what main reads of fq2 (:1265-1291): the phase calls below take it as it is
    coder = eref.random_coder(31)
    _, whole = eref.run_eref(f1, f2, db.text, coder=coder, return_results=True)
    _, three = eref.run_eref(f1, f2, db.text, coder=coder, devices=(0, 0, 0), return_results=True)
    assert (whole == three).all()
    with eref.ErefGPU(coder) as g:
        g.query_add(f1); g.query_add(f2); g.ref_set(db.text)
        a = g.scan().copy(); b = g.scan().copy()
        assert (a == b).all() and (a == whole).all()
        # counting the same reads again cannot remove hits: every el stays or grows
        g.query_add(f1); g.query_add(f2)
        c = g.scan()
        assert (c["el"] >= a["el"]).all()


@pytest.mark.parametrize("chunks", [1, 3, 8, 64])
def test_pipelined_run_host_matches_phase_api(golden, chunks):
    """eref_run_host (copies overlapped, DB cut into pieces) == the phase-by-phase calls, record for record"""
    name, meta, fa, f1, f2 = golden
    _, a = eref.run_eref(f1, f2, fa, coder=meta["coder"], return_results=True)
    text, b = eref.run_eref(f1, f2, fa, coder=meta["coder"], pipelined=True, chunks=chunks, return_results=True)
    assert a.size == b.size and (a == b).all()
    assert text == meta["stdout"]["0.9,0.85"]


def test_run_host_against_a_resident_database(golden):
    """eref_run_host(fq1, fq2, NULL): the sample against the database set before, several samples in a row"""
    name, meta, fa, f1, f2 = golden
    with eref.ErefGPU(meta["coder"]) as g:
        with pytest.raises(eref.ErefError):
            g.run_host(f1, f2, None)                       # nothing set yet
        g.ref_set(fa)
        for a, b in ((f1, f2), (f2, f1), (f1, f2)):
            # the table depends on the order of the files only through main's byte budget: the second file is read up to
            # the first one's size (:1265-1291); the phase calls count exactly the text they are given
            g.query_reset()
            g.query_add(a); g.query_add(b[:eref.fastq_limit(b, a.size)])
            want = g.scan().copy()
            got = g.run_host(a, b, None).copy()
            assert (got == want).all(), name
        with pytest.raises(eref.ErefError):
            g.run_host(f1[:0], f2, None)                   # empty fq1: the reference divides by zero (:1145)
        assert (g.run_host(f1, f2, None) == want).all()    # and the database is still set


def test_run_host_takes_the_query_from_device_memory(golden):
    """the FASTQ texts may already be in HBM (a query all-gathered over NVLink): same records"""
    import torch
    name, meta, fa, f1, f2 = golden
    d1, d2 = torch.from_numpy(np.array(f1)).cuda(), torch.from_numpy(np.array(f2)).cuda()
    with eref.ErefGPU(meta["coder"]) as g:
        want = g.run_host(f1, f2, fa).copy()
        got = g.run_host(eref.DeviceBytes(d1.data_ptr(), d1.numel()), eref.DeviceBytes(d2.data_ptr(), d2.numel()), fa)
        assert (got == want).all(), name


def test_run_host_refuses_empty_fq1():
    coder = eref.random_coder(3)
    fa = np.frombuffer(b">a\n" + b"ACGT" * 20 + b"\n", np.uint8)
    fq = np.frombuffer(b"@r\n" + b"ACGT" * 10 + b"\n+\n" + b"I" * 40 + b"\n", np.uint8)
    with eref.ErefGPU(coder) as g:
        with pytest.raises(eref.ErefError):
            g.run_host(np.zeros(0, np.uint8), fq, fa, cap=4)
        assert g.run_host(fq, fq, fa, cap=4).size == 1           # and the context is still usable


def test_fq2_is_read_only_up_to_fq1s_size(tmp_path):
    """main hands file_size(fq1) to read_fastq as the byte budget of BOTH files (:1265-1291, :942-945): with fq2 twice
    as long as fq1 the reference reads only its head.  Expected stdout: the unmodified reference on the same inputs
    (tests/golden/fq2_limit.json) — through run_eref by phases and pipelined, eref_run_host with the query in device
    memory (the caller passes the limited size), and the executable."""
    import torch
    from tests.conftest import fq2_limit_inputs
    meta, fa, f1, f2 = fq2_limit_inputs()
    lim = eref.fastq_limit(f2, f1.size)
    assert lim == meta["fq2_bytes_read"] < f2.size
    for key, expect in meta["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        for pipelined in (False, True):
            assert eref.run_eref(f1, f2, fa, None, hr, phr, 1, coder=meta["coder"], pipelined=pipelined) == expect, (key, pipelined)
    assert eref.run_eref(f1, f2, fa, None, 0.5, 0.3, 1, coder=meta["coder"], devices=(0, 0), shard_table=True,
                         share_reads=True) == meta["stdout"]["0.5,0.3"]
    d1, d2 = torch.from_numpy(np.array(f1)).cuda(), torch.from_numpy(np.array(f2)).cuda()
    with eref.ErefGPU(meta["coder"], 0.5, 0.3) as g:
        host = g.run_host(f1, f2, fa).copy()                      # host fq2: the call applies the budget itself
        dev = g.run_host(eref.DeviceBytes(d1.data_ptr(), d1.numel()), eref.DeviceBytes(d2.data_ptr(), lim), fa).copy()
        assert (host == dev).all() and eref.format_results(host) == meta["stdout"]["0.5,0.3"]
        whole = g.run_host(eref.DeviceBytes(d1.data_ptr(), d1.numel()), eref.DeviceBytes(d2.data_ptr(), d2.numel()), fa)
        assert eref.format_results(whole) != meta["stdout"]["0.5,0.3"]      # (and the budget is what makes the difference)
    fa.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    hdr = np.zeros(100, "<u4"); hdr[:96] = np.asarray(meta["coder"], np.uint32)
    hdr.tofile(tmp_path / "db.fa.k32.index.dat")
    for gpus in ("1", "2"):
        if gpus == "2" and eref.load_library().eref_device_count() < 2:
            continue
        for key, expect in meta["stdout"].items():
            hr, phr = key.split(",")
            r = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "4"], cwd=tmp_path,
                               capture_output=True, env=dict(os.environ, EREF_GPUS=gpus))
            assert r.returncode == 0, r.stderr.decode()
            assert r.stdout.decode() == expect, (key, gpus)


def test_strand_and_order_invariance():
    """reverse-complementing every read, or swapping fq1/fq2, leaves the table and the result unchanged"""
    db, f1, f2, _ = _case(41, n_genomes=40, db_bp=1_000_000, q_bp=300_000)
    coder = eref.random_coder(41)
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")

    def rc_fastq(fq):
        lines = fq.tobytes().split(b"\n")
        for i in range(1, len(lines), 4):
            lines[i] = lines[i].translate(comp)[::-1]
        return np.frombuffer(b"\n".join(lines), np.uint8)

    _, a = eref.run_eref(f1, f2, db.text, coder=coder, return_results=True)
    _, b = eref.run_eref(rc_fastq(f1), rc_fastq(f2), db.text, coder=coder, return_results=True)
    _, c = eref.run_eref(f2, f1, db.text, coder=coder, return_results=True)
    assert (a == b).all() and (a == c).all()
    assert (a["el"] > 0).any()


def test_present_genomes_are_selected_at_depth():
    """a genome tiled at 8x by error-free reads is reported with el_ratio ~ 1; absent genomes are not"""
    db = synth.make_db(50, 1_000_000, 51, frac_mutated=0.0, frac_n=0.0, frac_lower=0.0, n_short=0)
    present = np.array([3, 17, 30])
    f1, f2, _ = synth.make_reads(db, 0, 52, cov_lo=8, cov_hi=8, sub_rate=0.0, n_rate=0.0, present=present)
    coder = eref.random_coder(51)
    text, res = eref.run_eref(f1, f2, db.text, coder=coder, return_results=True)
    sel = {int(l.split("\t")[1]) for l in text.splitlines()}
    assert sel == {int(p) + 1 for p in present}
    absent = np.setdiff1d(np.arange(50), present)
    assert (res["el"][absent] == 0).all()


def test_eref_binary_end_to_end(golden, tmp_path):
    """the drop-in executable: same argv, stdout and side-effect files as the reference"""
    name, meta, fa, f1, f2 = golden
    if not os.path.exists(eref.EREF_BIN):
        pytest.fail("palace_b200/bin/eref not built")
    fa.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    # a reference-format index header next to the DB carries the coder (what a shared DB directory holds)
    hdr = np.zeros(100, "<u4"); hdr[:96] = np.asarray(meta["coder"], np.uint32)
    hdr.tofile(tmp_path / "db.fa.k32.index.dat")
    for key, expect in meta["stdout"].items():
        hr, phr = key.split(",")
        for gpus in ("1", "2"):
            env = dict(os.environ, EREF_GPUS=gpus)
            if gpus == "2" and eref.load_library().eref_device_count() < 2:
                continue
            r = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "4"], cwd=tmp_path,
                               capture_output=True, env=env)
            assert r.returncode == 0, r.stderr.decode()
            assert r.stdout.decode() == expect, f"{name} {key} gpus={gpus}"
            assert (tmp_path / "tmp_out.txt").exists() and (tmp_path / "tmp_out.txt").stat().st_size == 0


# ---- framing stress: line / record / tile boundaries in every position ----------------------------

def _adversarial_fasta(rng, n_records, big=False):
    """Records with random line widths (empty lines, lines longer than an 8 KiB tile, CRLF), headers of any
    length (also longer than a tile), records of 0..40 bases around the len > 32 rule, '>' inside lines,
    lower case, N and junk bytes, optional missing final newline and leading header-less sequence."""
    alphabet = np.frombuffer(b"ACGTACGTACGTACGTacgtNn>-\r *", np.uint8)
    parts = []
    if rng.random() < 0.5:
        parts.append(bytes(rng.choice(alphabet[:16], int(rng.integers(1, 200)))) + b"\n")    # record 0: no header
    for r in range(n_records):
        kind = rng.random()
        hdr_len = int(rng.integers(0, 30)) if kind < 0.9 else int(rng.integers(8000, 20000))
        parts.append(b">" + bytes(rng.choice(alphabet[:20], hdr_len)).replace(b"\n", b"") + b"\n")
        u = rng.random()
        if u < 0.25:
            length = int(rng.integers(0, 70))
        elif u < 0.9 or not big:
            length = int(rng.integers(70, 3000))
        else:
            length = int(rng.integers(9000, 40000))
        seq = rng.choice(alphabet if rng.random() < 0.3 else alphabet[:16], length)
        seq = bytes(seq).replace(b"\n", b"")
        pos = 0
        while pos < len(seq):
            w = int(rng.choice([1, 2, 31, 32, 33, 60, 70, 80, 255, 256, 257, 9000])) if rng.random() < 0.5 else 70
            line = seq[pos:pos + w]
            if line[:1] == b">":                      # a sequence line must not start with '>' (it would be a header)
                line = b"A" + line[1:]
            parts.append(line + (b"\r\n" if rng.random() < 0.05 else b"\n"))
            pos += w
            if rng.random() < 0.05:
                parts.append(b"\n")
    text = b"".join(parts)
    if rng.random() < 0.5 and text.endswith(b"\n"):
        text = text[:-1]
    return np.frombuffer(text, np.uint8)


@pytest.mark.parametrize("seed", [101, 102, 103, 104, 105, 106])
def test_fasta_framing_stress(oracle, seed):
    """records, lengths, ordinals and every code of every kept record == the oracle's getline framing (:686-756)"""
    rng = np.random.default_rng(seed)
    fa = _adversarial_fasta(rng, int(rng.integers(5, 400)), big=seed % 2 == 0)
    coder = eref.random_coder(seed)
    seqs, n_records = oracle.parse_fasta(fa)
    with eref.ErefGPU(coder) as g:
        g.ref_set(fa)
        kept, recs, bases = g.ref_count()
        assert kept == len(seqs)
        assert bases == sum(len(s) for s in seqs)
        raw = fa.tobytes()
        for i in rng.permutation(len(seqs))[:60]:
            off, hl, ln = g.ref_record(int(i) + 1)
            assert ln == len(seqs[i])
            if hl or off:
                assert raw[off:off + 1] == b">" and b"\n" not in raw[off:off + hl]
                assert off + hl == len(raw) or raw[off + hl:off + hl + 1] == b"\n"
            want = oracle.record_codes(coder, seqs[i])
            got = g.get_codes(int(i) + 1)
            assert got.size == want.size and (got == want).all(), f"record {i}"


def test_fasta_many_tiny_records_fill_a_tile(oracle):
    """33-base records with 2-byte headers: the densest packing of plane words per text tile (ring bound)"""
    rng = np.random.default_rng(7)
    recs = [b">\n" + bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), 33)) + b"\n" for _ in range(3000)]
    fa = np.frombuffer(b"".join(recs), np.uint8)
    coder = eref.random_coder(9)
    seqs, _ = oracle.parse_fasta(fa)
    with eref.ErefGPU(coder) as g:
        g.ref_set(fa)
        assert g.ref_count()[0] == 3000 == len(seqs)
        for i in (0, 1, 226, 227, 228, 1499, 2999):
            assert (g.get_codes(i + 1) == oracle.record_codes(coder, seqs[i])).all()


@pytest.mark.parametrize("seed", [201, 202, 203])
def test_fastq_framing_stress(oracle, seed):
    """FASTQ texts with empty lines, very long lines, CRLF and a missing final newline: the sequence-line byte
    sum (:1138-1140) and the table levels equal an independent count over the lines with index % 4 == 1"""
    import collections
    rng = np.random.default_rng(seed)
    alphabet = np.frombuffer(b"ACGTACGTACGTACGTacgtN", np.uint8)
    lines = []
    for _ in range(int(rng.integers(50, 1500))):
        u = rng.random()
        n = 0 if u < 0.1 else int(rng.integers(1, 400)) if u < 0.95 else int(rng.integers(8000, 30000))
        lines.append(bytes(rng.choice(alphabet, n)) + (b"\r" if rng.random() < 0.03 else b""))
    text = b"\n".join(lines) + (b"\n" if rng.random() < 0.5 else b"")
    fq = np.frombuffer(text, np.uint8)
    coder = eref.random_coder(seed)
    cnt = collections.Counter()
    seq_lines = text.split(b"\n")[1::4]
    for s in seq_lines:
        if len(s) >= 32:
            c = oracle.record_codes(coder, s).reshape(-1, 3)
            cnt.update(c[c.any(axis=1)].ravel().tolist())
    with eref.ErefGPU(coder) as g:
        assert g.query_add(fq) == sum(len(s) for s in seq_lines) == oracle.fastq_seq_bytes(fq)
        for level in (1, 2, 3):
            assert g.query_popcount(level) == sum(1 for v in cnt.values() if v >= level)
        # the same file again: every count doubles
        g.query_add(fq)
        for level in (1, 2, 3):
            assert g.query_popcount(level) == sum(1 for v in cnt.values() if 2 * v >= level)


# ---- the reference's on-disk index, produced on the GPU ---------------------

def test_index_file_matches_the_reference(golden, tmp_path):
    """<db>.k32.index.dat body (every code of every position) and <db>.genome.len.txt are byte-identical to
    what the UNMODIFIED reference wrote for the same FASTA and coder header (tests/golden/make_golden.py)"""
    import hashlib
    name, meta, fa, f1, f2 = golden
    idx, ln = str(tmp_path / "db.fa.k32.index.dat"), str(tmp_path / "db.fa.genome.len.txt")
    words = eref.write_index(fa, meta["coder"], idx, ln)
    assert words == meta["index_words"]
    w = np.fromfile(idx, "<u4")
    assert w.size == meta["index_words"]
    assert [int(x) for x in (w[:96] & 0xFFFF)] == meta["coder"]          # what saved_random_coder reads back (:1104-1122)
    assert hashlib.sha256(w[100:].tobytes()).hexdigest() == meta["index_body_sha256"]
    assert open(ln).read() == meta["genome_len_txt"]
    assert (eref.read_coder(idx) == np.asarray(meta["coder"])).all()


def test_index_body_chunking(oracle):
    """chunks of any size concatenate to the same body; a chunk smaller than one record is refused loudly"""
    db, _, _, _ = _case(41, n_genomes=30, db_bp=400_000, q_bp=50_000)
    coder = eref.random_coder(41)
    with eref.ErefGPU(coder) as g:
        g.ref_set(db.text)
        whole = np.concatenate(list(g.index_body(1 << 24)))
        small = np.concatenate(list(g.index_body(int(3 * db.lengths.max()) + 8)))
        assert whole.size == small.size and (whole == small).all()
        seqs, _ = oracle.parse_fasta(db.text)
        assert whole[0] == len(seqs[0])
        assert (whole[1:1 + 3 * (len(seqs[0]) - 31)] == oracle.record_codes(coder, seqs[0])).all()
        buf = np.zeros(16, "<u4")
        import ctypes
        n, nxt = ctypes.c_size_t(), ctypes.c_uint32()
        rc = g.lib.eref_index_body(g._h, 1, ctypes.c_void_p(buf.ctypes.data), buf.size, ctypes.byref(n), ctypes.byref(nxt))
        assert rc == -4


def test_eref_binary_writes_a_reference_index(golden, tmp_path):
    """EREF_WRITE_INDEX=1: the executable leaves the reference's index files next to the DB (coder from the sidecar)"""
    import hashlib
    name, meta, fa, f1, f2 = golden
    fa.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    with open(tmp_path / "db.fa.k32.coder", "w") as f:
        f.write(" ".join(str(x) for x in meta["coder"]) + "\n")
    key, expect = next(iter(meta["stdout"].items()))
    hr, phr = key.split(",")
    r = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "1"], cwd=tmp_path,
                       capture_output=True, env=dict(os.environ, EREF_WRITE_INDEX="1", EREF_GPUS="1"))
    assert r.returncode == 0, r.stderr.decode()
    assert r.stdout.decode() == expect
    w = np.fromfile(tmp_path / "db.fa.k32.index.dat", "<u4")
    assert w.size == meta["index_words"] and [int(x) for x in (w[:96] & 0xFFFF)] == meta["coder"]
    assert hashlib.sha256(w[100:].tobytes()).hexdigest() == meta["index_body_sha256"]
    assert open(tmp_path / "db.fa.genome.len.txt").read() == meta["genome_len_txt"]
    # a second run finds the index, takes the coder from it and prints the same lines
    r2 = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "1"], cwd=tmp_path,
                        capture_output=True, env=dict(os.environ, EREF_GPUS="1"))
    assert r2.returncode == 0 and r2.stdout.decode() == expect


# ---- the reference's rand()-driven down-sampling -----------------------------

def _downsample_fixture():
    import json
    from tests.conftest import load_golden, GOLDEN
    ds = json.load(open(os.path.join(GOLDEN, "downsample.json")))
    return (ds,) + load_golden(ds["inputs"])[1:]


@pytest.mark.parametrize("pipelined", [False, True])
def test_downsampling_matches_the_reference(oracle, pipelined):
    """stdout of the diagnostic reference variant (down_sampling_size lowered so that the small fixture down-samples)"""
    ds, fa, f1, f2 = _downsample_fixture()
    for key, expect in ds["stdout"].items():
        hr, phr = (float(x) for x in key.split(","))
        text, got = eref.run_eref(f1, f2, fa, None, hr, phr, 1, coder=ds["coder"], pipelined=pipelined,
                                  down_sampling_size=ds["down_sampling_size"], return_results=True)
        assert text == expect, key
        want = oracle.run(f1, f2, fa, ds["coder"], hr, phr, threads=2, down_sampling_size=ds["down_sampling_size"])
        assert (got == want).all()
    # two shards: the draws are made once and shared
    text2 = eref.run_eref(f1, f2, fa, None, 0.9, 0.85, 1, coder=ds["coder"], devices=(0, 0),
                          down_sampling_size=ds["down_sampling_size"])
    assert text2 == ds["stdout"]["0.9,0.85"]


def test_downsampling_table_counts_only_the_drawn_reads(oracle):
    import collections
    ds, fa, f1, f2 = _downsample_fixture()
    coder = ds["coder"]
    with eref.ErefGPU(coder) as g:
        seq1, l1 = g.query_census(f1)
        assert seq1 == oracle.fastq_seq_bytes(f1) and l1 == len(f1.tobytes().split(b"\n")[1::4])
        ratio = eref.sample_ratio(seq1, ds["down_sampling_size"])
        assert ratio == ds["ratio"]
        k1 = eref.draw_keep(ratio, l1)
        assert g.query_add_masked(f1, k1, l1) == seq1          # cal_sam_ratio's sum does not depend on the draws
        cnt = collections.Counter()
        for i, s in enumerate(f1.tobytes().split(b"\n")[1::4]):
            if (int(k1[i >> 5]) >> (i & 31)) & 1 and len(s) >= 32:
                c = oracle.record_codes(coder, s).reshape(-1, 3)
                cnt.update(c[c.any(axis=1)].ravel().tolist())
        for level in (1, 2, 3):
            assert g.query_popcount(level) == sum(1 for v in cnt.values() if v >= level)


def test_eref_binary_downsamples_like_the_reference(tmp_path):
    ds, fa, f1, f2 = _downsample_fixture()
    fa.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    hdr = np.zeros(100, "<u4"); hdr[:96] = np.asarray(ds["coder"], np.uint32)
    hdr.tofile(tmp_path / "db.fa.k32.index.dat")           # a warm run: the coder comes from the index, rand() from srand(1)
    for gpus in ("1", "2"):
        if gpus == "2" and eref.load_library().eref_device_count() < 2:
            continue
        for key, expect in ds["stdout"].items():
            hr, phr = key.split(",")
            r = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "1"], cwd=tmp_path,
                               capture_output=True,
                               env=dict(os.environ, EREF_GPUS=gpus, EREF_DOWN_SAMPLING_SIZE=str(ds["down_sampling_size"])))
            assert r.returncode == 0, r.stderr.decode()
            assert r.stdout.decode() == expect, f"{key} gpus={gpus}"


def test_eref_binary_reads_gzip_inputs(golden, tmp_path):
    """gzip-compressed FASTQ / FASTA inputs give the stdout of the plain files"""
    import shutil
    name, meta, fa, f1, f2 = golden
    from tests.conftest import GOLDEN
    for ext, dst in (("fa", "db.fa.gz"), ("r1.fq", "r1.fq.gz"), ("r2.fq", "r2.fq.gz")):
        shutil.copy(os.path.join(GOLDEN, f"{name}.{ext}.gz"), tmp_path / dst)
    with open(tmp_path / "db.fa.gz.k32.coder", "w") as f:
        f.write(" ".join(str(x) for x in meta["coder"]) + "\n")
    key, expect = next(iter(meta["stdout"].items()))
    hr, phr = key.split(",")
    r = subprocess.run([eref.EREF_BIN, "r1.fq.gz", "r2.fq.gz", "db.fa.gz", "tmp_out.txt", hr, phr, "1"], cwd=tmp_path,
                       capture_output=True, env=dict(os.environ, EREF_GPUS="1"))
    assert r.returncode == 0, r.stderr.decode()
    assert r.stdout.decode() == expect


# ---- candidate filter on ragged databases: a property that needs no oracle --------------------------

def _ragged_case(seed, n_records=3000):
    """records of 33..700 bases (plus a few long ones) packed tightly, so that the 17-word look-back / look-ahead of
    the candidate filter crosses record boundaries almost everywhere; a third of the records is covered 4x by reads"""
    rng = np.random.default_rng(seed)
    lens = rng.integers(33, 700, n_records)
    lens[rng.integers(0, n_records, 20)] = rng.integers(3000, 20000, 20)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    fa, reads = [], []
    for i, n in enumerate(lens):
        seq = acgt[rng.integers(0, 4, int(n))]
        fa.append(b">r%d\n" % i + seq.tobytes() + b"\n")
        if i % 3 == 0:
            for _ in range(4):
                reads.append(seq.tobytes())
                if n > 400:                                   # the reference's line buffer holds 400 characters
                    reads[-1] = None
                    for s in range(0, int(n) - 32, 150):
                        reads.append(seq[s:s + 300].tobytes())
    reads = [r for r in reads if r]
    order = rng.permutation(len(reads))
    def fastq(idx):
        return np.frombuffer(b"".join(b"@q\n" + reads[k] + b"\n+\n" + b"I" * len(reads[k]) + b"\n" for k in idx), np.uint8)
    return np.frombuffer(b"".join(fa), np.uint8), fastq(order[0::2]), fastq(order[1::2])


@pytest.mark.parametrize("seed", [301, 302, 303])
def test_pruned_scan_equals_full_scan_on_ragged_records(seed):
    fa, f1, f2 = _ragged_case(seed)
    coder = eref.random_coder(seed)
    for ratios in ((0.9, 0.85), (0.99, 0.97), (0.95, 0.9), (0.5, 0.45)):
        _, full = eref.run_eref(f1, f2, fa, None, *ratios, 1, coder=coder, prune=False, return_results=True)
        for pipelined in (False, True):
            _, got = eref.run_eref(f1, f2, fa, None, *ratios, 1, coder=coder, prune=True, pipelined=pipelined,
                                   chunks=3, return_results=True)
            assert (got == full).all(), f"seed {seed} ratios {ratios} pipelined={pipelined}"
        if ratios[0] == 0.9:
            assert (full["el"] > 0).sum() > 100


def test_ragged_records_match_the_oracle(oracle):
    fa, f1, f2 = _ragged_case(304, n_records=800)
    coder = eref.random_coder(304)
    want = oracle.run(f1, f2, fa, coder, 0.9, 0.85, threads=4)
    _, got = eref.run_eref(f1, f2, fa, None, 0.9, 0.85, 1, coder=coder, return_results=True)
    assert (got == want).all()


# ---- sharded table build (the table counted by N GPUs, exchanged through peer pointers / a hook) --------------

def test_thresholds_at_the_ends_of_their_range(oracle):
    """int(500 * ratio) of :513-514 at 0, 1, 500 and with the perfect threshold above the plain one, on records from 33
    bases up (shorter than a window, clamped paddings): every record equals the oracle's (which the live fuzz compares
    with the reference binary at the same pairs, oracle/fuzz_vs_reference.py ... extremes)"""
    fa, f1, f2 = _ragged_case(305, n_records=600)
    coder = eref.random_coder(305)
    for ratios in ((0.0, 0.0), (1.0, 1.0), (0.002, 0.002), (0.1, 0.9)):
        want = oracle.run(f1, f2, fa, coder, *ratios, threads=4)
        text, got = eref.run_eref(f1, f2, fa, None, *ratios, 1, coder=coder, return_results=True)
        assert got.size == want.size and (got == want).all(), ratios
        assert text == oracle.format(want), ratios


@pytest.mark.parametrize("parts", [2, 4, 3])
def test_sharded_table_build_equals_replicated(oracle, parts):
    """N contexts on one GPU emulate an N-GPU box: context i counts slices [8192 i/N, 8192 (i+1)/N) and stores them into
    every context's table; the records equal the replicated build's and the oracle's, with and without the filter"""
    db, f1, f2, _ = _case(71 + parts, n_genomes=90, db_bp=2_500_000, q_bp=700_000)
    coder = eref.random_coder(71)
    want = oracle.run(f1, f2, db.text, coder, 0.9, 0.85, threads=4)
    for prune in (True, False):
        _, rep = eref.run_eref(f1, f2, db.text, coder=coder, devices=(0,) * parts, prune=prune, return_results=True)
        _, sh = eref.run_eref(f1, f2, db.text, coder=coder, devices=(0,) * parts, prune=prune, shard_table=True,
                              return_results=True)
        assert rep.size == sh.size == want.size
        assert (sh == rep).all() and (sh == want).all()
    assert (want["el"] > 0).any()


def test_sharded_table_levels_sum_to_the_replicated_ones():
    """every level's population, summed over the parts, equals the replicated build's; every part's table is the same"""
    db, f1, f2, _ = _case(75, n_genomes=30, db_bp=600_000, q_bp=500_000)
    coder = eref.random_coder(75)
    with eref.ErefGPU(coder, level_pop=True) as g:
        g.query_add(f1); g.query_add(f2)
        whole = [g.query_popcount(l) for l in (1, 2, 3)]
    ctxs = [eref.ErefGPU(coder, level_pop=True) for _ in range(4)]
    try:
        eref.wire_table_peers(ctxs)
        for g in ctxs:
            g.query_add(f1); g.query_add(f2)
        for g in ctxs:
            g.ref_set(db.text)                             # queues the build
        for g in ctxs:
            g.sync()
        part_pops = [[g.query_popcount(l) for l in (1, 2, 3)] for g in ctxs]
        assert [sum(p[l] for p in part_pops) for l in range(3)] == whole
        assert all(p[0] > 0 for p in part_pops)
        recs = [g.scan().copy() for g in ctxs]
        assert all((r == recs[0]).all() for r in recs[1:])
    finally:
        for g in ctxs:
            g.close()


def test_sharded_table_pipelined_call_side_by_side(golden):
    """eref_run_host on two contexts that share the table build, one host thread each (as the executable runs its GPUs)"""
    name, meta, fa, f1, f2 = golden
    _, a = eref.run_eref(f1, f2, fa, coder=meta["coder"], devices=(0, 0), return_results=True)
    for _ in range(2):
        text, b = eref.run_eref(f1, f2, fa, coder=meta["coder"], devices=(0, 0), shard_table=True, pipelined=True, chunks=2,
                                return_results=True)
        assert a.size == b.size and (a == b).all()
        assert text == meta["stdout"]["0.9,0.85"]


def test_sharded_table_completion_hook(golden):
    """without peer pointers the caller's hook completes the table on the context's stream (here: device copies of the
    other context's part; in bench.py an in-place NCCL all-gather)"""
    try:
        from cuda.bindings import runtime as cudart
    except ImportError:
        from cuda import cudart
    name, meta, fa, f1, f2 = golden
    ctxs = [eref.ErefGPU(meta["coder"]) for _ in range(2)]
    try:
        ptrs = [g.table_export() for g in ctxs]
        for i, g in enumerate(ctxs):
            g.table_set_part(i, 2)
        called = []

        def make_hook(i):
            def hook(stream):
                j = 1 - i                                  # fetch part j from context j's table
                toff, tbytes, soff, sbytes = ctxs[i].table_part_range(j)    # (the parts are re-cut for every query)
                assert (toff, tbytes, soff, sbytes) == ctxs[j].table_part_range()
                for k, (off, nb) in ((1, (toff, tbytes)), (2, (soff, sbytes))):
                    err, = cudart.cudaMemcpyAsync(ptrs[i][k] + off, ptrs[j][k] + off, nb,
                                                  cudart.cudaMemcpyKind.cudaMemcpyDeviceToDevice, stream)
                    assert err == cudart.cudaError_t.cudaSuccess
                called.append(i)
            return hook
        for i, g in enumerate(ctxs):
            g.table_set_hook(make_hook(i))
            g.query_add(f1); g.query_add(f2)
        for g in ctxs:
            g.ref_set(fa)
        for g in ctxs:
            g.sync()                                       # both local parts are built before either hook copies
        recs = [g.scan().copy() for g in ctxs]
        assert sorted(called) == [0, 1]
        _, want = eref.run_eref(f1, f2, fa, coder=meta["coder"], return_results=True)
        assert all((r == want).all() for r in recs)
        assert eref.format_results(recs[0]) == meta["stdout"]["0.9,0.85"]
    finally:
        for g in ctxs:
            g.close()


# ---- SURVEY.md §8(f) row f1 fused into the executable ---------------------------------------------------------

def _f1_fixture_cases():
    import json
    from tests.conftest import GOLDEN
    golden = json.load(open(os.path.join(GOLDEN, "f1.json")))
    return [(k, v) for k, v in golden.items() if "fixture" in v]


@pytest.mark.parametrize("case", [k for k, _ in _f1_fixture_cases()])
def test_eref_binary_writes_the_next_steps_outputs(case, tmp_path):
    """EREF_REFS_FASTA / EREF_REFS_PERCENT: the drop-in writes phage_refs.fasta and {prefix}_ref_percent.txt itself;
    expected = what the UNMODIFIED get_ref_by_index.py made of the REFERENCE eref's stdout on the same fixture
    (tests/golden/f1.json, tests/golden/make_golden_f1.py), including the index drift past records of <= 32 bases"""
    import hashlib
    c = dict(_f1_fixture_cases())[case]
    meta, fa, f1, f2 = load_golden(c["fixture"])
    fa.tofile(tmp_path / "db.fa"); f1.tofile(tmp_path / "r1.fq"); f2.tofile(tmp_path / "r2.fq")
    hdr = np.zeros(100, "<u4"); hdr[:96] = np.asarray(meta["coder"], np.uint32)
    hdr.tofile(tmp_path / "db.fa.k32.index.dat")
    hr, phr = c["ratios"].split(",")
    env = dict(os.environ, EREF_GPUS="1", EREF_REFS_FASTA=str(tmp_path / "refs.fa"), EREF_REFS_PERCENT=str(tmp_path / "pct.txt"))
    r = subprocess.run([eref.EREF_BIN, "r1.fq", "r2.fq", "db.fa", "tmp_out.txt", hr, phr, "4"], cwd=tmp_path,
                       capture_output=True, env=env)
    assert r.returncode == 0, r.stderr.decode()
    assert r.stdout.decode() == c["names"]
    assert (tmp_path / "pct.txt").read_text() == c["out_percent"]
    out_fa = (tmp_path / "refs.fa").read_bytes()
    assert out_fa.count(b">") == c["out_fasta_records"]
    assert hashlib.sha256(out_fa).hexdigest() == c["out_fasta_sha256"]


def test_eref_binary_reads_bgzf_and_multi_member_gzip(golden, tmp_path):
    """BGZF inputs (blocks inflated by several threads) and concatenated gzip members give the stdout of the plain files"""
    import gzip
    import struct
    import zlib
    name, meta, fa, f1, f2 = golden

    def bgzf(data: bytes, block=30000) -> bytes:
        out = []
        for off in list(range(0, len(data), block)) + [len(data)]:      # the last (empty) block is BGZF's EOF marker
            chunk = data[off:off + block] if off < len(data) else b""
            co = zlib.compressobj(6, zlib.DEFLATED, -15)
            body = co.compress(chunk) + co.flush()
            bsize = len(body) + 25
            out.append(b"\x1f\x8b\x08\x04\x00\x00\x00\x00\x00\xff\x06\x00BC\x02\x00" + struct.pack("<H", bsize) + body +
                       struct.pack("<II", zlib.crc32(chunk), len(chunk)))
        return b"".join(out)

    raw1 = f1.tobytes()
    (tmp_path / "db.fa.gz").write_bytes(bgzf(fa.tobytes()))
    (tmp_path / "r1.fq.gz").write_bytes(gzip.compress(raw1[: len(raw1) // 3]) + gzip.compress(raw1[len(raw1) // 3:]))
    (tmp_path / "r2.fq.gz").write_bytes(bgzf(f2.tobytes(), block=65000))
    with open(tmp_path / "db.fa.gz.k32.coder", "w") as f:
        f.write(" ".join(str(x) for x in meta["coder"]) + "\n")
    key, expect = next(iter(meta["stdout"].items()))
    hr, phr = key.split(",")
    r = subprocess.run([eref.EREF_BIN, "r1.fq.gz", "r2.fq.gz", "db.fa.gz", "tmp_out.txt", hr, phr, "8"], cwd=tmp_path,
                       capture_output=True, env=dict(os.environ, EREF_GPUS="1", EREF_TIMING="1"))
    assert r.returncode == 0, r.stderr.decode()
    assert r.stdout.decode() == expect
    assert "inflated" in r.stderr.decode()


def test_a_query_that_outgrows_the_optimistic_buffers_is_retried(oracle):
    """the code buffers are sized from the text length (a FASTQ spends as many bytes on qualities as on bases) so that
    nothing waits for the count; a file whose quality lines are shorter than its sequence lines breaks that rule: the
    phase API reports EREF_ERR_RETRY once, the pipelined call retries by itself, and both then match the oracle"""
    db, f1, f2, _ = _case(81, n_genomes=40, db_bp=800_000, q_bp=600_000)
    coder = eref.random_coder(81)
    # the same reads with one-character quality lines, repeated until the file is as long as the original: nearly
    # twice the sequence bytes per text byte
    def squeeze(fq, target):
        lines = fq.tobytes().split(b"\n")
        recs = [b"\n".join([lines[k], lines[k + 1], b"+", b"I"]) + b"\n" for k in range(0, len(lines) - 3, 4)]
        out, size, k = [], 0, 0
        while size < target:
            out.append(recs[k % len(recs)]); size += len(out[-1]); k += 1
        return np.frombuffer(b"".join(out), np.uint8)
    s1, s2 = squeeze(f1, int(f1.size * 0.98)), squeeze(f2, int(f2.size * 0.98))
    want = oracle.run(s1, s2, db.text, coder, 0.9, 0.85, threads=4)
    with eref.ErefGPU(coder) as g:
        g.query_add(f1); g.query_add(f2); g.ref_set(db.text); g.scan()     # sizes the buffers for files of this size
        g.query_reset()
        g.query_add(s1); g.query_add(s2)
        try:
            got = g.scan().copy()
            retried = False
        except eref.ErefError as e:
            assert "(-6)" in str(e)
            retried = True
            g.query_reset(); g.query_add(s1); g.query_add(s2)
            got = g.scan().copy()
        assert (got == want).all()
    with eref.ErefGPU(coder) as g:
        g.run_host(f1, f2, db.text)
        got = g.run_host(s1, s2, db.text).copy()
        assert (got == want).all()
    assert retried


def test_small_queries_are_not_shared_out(golden, monkeypatch):
    """below EREF_SHARD_MIN_BYTES of query text the exchange of the table would cost more than counting it everywhere:
    every context of a sharded setup then builds the whole table for itself (same decision on every GPU: it depends
    only on the file sizes) and the records are the same"""
    name, meta, fa, f1, f2 = golden
    monkeypatch.delenv("EREF_SHARD_MIN_BYTES", raising=False)      # the library's default: 160 MB of FASTQ text
    _, want = eref.run_eref(f1, f2, fa, coder=meta["coder"], return_results=True)
    ctxs = [eref.ErefGPU(meta["coder"], level_pop=True) for _ in range(2)]
    try:
        eref.wire_table_peers(ctxs)
        with eref.ErefGPU(meta["coder"], level_pop=True) as one:
            one.query_add(f1); one.query_add(f2)
            whole = [one.query_popcount(l) for l in (1, 2, 3)]
        for g in ctxs:
            g.query_add(f1); g.query_add(f2); g.ref_set(fa)
        for g in ctxs:
            g.sync()
        for g in ctxs:
            assert [g.query_popcount(l) for l in (1, 2, 3)] == whole      # not a part: the whole table
            assert (g.scan() == want).all()
        assert ctxs[0].table_part_range() == (0, 1 << 29, 0, 1 << 25)
    finally:
        for g in ctxs:
            g.close()


# ---- shared reads: every GPU partitions 1/N of each query file; a build pulls its table slices out of the peers' lists ----

def test_split_fastq_cuts_at_record_boundaries():
    db, f1, f2, _ = _case(90, n_genomes=10, db_bp=100_000, q_bp=60_000)
    for shares in (1, 2, 3, 8):
        parts = eref.split_fastq(f1, shares)
        assert np.concatenate(parts).tobytes() == f1.tobytes()
        for p in parts:
            assert p.size == 0 or (p[0] == ord("@") and p[-1] == 10 and np.count_nonzero(p == 10) % 4 == 0)


@pytest.mark.parametrize("parts", [2, 3, 4, 8])
def test_shared_reads_equal_the_replicated_build(oracle, parts):
    """N contexts on one GPU, each adding 1/N of the reads: the tables every context ends up with, hence the records, equal
    those of one context that adds everything, and the oracle's"""
    db, f1, f2, _ = _case(91 + parts, n_genomes=90, db_bp=2_500_000, q_bp=700_000)
    coder = eref.random_coder(91)
    want = oracle.run(f1, f2, db.text, coder, 0.9, 0.85, threads=4)
    for prune in (True, False):
        _, sh = eref.run_eref(f1, f2, db.text, coder=coder, devices=(0,) * parts, prune=prune, shard_table=True, share_reads=True,
                              return_results=True)
        assert sh.size == want.size and (sh == want).all()
    assert (want["el"] > 0).any()


def test_shared_reads_levels_and_repeated_builds(golden):
    """level populations summed over the parts equal one context's; a second sample through the same wiring (new epoch,
    other files, ragged shares: one context gets an empty piece) is right too; run_host is refused while sharing"""
    name, meta, fa, f1, f2 = golden
    with eref.ErefGPU(meta["coder"], level_pop=True) as one:
        one.query_add(f1); one.query_add(f2)
        whole = [one.query_popcount(l) for l in (1, 2, 3)]
    _, want = eref.run_eref(f1, f2, fa, coder=meta["coder"], return_results=True)
    _, want_swapped = eref.run_eref(f2, f2, fa, coder=meta["coder"], return_results=True)
    ctxs = [eref.ErefGPU(meta["coder"], level_pop=True) for _ in range(3)]
    try:
        eref.wire_table_peers(ctxs, share_file_bytes=max(f1.size, f2.size))
        for sample, expect in (((f1, f2), want), ((f2, f2), want_swapped)):
            a, b = eref.split_fastq(sample[0], 3), eref.split_fastq(sample[1], 2) + [np.zeros(0, np.uint8)]
            for i, g in enumerate(ctxs):
                g.query_reset(); g.query_add(a[i]); g.query_add(b[i])
            for g in ctxs:
                g.ref_set(fa)
            for g in ctxs:
                g.sync()
            if sample[0] is f1:
                pops = [[g.query_popcount(l) for l in (1, 2, 3)] for g in ctxs]
                assert [sum(p[l] for p in pops) for l in range(3)] == whole
            for g in ctxs:
                assert (g.scan() == expect).all()
        with pytest.raises(eref.ErefError):
            ctxs[0].run_host(f1, f2, fa)
    finally:
        for g in ctxs:
            g.close()
