"""CPU suite, part 3: host-side logic — sharding at record boundaries, the down-sampling guard, and
the N>1 merge (allgather of padded fixed-size records, ordinal offsets) over gloo with world_size 2."""
import os
import socket

import numpy as np
import pytest

from palace_b200 import eref, synth
from palace_b200.dist import merge_gathered, pad_records


def test_shard_offsets_start_at_headers():
    db = synth.make_db(30, 200_000, 3, min_len=2000, max_len=20000)
    for shards in (1, 2, 3, 8):
        cuts = eref.shard_offsets(db.text, shards)
        assert cuts[0] == 0 and cuts[-1] == db.text.size and len(cuts) == shards + 1
        assert all(a <= b for a, b in zip(cuts, cuts[1:]))
        for c in cuts[1:-1]:
            assert c == db.text.size or (db.text[c] == ord(">") and db.text[c - 1] == ord("\n"))


def test_shards_cover_all_records(oracle):
    db = synth.make_db(30, 200_000, 4, min_len=2000, max_len=20000, n_short=4)
    whole, _ = oracle.parse_fasta(db.text)
    for shards in (2, 5):
        cuts = eref.shard_offsets(db.text, shards)
        parts = []
        for i in range(shards):
            parts += oracle.parse_fasta(db.text[cuts[i]:cuts[i + 1]])[0]
        assert parts == whole


def test_sample_ratio_and_draws():
    """cal_sam_ratio's integer arithmetic (:1144-1145) and the rand() % 100 replay (:955-957) — host-only entry points"""
    import ctypes
    assert eref.sample_ratio(50_000_000) == 2000
    assert eref.sample_ratio(1_000_000_000) == 100
    assert eref.sample_ratio(1_010_000_001) == 99                 # the reference starts to down-sample here
    assert eref.sample_ratio(450_000, 400_000) == 44
    with pytest.raises(eref.ErefError):
        eref.sample_ratio(0)                                      # the reference divides by zero (:1145)
    bits = eref.draw_keep(88, 1000)
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(ctypes.c_uint(1))
    want = [libc.rand() % 100 < 88 for _ in range(1000)]
    got = [bool((int(bits[i >> 5]) >> (i & 31)) & 1) for i in range(1000)]
    assert got == want and 800 < sum(got) < 950
    nxt = [libc.rand() % 100 < 88 for _ in range(10)]
    eref.draw_keep(88, 1000)
    more = eref.draw_keep(88, 10, reseed=False)                   # fq2 continues fq1's stream
    assert [bool((int(more[0]) >> i) & 1) for i in range(10)] == nxt


def test_merge_gathered_offsets_ordinals():
    a = np.zeros(3, eref.RESULT_DTYPE); a["ordinal"] = [1, 2, 3]; a["el"] = [5, 0, 7]
    b = np.zeros(2, eref.RESULT_DTYPE); b["ordinal"] = [1, 2]; b["el"] = [9, 1]
    slot = 3
    flat = np.concatenate([pad_records(a, slot), pad_records(b, slot)])
    out = merge_gathered(flat, [3, 2], slot)
    assert out["ordinal"].tolist() == [1, 2, 3, 4, 5]
    assert out["el"].tolist() == [5, 0, 7, 9, 1]


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from palace_b200.dist import allgather_records
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 4 if rank == 0 else 2                                  # ragged shards
        rec = np.zeros(n, eref.RESULT_DTYPE)
        rec["ordinal"] = np.arange(1, n + 1); rec["el"] = 100 * (rank + 1) + np.arange(n); rec["ref_len"] = 1000
        local = torch.from_numpy(rec.view(np.uint32).reshape(n, 4).astype(np.int32))
        merged = allgather_records(local, n)
        body, valid = allgather_records(local, n, compact=False)   # padded form: no count is read back
        assert body.shape == (world, 4, 4) and body[valid].tolist() == merged.tolist()
        q.put((rank, merged.tolist()))
    finally:
        dist.destroy_process_group()


def test_allgather_records_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == got[1]                                        # allgather: every rank holds the merge
    ords = [r[0] for r in got[0]]
    assert ords == [1, 2, 3, 4, 5, 6]
    assert [r[2] for r in got[0]] == [100, 101, 102, 103, 200, 201]


def test_get_ref_by_index_mirror(tmp_path):
    """the consumer of eref's stdout (share/palace/scripts/get_ref_by_index.py), hand-computed:
    index -> .fai line -> name -> single-line record; ascending index; the reference's index drift past short records"""
    from palace_b200 import get_ref_by_index as g
    fa = (b">p1 first phage\nACGT\nacgt \r\n>short\nAC\n>p2/1\tx\nTTTT\n\nGG\n>p3\nCCCC")
    (tmp_path / "db.fa").write_bytes(fa)
    (tmp_path / "db.fa.fai").write_text("p1\t8\t16\t4\t5\nshort\t2\t30\t2\t3\np2/1\t6\t41\t4\t5\np3\t4\t54\t4\t5\n")
    # eref skipped `short` (<= 32 bases do not consume an ordinal), so its ordinal 2 is p2/1 — but the script maps
    # ordinal 2 to the SECOND .fai line, as the reference does
    (tmp_path / "names.txt").write_text("noise\nref_index\t3\t1\t3900\t4000\t0.975\nref_index\t1\t2\t10\t12\t0.833333\n"
                                        "ref_index\t9\t1\t1\t1\t1\n")
    g.main(["x", str(tmp_path / "db.fa"), str(tmp_path / "db.fa.fai"), str(tmp_path / "names.txt"),
            str(tmp_path / "out.fa"), str(tmp_path / "out.txt")])
    assert (tmp_path / "out.fa").read_text() == ">p1\nACGTacgt\n>p2/1\nTTTTGG\n"
    assert (tmp_path / "out.txt").read_text() == "p1\t0.833333\np2/1\t0.975\n"
    (tmp_path / "dup.fa").write_bytes(b">a x\nAC\n>a y\nGT\n")
    with pytest.raises(ValueError):
        g.fasta_records((tmp_path / "dup.fa").read_bytes())


def test_get_ref_by_index_against_the_reference_script(tmp_path, capsys):
    """f1 pinned: tests/golden/f1.json holds what the UNMODIFIED reference script wrote (files and messages) on these
    inputs, generated by tests/golden/make_golden_f1.py with a stand-in for the two Bio.SeqIO calls"""
    import json
    from palace_b200 import get_ref_by_index as g
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "f1.json")))
    assert set(golden) >= {"hand", "missing", "edge", "empty_names"}
    import gzip
    import hashlib
    for name, c in golden.items():
        d = tmp_path / name.replace(":", "_").replace(",", "_")
        d.mkdir()
        if "fixture" in c:                                 # the eref fixtures: the reference eref's stdout on their FASTA
            fa = gzip.open(os.path.join(os.path.dirname(__file__), "golden", c["fixture"] + ".fa.gz"), "rb").read()
            (d / "db.fa").write_bytes(fa)
            (d / "db.fa.fai").write_text("".join((l[1:].split()[0].decode() if l[1:].split() else "") + "\t0\t0\t0\t0\n"
                                                 for l in fa.split(b"\n") if l.startswith(b">")))
            (d / "names.txt").write_text(c["names"])
            capsys.readouterr()
            g.main(["x", str(d / "db.fa"), str(d / "db.fa.fai"), str(d / "names.txt"), str(d / "out.fa"), str(d / "out.txt")])
            assert capsys.readouterr().out == c["log"], name
            assert hashlib.sha256((d / "out.fa").read_bytes()).hexdigest() == c["out_fasta_sha256"], name
            assert (d / "out.txt").read_text() == c["out_percent"], name
            continue
        (d / "db.fa").write_bytes(c["fasta"].encode("latin-1"))
        (d / "db.fa.fai").write_text(c["fai"])
        (d / "names.txt").write_text(c["names"])
        capsys.readouterr()
        g.main(["x", str(d / "db.fa"), str(d / "db.fa.fai"), str(d / "names.txt"), str(d / "out.fa"), str(d / "out.txt")])
        assert capsys.readouterr().out == c["log"], name
        assert (d / "out.fa").read_text() == c["out_fasta"], name
        assert (d / "out.txt").read_text() == c["out_percent"], name


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the driver's CPU arm) end to end without a GPU: forced onto the oracle port and the
    small sample so that it takes seconds; the line must carry the contract's keys with zero copy bytes"""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, timeout=300,
                       env=dict(os.environ, EREF_BENCH_CPU_SAMPLE="small", EREF_BENCH_CPU_KIND="port"))
    assert r.returncode == 0, r.stderr.decode()[-400:]
    lines = [l for l in r.stdout.decode().splitlines() if l.strip()]
    assert len(lines) == 1                                         # stdout carries the one JSON line and nothing else
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Gbp/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    # the labels say what ran: the development sample, the port — not the headline workload
    assert "development" in d["config"]["workload"] and "oracle port" in d["cpu_baseline"]["sample"]
    assert d["cpu_baseline"]["sample_db_bases"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


@pytest.mark.skipif(not os.path.exists("/root/reference/share/palace/scripts/get_ref_by_index.py"),
                    reason="needs the reference checkout (build container only)")
def test_get_ref_by_index_fuzz_against_the_reference_script(tmp_path, capsys):
    """f1, randomised: the unmodified reference script (Bio.SeqIO resolved to tests/golden/bio_standin) and the mirror on
    the same seeded inputs — ragged line lengths, blanks and CRs inside sequences, odd header white space, records the
    .fai does not know, indices past its end, repeated and malformed ref_index lines"""
    import importlib.util
    from palace_b200 import get_ref_by_index as g
    spec = importlib.util.spec_from_file_location("make_golden_f1", os.path.join(os.path.dirname(__file__), "golden", "make_golden_f1.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    rng = np.random.default_rng(4242)
    for case in range(40):
        n = int(rng.integers(1, 12))
        fa, names_in_fa = b"", []
        if rng.random() < 0.3:
            fa += b"junk before the first header\n"
        for r in range(n):
            name = f"s{case}_{r}" + ("|x" if rng.random() < 0.3 else "")
            sep = [b" ", b"\t", b"  ", b""][int(rng.integers(0, 4))]
            fa += b">" + name.encode() + (sep + b"desc " + str(r).encode() if sep else b"") + [b"\n", b"\r\n", b" \n"][int(rng.integers(0, 3))]
            names_in_fa.append(name)
            for _ in range(int(rng.integers(0, 5))):
                line = bytes(rng.choice(np.frombuffer(b"ACGTNacgt", np.uint8), int(rng.integers(0, 30))))
                if rng.random() < 0.2:
                    line = line[:3] + b" " + line[3:]
                fa += line + [b"\n", b"\r\n", b"  \n", b"\n\n"][int(rng.integers(0, 4))]
        if rng.random() < 0.5:
            fa = fa.rstrip(b"\n")
        fai_names = list(names_in_fa)
        if rng.random() < 0.4:
            fai_names.insert(int(rng.integers(0, len(fai_names) + 1)), "ghost")
        fai = "".join(f"{nm}\t1\t2\t3\t4\n" for nm in fai_names)
        lines = []
        for _ in range(int(rng.integers(0, 10))):
            k = int(rng.integers(0, len(fai_names) + 3))
            pct = ["0.9", "1", "0.7512", "7.5e-01", "nan", "x"][int(rng.integers(0, 6))]
            lines.append([f"ref_index\t{k}\t1\t10\t12\t{pct}", f"ref_index {k} {pct}", f"other\t{k}\t{pct}", f" ref_index\t{k}\t{pct} ",
                          f"ref_index\tx{k}\t{pct}"][int(rng.integers(0, 5))])
        names = "\n".join(lines) + ("\n" if lines else "")
        want = ref.run_reference(fa, fai, names)
        d = tmp_path / f"c{case}"
        d.mkdir()
        (d / "db.fa").write_bytes(fa); (d / "db.fa.fai").write_text(fai); (d / "names.txt").write_text(names)
        capsys.readouterr()
        g.main(["x", str(d / "db.fa"), str(d / "db.fa.fai"), str(d / "names.txt"), str(d / "out.fa"), str(d / "out.txt")])
        got = ((d / "out.fa").read_text(), (d / "out.txt").read_text(), capsys.readouterr().out)
        assert got == want, (case, fa, fai, names)


def test_fastq_limit_is_the_references_getline_budget(oracle):
    """eref_fastq_limit (host arithmetic of the library: block-wise newline counts with an early exit) against the
    oracle's literal getline loop (:942-945) on random texts with every line-length mix, and every kind of budget"""
    rng = np.random.default_rng(1)
    cut = 0
    for _ in range(1500):
        n = int(rng.integers(0, 400))
        a = np.where(rng.random(n) < rng.choice([0.02, 0.1, 0.3, 0.9]), 10, 65).astype(np.uint8)
        for end in [0, 1, 2, n // 2, max(n - 3, 0), max(n - 1, 0), n, n + 1] + [int(x) for x in rng.integers(0, n + 2, 4)]:
            want = oracle.fastq_limit(a, end)
            assert eref.fastq_limit(a, end) == want, (bytes(a), end)
            cut += want < n
    assert cut > 5000
    # blocks of 64 KiB, a budget deep inside a long text, both outcomes
    db = synth.make_db(20, 400_000, 5)
    f1, f2, _ = synth.make_reads(db, 900_000, 6)
    for end in (f2.size - 100, f2.size - 3000, f2.size - 20_000, f2.size // 2, 70_000, f1.size):
        assert eref.fastq_limit(f2, end) == oracle.fastq_limit(f2, end), end
    assert eref.fastq_limit(f2, f2.size - 20_000) < f2.size and eref.fastq_limit(f2, f2.size - 100) == f2.size


def test_candidate_filter_bound_holds_for_every_stride():
    """DESIGN.md §4.2 as arithmetic, without a GPU: a window is good only if >= three_min of its positions are hit by all
    three streams (:556-561), hence by stream 0; sampling stream 0 at the plane positions that are multiples of s leaves at
    most 500 - floor(500/s) positions of any window unsampled, so a window with fewer than three_min - (500 - 500 // s)
    sampled hits cannot be good.  cand[w] sums the sampled hits of words w-16..w of the record, need[w] = a candidate in
    w..w+16: every position of every good window must lie in a needed word, whatever the record's offset in the planes."""
    rng = np.random.default_rng(42)
    checked = good_total = 0
    for trial in range(60):
        length = int(rng.integers(40, 6000))
        p0 = int(rng.integers(0, 64)) + 32 * int(rng.integers(0, 3))         # the record's first plane position
        phr = float(rng.choice([0.85, 0.6, 0.97, 0.3]))
        hr = float(rng.choice([0.9, 0.5, 0.99]))
        one_min, three_min = int(np.float32(500) * np.float32(hr)), int(np.float32(500) * np.float32(phr))
        # hit patterns with long dense stretches (so that good windows exist) and noisy borders
        dense = np.zeros(length, bool)
        for _ in range(int(rng.integers(0, 4))):
            a = int(rng.integers(0, length)); dense[a: a + int(rng.integers(100, 2500))] = True
        h = [(rng.random(length) < np.where(dense, rng.uniform(0.8, 1.0), rng.uniform(0.0, 0.3))) for _ in range(3)]
        for hi in h:
            hi[max(length - 31, 0):] = False                                   # no 32-mer starts in the last 31 positions
        trio, single = h[0] & h[1] & h[2], h[0] | h[1] | h[2]
        ct, cs = np.concatenate([[0], np.cumsum(trio)]), np.concatenate([[0], np.cumsum(single)])
        j = np.arange(length)
        lo = np.maximum(j - 499, 0)
        good = (cs[j + 1] - cs[lo] >= one_min) & (ct[j + 1] - ct[lo] >= three_min)    # :548-561
        good_total += int(good.sum())
        pos = p0 + j                                                           # plane positions, word = pos >> 5
        w0, w1 = p0 >> 5, (p0 + length - 1) >> 5
        for s in range(4, 9):
            min_hits = three_min - (500 - 500 // s)
            if min_hits <= 0:
                continue                                                       # the filter is vacuous: everything is probed
            sampled_hit = h[0] & (pos % s == 0)
            per_word = np.bincount((pos >> 5) - w0, weights=sampled_hit, minlength=w1 - w0 + 1)
            cw = np.concatenate([[0], np.cumsum(per_word)])
            w = np.arange(w1 - w0 + 1)
            cand = cw[w + 1] - cw[np.maximum(w - 16, 0)] >= min_hits
            cc = np.concatenate([[0], np.cumsum(cand)])
            need = cc[np.minimum(w + 17, w.size)] - cc[w] > 0
            for jj in np.flatnonzero(good):
                words = np.arange((p0 + lo[jj]) >> 5, ((p0 + jj) >> 5) + 1) - w0
                assert need[words].all(), (trial, s, int(jj))
            checked += 1
    assert checked > 100 and good_total > 1000
    # the worst case for the bound: ONE window that is just good, its three_min trio hits placed on the unsampled
    # positions first, nothing else in the record; every word under the window must still be needed
    tight = 0
    for trial in range(200):
        length = int(rng.integers(600, 3000))
        p0 = int(rng.integers(0, 96))
        phr = float(rng.choice([0.85, 0.6, 0.97]))
        three_min = int(np.float32(500) * np.float32(phr))
        jj = int(rng.integers(499, length - 31))
        window = np.arange(jj - 499, jj + 1)
        for s in range(4, 9):
            min_hits = three_min - (500 - 500 // s)
            if min_hits <= 0:
                continue
            on_sample = (p0 + window) % s == 0
            order = np.concatenate([rng.permutation(window[~on_sample]), rng.permutation(window[on_sample])])
            h0 = np.zeros(length, bool)
            h0[order[:three_min]] = True                                      # stream 0 = the trio hits, and nothing more
            assert int(h0[window].sum()) == three_min                         # the window is good, by exactly one hit
            pos = p0 + np.arange(length)
            w0, w1 = p0 >> 5, (p0 + length - 1) >> 5
            per_word = np.bincount((pos >> 5) - w0, weights=h0 & (pos % s == 0), minlength=w1 - w0 + 1)
            cw = np.concatenate([[0], np.cumsum(per_word)])
            w = np.arange(w1 - w0 + 1)
            cand = cw[w + 1] - cw[np.maximum(w - 16, 0)] >= min_hits
            cc = np.concatenate([[0], np.cumsum(cand)])
            need = cc[np.minimum(w + 17, w.size)] - cc[w] > 0
            words = np.arange((p0 + jj - 499) >> 5, ((p0 + jj) >> 5) + 1) - w0
            assert need[words].all(), (trial, s, jj)
            assert min_hits <= int(per_word[words].sum()) <= min_hits + 1      # (the bound is tight: nothing to spare)
            tight += 1
    assert tight > 500
