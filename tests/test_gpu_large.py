"""GPU suite, maximum sizes (-m gpu): BASELINE configs[3]-sized queries, the real 1 Gbase down-sampling trigger
(bin/extract_ref.cpp:955-960, :1134-1146) and a FASTA shard close to the 4 GiB limit.  The oracle cannot run
these in seconds, so every check is a size-independent property that reduces the big input to a small one the
oracle does run: a read repeated K >= 3 times saturates its codes (:996-997), a database repeated K times
repeats its records."""
import numpy as np
import pytest

from palace_b200 import eref, synth

pytestmark = pytest.mark.gpu


def _small(seed, db_bp=20_000_000, n_genomes=400, q_bp=10_000_000):
    db = synth.make_db(n_genomes, db_bp, seed)
    f1, f2, _ = synth.make_reads(db, q_bp, seed + 1, cov_lo=1.0, cov_hi=4.0)
    return db, f1, f2


def _levels(g):
    return [g.query_popcount(level) for level in (1, 2, 3)]


def test_config4_sized_query_saturates_like_a_tripled_small_one(oracle):
    """2 x ~250 Mbp of reads (configs[3]: 500 Mbp query) = a 10 Mbp query repeated 50 times: every code that
    occurs at all occurs >= 3 times, so all three levels equal level 1 of the small query and the records equal
    those of the small query tripled, which the oracle computes"""
    db, f1, f2 = _small(31)
    coder = eref.random_coder(31)
    K = 50
    big1, big2 = np.tile(f1, K), np.tile(f2, K)
    assert big1.size + big2.size > 1_000_000_000
    with eref.ErefGPU(coder, 0.9, 0.85) as g:
        g.query_add(f1); g.query_add(f2)
        small_levels = _levels(g)
    with eref.ErefGPU(coder, 0.9, 0.85) as g:
        assert eref.sample_ratio(g.query_add(big1)) >= 100      # below the down-sampling trigger
        g.query_add(big2)
        assert _levels(g) == [small_levels[0]] * 3
        g.ref_set(db.text)
        got = g.scan().copy()
    tripled1, tripled2 = np.tile(f1, 3), np.tile(f2, 3)
    want = oracle.run(tripled1, tripled2, db.text, coder, 0.9, 0.85, threads=4)
    assert (got == want).all()
    assert (want["el"] > 0).any()


def test_the_real_downsampling_trigger(oracle):
    """fq1 above 1 Gbase: the ratio of cal_sam_ratio and the rand() replay decide which copies of which read
    are counted; with ~180 copies of every read and a ratio near 90 every read keeps >= 3 copies (checked from
    the drawn masks), so the table and the records must equal the small query tripled"""
    db, f1, f2 = _small(33, q_bp=12_000_000)
    coder = eref.random_coder(33)
    with eref.ErefGPU(coder) as g:
        per_copy, lines1 = g.query_census(f1)
        _, lines2 = g.query_census(f2)
    K = eref.DOWN_SAMPLING_SIZE // 2 // per_copy + 12
    big1, big2 = np.tile(f1, K), np.tile(f2, K)
    ratio = eref.sample_ratio(per_copy * K)
    assert 50 <= ratio < 100
    k1 = eref.draw_keep(ratio, lines1 * K)
    k2 = eref.draw_keep(ratio, lines2 * K, reseed=False)
    kept1 = np.unpackbits(k1.view(np.uint8), bitorder="little")[:lines1 * K].reshape(K, lines1).sum(0)
    kept2 = np.unpackbits(k2.view(np.uint8), bitorder="little")[:lines2 * K].reshape(K, lines2).sum(0)
    assert kept1.min() >= 3 and kept2.min() >= 3
    _, got = eref.run_eref(big1, big2, db.text, None, 0.9, 0.85, 1, coder=coder, pipelined=True, return_results=True)
    want = oracle.run(np.tile(f1, 3), np.tile(f2, 3), db.text, coder, 0.9, 0.85, threads=4)
    assert (got == want).all()
    assert (want["el"] > 0).any()
    # and by phases, with the table counts
    with eref.ErefGPU(coder) as g:
        g.query_add(f1); g.query_add(f2)
        level1 = g.query_popcount(1)
        g.query_reset()
        g.query_add_masked(big1, k1, lines1 * K)
        g.query_add_masked(big2, k2, lines2 * K)
        assert _levels(g) == [level1] * 3


def test_a_shard_close_to_4_GiB_repeats_its_records():
    """a 3.9 GB FASTA = a 30 MB database repeated: record r + i*n has the numbers of record r, ordinals run on,
    text offsets stay exact above 2^31; one byte more than the limit is refused before anything is copied"""
    db, f1, f2 = _small(35, db_bp=30_000_000, n_genomes=600, q_bp=4_000_000)
    coder = eref.random_coder(35)
    K = int(3_900_000_000 // db.text.size)
    big = np.tile(db.text, K)
    assert big.size > 3_700_000_000
    with eref.ErefGPU(coder) as g:
        g.query_add(f1); g.query_add(f2)
        g.ref_set(db.text)
        one = g.scan().copy()
        off0, hl0, len0 = g.ref_record(one.size)
        g.ref_set(big)
        kept, recs, bases = g.ref_count()
        assert kept == one.size * K
        got = g.scan().copy()
        assert (got["ordinal"] == np.arange(1, kept + 1)).all()
        for field in ("frag", "el", "ref_len"):
            assert (got[field].reshape(K, one.size) == one[field][None, :]).all(), field
        off, hl, ln = g.ref_record(kept)                        # header offset of the very last record
        assert (hl, ln) == (hl0, len0) and off == off0 + (K - 1) * db.text.size and off > 1 << 31
        assert big[off] == ord(">")
        assert (g.get_codes(kept) == g.get_codes(kept - one.size)).all()
        with pytest.raises(eref.ErefError):
            g.ref_set(np.zeros(0xFFFFFF00, np.uint8))           # calloc'ed, never touched
    assert (one["el"] > 0).any()
