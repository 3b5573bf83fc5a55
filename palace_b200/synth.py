"""Seeded synthetic inputs for the `eref` path.

A phage-style FASTA database and a paired FASTQ query in the only formats the reference
reads (bin/extract_ref.cpp:686-756 FASTA, :940-1004 four-line FASTQ).  Everything is
vectorised numpy so the 500 Mbp / 50 Mbp bench configuration is generated in seconds.

The generator stays clear of the reference's undefined behaviour:
sequence lines <= 400 chars (stack buffers :927-928), no records of 450..499 bp (:519),
non-empty fq1 (:1145), quality strings without '@' (:917).
"""
from __future__ import annotations

import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_NL = 10


def _wrap(seq: np.ndarray, width: int) -> np.ndarray:
    """ASCII sequence -> FASTA body wrapped at `width`, with trailing newline."""
    n = seq.size
    if n == 0:
        return np.empty(0, np.uint8)
    nfull, rem = divmod(n, width)
    parts = []
    if nfull:
        body = np.empty((nfull, width + 1), np.uint8)
        body[:, :width] = seq[: nfull * width].reshape(nfull, width)
        body[:, width] = _NL
        parts.append(body.ravel())
    if rem:
        parts.append(seq[nfull * width:])
        parts.append(np.array([_NL], np.uint8))
    return np.concatenate(parts)


class SynthDB:
    """text: FASTA bytes (uint8). codes: concatenated 2-bit genome array (for read sampling).
    starts/lengths: per *kept-for-sampling* genome offsets into `codes`."""

    def __init__(self, text, codes, starts, lengths, n_records):
        self.text = text
        self.codes = codes
        self.starts = starts
        self.lengths = lengths
        self.n_records = n_records


def make_db(n_genomes: int, total_bp: int, seed: int, *, width: int = 70,
            min_len: int = 5000, max_len: int = 200000, frac_mutated: float = 0.02,
            frac_n: float = 0.01, frac_lower: float = 0.005, n_short: int = 5,
            crlf: bool = False) -> SynthDB:
    rng = np.random.default_rng(seed)
    lens = np.exp(rng.uniform(np.log(min_len), np.log(max_len), n_genomes))
    lens = np.maximum((lens * (total_bp / lens.sum())).astype(np.int64), 600)
    # mutated copies of an earlier genome (partial hits / multi-fragment results)
    n_mut = int(round(frac_mutated * n_genomes))
    mut_idx = np.sort(rng.choice(np.arange(1, n_genomes), size=min(n_mut, max(n_genomes - 1, 0)),
                                 replace=False)) if n_genomes > 1 and n_mut else np.empty(0, np.int64)
    donors = {}
    mut_set = set(int(i) for i in mut_idx)
    for i in mut_idx:
        d = int(rng.integers(0, i))
        while d in mut_set:                # donor must be an original genome
            d = int(rng.integers(0, i))
        donors[int(i)] = d
        lens[i] = lens[d]
    starts = np.zeros(n_genomes, np.int64)
    starts[1:] = np.cumsum(lens)[:-1]
    total = int(lens.sum())
    codes = rng.integers(0, 4, size=total, dtype=np.uint8)
    for i, d in donors.items():
        src = codes[starts[d]: starts[d] + lens[d]].copy()
        rate = rng.uniform(0.01, 0.10)
        m = rng.random(src.size) < rate
        src[m] = (src[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) & 3
        codes[starts[i]: starts[i] + lens[i]] = src
    ascii_all = _ACGT[codes]
    # N runs and lower-case stretches (in the FASTA text only; reads are drawn from `codes`)
    for i in rng.choice(n_genomes, size=int(round(frac_n * n_genomes)), replace=False):
        L = int(lens[i]); run = int(rng.integers(10, 200)); p = int(rng.integers(0, max(L - run, 1)))
        ascii_all[starts[i] + p: starts[i] + p + run] = ord("N")
    for i in rng.choice(n_genomes, size=int(round(frac_lower * n_genomes)), replace=False):
        L = int(lens[i]); run = int(rng.integers(100, 2000)); p = int(rng.integers(0, max(L - run, 1)))
        sl = slice(starts[i] + p, starts[i] + min(p + run, L))
        ascii_all[sl] = ascii_all[sl] | 0x20
    # a handful of records <= 32 bp: dropped by the reference without consuming an ordinal (:697)
    short_after = set(int(x) for x in rng.choice(n_genomes, size=min(n_short, n_genomes), replace=False)) \
        if n_short else set()
    parts = []
    n_records = 0
    eol = np.frombuffer(b"\r\n" if crlf else b"\n", np.uint8)
    for i in range(n_genomes):
        hdr = (">phage_%06d synthetic len=%d" % (i, lens[i])).encode()
        parts.append(np.frombuffer(hdr, np.uint8)); parts.append(eol)
        body = _wrap(ascii_all[starts[i]: starts[i] + lens[i]], width)
        if crlf:
            body = np.frombuffer(body.tobytes().replace(b"\n", b"\r\n"), np.uint8)
        parts.append(body)
        n_records += 1
        if i in short_after:
            sl = int(rng.integers(1, 33))
            parts.append(np.frombuffer((">short_%06d tiny" % i).encode(), np.uint8)); parts.append(eol)
            parts.append(_ACGT[rng.integers(0, 4, size=sl)]); parts.append(eol)
            n_records += 1
    text = np.concatenate(parts)
    return SynthDB(text, codes, starts, lens, n_records)


def make_reads(db: SynthDB, total_bp: int, seed: int, *, min_len: int = 100, max_len: int = 400,
               cov_lo: float = 2.0, cov_hi: float = 6.0, sub_rate: float = 0.005,
               n_rate: float = 0.0005, present: np.ndarray | None = None):
    """Paired FASTQ (fq1_bytes, fq2_bytes, present_genomes). Reads tile a random subset of the
    genomes at 2-6x so some 32-mers reach the reference's count of 3 (:995-997) and some do not."""
    rng = np.random.default_rng(seed)
    G = db.lengths.size
    mean_len = 0.5 * (min_len + max_len)
    if present is None:
        order = rng.permutation(G)
        chosen, acc = [], 0.0
        covs = []
        for g in order:
            c = rng.uniform(cov_lo, cov_hi)
            chosen.append(int(g)); covs.append(c)
            acc += c * db.lengths[g]
            if acc >= total_bp:
                break
        present = np.array(chosen, np.int64); covs = np.array(covs)
    else:
        present = np.asarray(present, np.int64)
        covs = rng.uniform(cov_lo, cov_hi, present.size)
    n_reads_g = np.maximum((covs * db.lengths[present] / mean_len).astype(np.int64), 1)
    n_reads = int(n_reads_g.sum())
    gi = np.repeat(np.arange(present.size), n_reads_g)
    rlen = rng.integers(min_len, max_len + 1, size=n_reads).astype(np.int64)
    glen = db.lengths[present][gi]
    rlen = np.minimum(rlen, glen)
    off = (rng.random(n_reads) * (glen - rlen + 1)).astype(np.int64)
    start = db.starts[present][gi] + off
    rev = rng.random(n_reads) < 0.5
    perm = rng.permutation(n_reads)                 # shuffle so fq order is not genome order
    start, rlen, rev = start[perm], rlen[perm], rev[perm]
    tot = int(rlen.sum())
    rstart = np.zeros(n_reads, np.int64); rstart[1:] = np.cumsum(rlen)[:-1]
    k = np.arange(tot, dtype=np.int64) - np.repeat(rstart, rlen)
    rl = np.repeat(rlen, rlen); st = np.repeat(start, rlen); rv = np.repeat(rev, rlen)
    idx = np.where(rv, st + rl - 1 - k, st + k)
    b = db.codes[idx]
    b = np.where(rv, 3 - b, b).astype(np.uint8)
    m = rng.random(tot) < sub_rate
    b[m] = (b[m] + rng.integers(1, 4, size=int(m.sum()), dtype=np.uint8)) & 3
    seq = _ACGT[b]
    seq[rng.random(tot) < n_rate] = ord("N")
    del k, rl, st, rv, idx, b, m
    out = []
    for mate in (0, 1):
        sel = np.arange(mate, n_reads, 2)
        out.append(_fastq_text(seq, rstart[sel], rlen[sel], sel, mate + 1))
    return out[0], out[1], present


def _fastq_text(seq, rstart, rlen, ids, mate) -> np.ndarray:
    """Four-line FASTQ records: '@r<9 digits>/<mate>', sequence, '+', constant 'I' qualities."""
    n = rlen.size
    if n == 0:
        return np.empty(0, np.uint8)
    H = 13                                             # "@r%09d/%d"
    rec = (H + 1) + (rlen + 1) + 2 + (rlen + 1)
    roff = np.zeros(n, np.int64); roff[1:] = np.cumsum(rec)[:-1]
    total = int(rec.sum())
    out = np.full(total, ord("I"), np.uint8)
    hdr = np.empty((n, H), np.uint8)
    hdr[:, 0] = ord("@"); hdr[:, 1] = ord("r")
    v = ids.astype(np.int64).copy()
    for d in range(9):
        hdr[:, 10 - d] = ord("0") + (v % 10); v //= 10
    hdr[:, 11] = ord("/"); hdr[:, 12] = ord("0") + mate
    out[(roff[:, None] + np.arange(H)[None, :]).ravel()] = hdr.ravel()
    out[roff + H] = _NL
    tot = int(rlen.sum())
    cs = np.zeros(n, np.int64); cs[1:] = np.cumsum(rlen)[:-1]
    within = np.arange(tot, dtype=np.int64) - np.repeat(cs, rlen)
    out[np.repeat(roff + H + 1, rlen) + within] = seq[np.repeat(rstart, rlen) + within]
    out[roff + H + 1 + rlen] = _NL
    out[roff + H + 2 + rlen] = ord("+")
    out[roff + H + 3 + rlen] = _NL
    out[roff + rec - 1] = _NL
    return out
