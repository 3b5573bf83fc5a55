#!/usr/bin/env python
"""Golden vectors for SURVEY.md §8(f) row f1: run the UNMODIFIED reference consumer script
(/root/reference/share/palace/scripts/get_ref_by_index.py) on small inputs and record what it wrote.

Biopython is absent from this image, so the script's `from Bio import SeqIO` resolves to the stand-in under
tests/golden/bio_standin (two functions restating Biopython's FASTA rules).  Everything else — ref_index line
parsing, .fai mapping, ordering, output formatting, warnings — is the reference's own code.  Run from the repo
root in the build container:

    python tests/golden/make_golden_f1.py        # writes tests/golden/f1.json
"""
import contextlib
import gzip
import io
import json
import os
import runpy
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
SCRIPT = "/root/reference/share/palace/scripts/get_ref_by_index.py"


def fai_of(fa: bytes) -> str:
    """what `samtools faidx` would list, as far as the consumer reads it: one line per record, name first"""
    out = []
    for line in fa.split(b"\n"):
        if line.startswith(b">"):
            words = line[1:].split()
            out.append((words[0].decode() if words else "") + "\t0\t0\t0\t0\n")
    return "".join(out)


def cases():
    hand_fa = (b"stray line before any header\n>p1 first phage\nACGT\nacgt \r\n>short\nAC\n>p2/1\tx\nTTTT\n\nGG\n"
               b">p3|x  two  blanks\nCC CC\n>lastNoNewline\nGATTACA")
    hand_names = ("noise\nref_index\t3\t1\t3900\t4000\t0.975\nref_index\t1\t2\t10\t12\t0.833333\n"
                  "ref_index\t9\t1\t1\t1\t1\nref_index x 5 1 7 7 1\nref_index\t4\tnan\nref_index\tnope\t0.5\n"
                  "  ref_index 2 1 1 2 5e-1\nref_index\t3\t1\t1\t1\t0.5\n")
    yield "hand", hand_fa, fai_of(hand_fa), hand_names
    # a .fai that names a record the FASTA does not hold, and an index past its end
    yield "missing", b">a\nAC\n>b\nGT\n", "a\t2\nzz\t2\nb\t2\n", "ref_index\t2\t1\t1\t1\t0.9\nref_index\t3\t1\t1\t1\t0.8\nref_index\t7\t1\t1\t1\t0.7\n"
    # the framing fixture of the eref goldens with the reference eref's own stdout
    edge_fa = gzip.open(os.path.join(HERE, "edge.fa.gz"), "rb").read()
    edge = json.load(open(os.path.join(HERE, "edge.json")))
    stdout = next(iter(edge["stdout"].values())) if isinstance(edge["stdout"], dict) else edge["stdout"]
    edge_fa = edge_fa[edge_fa.index(b">"):] if not edge_fa.startswith(b">") else edge_fa
    yield "edge", edge_fa, fai_of(edge_fa), stdout
    yield "empty_names", b">a\nAC\n", "a\t2\n", ""


def run_reference(fa: bytes, fai: str, names: str):
    """(out_fasta, out_percent, messages) of the unmodified reference script on these inputs"""
    standin = os.path.join(HERE, "bio_standin")
    if standin not in sys.path:
        sys.path.insert(0, standin)
    with tempfile.TemporaryDirectory() as d:
        p = {k: os.path.join(d, k) for k in ("db.fa", "db.fa.fai", "names.txt", "out.fa", "out.txt")}
        open(p["db.fa"], "wb").write(fa)
        open(p["db.fa.fai"], "w").write(fai)
        open(p["names.txt"], "w").write(names)
        argv, sys.argv = sys.argv, ["get_ref_by_index.py"] + [p[k] for k in ("db.fa", "db.fa.fai", "names.txt", "out.fa", "out.txt")]
        log = io.StringIO()
        try:
            with contextlib.redirect_stdout(log):
                runpy.run_path(SCRIPT, run_name="__main__")
        finally:
            sys.argv = argv
        return open(p["out.fa"]).read(), open(p["out.txt"]).read(), log.getvalue()


def fixture_cases():
    """the eref fixtures themselves: the REFERENCE eref's stdout (every threshold pair of <fixture>.json) through the
    reference script, on the fixture's own FASTA — what `EREF_REFS_FASTA/EREF_REFS_PERCENT` of the drop-in must write.
    The FASTA is not repeated in f1.json (it is tests/golden/<fixture>.fa.gz); the sequence file is pinned by sha256."""
    for fixture in ("basic", "short_reads", "edge"):
        fa = gzip.open(os.path.join(HERE, fixture + ".fa.gz"), "rb").read()
        meta = json.load(open(os.path.join(HERE, fixture + ".json")))
        for key, stdout in meta["stdout"].items():
            yield fixture, key, fa, fai_of(fa), stdout


def main():
    import hashlib
    out = {}
    for name, fa, fai, names in cases():
        out_fa, out_txt, log = run_reference(fa, fai, names)
        out[name] = {"fasta": fa.decode("latin-1"), "fai": fai, "names": names,
                     "out_fasta": out_fa, "out_percent": out_txt, "log": log}
    for fixture, key, fa, fai, names in fixture_cases():
        out_fa, out_txt, log = run_reference(fa, fai, names)
        out[f"fixture:{fixture}:{key}"] = {"fixture": fixture, "ratios": key, "names": names, "out_percent": out_txt,
                                          "out_fasta_sha256": hashlib.sha256(out_fa.encode("latin-1")).hexdigest(),
                                          "out_fasta_records": out_fa.count(">"), "log": log}
    json.dump(out, open(os.path.join(HERE, "f1.json"), "w"), indent=1)
    for k, v in out.items():
        print(k, len(v.get("out_fasta", "")), v.get("out_fasta_records"), repr(v["out_percent"][:80]))


if __name__ == "__main__":
    main()
