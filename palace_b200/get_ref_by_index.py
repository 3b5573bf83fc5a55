#!/usr/bin/env python
"""Drop-in for share/palace/scripts/get_ref_by_index.py (palace:482-498).

    get_ref_by_index.py <phagedb.fasta> <phagedb.fasta.fai> <ref_names.txt> <out.fasta> <out_percent.txt>

Takes the `ref_index ...` lines `eref` printed (bin/extract_ref.cpp:617), maps every index to a sequence name
through the .fai (index = 1-based line number, get_ref_by_index.py:40-49 — including the reference's own drift
when the database holds records of <= 32 bases), and writes the named records as single-line
FASTA plus `name<TAB>percentage`, in ascending index order (:74).

The reference loads the whole database into a Biopython dictionary first (:62); this version needs no
Biopython: it finds the header lines with one vectorised pass over the bytes and touches only the selected
records.  Semantics kept from Biopython's FASTA parser: the record id is the header up to the first white
space, the sequence is the record's lines joined with trailing white space, blanks and '\\r' removed, and a
duplicated id is an error.  Parity: pinned to the UNMODIFIED reference script's outputs and messages
(tests/golden/f1.json, made by tests/golden/make_golden_f1.py); Biopython itself is not in the build image, so
that run used a stand-in for SeqIO.parse/to_dict restating Biopython's FASTA rules — the Biopython half stays
unpinned, the script's own logic does not.
"""
import sys

import numpy as np


def parse_ref_file(ref_file):
    """index -> percentage from the lines tagged ref_index: first all-digit token after the tag, last token that
    parses as a float (get_ref_by_index.py:6-37)"""
    ref_data = {}
    with open(ref_file, "r") as f:
        for line in f:
            line = line.strip()
            if not line.startswith("ref_index"):
                continue
            parts = line.split()
            index = next((int(p) for p in parts[1:] if p.isdigit()), None)
            percentage = None
            for p in reversed(parts):
                try:
                    percentage = float(p)
                    break
                except ValueError:
                    continue
            if index is not None and percentage is not None:
                ref_data[index] = percentage
    return ref_data


def load_fai_index(fai_file):
    with open(fai_file, "r") as f:
        return {idx: line.split("\t")[0] for idx, line in enumerate(f, 1)}


def fasta_records(raw: bytes):
    """id -> (first sequence byte, end of record) for every record; raises on a duplicated id like SeqIO.to_dict"""
    a = np.frombuffer(raw, np.uint8)
    gt = np.flatnonzero(a == 62)
    starts = gt[(gt == 0) | (a[np.maximum(gt, 1) - 1] == 10)]             # '>' at the start of a line
    records = {}
    for k, s in enumerate(starts):
        s = int(s)
        e = raw.find(b"\n", s)
        e = len(raw) if e < 0 else e
        title = raw[s + 1:e].rstrip()
        words = title.split(None, 1)
        rid = words[0].decode("latin-1") if words else ""
        if rid in records:
            raise ValueError(f"Duplicate key '{rid}'")
        records[rid] = (min(e + 1, len(raw)), int(starts[k + 1]) if k + 1 < len(starts) else len(raw))
    return records


def record_sequence(raw: bytes, span) -> str:
    lines = raw[span[0]:span[1]].split(b"\n")
    return b"".join(l.rstrip() for l in lines).replace(b" ", b"").replace(b"\r", b"").decode("latin-1")


def main(argv=None):
    argv = sys.argv if argv is None else argv
    fasta_file, fai_file, ref_file, out_fasta, out_percent = argv[1:6]
    print("Loading FASTA sequences...")
    with open(fasta_file, "rb") as f:
        raw = f.read()
    records = fasta_records(raw)
    print("Loading FAI index...")
    index_to_name = load_fai_index(fai_file)
    print("Processing reference file...")
    ref_data = parse_ref_file(ref_file)
    print("Writing output files...")
    with open(out_fasta, "w") as fasta_out, open(out_percent, "w") as percent_out:
        for index, percentage in sorted(ref_data.items()):
            if index in index_to_name:
                seq_name = index_to_name[index]
                if seq_name in records:
                    fasta_out.write(f">{seq_name}\n")
                    fasta_out.write(f"{record_sequence(raw, records[seq_name])}\n")
                    percent_out.write(f"{seq_name}\t{percentage}\n")
                else:
                    print(f"Warning: Sequence '{seq_name}' not found in FASTA file")
            else:
                print(f"Warning: Index {index} not found in FAI file")
    print("Processing complete!")


if __name__ == "__main__":
    main()
