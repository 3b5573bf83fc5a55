"""SeqIO.parse(handle_or_path, "fasta") and SeqIO.to_dict, restating Biopython's published FASTA rules
(Bio.SeqIO.FastaIO.SimpleFastaParser / FastaIterator): lines before the first '>' are skipped; the title is the
header line without '>' and trailing white space; id = first white-space-delimited word of the title ('' if
none); the sequence is the record's lines, each right-stripped, joined, with blanks and '\\r' removed;
to_dict raises ValueError on a repeated id."""


class _Record:
    def __init__(self, title, seq):
        words = title.split(None, 1)
        self.id = self.name = words[0] if words else ""
        self.description = title
        self.seq = seq


def parse(source, fmt):
    if fmt != "fasta":
        raise ValueError("stand-in knows FASTA only")
    handle = open(source, "r") if isinstance(source, str) else source
    try:
        title, lines = None, []
        for line in handle:
            if line[0:1] == ">":
                if title is not None:
                    yield _Record(title, "".join(lines).replace(" ", "").replace("\r", ""))
                title, lines = line[1:].rstrip(), []
            elif title is not None:
                lines.append(line.rstrip())
        if title is not None:
            yield _Record(title, "".join(lines).replace(" ", "").replace("\r", ""))
    finally:
        if isinstance(source, str):
            handle.close()


def to_dict(records):
    out = {}
    for r in records:
        if r.id in out:
            raise ValueError(f"Duplicate key '{r.id}'")
        out[r.id] = r
    return out
