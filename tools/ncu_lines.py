"""Development helper: per-CUDA-source-line share of executed instructions and stall samples of one kernel, from
`ncu -i REP --page source --csv --print-source cuda,sass --kernel-name K > file.csv`.
usage: python tools/ncu_lines.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname, hdr, lines = "", None, {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].rsplit("/", 1)[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if hdr is None or len(r) <= ie or not r[0].isdigit() or not r[ie].isdigit():
        continue
    k = (fname, int(r[0]))
    v = lines.setdefault(k, [0, 0, r[1].strip()[:100]])
    v[0] += int(r[ie]); v[1] += int(r[ss]) if r[ss].isdigit() else 0
tot = sum(v[0] for v in lines.values()) or 1
tots = sum(v[1] for v in lines.values()) or 1
print(f"total warp instructions {tot}, samples {tots}")
for (f, l), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * v[0] / tot:5.1f}% inst {100 * v[1] / tots:5.1f}% smp  {f}:{l:<4} {v[2]}")
