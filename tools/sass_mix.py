#!/usr/bin/env python
"""SASS evidence for the instruction-bound claims ("no SASS listing is committed"): per-kernel opcode
histograms of liberef_gpu.so from `cuobjdump -sass`, plus the listing of the inner code of the kernels named on the
command line.  Usage: python tools/sass_mix.py [kernel-substring ...] > profiles/sass_r02.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "palace_b200", "liberef_gpu.so")


def main():
    want = sys.argv[1:] or ["k_pack_fasta", "k_fasta_tiles", "k_pack_fastq", "k_probe_part", "k_probe_sampled",
                            "k_count_slices_persistent", "k_fine_partition", "k_coarse_scatter_staged"]
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = re.split(r"\n\s*Function : ", txt)[1:]
    print("# cuobjdump -sass palace_b200/liberef_gpu.so (sm_100a): opcode histograms per kernel\n")
    listings = []
    for k in kernels:
        name = k.split("\n", 1)[0].strip()
        dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        short = re.sub(r"\(.*", "", dem.replace("(anonymous namespace)::", ""))
        if not any(w in short for w in want):
            continue
        ops = collections.Counter()
        lines = []
        for line in k.splitlines():
            m = re.match(r"\s*/\*([0-9a-f]{4})\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if m:
                ops[m.group(3).split(".")[0]] += 1
                lines.append(line.rstrip())
        total = sum(ops.values())
        print(f"## {short}: {total} SASS instructions")
        print("   " + ", ".join(f"{o} {n}" for o, n in ops.most_common(18)))
        groups = {"integer/logic (LOP3, IADD3, SHF, IMAD, LEA, POPC, FLO, BREV, PRMT, ISETP, SEL, VIADD...)":
                  sum(n for o, n in ops.items() if o in ("LOP3", "IADD3", "IADD", "SHF", "IMAD", "LEA", "POPC", "FLO", "BREV", "PRMT",
                                                        "ISETP", "SEL", "VIADD", "IABS", "VIMNMX", "IMNMX", "MOV", "UMOV", "UIADD3",
                                                        "ULOP3", "USHF", "UISETP", "UIMAD", "ULEA", "R2UR", "S2R", "S2UR", "CS2R", "SGXT",
                                                        "BMSK", "LOP", "UFLO", "UPOPC", "USEL", "UPRMT", "VOTE", "VOTEU", "SHFL", "REDUX",
                                                        "PLOP3", "UPLOP3", "P2R", "R2P")),
                  "global/local memory (LDG, STG, LD, ST, ATOMG, RED, LDC, LDCU)":
                  sum(n for o, n in ops.items() if o in ("LDG", "STG", "LD", "ST", "ATOMG", "RED", "LDC", "LDCU", "ULDC", "LDL", "STL", "ATOM")),
                  "shared memory (LDS, STS, ATOMS)": sum(n for o, n in ops.items() if o in ("LDS", "STS", "ATOMS", "LDSM")),
                  "control (BRA, BAR, BSSY, BSYNC, EXIT, WARPSYNC, ...)":
                  sum(n for o, n in ops.items() if o in ("BRA", "BAR", "BSSY", "BSYNC", "EXIT", "WARPSYNC", "CALL", "RET", "BRX", "NOP",
                                                        "YIELD", "MEMBAR", "ERRBAR", "DEPBAR", "NANOSLEEP", "BREAK", "CCTL", "FENCE")),
                  "tensor / TMA (HMMA, UTC*, UTMA*, UBLKCP): none expected on this path":
                  sum(n for o, n in ops.items() if o.startswith(("HMMA", "UTC", "UTMA", "UBLKCP", "IMMA", "QMMA", "TCGEN")))}
        for g, n in groups.items():
            print(f"     {n:5d}  {g}")
        print()
        listings.append((short, lines))
    for short, lines in listings[:2]:
        print(f"# listing: {short} (first 160 instructions)\n")
        print("\n".join(lines[:160]))
        print()


if __name__ == "__main__":
    main()
