"""Summarise an .ncu-rep (raw page) into one line per captured launch. usage: ncu_summary.py rep [more reps]"""
import csv, io, subprocess, sys
WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__t_sector_hit_rate.pct", "l2hit%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "inst"),
        ("lts__t_sectors_srcunit_tex_op_atom.sum", "atom_sectors"), ("launch__grid_size", "grid")]
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {rep}")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("<unnamed>::", "")
        parts = [name]
        for key, short in WANT:
            if key in hdr:
                v = r[hdr.index(key)]; u = units[hdr.index(key)]
                try:
                    v = f"{float(v):.4g}"
                except ValueError:
                    pass
                parts.append(f"{short}={v}{u if u not in ('%', 'register/thread', 'inst', 'sector') else ''}")
        print("  " + "  ".join(parts))
