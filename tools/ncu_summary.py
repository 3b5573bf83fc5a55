"""Summarise an .ncu-rep (raw page) into one line per captured launch. usage: ncu_summary.py [--json out.json] rep [more reps]
--json also writes per-kernel averages (duration, DRAM bytes) that bench.py reads for roofline.traffic."""
import csv, io, subprocess, sys
WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__t_sector_hit_rate.pct", "l2hit%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "inst"),
        ("lts__t_sectors_srcunit_tex_op_atom.sum", "atom_sectors"), ("launch__grid_size", "grid")]
import json, re
args = sys.argv[1:]
json_out = None
if args and args[0] == "--json":
    json_out, args = args[1], args[2:]
agg = {}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
for rep in args:
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
        def val(key):
            return float(r[hdr.index(key)]) * UNIT.get(units[hdr.index(key)], 1) if key in hdr else 0.0
        base = re.sub(r"^void ", "", name).split("<")[0].strip()
        a = agg.setdefault(base, {"launches": 0, "time_us": 0.0, "dram": 0.0})
        a["launches"] += 1; a["time_us"] += val("gpu__time_duration.sum")
        a["dram"] += val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
if json_out:
    out = {"source": [os.path.basename(x) for x in args] if (os := __import__("os")) else None,
           "how": "ncu --set full --clock-control none (cold L2 per launch); averages over the captured launches",
           "kernels": {k: {"launches_captured": v["launches"], "time_us": round(v["time_us"] / v["launches"], 2),
                           "dram_bytes_per_launch": int(v["dram"] / v["launches"])} for k, v in sorted(agg.items())}}
    json.dump(out, open(json_out, "w"), indent=1)
