"""Development helper: per-phase CUDA-event times of the eref path on a synthetic workload.
usage: python tools/phase_times.py [db_genomes db_bp query_bp reps]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from palace_b200 import eref, synth

G, R, Q, reps = (int(x) for x in (sys.argv[1:5] + ["2000", "100000000", "10000000", "3"][len(sys.argv) - 1:]))
t = time.time()
db = synth.make_db(G, R, 1000)
f1, f2, present = synth.make_reads(db, Q, 2000)
print(f"generated DB {db.text.size} B, fq {f1.size}+{f2.size} B, present {len(present)} in {time.time() - t:.1f}s", flush=True)
coder = eref.random_coder(1)
for prune, bucketed in ((True, True), (True, False), (False, True)):
    with eref.ErefGPU(coder, prune=prune, bucketed=bucketed) as g:
        for rep in range(reps):
            g.query_reset()
            t0 = time.perf_counter()
            g.query_add(f1); ta = g.timings()
            g.query_add(f2); tb = g.timings()
            g.ref_set(db.text); tc = g.timings()
            res = g.scan(); td = g.timings()
            wall = time.perf_counter() - t0
            out = {k: round(v, 3) if isinstance(v, float) else v for k, v in td.items()}
            out["fq1_pack_ms"] = round(ta["query_pack_ms"], 3)
            out["fq1_bucket_ms"] = round(ta["query_bucket_ms"], 3)
            out["fq1_h2d_ms"] = round(ta["query_h2d_ms"], 3)
            out["wall_ms"] = round(wall * 1e3, 2); out["prune"] = prune; out["bucketed"] = bucketed; out["selected"] = int((res["el"] > 0).sum())
            print(json.dumps(out), flush=True)
        print("table popcounts", [g.query_popcount(l) for l in (1, 2, 3)], flush=True)
