"""Development helper: e2e (eref_run_host) time vs number of database pieces on the bench workload."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from palace_b200 import eref
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
fa, f1, f2, bases = bench.gen_inputs(wl, 0, 1)
fa_h, f1_h, f2_h = (torch.from_numpy(x).pin_memory().numpy() for x in (fa, f1, f2))
cap = int(np.count_nonzero(fa == 62)) + 16
g = eref.ErefGPU(eref.random_coder(1))
for chunks in (1, 2, 4, 6, 8, 12, 16):
    for _ in range(2):
        g.run_host(f1_h, f2_h, fa_h, chunks=chunks, cap=cap)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5):
        r = g.run_host(f1_h, f2_h, fa_h, chunks=chunks, cap=cap)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    t = g.timings()
    print(f"chunks={chunks:2d}  {dt*1e3:7.3f} ms  {bases/1e9/dt:6.2f} Gbp/s  records={r.size} selected={(r['el']>0).sum()} "
          f"frame={t['ref_frame_ms']:.2f} pack={t['ref_pack_ms']:.2f} count={t['query_count_ms']:.2f}", flush=True)
