"""Development helper: scan phase times vs EREF_PROBE_PARTS on one GPU for a bench workload (default c4: dense table)."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from palace_b200 import eref
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
t0 = time.time()
fa, f1, f2, bases = bench.gen_inputs(wl, 0, 1)
print(f"{wl}: db {fa.size/1e6:.0f} MB, fq {f1.size/1e6:.0f}+{f2.size/1e6:.0f} MB, generated in {time.time()-t0:.0f} s", flush=True)
fa_d, f1_d, f2_d = (torch.from_numpy(x).cuda() for x in (fa, f1, f2))
lpw = os.environ.get('EREF_PROBE_LPW', '4')
for parts in (int(x) for x in os.environ.get('PARTS', '1,4,8,16').split(',')):
    os.environ["EREF_PROBE_PARTS"] = str(parts)
    with eref.ErefGPU(eref.random_coder(1)) as g:
        for it in range(3):
            g.query_reset()
            g.query_add_device(f1_d.data_ptr(), f1_d.numel(), want_bytes=False)
            g.query_add_device(f2_d.data_ptr(), f2_d.numel(), want_bytes=False)
            g.ref_set_device(fa_d.data_ptr(), fa_d.numel())
            g.scan_device()
            t = g.timings()
        print(f"parts={parts:2d} lpw={lpw} sample={t['scan_sample_ms']:.3f} need={t['scan_cand_ms']:.3f} probe={t['scan_probe_ms']:.3f} "
              f"window={t['scan_window_ms']:.3f} need_frac={t['ref_need_words']/max(t['ref_words'],1):.3f} "
              f"bucket={t['query_bucket_ms']:.2f} count={t['query_count_ms']:.2f} codes={t['query_codes']/1e6:.0f}M", flush=True)
