"""Development helper: whole-step and table-update time vs the knobs of the table build on the bench workload
(EREF_BUCKET_BITS, EREF_COUNT_STREAMS, EREF_COUNT_UNITS, EREF_COUNT_CTAS).  usage: count_sweep.py [workload] [quick]"""
import itertools, os, sys
import numpy as np
import torch
sys.path.insert(0, ".")
from palace_b200 import eref
import bench
fa, f1, f2, bases = bench.gen_inputs(sys.argv[1] if len(sys.argv) > 1 else "c2", 0, 1)
d1, d2, dfa = (torch.from_numpy(x).cuda() for x in (f1, f2, fa))
quick = len(sys.argv) > 2
grid = [(5, 2, 2, 8), (5, 2, 2, 4), (5, 2, 2, 2), (5, 2, 2, 1), (6, 2, 2, 4), (6, 2, 2, 2), (6, 4, 2, 2), (7, 4, 2, 2), (7, 4, 1, 2),
        (6, 2, 1, 2), (6, 2, 1, 4), (6, 4, 1, 1), (7, 4, 1, 1), (6, 2, 4, 1), (6, 3, 2, 2)] if quick else \
    list(itertools.product((5, 6, 7), (2, 4), (1, 2), (1, 2, 4, 8)))
ref = None
stream = torch.cuda.Stream()
for bits, streams, units, ctas in grid:
    os.environ.update(EREF_COUNT_STREAMS=str(streams), EREF_COUNT_UNITS=str(units), EREF_BUCKET_BITS=str(bits),
                      EREF_COUNT_CTAS=str(ctas))
    with eref.ErefGPU(eref.random_coder(1)) as g, torch.cuda.stream(stream):
        g.set_stream(stream.cuda_stream)
        ts, steps = [], []
        for rep in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            g.query_reset()
            g.query_add_device(d1.data_ptr(), d1.numel()); g.query_add_device(d2.data_ptr(), d2.numel())
            g.ref_set_device(dfa.data_ptr(), dfa.numel())
            r = g.scan()
            e1.record(stream); e1.synchronize()
            t = g.timings()
            ts.append(t["query_count_ms"]); steps.append(e0.elapsed_time(e1))
        pops = [g.query_popcount(l) for l in (1, 2, 3)]
        if ref is None:
            ref = (pops, r.copy())
        assert pops == ref[0] and (r == ref[1]).all()
        print(f"bits={bits} streams={streams} units={units} ctas={ctas}  step_ms min={min(steps[2:]):.3f} med={sorted(steps[2:])[2]:.3f}  "
              f"count_ms med={sorted(ts[2:])[2]:.3f} frame={t['ref_frame_ms']:.3f} pack={t['ref_pack_ms']:.3f} bucket={t['query_bucket_ms']:.3f}", flush=True)
