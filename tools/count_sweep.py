"""Development helper: table-update time vs EREF_COUNT_STREAMS / EREF_COUNT_UNITS on the bench workload."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, ".")
from palace_b200 import eref
import bench
fa, f1, f2, bases = bench.gen_inputs(sys.argv[1] if len(sys.argv) > 1 else "c2", 0, 1)
d1, d2, dfa = (torch.from_numpy(x).cuda() for x in (f1, f2, fa))
ref = None
for streams in (1, 2, 3, 4):
    for units in (1, 2, 4):
        os.environ["EREF_COUNT_STREAMS"] = str(streams); os.environ["EREF_COUNT_UNITS"] = str(units)
        with eref.ErefGPU(eref.random_coder(1)) as g:
            ts = []
            for rep in range(6):
                g.query_reset()
                g.query_add_device(d1.data_ptr(), d1.numel()); g.query_add_device(d2.data_ptr(), d2.numel())
                g.ref_set_device(dfa.data_ptr(), dfa.numel())
                r = g.scan()
                ts.append(g.timings()["query_count_ms"])
            pops = [g.query_popcount(l) for l in (1, 2, 3)]
            if ref is None:
                ref = (pops, r.copy())
            assert pops == ref[0] and (r == ref[1]).all()
            print(f"streams={streams} units={units}  count_ms min={min(ts[2:]):.3f} med={sorted(ts[2:])[2]:.3f}", flush=True)
