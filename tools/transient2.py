"""GPU probe: kernel time of the first launches of a fresh handle (no histogram), with the split the scheduler chose."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
C = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
pieces = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [8, 92, 100, 100, 100, 400]
ds, xm, xp = make_workload(100, 8, 4)
for rep in range(2):
    with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=100) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(C, seed=rep)
        prev = s.stats()
        for piece in pieces:
            s.update_loop(piece)
            st = s.stats()
            cyc, mk, sp = s.vertex_cost()
            o = np.argsort(-cyc)[:6]
            print("rep %d +%4d sweeps: kernel %.2f ms, %d launches | split hist %s | top vertices (cycles/1e6, mean k, split): %s"
                  % (rep, piece, st.kernel_ms - prev.kernel_ms, st.mh_launches - prev.mh_launches, np.bincount(sp, minlength=33)[[1, 2, 4, 8, 16, 32]].tolist(),
                     [(round(cyc[v] / 1e6, 1), round(float(mk[v]), 2), int(sp[v])) for v in o]), flush=True)
            prev = st
