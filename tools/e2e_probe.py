import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
ds, xm, xp = make_workload(100, 8, 4)
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 6):
    t0 = time.perf_counter()
    s = Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=100)
    s.set_data(xm, xp, ds.prior_graph)
    t1 = time.perf_counter()
    s.generate(4096, seed=rep)
    t2 = time.perf_counter()
    s.update_loop(400)
    t3 = time.perf_counter()
    st = s.stats()
    s.close()
    print("rep %d: data %.1f init %.1f sweeps %.1f ms (kernel %.1f ms, %d launches)" % (rep, 1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), st.kernel_ms, st.mh_launches), file=sys.stderr, flush=True)
