"""GPU probe for metric M2: max edge split-PSRF versus sweeps (V=100 metric point)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
C = int(sys.argv[1]) if len(sys.argv) > 1 else 512
cap_s = float(sys.argv[2]) if len(sys.argv) > 2 else 120.0
check = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
ds, xm, xp = make_workload(100, 8, 4)
t0 = time.perf_counter()
with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=check // 4) as s:
    s.set_data(xm, xp, ds.prior_graph)
    s.generate(C, seed=18)
    done = 0; nxt = check
    while time.perf_counter() - t0 < cap_s:
        s.update_loop(check); done += check
        mx, edge = s.psrf(0.5)
        if done >= nxt or mx < 1.1:
            e = edge[np.isfinite(edge)]
            print("sweeps %8d  t %7.2fs  max psrf %.4f  #edges>1.1: %d  >1.05: %d" % (done, time.perf_counter() - t0, mx, (e > 1.1).sum(), (e > 1.05).sum()), flush=True)
            nxt = int(nxt * 1.5)
        if mx < 1.1: break
