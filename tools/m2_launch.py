"""A few MH launches in the time-to-convergence regime (few chains: every heavy micro-chain on a warp of its own), for ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_workload
from ssps_b200 import Sampler
C = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ds, xm, xp = make_workload(100, 8, 4)
with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=1000) as s:
    s.set_data(xm, xp, ds.prior_graph)
    s.generate(C, seed=18)
    for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 6):
        s.update_loop(1000)
    print(s.stats().kernel_ms)
