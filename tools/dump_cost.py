import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
ds, xm, xp = make_workload(100, 8, 4)
with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=1000) as s:
    s.set_data(xm, xp, ds.prior_graph)
    s.generate(128, seed=1)
    out = []
    for rep in range(6):
        s.update_loop(1000)
        cost, mk, sp = s.vertex_cost()
        out.append((cost.copy(), sp.copy()))
        print(rep, "max task", (cost / (4 * sp)).max() / 1e6, "argmax", int(np.argmax(cost / (4 * sp))), "split of argmax", int(sp[np.argmax(cost / (4 * sp))]), "splits>1:", {int(v): int(sp[v]) for v in np.nonzero(sp > 1)[0]})
    np.save("/root/repo/gpurun_out/cost.npy", np.array([o[0] for o in out])); np.save("/root/repo/gpurun_out/split.npy", np.array([o[1] for o in out]))
