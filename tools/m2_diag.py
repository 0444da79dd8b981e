"""GPU probe, time-to-convergence regime (few chains): where the cycles of the slowest task of a launch go."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
C = int(sys.argv[1]) if len(sys.argv) > 1 else 16
S = int(sys.argv[2]) if len(sys.argv) > 2 else 4000
ds, xm, xp = make_workload(100, 8, 4)
with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=S // 4, collect_khist=bool(os.environ.get("DIAG_KHIST"))) as s:
    s.set_data(xm, xp, ds.prior_graph)
    s.generate(C, seed=18)
    prev, praw = s.stats(), s.raw_stats(48)
    for rep in range(5):
        s.update_loop(S)
        st, raw = s.stats(), s.raw_stats(48)
        cyc, mk, sp = s.vertex_cost()
        n_slabs = (C + 31) // 32
        task = cyc / (n_slabs * sp)             # mean cycles per task of the vertex in the LAST launch (S / 4 sweeps)
        o = np.argsort(-task)[:5]
        dms = st.kernel_ms - prev.kernel_ms
        dp = st.n_proposals - prev.n_proposals
        print("rep %d: %.2f us per sweep (kernel), %d launches | MID lane-steps %.4f, BIG %.5f per proposal | BIG %.0f cycles per lane-step, MID %.0f per pass"
              % (rep, 1e3 * dms / S, st.mh_launches - prev.mh_launches, (raw[14] - praw[14]) / dp, (raw[13] - praw[13]) / dp,
                 (raw[11] - praw[11]) / max(raw[13] - praw[13], 1), (raw[15] - praw[15]) / max(raw[39] - praw[39], 1)))
        print("       slowest tasks (vertex, cycles per sweep, split, mean k): %s | median task %.0f cycles per sweep, tasks %d"
              % ([(int(v), int(task[v] / (S // 4)), int(sp[v]), round(float(mk[v]), 2)) for v in o], np.median(task) / (S // 4), int((n_slabs * sp).sum())), flush=True)
        prev, praw = st, raw
