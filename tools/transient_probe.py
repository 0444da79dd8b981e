"""GPU probe: cost of the first launches of a fresh handle (chains drawn from the prior) by storage tier."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler
C = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ds, xm, xp = make_workload(100, 8, 4)
for rep in range(2):
    with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=100, collect_khist=True) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(C, seed=rep)
        prev, praw, pk = s.stats(), s.raw_stats(32), s.khist(40)
        for piece in (8, 92, 100, 100, 100, 400):
            s.update_loop(piece)
            st, raw, kh = s.stats(), s.raw_stats(32), s.khist(40)
            dp = st.n_proposals - prev.n_proposals
            dk = kh - pk
            print("rep %d +%4d sweeps: kernel %.2f ms (%.3e prop/s) | mean k %.2f, k>5 %.3f, k>15 %.4f | mid %.4f big %.5f per proposal | big %.0f cyc/step, mid %.0f cyc/step"
                  % (rep, piece, st.kernel_ms - prev.kernel_ms, dp / ((st.kernel_ms - prev.kernel_ms) * 1e-3), (st.sum_k - prev.sum_k) / dp,
                     dk[6:].sum() / dk.sum(), dk[16:].sum() / dk.sum(), (raw[14] - praw[14]) / dp, (raw[13] - praw[13]) / dp,
                     (raw[11] - praw[11]) / max(raw[13] - praw[13], 1), (raw[15] - praw[15]) / max(raw[14] - praw[14], 1)), flush=True)
            prev, praw, pk = st, raw, kh
