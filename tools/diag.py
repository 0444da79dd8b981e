"""GPU diagnostic: tier statistics and throughput of the MH kernel on the bench workload."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from bench import make_workload
from ssps_b200 import Sampler

V = int(sys.argv[1]) if len(sys.argv) > 1 else 100
C = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
S = int(sys.argv[3]) if len(sys.argv) > 3 else 400
verify = len(sys.argv) > 4 and sys.argv[4] == "verify"
T = int(os.environ.get("DIAG_T", "8"))
ds, xm, xp = make_workload(V, T, 4)
with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=max(S // 4, 1), verify=verify, collect_khist=bool(os.environ.get('DIAG_KHIST'))) as s:
    s.set_data(xm, xp, ds.prior_graph)
    s.generate(C, seed=1)
    prev = s.stats()
    for rep in range(4):
        s.update_loop(S)
        st = s.stats()
        dp = st.n_proposals - prev.n_proposals
        dms = st.kernel_ms - prev.kernel_ms
        print("rep %d: %.3e prop/s kernel | t1 %.2e big %.2e lamx %.2e (per proposal) acc %.3f lam_acc %.3f mism %d"
              % (rep, dp / (dms * 1e-3), (st.n_tier1 - prev.n_tier1) / dp, (st.n_big_steps - prev.n_big_steps) / dp,
                 (st.n_lambda_exact - prev.n_lambda_exact) / dp, (st.n_accepted - prev.n_accepted) / dp,
                 (st.n_lambda_accepted - prev.n_lambda_accepted) / dp, st.n_mismatch))
        print("      big: %.0f cycles per big lane-step, %.3f of all task cycles" % ((st.big_cycles - prev.big_cycles) / max(st.n_big_steps - prev.n_big_steps, 1),
              (st.big_cycles - prev.big_cycles) / max(float(s.vertex_cost()[0].sum()) * 4, 1.0)))
        prev = st
        r = s.raw_stats(32)
        print("      mid: %d lane-steps (%.2e per proposal), %.0f cycles per lane-step, %.3f of all task cycles (cumulative; cycles only in the profiling build)"
              % (r[14], r[14] / max(st.n_proposals, 1), r[15] / max(r[14], 1), r[15] / max(float(s.vertex_cost()[0].sum()) * (rep + 1), 1.0)))
    cost, mk, sp = s.vertex_cost()
    n_slabs = (C + 31) // 32
    task = cost / (n_slabs * sp)
    o = np.argsort(-task)
    print("kernel ms last rep:", dms)
    print("(vertex, Mcycles per task, split, mean indeg) top 12:", [(int(i), round(task[i] / 1e6, 3), int(sp[i]), round(float(mk[i]), 2)) for i in o[:12]])
    print("sum task Mcycles %.1f, tasks %d, max task Mcycles %.3f, median %.3f, split hist %s" % (cost.sum() / 1e6, int((n_slabs * sp).sum()), task.max() / 1e6, np.median(task) / 1e6, np.bincount(sp).tolist()))
    print("mean indeg hist:", np.histogram(mk, bins=[0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 16, 32, 64])[0])
    if os.environ.get('DIAG_KHIST'):
        print("khist:", s.khist(40).tolist())
    if os.environ.get('DIAG_PROF'):
        r = s.raw_stats(48)
        if r[39]:
            print("MID: %d passes (two lanes each at most), %.0f cycles per pass, %.0f per lane-step | per pass: copy-in %.0f, scans+prefix %.0f, rows %.0f (%.1f rows), accept %.0f, commit %.0f"
                  % (r[39], r[15] / r[39], r[15] / max(r[14], 1), r[32] / r[39], r[33] / r[39], r[35] / r[39], r[38] / r[39], r[36] / r[39], r[37] / r[39]))
        names = ["select+bcast", "stage_load", "scans", "remove (+q prefix)", "append", "accept (src lane)", "commit", "promote+tail"]
        tot = float(r[16:24].sum()) or 1.0
        print("BIG phase cycles by section (profiling build): " + ", ".join("%s %.1f%%" % (n, 100 * r[16 + e] / tot) for e, n in enumerate(names)),
              "| per big lane-step %.0f" % (tot / max(st.n_big_steps, 1)))
        if r[31]:
            print("fac_remove: %d calls, mean trailing rows %.1f, partial sums %.0f cycles, trailing loop %.0f cycles per call (%.0f per pivot)"
                  % (r[31], r[30] / r[31], r[28] / r[31], r[29] / r[31], r[29] / max(r[30], 1)))
