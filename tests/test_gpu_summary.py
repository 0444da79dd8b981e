"""f1 on the GPU: the device-emitted post-processed summary (K5, N_eff included) against the host twin of the reference's
scripts/postprocess_samples.jl, and the reference's ONE test (tests/ssps_csv.sh) reproduced end to end through the CLI."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

REF_KEYS = {"stop_points", "summary", "conv_stats", "parent_sets", "edge_conf_key", "t_elapsed", "n"}     # collect_stats :223-231


def _close(a, b, rel):
    if a is None or b is None:
        return a is None and b is None
    return abs(a - b) <= rel * max(abs(a), abs(b), 1e-300)


def _compare_summaries(got, want, V, rel=1e-9):
    assert set(got) == REF_KEYS | {"ssps_b200"}
    assert got["stop_points"] == want["stop_points"] and got["n"] == want["n"] and got["edge_conf_key"] == "parent_sets"
    worst = 0.0
    n_null = 0
    for i in range(V):
        assert _close(got["summary"]["lambda"][i], want["summary"]["lambda"][i], 1e-12)
        for sp in range(len(want["stop_points"])):
            for k in (0, 1):
                assert _close(got["conv_stats"]["lambda"][i][sp][k], want["conv_stats"]["lambda"][i][sp][k], rel), ("lambda", i, sp, k)
        for j in range(V):
            assert _close(got["parent_sets"][i][j], want["parent_sets"][i][j], 1e-12)
            for sp in range(len(want["stop_points"])):
                g, w = got["conv_stats"]["parent_sets"][i][j][sp], want["conv_stats"]["parent_sets"][i][j][sp]
                n_null += g[0] is None
                for k in (0, 1):
                    assert _close(g[k], w[k], rel), ("edge", i, j, sp, k, g, w)
                    if g[k] is not None:
                        worst = max(worst, abs(g[k] - w[k]) / max(abs(w[k]), 1e-300))
    return worst, n_null


def test_device_summary_matches_the_reference_script_twin(c1, tmp_path):
    """PSRF and N_eff of every edge and every lambda_i, the pooled posterior and the lambda means: device (ssps_convergence,
    from the change-point log) vs postprocess.collect_stats on the per-chain JSON files of the same chains (dense sequences,
    the reference's algorithm line by line), at two stop points."""
    from ssps_b200 import Sampler, postprocess
    V, n_chains, n_sweeps = 46, 5, 600
    with Sampler(large_indeg=20.0, lambda_prop_std=3.0, record_changepoints=True) as s:
        s.set_data(c1["Xminus"], c1["Xplus"], c1["prior"])
        s.generate(n_chains, seed=41)
        s.update_loop(n_sweeps)
        files = []
        for c in range(n_chains):
            f = tmp_path / ("chain%d.json" % c)
            f.write_text(json.dumps(s.results(c, t_elapsed=1.5)))
            files.append(str(f))
        got = postprocess.summarize_sampler(s, burnin=0.5, t_elapsed=1.5, stop_points=[301, 600])
        got_b = postprocess.summarize_sampler(s, burnin=0.3, t_elapsed=1.5)
    want = postprocess.collect_stats([postprocess.load_samples(f) for f in files], [301, 600], 0.5)
    want_b = postprocess.collect_stats([postprocess.load_samples(f) for f in files], [600], 0.3)
    worst, n_null = _compare_summaries(got, want, V)
    worst_b, _ = _compare_summaries(got_b, want_b, V)
    n_eff = np.array([[e[1][1] for e in row] for row in got["conv_stats"]["parent_sets"]], dtype=float)
    print("device summary vs reference-script twin: worst relative difference %.2e / %.2e (burn-in 0.5 / 0.3); %d constant edges (PSRF null); "
          "N_eff range %.1f .. %.1f" % (worst, worst_b, n_null, np.nanmin(n_eff), np.nanmax(n_eff)))
    assert got["ssps_b200"]["burnin_index"] == [150, 300] and got["ssps_b200"]["split_index"] == [225, 450]
    assert n_null > 0                                   # edges that never appear: W = 0 -> PSRF null, N_eff = M n exactly
    psrf = np.array([[e[1][0] for e in row] for row in got["conv_stats"]["parent_sets"]], dtype=float)
    const = n_eff[np.isnan(psrf)]       # W = 0: never / always present in every half (N_eff = M n exactly), or constant per half at different values
    assert (const == 2 * n_chains * 150.0).sum() > 0.9 * const.size and (const <= 2 * n_chains * 150.0).all()


def test_tally_summary_without_the_log_reports_what_it_used():
    from conftest import make_synth
    from ssps_b200 import Sampler, postprocess
    ds, xm, xp = make_synth(30, seed=2)
    with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=50) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(64, seed=3)
        s.update_loop(400)
        out = postprocess.summarize_sampler(s, burnin=0.5, t_elapsed=0.1)
        mx, edge = s.psrf(0.5)
    assert set(out) == REF_KEYS | {"ssps_b200"}
    assert out["stop_points"] == [400] and out["ssps_b200"]["burnin_index"] == [200]
    assert out["conv_stats"]["parent_sets"][0][0][0][1] is None                 # N_eff needs the log
    assert out["ssps_b200"]["max_edge_psrf"] == pytest.approx(mx, rel=1e-12)
    json.dumps(out)


def _write_example_inputs(c1, d):
    """run_ssps/example_timeseries.csv (TAB separated, header timeseries / timestep / variables) and example_prior.csv
    (comma separated, no header) rebuilt from the committed fixture."""
    ts = os.path.join(d, "example_timeseries.csv")
    with open(ts, "w") as f:
        f.write("timeseries\ttimestep\t" + "\t".join(str(v) for v in c1["var_names"]) + "\n")
        for r in range(c1["raw_table"].shape[0]):
            f.write("%s\t%r\t%s\n" % (c1["raw_series"][r], float(c1["raw_timestep"][r]), "\t".join(repr(float(x)) for x in c1["raw_table"][r])))
    pr = os.path.join(d, "example_prior.csv")
    np.savetxt(pr, c1["prior"], delimiter=",", fmt="%.17g")
    return ts, pr


def test_reference_csv_test_reproduced_on_the_gpu(c1, tmp_path):
    """tests/ssps_csv.sh:9-20: run the run_ssps example through the CLI (4 chains, the Snakefile's flags, run_ssps/Snakefile:81-85),
    post-process (postprocess_samples.jl's CLI twin), convert to the genie edge list and compare the (node1, node2) key set with
    tests/baseline_predictions.genie -- csvdiff on the key columns, scores ignored.  Also: --resume-sampling continues a chain,
    and an expiring timeout still writes a valid file."""
    from ssps_b200 import postprocess
    d = str(tmp_path)
    ts, pr = _write_example_inputs(c1, d)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    chains = []
    for c in range(4):
        out = os.path.join(d, "%d.json" % c)
        cmd = [sys.executable, "-m", "ssps_b200.ssps_wrapper", ts, pr, out, "60", "--regression-deg", "1", "--n-steps", "3000",
               "--lambda-prop-std", "3.0", "--large-indeg", "20", "--multithread", "--seed", str(100 + c)]
        subprocess.check_call(cmd, env=env, cwd=d)
        chains.append(out)
    pred = os.path.join(d, "predictions.json")
    subprocess.check_call([sys.executable, "-m", "ssps_b200.postprocess", "--chain-samples"] + chains + ["--output-file", pred], env=env, cwd=d)
    res = json.load(open(pred))
    assert res["n"] == [3001] * 4 and res["stop_points"] == [3001]                # quirk Q3: n_steps is inclusive
    genie = os.path.join(d, "example_predictions.genie")
    postprocess.write_genie(genie, res["parent_sets"], [str(x) for x in c1["node_names"]])
    lines = open(genie).read().strip().split("\n")
    assert lines[0] == "node1,node2,score"
    keys = {tuple(l.split(",")[:2]) for l in lines[1:]}
    want = set(zip((str(x) for x in c1["genie_node1"]), (str(x) for x in c1["genie_node2"])))
    assert len(lines) - 1 == len(want) == 2116 and keys == want                   # "files are identical" on the key columns
    # the scores are not part of the reference's test; they still have to look like its artefact
    p = np.array(res["parent_sets"])
    assert p[21, 21] > 0.9 and p[30, 21] > 0.9 and abs(p.mean() - c1["genie_scores"].mean()) < 0.03

    # --resume-sampling: continues from the last state of a chain file (recover_state!, SSPS.jl:24-43)
    first = json.load(open(chains[0]))
    out2 = os.path.join(d, "resumed.json")
    subprocess.check_call([sys.executable, "-m", "ssps_b200.ssps_wrapper", ts, pr, out2, "60", "--regression-deg", "1", "--n-steps", "10",
                           "--lambda-prop-std", "3.0", "--large-indeg", "20", "--resume-sampling", chains[0], "--seed", "7"], env=env, cwd=d)
    r2 = json.load(open(out2))
    last_sets = [sorted(int(k) for k, v in ps.items() if v[-1][1] == 1) for ps in first["parent_sets"]]
    init_sets = [sorted(int(k) for k, v in ps.items() if v[0] == [0, 1]) for ps in r2["parent_sets"]]
    assert init_sets == last_sets and r2["n"] == 11
    assert [lv[0][1] for lv in r2["lambda"]] == [lv[-1][1] for lv in first["lambda"]]

    # timeout expiry: no iteration limit, 2 s budget -> stops on the clock and writes what it has (mcmc_inference.jl:58)
    out3 = os.path.join(d, "timed.json")
    subprocess.check_call([sys.executable, "-m", "ssps_b200.ssps_wrapper", ts, pr, out3, "2", "--regression-deg", "1",
                           "--lambda-prop-std", "3.0", "--large-indeg", "20", "--seed", "8"], env=env, cwd=d)
    r3 = json.load(open(out3))
    assert r3["n"] > 100 and 1.5 < r3["t_elapsed"] < 30.0
    assert len(r3["parent_sets"]) == 46 and len(r3["lambda"]) == 46


def test_default_cli_flags_run_the_shipped_example(c1, tmp_path):
    """`ssps_wrapper ts prior out timeout` with the reference's DEFAULT flags (--regression-deg -1 = the empty design of quirk
    Q2, --large-indeg 20: initial in-degrees around 17): round 1 sent this through a 12-parent path and failed at once."""
    d = str(tmp_path)
    ts, pr = _write_example_inputs(c1, d)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    out = os.path.join(d, "default.json")
    subprocess.check_call([sys.executable, "-m", "ssps_b200.ssps_wrapper", ts, pr, out, "30", "--n-steps", "500", "--seed", "5",
                           "--posterior-json", os.path.join(d, "post.json")], env=env, cwd=d)
    r = json.load(open(out))
    assert r["n"] == 501
    k0 = [sum(1 for v in ps.values() if v[0] == [0, 1]) for ps in r["parent_sets"]]
    assert max(k0) > 12                                        # the initial state alone is beyond round 1's limit
    post = json.load(open(os.path.join(d, "post.json")))
    assert set(post) == REF_KEYS | {"ssps_b200"}
