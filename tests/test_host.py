"""Host-side logic (no GPU): input-file ingest, CLI argument table, resume parsing, simulator."""
import json
import os

import numpy as np
import pytest

REF_EXAMPLE = "/root/reference/run_ssps"


def _write_example(tmp_path, c1):
    """Re-create the reference's example input files from the committed fixture."""
    ts = tmp_path / "ts.csv"
    with open(ts, "w") as f:
        f.write("\t".join(["timeseries", "timestep"] + list(c1["var_names"])) + "\n")
        for r in range(c1["raw_table"].shape[0]):
            f.write("\t".join([str(c1["raw_series"][r]), repr(float(c1["raw_timestep"][r]))] +
                              [repr(float(x)) for x in c1["raw_table"][r]]) + "\n")
    pr = tmp_path / "prior.csv"
    np.savetxt(pr, c1["prior"], delimiter=",", fmt="%.17g")
    return str(ts), str(pr)


def test_ingest_round_trip(tmp_path, c1):
    from ssps_b200.preprocess import combine_X, load_formatted_data
    ts, pr = _write_example(tmp_path, c1)
    series, prior, names = load_formatted_data(ts, pr)
    assert [s.shape[0] for s in series] == list(c1["series_lengths"])
    assert names == list(c1["var_names"])
    xm, xp = combine_X(series)
    assert np.array_equal(xm, c1["Xminus"]) and np.array_equal(xp, c1["Xplus"])
    assert np.array_equal(prior, c1["prior"])


def test_ingest_sorts_by_timestep_and_keeps_series_order(tmp_path):
    from ssps_b200.preprocess import ingest_timeseries
    p = tmp_path / "t.tsv"
    p.write_text("timeseries\ttimestep\ta\tb\nB\t2\t3\t4\nB\t1\t1\t2\nA\t1\t5\t6\nB\t3\t7\t8\nA\t2\t9\t10\n")
    series, names = ingest_timeseries(str(p))
    assert names == ["a", "b"]
    assert np.array_equal(series[0], [[1, 2], [3, 4], [7, 8]])      # "B" appeared first
    assert np.array_equal(series[1], [[5, 6], [9, 10]])


@pytest.mark.skipif(not os.path.isdir(REF_EXAMPLE), reason="reference checkout not present (GPU box)")
def test_ingest_reference_files_directly(c1):
    from ssps_b200.preprocess import combine_X, load_formatted_data
    series, prior, _ = load_formatted_data(os.path.join(REF_EXAMPLE, "example_timeseries.csv"),
                                           os.path.join(REF_EXAMPLE, "example_prior.csv"))
    xm, xp = combine_X(series)
    assert np.array_equal(xm, c1["Xminus"]) and np.array_equal(xp, c1["Xplus"]) and np.array_equal(prior, c1["prior"])


def test_cli_argument_table_matches_reference():
    from ssps_b200.ssps_wrapper import parse_script_arguments
    a = parse_script_arguments(["ts", "ref", "out.json", "60"])
    # defaults of SSPS.jl:50-100
    assert (a.timeout, a.thinning, a.large_indeg, a.lambda_max, a.regression_deg, a.lambda_prop_std, a.n_steps,
            a.proposal, a.resume_sampling, a.multithread) == (60.0, 1, 20.0, 15.0, -1, 0.25, -1, "sparse", "", False)
    # the shell line of run_ssps/Snakefile:81-85
    a = parse_script_arguments("ts ref out 60 --regression-deg 1 --n-steps 100000 --lambda-prop-std 3.0 --large-indeg 20 --multithread".split())
    assert (a.regression_deg, a.n_steps, a.lambda_prop_std, a.large_indeg, a.multithread) == (1, 100000, 3.0, 20.0, True)
    with pytest.raises(SystemExit):
        parse_script_arguments(["ts", "ref", "out", "60", "--proposal", "bogus"])


def test_recover_state():
    from ssps_b200.ssps_wrapper import recover_state
    samples = {"lambda": [[[0, 1.5], [7, 2.5]], [[0, 3.0]]],
               "parent_sets": [{"2": [[0, 1], [3, 0]], "1": [[4, 1]]}, {"2": [[0, 1]], "1": [[1, 1], [2, 0], [9, 1]]}]}
    psets, lam = recover_state(samples)
    assert psets == [[0], [0, 1]] and lam == [2.5, 3.0]


def test_unsupported_modes_exit_loudly():
    from ssps_b200.ssps_wrapper import perform_inference
    with pytest.raises(SystemExit):     # the deprecated uniform proposal exists for regression_deg <= 1 only
        perform_inference("a", "b", "c", 1.0, 10, 1, 20.0, 15.0, 2, 0.25, "uniform", "", False)
    with pytest.raises(SystemExit):
        perform_inference("a", "b", "c", 1.0, 10, 1, 20.0, 15.0, 1, 0.25, "edge-reversal", "", False)


def test_simulator_shapes_and_prior_edit_counts():
    from ssps_b200.simulate import n_coeffs, simulate
    ds = simulate(40, 8, 4, seed=3)
    assert len(ds.series) == 4 and ds.series[0].shape == (8, 40)
    xm, xp = ds.lagged()
    assert xm.shape == (28, 40) and np.array_equal(xm[1:7], xp[0:6])
    E = int(ds.true_graph.sum())
    kept = int((ds.true_graph & (ds.prior_graph > 0)).sum())
    added = int((~ds.true_graph & (ds.prior_graph > 0)).sum())
    assert kept == int(np.round(0.5 * E)) and added == int(np.round(0.5 * E))     # modify_digraph!, remove = add = 0.5
    assert max(float(s.max()) for s in ds.series) < 100.0
    assert n_coeffs(3, 3) == 8 and n_coeffs(5, 3) == 26 and n_coeffs(0, 3) == 1    # simulate_data.jl:105-107
    ds2 = simulate(40, 8, 4, seed=3)
    assert np.array_equal(ds.series[2], ds2.series[2])


def test_simulator_file_writers(tmp_path):
    from ssps_b200.preprocess import load_formatted_data
    from ssps_b200.simulate import save_dataset, save_graph, simulate
    ds = simulate(12, 5, 3, seed=1)
    save_dataset(ds.series, str(tmp_path / "ts.tsv"))
    save_graph(ds.prior_graph, str(tmp_path / "prior.csv"))
    series, prior, names = load_formatted_data(str(tmp_path / "ts.tsv"), str(tmp_path / "prior.csv"))
    assert names[0] == "var01" and len(series) == 3
    assert np.array_equal(series[1], ds.series[1]) and np.array_equal(prior, ds.prior_graph)


def test_bench_workload_presets(monkeypatch, capsys):
    """bench.py's presets: the metric point (weak scaling), C4 (1024 chains IN TOTAL: strong scaling splits them over the ranks), C5
    (regression_deg 'full' = 64); the reference arm cannot exist for C5 and must say so in one JSON line."""
    import importlib, json, sys
    bench = importlib.import_module("bench")
    seen = {}

    def fake_ours(args, rank, world, local_rank):
        seen["args"] = args

    monkeypatch.setattr(bench, "run_ours", fake_ours)
    monkeypatch.setattr(bench, "protect_stdout", lambda: None)
    monkeypatch.setenv("WORLD_SIZE", "1")
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    bench.main()
    a = seen["args"]
    assert (a.V, a.T, a.deg, a.chains, a.sweeps, a.scaling, a.no_psrf) == (100, 8, 1, 4096, 400, "weak", False)
    assert a.metric == bench.METRIC
    monkeypatch.setattr(sys, "argv", ["bench.py", "--workload", "c4"])
    bench.main()
    a = seen["args"]
    assert (a.V, a.T, a.deg, a.chains, a.scaling, a.no_psrf) == (200, 100, 1, 1024, "strong", True) and "V=200" in a.metric
    monkeypatch.setattr(sys, "argv", ["bench.py", "--workload", "c5"])
    bench.main()
    a = seen["args"]
    assert (a.V, a.T, a.deg, a.deg_label, a.chains, a.scaling, a.no_cpu_baseline) == (1000, 100, 64, "full", 512, "weak", True)
    capsys.readouterr()
    monkeypatch.setattr(sys, "argv", ["bench.py", "--workload", "c5", "--impl", "reference"])
    bench.main()
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line
