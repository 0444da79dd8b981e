"""GPU tests of the deprecated uniform graph proposal (--proposal uniform: uniform_update_loop / uniform_proposal /
uniform_involution, state_updates.jl:98-112, dbn_proposals.jl:142-230) against the oracle's restatement
(ORC_FLAG_UNIFORM, whose stationary law is checked on the CPU in tests/test_oracle.py)."""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from conftest import make_synth  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("thinning", [1, 2])
def test_l2_uniform_replay_bit_exact_c1(c1, c1_problem, oracle, thinning):
    """Replayed stream on the shipped example (V = 46, n = 18: rank-deficient sets, dense starts): identical decisions,
    parent sets, lambdas, tallies and change-point log; add, remove AND reverse moves must have been accepted."""
    from ssps_b200 import Sampler
    n_chains, n_sweeps, V = 5, 40, c1_problem.V
    rng = np.random.default_rng(300 + thinning)
    draws = rng.random((n_sweeps, n_chains, V, 6))
    draws[..., 0] = rng.standard_normal((n_sweeps, n_chains, V))
    draws[::3, :, ::2, 2] = 1.0 - 1e-4 * rng.random(draws[::3, :, ::2, 2].shape)     # many draws land on the reverse-edge range
    with Sampler(large_indeg=20.0, lambda_max=15.0, lambda_prop_std=3.0, record_changepoints=True, proposal="uniform") as s:
        s.set_data(c1["Xminus"], c1["Xplus"], c1["prior"])
        s.generate(n_chains, seed=5)
        s.replay(draws, thinning=thinning)
        cnt = s.chain_edge_counts()
        n_rev = 0
        for c in range(n_chains):
            ch = oracle.Chain(c1_problem, scorer=oracle.SCORER_CHOL, flags=oracle.FLAG_UNIFORM)
            ch.init(5, c)
            ch.recorder_enable()
            for t in range(n_sweeps):
                acc = ch.sweep(draws[t, c])
                n_rev += int(((acc[:, 1] == 1) & (acc[:, 2] == 2)).sum())
                if (t + 1) % thinning == 0:
                    ch.record()
            sets, lam, lml = s.state(c)
            assert sets == ch.parent_sets()
            assert np.array_equal(lam, ch.lambdas())
            np.testing.assert_allclose(lml, ch.lmls(), rtol=1e-12)
            assert np.array_equal(cnt[c].astype(np.int64), ch.edge_counts())
            got, want = s.results(c), ch.results_dict()
            assert got["n"] == want["n"] == n_sweeps // thinning
            assert got["parent_sets"] == want["parent_sets"]
        st = s.stats()
        assert st.n_proposals == n_chains * V * n_sweeps and st.n_accepted > 0
    assert n_rev > 0, "no reverse-edge move was accepted: the test does not exercise the two-vertex update"


def test_uniform_philox_run_matches_oracle_chain(oracle):
    """The library's own Philox streams (ssps_run): ssps_export_draws returns the very numbers the kernel used (the normal is
    formed with fp32 MUFU approximations on the device); the oracle chain fed with them must visit the same graphs."""
    from ssps_b200 import Sampler
    ds, xm, xp = make_synth(20, seed=8)
    P = oracle.Problem(xm, xp, ds.prior_graph, large_indeg=15.0, lambda_prop_std=3.0)
    n_chains, n_sweeps = 4, 50
    with Sampler(large_indeg=15.0, lambda_prop_std=3.0, proposal="uniform") as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(n_chains, seed=21, chain_id_offset=2)
        s.update_loop(30)
        s.update_loop(n_sweeps - 30)                # two runs: the launch-start rebuild of the adjacency state must be seamless
        draws = s.export_draws(0, n_sweeps)
        for c in range(n_chains):
            ch = oracle.Chain(P, scorer=oracle.SCORER_CHOL, flags=oracle.FLAG_UNIFORM)
            ch.init(21, 2 + c)
            for t in range(n_sweeps):
                ch.sweep(draws[t, c])
            sets, lam, _ = s.state(c)
            assert sets == ch.parent_sets()
            assert np.array_equal(lam, ch.lambdas())


def test_uniform_and_sparse_proposals_sample_the_same_posterior():
    """L3: two different MH kernels for one target.  Edge posteriors of a uniform-proposal run and of a sparse-proposal run on the
    same data agree within Monte-Carlo error (V = 12: the uniform proposal mixes slowly, one graph move per vertex and sweep)."""
    from ssps_b200 import Sampler
    ds, xm, xp = make_synth(12, seed=31, T=10, M=4)
    post = {}
    for prop, sweeps in (("uniform", 30000), ("sparse", 6000)):
        with Sampler(large_indeg=4.0, lambda_prop_std=3.0, psrf_grid=500, proposal=prop) as s:
            s.set_data(xm, xp, ds.prior_graph)
            s.generate(256, seed=3)
            s.update_loop(sweeps)
            post[prop] = s.edge_posterior(0.5)[0]
    d = np.abs(post["uniform"] - post["sparse"])
    print("uniform vs sparse proposal: max |dp| %.4f, mean |dp| %.4f, Pearson %.4f" % (d.max(), d.mean(), np.corrcoef(post["uniform"].ravel(), post["sparse"].ravel())[0, 1]))
    assert d.max() < 0.05 and np.corrcoef(post["uniform"].ravel(), post["sparse"].ravel())[0, 1] > 0.995


def test_cli_uniform_proposal_writes_the_reference_json(c1):
    """`ssps_wrapper ts prior out timeout --proposal uniform` (without --multithread, which would override it, SSPS.jl:98)."""
    from test_gpu_summary import _write_example_inputs
    d = tempfile.mkdtemp()
    ts, pr = _write_example_inputs(c1, d)
    out = os.path.join(d, "chain.json")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    subprocess.check_call([sys.executable, "-m", "ssps_b200.ssps_wrapper", ts, pr, out, "60", "--regression-deg", "1", "--n-steps", "300",
                           "--lambda-prop-std", "3.0", "--large-indeg", "20", "--proposal", "uniform", "--seed", "11"], env=env, cwd=d)
    res = json.load(open(out))
    assert res["n"] == 301 and len(res["parent_sets"]) == 46 and len(res["lambda"]) == 46
