"""BASELINE.json's configurations as GPU parity / property cases (C2..C5; C1 is tests/test_gpu_parity.py).
Everything goes through the C ABI (ssps_b200.Sampler is a ctypes shim over libssps_b200.so)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conftest import make_synth  # noqa: E402

pytestmark = pytest.mark.gpu


def _dream_like(V=46, n_series=4, T=6, seed=3):
    """C3-shaped input: V ~ 46 proteins, 4 series x 6 time points (n = 20 lagged rows), prior confidences
    U(0, 0.3) with 30 % zeros (the DREAM priors are real-valued, unlike the simulator's 0/1 graphs)."""
    rng = np.random.default_rng(seed)
    series = [np.cumsum(rng.standard_normal((T, V)), axis=0) + rng.standard_normal(V) for _ in range(n_series)]
    xm = np.concatenate([s[:-1] for s in series])
    xp = np.concatenate([s[1:] for s in series])
    prior = rng.uniform(0.0, 0.3, (V, V)) * (rng.random((V, V)) > 0.3)
    return xm, xp, prior


def test_c2_posterior_agrees_with_oracle_chains_within_monte_carlo_error(oracle):
    """C2: V=40, T=8 x 4 series (n=28), 64 chains on one GPU.  Level-3 parity: pooled post-burn-in edge
    frequencies of GPU chains vs INDEPENDENT oracle chains (different streams) agree within Monte-Carlo error,
    and both rank the true edges equally well."""
    from ssps_b200 import Sampler
    from ssps_b200 import scoring
    ds, xm, xp = make_synth(40, seed=20200702)
    P = oracle.Problem(xm, xp, ds.prior_graph, large_indeg=15.0, lambda_prop_std=3.0)
    n_sweeps = 20000
    with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=1000) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(64, seed=11)
        s.update_loop(n_sweeps)
        g = s.chain_edge_counts().astype(np.float64) / n_sweeps          # per chain, whole run
    # oracle: 8 independent chains, kernel-faithful scorer
    n_cpu = 8
    o = []
    for c in range(n_cpu):
        ch = oracle.Chain(P, scorer=oracle.SCORER_CHOL)
        ch.init(77, 1000 + c)
        ch.recorder_enable()
        ch.run(n_sweeps)
        o.append(ch.edge_counts() / n_sweeps)
    o = np.array(o)
    p_gpu, p_cpu = g.mean(axis=0), o.mean(axis=0)
    # Welch statistic per edge from the chain-to-chain scatter of BOTH arms (chains stuck in different modes widen it)
    se = np.sqrt(g.var(axis=0, ddof=1) / 64 + o.var(axis=0, ddof=1) / n_cpu)
    z = np.abs(p_gpu - p_cpu) / np.maximum(se, 2e-3)
    assert np.mean(z > 5.0) < 0.01, (np.mean(z > 5.0), z.max())
    assert np.abs(p_gpu - p_cpu).mean() < 0.01
    truth = ds.true_graph.T.astype(bool)          # simulator: [parent, child]; posterior: [child][parent]
    if True:
        a_gpu, a_cpu = scoring.aucpr(p_gpu, truth), scoring.aucpr(p_cpu, truth)
        assert abs(a_gpu - a_cpu) < 0.05 and a_gpu > 2 * truth.mean()        # far better than chance


def test_c3_dream_like_priors_replay_bit_exact(oracle):
    """C3 shape: real-valued confidences (many distinct 1 - c groups in the lambda normaliser), n = 20 < V = 46,
    16 chains.  Replayed stream: parent sets and lambdas bit-identical to the oracle."""
    from ssps_b200 import Sampler
    xm, xp, prior = _dream_like()
    P = oracle.Problem(xm, xp, prior, large_indeg=20.0, lambda_prop_std=3.0)
    n_chains, n_sweeps, V = 16, 40, 46
    rng = np.random.default_rng(5)
    draws = rng.random((n_sweeps, n_chains, V, 6))
    draws[..., 0] = rng.standard_normal((n_sweeps, n_chains, V))
    with Sampler(large_indeg=20.0, lambda_prop_std=3.0, verify=True) as s:
        s.set_data(xm, xp, prior)
        s.generate(n_chains, seed=2)
        s.replay(draws)
        for c in range(n_chains):
            ch = oracle.Chain(P, scorer=oracle.SCORER_CHOL)
            ch.init(2, c)
            for t in range(n_sweeps):
                ch.sweep(draws[t, c])
            sets, lam, _ = s.state(c)
            assert sets == ch.parent_sets()
            assert np.array_equal(lam, ch.lambdas())
        assert s.stats().n_mismatch == 0


def test_c4_full_size_properties():
    """C4: V=200, T=100 x 4 series (n=396), 1024 chains: size-independent properties at full size
    (every cached-factor score equals its from-scratch score; tallies, counters and bounds are consistent;
    two handles with the same seed give identical chains whatever the task schedule does)."""
    from ssps_b200 import Sampler
    ds, xm, xp = make_synth(200, T=100, M=4, seed=20200704)
    assert xm.shape == (396, 200)
    n_chains, n_sweeps = 1024, 40

    def run(verify):
        with Sampler(large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=10, verify=verify) as s:
            s.set_data(xm, xp, ds.prior_graph)
            s.generate(n_chains, seed=9)
            s.update_loop(n_sweeps)        # 4 launches: the scheduler re-plans from measured cycles after the first
            return s.stats(), s.chain_edge_counts(), s.state(1023), s.psrf(0.5)[0]

    st, cnt, (sets, lam, lml), mx = run(True)
    assert st.n_mismatch == 0 and st.n_proposals == n_chains * 200 * n_sweeps
    assert cnt.max() <= n_sweeps and cnt.sum() > 0
    assert all(s_ == sorted(set(s_)) for s_ in sets) and np.isfinite(lml).all() and mx > 1.0
    st2, cnt2, (sets2, lam2, _), _ = run(False)
    assert np.array_equal(cnt, cnt2) and sets == sets2 and np.array_equal(lam, lam2)
    assert st2.n_accepted == st.n_accepted


def test_c5_interaction_design_at_v1000(synth_v1000):
    """C5: V=1000, T=100, M=4 (n=396), regression_deg 'full' read as deg >= |S| (all non-empty product subsets, quirk Q2 /
    Hill's 'full'), chains initialised FROM THE MODEL (ConfParentPrior.random with lambda ~ U(log(1000/15 - 1), 15): 4 parents
    per vertex on average, up to about 20: 2^k - 1 design columns, saturated and closed form from k = 9).  The chains shed
    parents through the warp-per-set tier (255 / 127 / 63 columns at k = 8 / 7 / 6) into the thread-per-set tier; no proposal may be refused
    for a scorer limit, the state must reproduce across handles and the kept lml must be the from-scratch score."""
    from ssps_b200 import Sampler
    ds, xm, xp = synth_v1000
    n_chains, n_sweeps, V = 64, 100, 1000
    out = []
    for rep in range(2):
        with Sampler(regression_deg=64, large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=-1) as s:
            s.set_data(xm, xp, ds.prior_graph)
            s.generate(n_chains, seed=4)
            k0 = np.array([len(x) for x in s.state(0)[0]])
            assert k0.max() > 12 and (k0 >= 6).sum() > 100   # the start IS beyond the old 12-parent / 48-column limits
            s.update_loop(n_sweeps if rep == 0 else 10)
            st = s.stats()
            sets, lam, lml = s.state(3)
            out.append((sets, lam.copy(), st.n_accepted))
            assert st.n_proposals == n_chains * V * (n_sweeps if rep == 0 else 10) and 0 < st.n_accepted < st.n_proposals
            assert st.n_indegree_rejects == 0
            assert np.isfinite(lml).all()
            relml, _ = s.score(list(range(V)), sets)
            np.testing.assert_array_equal(lml, relml)       # one definition, whatever tier scored it last
            if rep == 0:
                k1 = np.array([len(x) for x in sets])
                print("C5: mean in-degree %.1f -> %.2f after %d sweeps (max %d); accepted %.3f" % (k0.mean(), k1.mean(), n_sweeps, k1.max(), st.n_accepted / st.n_proposals))
                # 2^k - 1 columns at -1/2 log(1 + n) each: every set above k = 9 sheds parents at once.  k = 9 itself (511 columns, the
                # first saturated size: a perfect fit) is a local mode of the reference's score against k = 8 (255 columns of RANDOM parents
                # explain 255/395 of y): chains that start above it settle there, the others walk down to the sparse sets
                assert k1.max() <= 9 and k1.mean() < k0.mean()
                first = (sets, lam.copy())
    # determinism: a second handle, 10 sweeps, then the same 10 sweeps again on a third
    with Sampler(regression_deg=64, large_indeg=15.0, lambda_prop_std=3.0, psrf_grid=-1) as s:
        s.set_data(xm, xp, ds.prior_graph)
        s.generate(n_chains, seed=4)
        s.update_loop(10)
        sets2, lam2, _ = s.state(3)
    assert out[1][0] == sets2 and np.array_equal(out[1][1], lam2)
