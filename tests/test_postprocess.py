"""Host-side twins of the reference's post-processing and scoring scripts: checked on CPU
against the oracle's dense C restatement, scikit-learn, and sample files written by the oracle chain."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conftest import make_synth  # noqa: E402
from ssps_b200 import postprocess as pp  # noqa: E402
from ssps_b200 import scoring  # noqa: E402


def _flip_seq(rng, n, p):
    x = np.zeros(n, dtype=np.uint8)
    v = int(rng.random() < 0.5)
    for t in range(n):
        if rng.random() < p:
            v = 1 - v
        x[t] = v
    return x


def _changepoints(x):
    cps = [[0, int(x[0])]]
    for t in range(1, len(x) + 1):
        if x[t - 1] != cps[-1][1]:
            cps.append([t, int(x[t - 1])])
    return cps


def test_dense_sequence_follows_changepointvec_indexing():
    # index 0 = initial state; the value at t is that of the last change point <= t; no change points -> 0
    assert pp.dense_sequence([], 4).tolist() == [0, 0, 0, 0]
    assert pp.dense_sequence([[0, 1], [3, 0]], 5).tolist() == [1, 1, 0, 0, 0]
    assert pp.dense_sequence([[2, 1]], 3).tolist() == [0, 1, 1]
    assert pp.dense_sequence([[0, 2.5], [1, 3.5]], 2).tolist() == [3.5, 3.5]


@pytest.mark.parametrize("stop", [400, 399, 250])
def test_psrf_neff_match_the_oracle_dense_restatement(oracle, stop):
    rng = np.random.default_rng(stop)
    seqs = [_flip_seq(rng, 400, 0.07) for _ in range(4)]
    cps = [_changepoints(x) for x in seqs]
    (psrf, neff), = pp.evaluate_convergence(cps, [400] * 4, [stop], 0.5)
    cut = [x[:stop] for x in seqs]
    if (stop - stop // 2) % 2 == 0:          # the oracle's dense routine takes equal halves
        assert psrf == pytest.approx(oracle.psrf_binary_dense(cut, 0.5), rel=1e-12)
        assert neff == pytest.approx(oracle.neff_binary_dense(cut, 0.5), rel=1e-12)
    assert 1.0 <= psrf < 1.5 and 0 < neff <= 8 * stop


def test_constant_sequences_give_null_psrf_like_the_reference():
    out = pp.evaluate_convergence([[], []], [50, 50], [50], 0.5)
    assert np.isnan(out[0][0]) and out[0][1] == 4 * 12      # W = 0 -> NaN -> JSON null; corr_sum 0 -> m n


def test_postprocess_sample_files_on_oracle_chains(tmp_path, oracle):
    ds, xm, xp = make_synth(12, T=8, M=3, seed=9)
    P = oracle.Problem(xm, xp, ds.prior_graph, large_indeg=4.0, lambda_prop_std=3.0)
    files, counts = [], []
    for c in range(3):
        ch = oracle.Chain(P, scorer=oracle.SCORER_CHOL)
        ch.init(5, c)
        ch.recorder_enable()
        ch.run(300)
        f = tmp_path / ("chain%d.json" % c)
        f.write_text(json.dumps(ch.results_dict(t_elapsed=1.0 + c)))
        files.append(str(f))
        counts.append(ch)
    out = tmp_path / "post.json"
    assert pp.main(["--chain-samples"] + files + ["--output-file", str(out), "--stop-points", "200", "300", "999"]) == 0
    res = json.loads(out.read_text())
    assert res["stop_points"] == [200, 300] and res["n"] == [300] * 3 and res["t_elapsed"] == [1.0, 2.0, 3.0]
    assert res["edge_conf_key"] == "parent_sets" and set(res["summary"]) == {"lambda"}
    V = 12
    probs = np.array(res["parent_sets"])
    assert probs.shape == (V, V) and probs.min() >= 0 and probs.max() <= 1
    # pooled post-burn-in frequency = (tallies over samples 151..300) / (3 * 150), from an independent replay of chain 0..2
    want = np.zeros((V, V))
    for c in range(3):
        ch = oracle.Chain(P, scorer=oracle.SCORER_CHOL)
        ch.init(5, c)
        ch.recorder_enable()
        ch.run(150)
        first = ch.edge_counts().copy()
        ch.run(150)
        want += ch.edge_counts() - first
    assert np.allclose(probs, want / 450.0, atol=1e-15)
    conv = res["conv_stats"]["parent_sets"]
    assert len(conv) == V and len(conv[0]) == V and len(conv[0][0]) == 2 and len(conv[0][0][0]) == 2
    assert len(res["conv_stats"]["lambda"]) == V and len(res["summary"]["lambda"]) == V
    from ssps_b200.postprocess import genie_edge_list
    rows = genie_edge_list(probs, ["n%d" % i for i in range(V)])
    assert len(rows) == V * V and all(rows[i][2] >= rows[i + 1][2] for i in range(len(rows) - 1))


def test_scores_match_sklearn_without_ties():
    from sklearn.metrics import average_precision_score, roc_auc_score
    rng = np.random.default_rng(1)
    truth = rng.random(500) < 0.1
    preds = rng.random(500) * 0.5 + truth * rng.random(500) * 0.5
    assert scoring.aucroc(preds, truth) == pytest.approx(roc_auc_score(truth, preds), rel=1e-12)
    assert scoring.average_precision(preds, truth) == pytest.approx(average_precision_score(truth, preds), rel=1e-12)
    assert 0.0 < scoring.aucpr(preds, truth) <= 1.0
    # perfect and inverted rankings
    assert scoring.aucroc([3, 2, 1, 0], [1, 1, 0, 0]) == 1.0 and scoring.aucroc([0, 1, 2, 3], [1, 1, 0, 0]) == 0.0
    assert scoring.aucpr([3, 2, 1, 0], [1, 1, 0, 0]) == 1.0
    # all tied: one step of the curve
    assert scoring.aucroc([1, 1, 1, 1], [1, 0, 1, 0]) == 0.5
