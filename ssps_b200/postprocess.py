"""Pooled edge posterior / convergence summary in the format of the reference's
scripts/postprocess_samples.jl (:212-233), produced directly from the device tallies so that
scoring.jl / score_dream.py / json_to_genie.py consume many-chain runs without per-chain JSON.

    {"parent_sets": V x V probabilities [child][parent], "edge_conf_key": "parent_sets",
     "conv_stats": {"parent_sets": [[ [[psrf, n_eff]] per parent ] per child ], "lambda": [...]},
     "summary": {"lambda": [...]}, "stop_points": [n], "t_elapsed": [...], "n": [...]}

PSRF follows compute_psrf_neff (:99-138) incl. its 1/M between-chain variance (quirk Q6); NaN
(constant indicator, W = 0) is written as null exactly like JSON.print does.  N_eff needs lagged
autocorrelations of the dense sequences (:58-94), which the tallies do not hold: it is written as
null here and is available per chain from the change-point JSON through the reference's own script.
"""
from __future__ import annotations

import math


def _num(x):
    x = float(x)
    return x if math.isfinite(x) else None


def summarize_sampler(sampler, burnin: float = 0.5, t_elapsed: float = 0.0) -> dict:
    V = sampler.V
    probs, lam_mean, b_used, n_used = sampler.edge_posterior(burnin)
    try:
        mx, edge_psrf, lam_psrf = sampler.psrf_finish(sampler.psrf_sums(burnin))
    except Exception:            # not enough snapshots yet: PSRF undefined
        mx, edge_psrf, lam_psrf = float("nan"), None, None
    conv_ps = [[[[_num(edge_psrf[i][j]) if edge_psrf is not None else None, None]] for j in range(V)] for i in range(V)]
    conv_lam = [[[_num(lam_psrf[i]) if lam_psrf is not None else None, None]] for i in range(V)]
    return {
        "parent_sets": [[float(probs[i][j]) for j in range(V)] for i in range(V)],
        "edge_conf_key": "parent_sets",
        "conv_stats": {"parent_sets": conv_ps, "lambda": conv_lam},
        "summary": {"lambda": [float(x) for x in lam_mean]},
        "stop_points": [int(n_used)],
        "burnin_index": int(b_used),
        "max_edge_psrf": _num(mx),
        "t_elapsed": [float(t_elapsed)] * sampler.n_chains,
        "n": [int(n_used)] * sampler.n_chains,
    }


def genie_edge_list(probs_child_major, node_names):
    """json_to_genie.py:9-30 restated without pandas' removed DataFrame.append: rows (node1 = parent,
    node2 = child, score), all V^2 pairs, sorted by score descending (stable, like sort_values)."""
    rows = []
    for i, child in enumerate(node_names):
        for j, parent in enumerate(node_names):
            rows.append((parent, child, float(probs_child_major[i][j])))
    rows.sort(key=lambda r: -r[2])
    return rows


def write_genie(path, probs_child_major, node_names):
    with open(path, "w") as f:
        f.write("node1,node2,score\n")
        for a, b, s in genie_edge_list(probs_child_major, node_names):
            f.write("%s,%s,%r\n" % (a, b, s))
