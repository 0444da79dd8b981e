"""Pooled edge posterior / convergence summary in the format of the reference's
scripts/postprocess_samples.jl (:212-233), produced directly on the device so that scoring.jl / score_dream.py /
json_to_genie.py consume many-chain runs without per-chain JSON files.  Exactly the reference's keys:

    {"parent_sets": V x V probabilities [child][parent], "edge_conf_key": "parent_sets",
     "conv_stats": {"parent_sets": [[ [[psrf, n_eff] per stop point] per parent ] per child ], "lambda": [...]},
     "summary": {"lambda": [...]}, "stop_points": [...], "t_elapsed": [...], "n": [...]}

Two sources:
  * the change-point log (record_changepoints = 1): ssps_convergence (K5, csrc/ssps_neff.cuh) evaluates compute_psrf_neff
    (:99-138) and correlation_sum (:58-91) per edge and per lambda_i at the reference's EXACT burn-in / split indices,
    N_eff included -- what tests/test_gpu_summary.py compares with this module's host twin of the reference's script;
  * the tallies only (thousands of chains, no log): PSRF from the snapshot-aligned tallies (K4), N_eff = null, and the
    indices actually used are reported under the extension key "ssps_b200" (a consumer of the reference's format ignores it).
PSRF follows the reference incl. its 1/M between-chain variance (quirk Q6); NaN (constant indicator, W = 0) is written
as null exactly like JSON.print does.
"""
from __future__ import annotations

import math
import warnings


def _num(x):
    x = float(x)
    return x if math.isfinite(x) else None


def _has_log(sampler) -> bool:
    if hasattr(sampler, "record_changepoints"):
        return bool(sampler.record_changepoints)
    return bool(sampler.cfg.record_changepoints)


def _n_recorded(sampler) -> int:
    return int(sampler.samples_recorded()) if hasattr(sampler, "samples_recorded") else int(sampler.stats().samples_recorded)


def summarize_sampler(sampler, burnin: float = 0.5, t_elapsed: float = 0.0, stop_points=()) -> dict:
    """collect_stats (:212-233) for the chains of a Sampler / MultiGpuSampler, computed on the device."""
    from ._lib import ERR_UNAVAILABLE, SspsError
    V, n = sampler.V, _n_recorded(sampler)
    stops = [int(sp) for sp in stop_points if sp <= n] if len(stop_points) else [n]
    if _has_log(sampler):
        full = sampler.convergence(burnin, n)                    # evaluate_summary ignores the stop points (:183-191)
        per_stop = [full if sp == n else sampler.convergence(burnin, sp) for sp in stops]
        conv_ps = [[[[_num(c["edge_psrf"][i][j]), _num(c["edge_neff"][i][j])] for c in per_stop] for j in range(V)] for i in range(V)]
        conv_lam = [[[_num(c["lambda_psrf"][i]), _num(c["lambda_neff"][i])] for c in per_stop] for i in range(V)]
        probs, lam_mean = full["edge_mean"], full["lambda_mean"]
        ext = {"source": "change-point log (K5)", "burnin_index": [int(c["burnin_index"]) for c in per_stop],
               "split_index": [int(c["split_index"]) for c in per_stop]}
    else:
        out = sampler.edge_posterior(burnin)
        probs, lam_mean, b_used, n_used = out[0], out[1], out[2], out[3]
        want = int(math.floor(burnin * n_used))
        if b_used != want:
            warnings.warn("edge posterior: burn-in index %d instead of floor(%g * %d) = %d (tally snapshots lie on a grid; "
                          "lower psrf_grid or keep the change-point log for the exact index)" % (b_used, burnin, n_used, want))
        try:
            res = sampler.psrf(burnin) if not hasattr(sampler, "psrf_sums") else sampler.psrf_finish(sampler.psrf_sums(burnin))
            mx, edge_psrf, lam_psrf = res[0], res[1], (res[2] if len(res) > 2 else None)
        except SspsError as e:       # not enough snapshots yet: PSRF undefined.  Anything else (CUDA, memory) is an error
            if e.code != ERR_UNAVAILABLE:
                raise
            mx, edge_psrf, lam_psrf = float("nan"), None, None
        stops = [int(n_used)]
        conv_ps = [[[[_num(edge_psrf[i][j]) if edge_psrf is not None else None, None]] for j in range(V)] for i in range(V)]
        conv_lam = [[[_num(lam_psrf[i]) if lam_psrf is not None else None, None]] for i in range(V)]
        ext = {"source": "tallies (K4): N_eff needs the change-point log", "burnin_index": [int(b_used)], "max_edge_psrf": _num(mx)}
    return {
        "stop_points": stops,
        "summary": {"lambda": [float(x) for x in lam_mean]},
        "conv_stats": {"parent_sets": conv_ps, "lambda": conv_lam},
        "parent_sets": [[float(probs[i][j]) for j in range(V)] for i in range(V)],
        "edge_conf_key": "parent_sets",
        "t_elapsed": [float(t_elapsed)] * sampler.n_chains,
        "n": [int(n)] * sampler.n_chains,
        "ssps_b200": ext,
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


# ---------------------------------------------------------------------------------------------------
# Twin of scripts/postprocess_samples.jl for per-chain sample files (the reference's or ours):
#     python -m ssps_b200.postprocess --chain-samples c1.json c2.json ... --output-file out.json
#                                     [--stop-points n1 n2 ...] [--burnin 0.5]
# Restates load_samples (:23-35), ChangepointVec (changepoint_vec.jl), compute_psrf_neff (:99-138),
# correlation_sum (:58-91), evaluate_convergence (:155-180), evaluate_summary (:183-191),
# collect_stats (:212-233) and postprocess_sample_files (:242-259), on dense sequences.
# ---------------------------------------------------------------------------------------------------
import numpy as _np


def load_samples(path: str) -> dict:
    """load_samples: parent_sets[i] becomes a list over ALL V parents (missing key -> no change points)."""
    import json
    with open(path) as f:
        d = json.load(f)
    V = len(d["parent_sets"])
    d["parent_sets"] = [[ps.get(str(p), []) for p in range(1, V + 1)] for ps in d["parent_sets"]]
    return d


def dense_sequence(changepoints, n: int) -> _np.ndarray:
    """ChangepointVec(changepoints, n) as a dense vector over positions 1..n: the value at t is that of the
    last change point with index <= t (index 0 = the initial state), default 0 (changepoint_vec.jl getindex)."""
    out = _np.zeros(n, dtype=_np.float64)
    for a, (t, v) in enumerate(changepoints):
        t_next = changepoints[a + 1][0] if a + 1 < len(changepoints) else n + 1
        lo, hi = max(int(t), 1), min(int(t_next) - 1, n)
        if hi >= lo:
            out[lo - 1:hi] = v
    return out


def _is_binary(changepoints) -> bool:
    return all(v == 0 or v == 1 for _, v in changepoints)


def correlation_sum(seqs, varplus: float, is_binary: bool) -> float:
    """correlation_sum (:58-91) incl. its accumulation rule: at lag t > 2 the correlation of lag t - 2 is
    added while corr(t) + corr(t - 1) >= 0.  seqs: list of 1-D arrays (lags run to len(seqs[0]) - 2; every
    sequence's variogram uses its own length, as seq_variogram does)."""
    m = len(seqs)
    s, prev2, prev1 = 0.0, 0.0, 0.0
    for t in range(1, len(seqs[0]) - 1):
        comb = 0.0
        for q in seqs:
            nq = len(q)
            if is_binary:
                comb += float((q[:nq - t] != q[t:]).sum()) / (nq - t)
            else:
                dq = q[:nq - t] - q[t:]
                comb += float((dq * dq).sum()) / (nq - t)
        comb /= m
        with _np.errstate(divide="ignore", invalid="ignore"):
            corr = float(1.0 - _np.float64(comb) / _np.float64(2.0 * varplus))
        if t > 2:
            if corr + prev1 >= 0.0:
                s += prev2
            else:
                break
        prev2, prev1 = prev1, corr
    return s


def compute_psrf_neff(half_seqs, is_binary: bool):
    """compute_psrf_neff (:99-138): half_seqs = list of 1-D arrays (m half-chains); n = len(half_seqs[0]);
    B uses 1/m (quirk Q6)."""
    m, n = len(half_seqs), len(half_seqs[0])
    means, variances = _np.zeros(m), _np.zeros(m)
    for a, q in enumerate(half_seqs):
        N = len(q)
        if is_binary:
            means[a] = float((q != 0).sum()) / N
            variances[a] = means[a] * (1.0 - means[a]) * N / (N - 1)
        else:
            means[a] = float(q.sum()) / N
            variances[a] = float(((q - means[a]) ** 2).sum()) / (N - 1)
    B = n * float(((means - means.sum() / m) ** 2).sum()) / m
    W = float(variances.sum()) / m
    vp = (n - 1) * W / n + B / n
    with _np.errstate(divide="ignore", invalid="ignore"):
        psrf = float(_np.sqrt(_np.float64(vp) / _np.float64(W)))
    corr_sum = correlation_sum(half_seqs, vp, is_binary)
    n_eff = m * n / (1.0 + 2.0 * corr_sum)
    return psrf, n_eff


def evaluate_convergence(whole_seqs, ns, stop_points, burnin=0.5, is_binary=None):
    """evaluate_convergence (:155-180): whole_seqs = change-point lists of ONE quantity, one per chain."""
    if is_binary is None:
        is_binary = _is_binary(whole_seqs[0])
    dense = [dense_sequence(cp, n) for cp, n in zip(whole_seqs, ns)]
    out = []
    for ln in stop_points:
        b = int(math.floor(burnin * ln))
        sp = b + (ln - b) // 2
        halves = []
        for d in dense:
            halves.append(d[b:sp]); halves.append(d[sp:ln])
        out.append(compute_psrf_neff(halves, is_binary))
    return out


def evaluate_summary(whole_seqs, ns, burnin=0.5) -> float:
    """evaluate_summary (:183-191): pooled mean after each chain's own burn-in (stop points are ignored there too)."""
    tot, cnt = 0.0, 0
    for cp, n in zip(whole_seqs, ns):
        b = int(math.floor(burnin * n))
        d = dense_sequence(cp, n)
        tot += float(d[b:].sum()); cnt += n - b
    return tot / cnt


def collect_stats(dict_vec, stop_points, burnin=0.5) -> dict:
    """collect_stats (:212-233): same tree as the reference's output JSON."""
    V = len(dict_vec[0]["parent_sets"])
    ns = [int(d["n"]) for d in dict_vec]

    def conv(seqs):
        return [[_num(p), _num(e)] for p, e in evaluate_convergence(seqs, ns, stop_points, burnin)]

    ps_sum = [[evaluate_summary([d["parent_sets"][i][j] for d in dict_vec], ns, burnin) for j in range(V)] for i in range(V)]
    ps_conv = [[conv([d["parent_sets"][i][j] for d in dict_vec]) for j in range(V)] for i in range(V)]
    lam_sum = [evaluate_summary([d["lambda"][i] for d in dict_vec], ns, burnin) for i in range(len(dict_vec[0]["lambda"]))]
    lam_conv = [conv([d["lambda"][i] for d in dict_vec]) for i in range(len(dict_vec[0]["lambda"]))]
    return {
        "stop_points": list(stop_points),
        "summary": {"lambda": lam_sum},
        "conv_stats": {"parent_sets": ps_conv, "lambda": lam_conv},
        "parent_sets": ps_sum,
        "edge_conf_key": "parent_sets",
        "t_elapsed": [d["t_elapsed"] for d in dict_vec],
        "n": ns,
    }


def postprocess_sample_files(sample_filenames, output_file, stop_points=(), burnin=0.5) -> dict:
    """postprocess_sample_files (:242-259)."""
    import json
    dict_vec = [load_samples(f) for f in sample_filenames]
    n_max = min(int(d["n"]) for d in dict_vec)
    stop_points = [n_max] if len(stop_points) == 0 else [sp for sp in stop_points if sp <= n_max]
    results = collect_stats(dict_vec, stop_points, burnin)
    with open(output_file, "w") as f:
        json.dump(results, f)
    print("Saved postprocessed MCMC results to", output_file)
    return results


def main(argv=None) -> int:
    import argparse
    ap = argparse.ArgumentParser(prog="postprocess_samples")
    ap.add_argument("--chain-samples", nargs="+", required=True, help="mcmc sample files for one or more chains")
    ap.add_argument("--stop-points", nargs="+", type=int, default=[], help="points in the sequence where convergence should be analyzed.")
    ap.add_argument("--burnin", type=float, default=0.5, help="fraction of sequence to discard as burnin.")
    ap.add_argument("--output-file", required=True, help="name of output JSON file containing convergence information")
    a = ap.parse_args(argv)
    postprocess_sample_files(a.chain_samples, a.output_file, a.stop_points, a.burnin)
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
