"""Synthetic DBN data generator -- a restatement of the reference's scripts/simulate_data.jl.

Used to build the inputs of BASELINE.json configs 2/4/5 and of the metric point (V=100, T=8,
M=4) without Julia.  Follows simulate_data.jl:26-327 step by step:

* true graph: every (child, parent) pair, self-edges included, is an edge with probability
  p = 5/V (generate_random_digraph :26-29, julia_main :296);
* regression coefficients ~ N(0, coeff_std^2), coeff_std = 1/sqrt(V), one per monomial of the
  child's parents up to `regression_deg`, constant term first (generate_reg_coeffs :117-127,
  n_coeffs :105-107, compute_B_data :142-160 -- monomials in Combinatorics.combinations order);
* each series: x_1 ~ N(0, s^2 I), x_{t+1,i} = <monomials(x_t[parents_i]), coeffs_i> + N(0, s^2),
  s = 1/sqrt(T)/1000 (generate_time_series :191-209, julia_main :304);
* the PRIOR graph is the true graph with round(remove * E) edges deleted and round(add * E)
  non-edges inserted (modify_digraph! :67-100; the data come from the TRUE graph :250-252);
* everything is redrawn until max(data) < 100 (julia_main :306-318).

The RNG is numpy's PCG64 (Julia's unseeded MersenneTwister stream cannot be reproduced), so
datasets are statistically -- not bit-wise -- those of the reference generator.
"""
from __future__ import annotations

import itertools
import json
import math
from dataclasses import dataclass

import numpy as np


@dataclass
class SimulatedDataset:
    true_graph: np.ndarray      # V x V bool, [parent, child]
    prior_graph: np.ndarray     # V x V float 0/1, [parent, child]  (what save_graph writes)
    series: list                # M arrays of shape (T, V)
    V: int
    T: int

    def lagged(self):
        """combine_X (dbn_preprocess.jl:70-87): Xminus = rows 1..T-1, Xplus = rows 2..T."""
        xm = np.concatenate([s[:-1] for s in self.series], axis=0)
        xp = np.concatenate([s[1:] for s in self.series], axis=0)
        return np.asfortranarray(xm), np.asfortranarray(xp)


def n_coeffs(num_vars: int, max_degree: int) -> int:
    return sum(math.comb(num_vars, i) for i in range(0, min(num_vars, max_degree) + 1))


def _julia_round(x: float) -> int:
    return int(np.round(x))  # numpy rounds half to even, like Julia's round()


def generate_random_digraph(V: int, p: float, rng) -> np.ndarray:
    """parent_sets[i][j] -> returned as G[j, i] (parent-major rows, child columns)."""
    ps = rng.random((V, V)) < p          # ps[i, j]: j is a parent of i
    return ps.T.copy()


def modify_digraph(true_graph: np.ndarray, remove: float, add: float, rng) -> np.ndarray:
    V = true_graph.shape[0]
    ps = true_graph.T.copy()             # ps[i, j] child-major as in the reference loops
    edges = [(i, j) for i in range(V) for j in range(V) if ps[i, j]]
    non_edges = [(i, j) for i in range(V) for j in range(V) if not ps[i, j]]
    n_edges = len(edges)
    keep = _julia_round((1.0 - remove) * n_edges)
    perm = rng.permutation(n_edges)
    for idx in perm[: max(n_edges - keep, 0)]:
        i, j = edges[idx]
        ps[i, j] = False
    n_add = _julia_round(add * n_edges)
    perm = rng.permutation(len(non_edges))
    for idx in perm[: min(n_add, len(non_edges))]:
        i, j = non_edges[idx]
        ps[i, j] = True
    return ps.T.copy()


def _monomials(x_used: np.ndarray, deg: int) -> np.ndarray:
    out = [1.0]
    for d in range(1, deg + 1):
        for comb in itertools.combinations(range(len(x_used)), d):
            out.append(float(np.prod(x_used[list(comb)])))
    return np.array(out[: n_coeffs(len(x_used), deg)])


def generate_dataset(true_graph: np.ndarray, T: int, M: int, coeff_std: float, regression_deg: int,
                     regression_std: float, rng) -> list:
    V = true_graph.shape[0]
    parents = [np.flatnonzero(true_graph[:, i]) for i in range(V)]
    coeffs = [coeff_std * rng.standard_normal(n_coeffs(len(parents[i]), regression_deg)) for i in range(V)]
    # index tables for the monomials of each child (combinations order, degree-major)
    combos = []
    for i in range(V):
        k = len(parents[i])
        cl = []
        for d in range(1, min(regression_deg, k) + 1):
            cl.extend(itertools.combinations(range(k), d))
        combos.append(cl)
    series = []
    for _ in range(M):
        x = np.zeros((T, V))
        x[0] = regression_std * rng.standard_normal(V)
        for t in range(1, T):
            prev = x[t - 1]
            for i in range(V):
                used = prev[parents[i]]
                val = coeffs[i][0]
                for c, comb in enumerate(combos[i]):
                    term = 1.0
                    for a in comb:
                        term *= used[a]
                    val += coeffs[i][c + 1] * term
                x[t, i] = val + regression_std * rng.standard_normal()
        series.append(x)
    return series


def simulate(V: int, T: int, M: int, remove: float = 0.5, add: float = 0.5, regression_deg: int = 3,
             seed: int = 20200701) -> SimulatedDataset:
    """julia_main of simulate_data.jl (:289-327)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    p = 5.0 / V
    coeff_std = 1.0 / math.sqrt(V)
    regression_std = 1.0 / math.sqrt(T) / 1000.0
    while True:
        true_graph = generate_random_digraph(V, p, rng)
        prior = modify_digraph(true_graph, remove, add, rng)
        series = generate_dataset(true_graph, T, M, coeff_std, regression_deg, regression_std, rng)
        if max(float(s.max()) for s in series) < 100.0:
            break
    return SimulatedDataset(true_graph=true_graph, prior_graph=prior.astype(np.float64), series=series, V=V, T=T)


def save_graph(graph: np.ndarray, file_name: str) -> None:
    """save_graph (:277-283): V x V, comma separated, no header, entry [parent, child] as 0/1."""
    np.savetxt(file_name, np.asarray(graph).astype(np.int64), fmt="%d", delimiter=",")


def save_dataset(series: list, file_name: str) -> None:
    """save_dataset (:258-272): TAB separated, header timeseries/timestep/var###, 1-based ids."""
    V = series[0].shape[1]
    ndigs = len(str(V))
    with open(file_name, "w") as f:
        f.write("\t".join(["timeseries", "timestep"] + ["var" + str(i + 1).zfill(ndigs) for i in range(V)]) + "\n")
        for m, ts in enumerate(series):
            for t in range(ts.shape[0]):
                f.write("\t".join([str(m + 1), str(t + 1)] + [repr(float(v)) for v in ts[t]]) + "\n")


def dream_like(V: int = 46, M: int = 4, T: int = 6, seed: int = 8) -> SimulatedDataset:
    """HPN-DREAM8-shaped stand-in (config C3 when the challenge files cannot be preprocessed):
    same simulator, but a continuous prior U(0, 0.3) with 30 % zeros."""
    ds = simulate(V, T, M, seed=seed)
    rng = np.random.Generator(np.random.PCG64(seed + 1))
    conf = rng.uniform(0.0, 0.3, size=(V, V))
    conf[rng.random((V, V)) < 0.3] = 0.0
    conf[ds.true_graph & (rng.random((V, V)) < 0.5)] += 0.5
    ds.prior_graph = np.clip(conf, 0.0, 1.0)
    return ds


def main(argv=None) -> int:
    """CLI twin: V T N remove add regression_deg ref_dg_file true_dg_file timeseries_file [seed]."""
    import sys
    a = list(sys.argv[1:] if argv is None else argv)
    if len(a) < 9:
        print(main.__doc__)
        return 2
    V, T, N = int(a[0]), int(a[1]), int(a[2])
    ds = simulate(V, T, N, remove=float(a[3]), add=float(a[4]), regression_deg=int(a[5]),
                  seed=int(a[9]) if len(a) > 9 else 20200701)
    save_graph(ds.prior_graph, a[6])
    save_graph(ds.true_graph, a[7])
    save_dataset(ds.series, a[8])
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
