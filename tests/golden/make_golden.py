"""Generate tests/golden/example_c1.npz from the reference's shipped example (config C1).

Run HERE (the container that has /root/reference); the GPU box only sees the committed .npz.
The reference's input files are parsed with pandas exactly as dbn_preprocess.jl:12-87 describes
(series = unique `timeseries` values in order of appearance, rows sorted by `timestep`, data =
columns 3..end; prior CSV has no header, entry [parent j, child i]).  Also stores the edge
scores of tests/baseline_predictions.genie (one unseeded 4x60 s reference run: a loose
Monte-Carlo sanity target, NOT a golden vector -- the reference's test ignores the score column,
tests/ssps_csv.sh:11) as a V x V [child, parent] matrix.
"""
import json
import os
import sys

import numpy as np
import pandas as pd

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "example_c1.npz")


def main():
    ts = pd.read_csv(os.path.join(REF, "run_ssps/example_timeseries.csv"), sep="\t", float_precision="round_trip")
    prior = pd.read_csv(os.path.join(REF, "run_ssps/example_prior.csv"), header=None, float_precision="round_trip").to_numpy(dtype=np.float64)
    names = json.load(open(os.path.join(REF, "run_ssps/example_node_names.json")))
    var_names = list(ts.columns[2:])
    series_names, series = [], []
    for name in pd.unique(ts["timeseries"]):
        sub = ts[ts["timeseries"] == name].sort_values("timestep", kind="stable")
        series_names.append(str(name))
        series.append(sub.iloc[:, 2:].to_numpy(dtype=np.float64))
    xm = np.concatenate([s[:-1] for s in series], axis=0)
    xp = np.concatenate([s[1:] for s in series], axis=0)
    V = xm.shape[1]
    genie = pd.read_csv(os.path.join(REF, "tests/baseline_predictions.genie"))
    idx = {n: i for i, n in enumerate(names)}
    scores = np.zeros((V, V))
    for a, b, s in zip(genie["node1"], genie["node2"], genie["score"]):
        scores[idx[b], idx[a]] = s  # node1 = parent, node2 = child (json_to_genie.py:17-22)
    np.savez_compressed(
        OUT,
        raw_table=ts.iloc[:, 2:].to_numpy(dtype=np.float64),
        raw_series=np.array([str(x) for x in ts["timeseries"]]),
        raw_timestep=ts["timestep"].to_numpy(dtype=np.float64),
        var_names=np.array(var_names), node_names=np.array(names),
        series_lengths=np.array([s.shape[0] for s in series]),
        Xminus=xm, Xplus=xp, prior=prior, genie_scores=scores,
        # the key columns of the reference's baseline file, in file order: what tests/ssps_csv.sh diffs (csvdiff on node1,node2)
        genie_node1=np.array([str(x) for x in genie["node1"]]), genie_node2=np.array([str(x) for x in genie["node2"]]))
    print("wrote", OUT, "V=%d n=%d series=%s" % (V, xm.shape[0], [s.shape[0] for s in series]))


if __name__ == "__main__":
    sys.exit(main())
