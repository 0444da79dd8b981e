"""File ingest for the reference's input formats (SSPS/src/dbn_preprocess.jl:12-67).

Only parsing lives here; lagging (combine_X), standardisation and the Gram matrix are built on the
device by ssps_set_timeseries / ssps_set_data.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def ingest_timeseries(path: str) -> Tuple[List[np.ndarray], List[str]]:
    """ingest_timeseries! (:12-25): TAB- (or comma-) separated table with header
    `timeseries, timestep, var...`; one matrix per unique `timeseries` value in order of first
    appearance, rows sorted by `timestep` (stable), data = columns 3..end."""
    with open(path, "r") as f:
        header = f.readline().rstrip("\r\n")
        sep = "\t" if "\t" in header else ","
        names = header.split(sep)
        if len(names) < 3:
            raise ValueError("%s: expected columns timeseries, timestep, <variables...>" % path)
        groups, order = {}, []
        for ln, line in enumerate(f, start=2):
            line = line.rstrip("\r\n")
            if not line:
                continue
            parts = line.split(sep)
            if len(parts) != len(names):
                raise ValueError("%s:%d: %d fields, header has %d" % (path, ln, len(parts), len(names)))
            key = parts[0]
            if key not in groups:
                groups[key] = []
                order.append(key)
            groups[key].append((float(parts[1]), [float(x) for x in parts[2:]]))
    series = []
    for key in order:
        rows = sorted(groups[key], key=lambda r: r[0])   # Python's sort is stable, like sort!(ts, [:timestep])
        series.append(np.array([r[1] for r in rows], dtype=np.float64))
    return series, names[2:]


def ingest_conf_adjacency(path: str) -> np.ndarray:
    """ingest_conf_adjacency! (:28-44): headerless comma-separated V x V matrix, entry
    [parent j, child i]; only entries > 0 carry a confidence (applied on the device)."""
    m = np.loadtxt(path, delimiter=",", dtype=np.float64, ndmin=2)
    if m.shape[0] != m.shape[1]:
        raise ValueError("%s: prior matrix must be square, got %s" % (path, m.shape))
    return m


def load_formatted_data(timeseries_data_path: str, reference_graph_path: str):
    """load_formatted_data (:57-67)."""
    series, var_names = ingest_timeseries(timeseries_data_path)
    prior = ingest_conf_adjacency(reference_graph_path)
    if prior.shape[0] != series[0].shape[1]:
        raise ValueError("prior is %s but the time series have %d variables" % (prior.shape, series[0].shape[1]))
    return series, prior, var_names


def combine_X(series: List[np.ndarray]):
    """combine_X (:70-87) on the host -- used by tests and by callers that already hold lagged
    matrices; the product path lags on the device (ssps_set_timeseries)."""
    xm = np.concatenate([s[:-1] for s in series], axis=0)
    xp = np.concatenate([s[1:] for s in series], axis=0)
    return np.asfortranarray(xm), np.asfortranarray(xp)
