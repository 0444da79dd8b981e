"""Edge-prediction scores of the reference's scripts/scoring.jl, restated for host-side checks of the
posterior the device produces: aucroc (:24-70), aucpr (:79-126), average_precision (:135-180).
Predictions = the V x V edge posterior [child][parent], truths = the true adjacency in the same layout.
Ties are handled the way the reference does: the curve only advances when the score changes.
"""
from __future__ import annotations

import numpy as np


def _sorted(preds, truths):
    p = np.asarray(preds, dtype=np.float64).reshape(-1)
    t = np.asarray(truths).reshape(-1).astype(bool)
    o = np.argsort(-p, kind="stable")        # sortperm(preds, order = Reverse) is stable too
    return p[o], t[o]


def aucroc(preds, truths) -> float:
    p, t = _sorted(preds, truths)
    N, npos = len(p), int(t.sum())
    nneg = N - npos
    area, tp, fp, prev_tp, prev_fp = 0.0, 0, 0, 0, 0
    for i in range(N):
        if i > 0 and p[i] != p[i - 1]:
            area += 0.5 * (prev_tp + tp) * (fp - prev_fp) / (npos * nneg)
            prev_tp, prev_fp = tp, fp
        if t[i]:
            tp += 1
        else:
            fp += 1
    return area + 0.5 * (prev_tp + tp) * (fp - prev_fp) / (npos * nneg)


def aucpr(preds, truths) -> float:
    p, t = _sorted(preds, truths)
    N, npos = len(p), int(t.sum())
    area, tp, fp, prev_p, prev_r = 0.0, 0, 0, 1.0, 0.0
    for i in range(N):
        if i > 0 and p[i] < p[i - 1]:
            recall, precision = tp / npos, tp / (tp + fp)
            area += 0.5 * (precision + prev_p) * (recall - prev_r)
            prev_r, prev_p = recall, precision
        if t[i]:
            tp += 1
        else:
            fp += 1
    recall, precision = tp / npos, tp / (tp + fp)
    return area + 0.5 * (precision + prev_p) * (recall - prev_r)


def average_precision(preds, truths) -> float:
    p, t = _sorted(preds, truths)
    N, npos = len(p), int(t.sum())
    s, tp, fp, prev_r = 0.0, 0, 0, 0.0
    for i in range(N):
        if i > 0 and p[i] < p[i - 1]:
            recall, precision = tp / npos, tp / (tp + fp)
            s += precision * (recall - prev_r)
            prev_r = recall
        if t[i]:
            tp += 1
        else:
            fp += 1
    recall, precision = tp / npos, tp / (tp + fp)
    return s + precision * (recall - prev_r)
