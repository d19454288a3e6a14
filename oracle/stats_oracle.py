"""CPU ORACLE (test infrastructure, NOT product code) for the group-level statistics that consume the
searchlight score maps: one-sample t-maps over subjects and threshold-free cluster
enhancement (TFCE) on the cortical mesh.

Only `tests/` may import this module.  Parity status: **pinned against the reference itself** --
`tests/golden/make_golden.py::stats_case` runs the unmodified `ttest_1samp_no_p` / `calc_image_t_values`
(`analyses/decoding/searchlight/searchlight_permutation_testing.py:213-258`) and `calc_clusters` /
`calc_tfce_values` (`analyses/cluster_analysis.py:15-75, 248-312`) on seeded synthetic maps and commits
the outputs as `tests/golden/group_stats.npz`; `tests/test_oracle.py` asserts this restatement equals them.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
from scipy.sparse.csgraph import connected_components


def ttest_1samp_no_p(x, sigma: float = 0.0, method: str = "relative"):
    """searchlight_permutation_testing.py:213-242 (adapted there from mne): t = mean / sqrt(var_ddof1 / n) per
    column, with the optional "hat" variance adjustment."""
    x = np.asarray(x, dtype=np.float64)
    var = np.var(x, axis=0, ddof=1)
    if sigma > 0:
        var = var + (sigma * np.max(var) if method == "relative" else sigma)
    return np.mean(x, axis=0) / np.sqrt(var / x.shape[0])


def calc_image_t_values(data, popmean: float, sigma: float = 0.0):
    """searchlight_permutation_testing.py:245-258: per vertex (column of `data`, subjects x vertices) the t-value
    of `values - popmean`; NaN where all subjects have the same value (`calc_t_value`, :245-251).  The reference
    calls ttest_1samp_no_p per vertex, so the "relative" sigma limit is that vertex's own variance."""
    data = np.asarray(data, dtype=np.float64)
    out = np.full(data.shape[1], np.nan)
    for v in range(data.shape[1]):
        col = data[:, v]
        if np.all(col == col[0]):
            continue
        out[v] = ttest_1samp_no_p(col - popmean, sigma=sigma)
    return out


def clusters_at(values, edges, threshold):
    """cluster_analysis.py:15-75 (`calc_clusters`): connected components of the graph whose edges join two vertices
    with value >= threshold.  Returns (labels (V,) with -1 outside clusters, sizes per label)."""
    values = np.asarray(values, dtype=np.float64)
    with np.errstate(invalid="ignore"):
        act = values >= threshold                        # NaN compares False, as in the reference
    e = edges[act[edges[:, 0]] & act[edges[:, 1]]]
    v = len(values)
    labels = np.full(v, -1, dtype=np.int64)
    if len(e) == 0:
        return labels, np.zeros(0, dtype=np.int64)
    g = sp.coo_matrix((np.ones(len(e)), (e[:, 0], e[:, 1])), shape=(v, v))
    _, comp = connected_components(g, directed=False)
    in_cluster = np.zeros(v, dtype=bool)
    in_cluster[e.ravel()] = True                         # a cluster needs at least one edge
    ids, inv = np.unique(comp[in_cluster], return_inverse=True)
    labels[in_cluster] = inv
    return labels, np.bincount(inv, minlength=len(ids))


def tfce(values, edges, h: float = 2.0, e: float = 1.0, dh=0.1):
    """cluster_analysis.py:248-312 (`calc_tfce_values` for one hemisphere, cluster extent = number of vertices):
    for thresholds t = dh, 2dh, ... (np.arange(step, max + step, step)) every vertex of a cluster gains
    size**e * t**h * dh, a supra-threshold (strictly greater) vertex outside clusters gains t**h * dh; float32
    accumulator, thresholds in ascending order."""
    values = np.asarray(values, dtype=np.float64)
    out = np.zeros(values.shape, dtype=np.float32)
    max_score = np.nanmax(values) if not np.all(np.isnan(values)) else np.nan
    if np.isnan(max_score) or np.isinf(max_score) or max_score <= 0:
        return out
    step = max_score / 100 if dh in (0, "auto") else dh
    for t in np.arange(step, max_score + step, step):
        labels, sizes = clusters_at(values, edges, t)
        inside = labels >= 0
        term = (sizes.astype(np.float64) ** e) * (t ** h) * step
        out[inside] = (out[inside].astype(np.float64) + term[labels[inside]]).astype(np.float32)
        with np.errstate(invalid="ignore"):
            single = (values > t) & ~inside
        out[single] = (out[single].astype(np.float64) + (1 ** e) * (t ** h) * step).astype(np.float32)
    return out
