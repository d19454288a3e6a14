"""Drop-in for the cluster / TFCE functions of the reference's `analyses/cluster_analysis.py`, backed by the CUDA
kernels of `csrc/stats.cu` (`mmd_tfce`, `mmd_cluster_labels`).  Same names, argument meaning and return layouts
as `cluster_analysis.py:15-75` (`calc_clusters`), `:155-178` (`calc_significance_cutoff`) and `:248-312`
(`calc_tfce_values`); nothing here computes statistics on the CPU except the final quantile of the null maxima.

Only the reference's default cluster-extent measure ("num_vertices") is built on the GPU; "edge_lengths" raises.
"""
from __future__ import annotations

import numpy as np

from ..utils import HEMIS


def _engine():
    from .decoding.searchlight.searchlight import get_engine
    return get_engine()


def edges_array(edge_lengths: dict) -> np.ndarray:
    """The keys of one hemisphere's `edge_lengths` dict ((v0, v1) -> length) as an (E, 2) int32 array."""
    if isinstance(edge_lengths, np.ndarray):
        return np.ascontiguousarray(edge_lengths, dtype=np.int32).reshape(-1, 2)
    return np.array(list(edge_lengths.keys()), dtype=np.int32).reshape(-1, 2)


def calc_clusters(scores, threshold, edge_lengths=None, return_clusters=True, return_cluster_edge_lengths=False,
                  return_agg_t_values=False, return_cluster_map=False):
    """cluster_analysis.py:15-75: connected components of the vertices with score >= threshold (a cluster needs at
    least one edge).  `edge_lengths`: dict {(v0, v1): length} or an (E, 2) array of vertex pairs."""
    scores = np.asarray(scores, dtype=np.float64)
    edges = edges_array(edge_lengths)
    labels = _engine().cluster_labels(scores, edges, threshold)
    order = np.argsort(labels, kind="stable")
    order = order[labels[order] >= 0]
    bounds = np.flatnonzero(np.diff(labels[order])) + 1
    groups = np.split(order, bounds) if len(order) else []
    result = dict()
    if return_clusters:
        result["clusters"] = [set(g.tolist()) for g in groups]
    if return_cluster_map:
        cluster_map = np.zeros_like(scores)
        for g in groups:
            cluster_map[g] = sum(scores[n] for n in g.tolist())
        result["cluster_map"] = cluster_map
    if return_agg_t_values:
        result["agg_t_values"] = [sum(scores[n] for n in g.tolist()) for g in groups]
    if return_cluster_edge_lengths:
        if isinstance(edge_lengths, dict):
            lengths = np.fromiter(edge_lengths.values(), dtype=np.float64, count=len(edge_lengths))
        else:
            raise ValueError("return_cluster_edge_lengths needs the {(v0, v1): length} dict")
        inside = (labels[edges[:, 0]] >= 0) & (labels[edges[:, 0]] == labels[edges[:, 1]])
        ids = {int(labels[g[0]]): i for i, g in enumerate(groups)}
        sums = np.zeros(len(groups))
        np.add.at(sums, [ids[int(l)] for l in labels[edges[inside, 0]]], lengths[inside])
        result["cluster_edge_lengths"] = sums.tolist()
    return result


def calc_tfce_values(t_values, edge_lengths_dicts, metric, h=2, e=1, dh=0.1, cluster_extents_measure="num_vertices",
                     use_tqdm=False):
    """cluster_analysis.py:248-312: {hemi: {metric: float32 TFCE map}}.  `t_values[hemi][metric]` may also be a
    (n_maps, V) array (all maps of a null distribution in one launch per chunk)."""
    if cluster_extents_measure != "num_vertices":
        raise RuntimeError("Unknown cluster extents measure on the GPU path: ", cluster_extents_measure)
    tfce_values = dict()
    for hemi in HEMIS:
        values = np.asarray(t_values[hemi][metric], dtype=np.float64)
        if values.ndim == 1:
            max_score = np.nan if np.isnan(values).all() else np.nanmax(values)
            if np.isnan(max_score) or np.isinf(max_score) or max_score <= 0:
                # the reference's early exits (NaN-only, inf, non-positive maps): zeros like the input
                tfce_values[hemi] = {metric: np.zeros_like(values)}
                continue
        step = 0.0 if dh in (0, "auto") else float(dh)
        tfce_values[hemi] = {metric: _engine().tfce(values, edges_array(edge_lengths_dicts[hemi]), h, e, step)}
    return tfce_values


def calc_significance_cutoff(null_distribution_tfce_values, metric, p_value_threshold=0.05):
    """cluster_analysis.py:155-178: quantile of the per-permutation maxima over both hemispheres."""
    null_distr = np.sort([
        np.nanmax(np.concatenate((n[HEMIS[0]][metric], n[HEMIS[1]][metric])))
        for n in null_distribution_tfce_values
    ])
    if p_value_threshold == 1 / len(null_distribution_tfce_values):
        significance_cutoff = np.max(null_distr)
    else:
        significance_cutoff = np.quantile(null_distr, 1 - p_value_threshold, method="closest_observation")
    return significance_cutoff, null_distr
