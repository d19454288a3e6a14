"""Drop-in for the t-map functions of the reference's
`analyses/decoding/searchlight/searchlight_permutation_testing.py` (`ttest_1samp_no_p` :213-242, `calc_t_value`
:245-251, `calc_image_t_values` :254-258), backed by `mmd_ttest_1samp` (`csrc/stats.cu`).  The TFCE side lives in
`analyses/cluster_analysis.py` like in the reference.  The file-system plumbing around them (per-subject pickles, h5py
null-distribution files, :102-180 and :349-556) belongs to the dataset and is not rebuilt here.
"""
from __future__ import annotations

import numpy as np


def _engine():
    from .searchlight import get_engine
    return get_engine()


def ttest_1samp_no_p(X, sigma=0, method="relative"):
    """One-sample t-values of the columns of X (subjects x vertices, or one vertex's subjects) against 0."""
    X = np.asarray(X, dtype=np.float64)
    one_d = X.ndim == 1
    data = X[:, None] if one_d else X
    if sigma > 0 and method == "relative" and not one_d:
        # the variance floor is sigma * max(var) over ALL columns (:238-240): one extra pass for the maximum
        var = data.shape[0] / np.square(_engine().ttest_1samp(data, 0.0, 0.0, constant_is_nan=False) / data.mean(0))
        t = _engine().ttest_1samp(data, 0.0, float(sigma) * np.nanmax(var), relative=False, constant_is_nan=False)
    else:
        t = _engine().ttest_1samp(data, 0.0, float(sigma), relative=(method == "relative"), constant_is_nan=False)
    return t[0] if one_d else t


def calc_t_value(values, popmean, sigma=0):
    values = np.asarray(values, dtype=np.float64)
    return _engine().ttest_1samp(values[:, None], popmean, float(sigma))[0]


def calc_image_t_values(data, popmean, use_tqdm=False, metric=None, sigma=0):
    """(subjects x vertices) -> (vertices,) t-values of `data - popmean`; NaN where all subjects agree."""
    return _engine().ttest_1samp(np.asarray(data, dtype=np.float64), popmean, float(sigma))
