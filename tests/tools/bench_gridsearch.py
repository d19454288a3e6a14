"""BASELINE.json configs[0]: ridge_regression_decoding's alpha grid search on synthetic betas (1 subject, ~10k voxels,
4000 train trials, 1024-d features, DEFAULT_ALPHAS, 5-fold CV): the GPU drop-in next to sklearn's GridSearchCV on the
host cores (the reference's own code path, ridge_regression_decoding.py:111-117).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from multimodal_decoding_b200.analyses.decoding import ridge_regression_decoding as rr  # noqa: E402
from multimodal_decoding_b200 import eval as ev  # noqa: E402

n, v, d = 4000, 10000, 1024
rng = np.random.default_rng(0)
z = rng.standard_normal((n, 32))
X = (0.3 * z @ rng.standard_normal((32, v)) + rng.standard_normal((n, v))).astype(np.float32)
Y = (z @ rng.standard_normal((32, d)) + 0.5 * rng.standard_normal((n, d))).astype(np.float32)
alphas = rr.DEFAULT_ALPHAS

rr.grid_search_ridge_gpu(X[:500], Y[:500], alphas)                 # warm-up (workspaces)
t0 = time.perf_counter()
best, mean, model = rr.grid_search_ridge_gpu(X, Y, alphas)
gpu_s = time.perf_counter() - t0

from sklearn.linear_model import Ridge  # noqa: E402
from sklearn.metrics import make_scorer  # noqa: E402
from sklearn.model_selection import GridSearchCV  # noqa: E402
from oracle import searchlight_oracle as orc  # noqa: E402  (CPU scorer of the baseline leg)
scorer = make_scorer(lambda yt, yp: orc.pairwise_accuracy(yt, yp), greater_is_better=True)
t0 = time.perf_counter()
clf = GridSearchCV(estimator=Ridge(fit_intercept=False), param_grid=dict(alpha=alphas), scoring=scorer,
                   cv=rr.NUM_CV_SPLITS, n_jobs=int(os.environ.get("N_JOBS", "8")), refit=True)
clf.fit(X, Y)
cpu_s = time.perf_counter() - t0
print(json.dumps({
    "metric": "alpha grid search wall time (configs[0]: n=4000, v=10000, d=1024, 6 alphas x 5 folds + refit)",
    "gpu_s": round(gpu_s, 3), "cpu_sklearn_s": round(cpu_s, 2), "cpu_n_jobs": clf.n_jobs,
    "best_alpha_gpu": best, "best_alpha_sklearn": clf.best_params_["alpha"],
    "max_abs_diff_mean_test_score": float(np.abs(mean - clf.cv_results_["mean_test_score"]).max()),
    "note": "GPU: float64 on the library's own kernels (csrc/dense.cu: DMMA GEMM, blocked Cholesky, triangular solves; csrc/pairwise.cu); "
            "sklearn: float32 inputs as in the reference"}))
