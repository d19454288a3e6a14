"""CPU ORACLE (test infrastructure, NOT product code) for the searchlight ridge-decoding path.

A float64 numpy/scipy restatement of what the reference computes per searchlight centre.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import this
module; the product (`multimodal_decoding_b200`) never does.

Parity status: **pinned against the reference itself**.  The reference has no tests or golden
vectors, so `tests/golden/make_golden.py` imports the unmodified reference from
`/root/reference` (with stub modules for nilearn/nibabel/matplotlib) in the build container, runs
its `custom_search_light`, `eval.*` and sklearn k-NN call on seeded synthetic inputs, and commits
the outputs under `tests/golden/`; `tests/test_oracle.py` asserts this file reproduces them.

The arithmetic the reference delegates to un-vendored third-party packages is restated from
their documented algorithms (reference pins: scikit-learn 1.0.2, scipy 1.11.4, numpy 1.26.2 —
`environment_files/environment_gpu.yml:85,176,232,233`; this container: 1.9.0 / 1.18.1 / 2.3.5):

* `sklearn.linear_model.Ridge(alpha, fit_intercept=False).fit` with k <= N_train: inputs are
  up-cast to float64, `coef = solve(XtX + alpha*I, XtY)` by Cholesky (`_ridge.py` `_solve_cholesky`;
  call site `searchlight.py:56`, estimator built at `searchlight.py:240`).
* `sklearn.preprocessing.StandardScaler().fit_transform`: per-feature (x - mean) / std with
  ddof=0 and scale 1 for zero-variance features (call site `eval.py:118`).
* `scipy.spatial.distance.cdist(a, b, "cosine")`: 1 - a.b / (|a| |b|) (call site `eval.py:104`).
* `sklearn.neighbors.NearestNeighbors().kneighbors`: exact Euclidean k-NN, ascending distance,
  the query point itself first (call site `searchlight.py:173`).
* `NearestNeighbors(radius=r).radius_neighbors_graph(coords)`: every point within Euclidean distance r,
  tested on the squared distance (call site `searchlight.py:164-165`).
* `sklearn.model_selection.KFold(5)` folds and GridSearchCV's first-best selection
  (call site `ridge_regression_decoding.py:111-117`).
"""
from __future__ import annotations

import numpy as np
import scipy.linalg

SCORE_KEYS = (
    "pairwise_acc_captions",                       # eval.py:11
    "pairwise_acc_images",                         # eval.py:12
    "pairwise_acc_cross_captions_to_images",       # eval.py:15
    "pairwise_acc_cross_images_to_captions",       # eval.py:14
    "pairwise_acc_imagery",                        # eval.py:42
    "pairwise_acc_imagery_whole_test_set",         # eval.py:43
    "pairwise_acc_imagery_no_std",                 # searchlight.py:87
    "pairwise_acc_imagery_whole_test_set_no_std",  # searchlight.py:87
)


# ------------------------------------------------------------------------------------------
# scoring (eval.py:103-181)
# ------------------------------------------------------------------------------------------
def zscore_rows_population(p: np.ndarray) -> np.ndarray:
    """StandardScaler().fit_transform(p) (eval.py:118): per column, over the rows of p."""
    p = np.asarray(p, dtype=np.float64)
    mu = p.mean(axis=0)
    sd = p.std(axis=0)
    sd = np.where(sd == 0.0, 1.0, sd)
    return (p - mu) / sd


def cosine_distance_matrix(pred: np.ndarray, cand: np.ndarray) -> np.ndarray:
    """cdist(pred, cand, 'cosine') (eval.py:103-104): rows = predictions, cols = candidates."""
    pred = np.asarray(pred, dtype=np.float64)
    cand = np.asarray(cand, dtype=np.float64)
    pn = np.sqrt((pred * pred).sum(1))
    cn = np.sqrt((cand * cand).sum(1))
    return 1.0 - (pred @ cand.T) / (pn[:, None] * cn[None, :])


def pairwise_count(dist: np.ndarray):
    """(#strictly-correct comparisons, #comparisons) of eval.py:107-113.

    The diagonal holds the first min(rows, cols) entries; every entry of row i is compared
    with dist[i, i] using a strict '<', the diagonal itself can never count.
    """
    diag = np.diagonal(dist)
    correct = int((diag[:, None] < dist[: len(diag)]).sum())
    return correct, dist.size - len(diag)


def pairwise_accuracy(latents, predictions, standardize_predictions=False):
    """eval.py:116-123 for metric='cosine', standardize_latents=False."""
    p = zscore_rows_population(predictions) if standardize_predictions else predictions
    c, n = pairwise_count(cosine_distance_matrix(p, latents))
    return c / n


def score_centre(y_test, pred_test, y_im, pred_im, is_caption) -> dict:
    """The eight scores one centre emits (searchlight.py:81-90, eval.py:126-181)."""
    cap, img = is_caption, ~is_caption
    out = {
        SCORE_KEYS[0]: pairwise_accuracy(y_test[cap], pred_test[cap], True),
        SCORE_KEYS[1]: pairwise_accuracy(y_test[img], pred_test[img], True),
        SCORE_KEYS[2]: pairwise_accuracy(y_test[img], pred_test[cap], True),
        SCORE_KEYS[3]: pairwise_accuracy(y_test[cap], pred_test[img], True),
    }
    whole = np.concatenate([y_im, y_test])                      # eval.py:176
    out[SCORE_KEYS[4]] = pairwise_accuracy(y_im, pred_im, True)
    out[SCORE_KEYS[5]] = pairwise_accuracy(whole, pred_im, True)
    out[SCORE_KEYS[6]] = pairwise_accuracy(y_im, pred_im, False)
    out[SCORE_KEYS[7]] = pairwise_accuracy(whole, pred_im, False)
    return out


def score_margins(y_test, pred_test, y_im, pred_im, is_caption) -> np.ndarray:
    """Smallest |diag - off-diag| gap behind each of the 8 scores (a count can only differ
    between two implementations where this gap is at rounding level)."""
    cap, img = is_caption, ~is_caption
    whole = np.concatenate([y_im, y_test])
    cases = [(y_test[cap], pred_test[cap], True), (y_test[img], pred_test[img], True),
             (y_test[img], pred_test[cap], True), (y_test[cap], pred_test[img], True),
             (y_im, pred_im, True), (whole, pred_im, True), (y_im, pred_im, False), (whole, pred_im, False)]
    res = []
    for lat, pr, std in cases:
        d = cosine_distance_matrix(zscore_rows_population(pr) if std else pr, lat)
        n = min(d.shape)
        gap = np.abs(d[:n] - np.diagonal(d)[:, None])
        gap[np.arange(n), np.arange(n)] = np.inf
        res.append(gap.min())
    return np.array(res)


# ------------------------------------------------------------------------------------------
# ridge fit / predict (searchlight.py:50-59 -> sklearn Ridge, cholesky solver)
# ------------------------------------------------------------------------------------------
def ridge_fit(x_train, y_train, alpha: float) -> np.ndarray:
    """Returns W (k x D) with predictions = X @ W; float64 throughout like sklearn."""
    x = np.asarray(x_train, dtype=np.float64)
    y = np.asarray(y_train, dtype=np.float64)
    n, k = x.shape
    if k <= n:                                    # primal normal equations
        a = x.T @ x
        a[np.diag_indices_from(a)] += alpha
        return scipy.linalg.solve(a, x.T @ y, assume_a="pos", overwrite_a=True)
    kmat = x @ x.T                                # dual form when features outnumber samples
    kmat[np.diag_indices_from(kmat)] += alpha
    return x.T @ scipy.linalg.solve(kmat, y, assume_a="pos", overwrite_a=True)


def train_and_test(x_nbh, y, train_ids, test_ids, imagery_ids, alpha, is_caption, return_preds=False):
    """One centre on its already-gathered neighbourhood columns (searchlight.py:38-90)."""
    w = ridge_fit(x_nbh[train_ids], y[train_ids], alpha)
    x64 = np.asarray(x_nbh, dtype=np.float64)
    y64 = np.asarray(y, dtype=np.float64)
    pred_test = x64[test_ids] @ w
    pred_im = x64[imagery_ids] @ w
    scores = score_centre(y64[test_ids], pred_test, y64[imagery_ids], pred_im, is_caption)
    if return_preds:
        return scores, pred_test, pred_im
    return scores


def default_is_caption(n_test: int) -> np.ndarray:
    """TEST_STIM_TYPES (data.py:182-187): first half captions, second half images."""
    m = np.zeros(n_test, dtype=bool)
    m[: n_test // 2] = True
    return m


def search_light(X, y, A, train_ids, test_ids, imagery_ids, alpha=1.0, is_caption=None,
                 centres=None) -> np.ndarray:
    """All centres (searchlight.py:93-149): returns an (n_centres, 8) float64 array in SCORE_KEYS order."""
    if is_caption is None:
        is_caption = default_is_caption(len(test_ids))
    train_ids, test_ids, imagery_ids = (np.asarray(i, dtype=np.int64) for i in (train_ids, test_ids, imagery_ids))
    centres = range(len(A)) if centres is None else centres
    out = np.empty((len(centres), len(SCORE_KEYS)))
    for r, c in enumerate(centres):
        s = train_and_test(X[:, np.asarray(A[c])], y, train_ids, test_ids, imagery_ids, alpha, is_caption)
        out[r] = [s[key] for key in SCORE_KEYS]
    return out


# ------------------------------------------------------------------------------------------
# adjacency (searchlight.py:152-178, k-NN branch)
# ------------------------------------------------------------------------------------------
def knn_adjacency(coords, k: int, block: int = 2048):
    """Exact k nearest neighbours of every point among all points, self first.

    Brute force in float64, distances squared as ((dx*dx + dy*dy) + dz*dz), stable ordering
    by (distance, index): reproduces sklearn's kd-tree result whenever distances are tie-free
    (SURVEY.md §8c).  Returns (distances (V,k) float64, indices (V,k) int64).
    """
    c = np.asarray(coords, dtype=np.float64)
    v = len(c)
    idx = np.empty((v, k), dtype=np.int64)
    dist = np.empty((v, k), dtype=np.float64)
    for s in range(0, v, block):
        q = c[s:s + block]
        dx = q[:, None, 0] - c[None, :, 0]
        dy = q[:, None, 1] - c[None, :, 1]
        dz = q[:, None, 2] - c[None, :, 2]
        d2 = (dx * dx + dy * dy) + dz * dz
        order = np.argsort(d2, axis=1, kind="stable")[:, :k]
        idx[s:s + block] = order
        dist[s:s + block] = np.sqrt(np.take_along_axis(d2, order, axis=1))
    return dist, idx


# ------------------------------------------------------------------------------------------
# null distribution (searchlight.py:61-79, data.py:540-561)
# ------------------------------------------------------------------------------------------
def radius_adjacency(coords, radius: float, block: int = 2048):
    """Radius branch of get_adjacency_matrix (searchlight.py:160-166): for every point the ascending indices of all
    points (itself included) with squared distance ((dx*dx + dy*dy) + dz*dz) <= radius*radius -- the kd-tree's
    reduced-distance test of sklearn's `radius_neighbors_graph`.  Returns a list of int64 arrays."""
    c = np.asarray(coords, dtype=np.float64)
    r2 = float(radius) * float(radius)
    out = []
    for s in range(0, len(c), block):
        q = c[s:s + block]
        dx = q[:, None, 0] - c[None, :, 0]
        dy = q[:, None, 1] - c[None, :, 1]
        dz = q[:, None, 2] - c[None, :, 2]
        inside = ((dx * dx + dy * dy) + dz * dz) <= r2
        out.extend(np.flatnonzero(row).astype(np.int64) for row in inside)
    return out


def shuffled_test_indices(n_permutations: int, n_test: int = 140):
    """create_null_distr_shuffled_indices (data.py:540-561): legacy numpy RandomState seeds
    0,1,2,...; each half permuted on its own; draws with any fixed point are skipped."""
    half = n_test // 2
    out, seed = [], 0
    while len(out) < n_permutations:
        rs = np.random.RandomState(seed)
        first = rs.choice(np.arange(half), size=half, replace=False)
        second = rs.choice(np.arange(half, n_test), size=half, replace=False)
        perm = np.concatenate([first, second])
        seed += 1
        if not np.any(perm == np.arange(n_test)):
            out.append(perm)
    return out


def null_scores_centre(y_test, pred_test, y_im, pred_im, is_caption, perms, imagery_perms):
    """The per-permutation score dicts of searchlight.py:61-77 (no cross-decoding scores).
    `imagery_perms[p]` is the 3-derangement used for permutation p (the reference draws it
    unseeded, searchlight.py:65; callers pass it in so results are reproducible)."""
    cap, img = is_caption, ~is_caption
    rows = []
    for perm, imp in zip(perms, imagery_perms):
        yt, yi = y_test[perm], y_im[np.asarray(imp)]
        whole = np.concatenate([yi, yt])
        rows.append({
            SCORE_KEYS[0]: pairwise_accuracy(yt[cap], pred_test[cap], True),
            SCORE_KEYS[1]: pairwise_accuracy(yt[img], pred_test[img], True),
            SCORE_KEYS[4]: pairwise_accuracy(yi, pred_im, True),
            SCORE_KEYS[5]: pairwise_accuracy(whole, pred_im, True),
            SCORE_KEYS[6]: pairwise_accuracy(yi, pred_im, False),
            SCORE_KEYS[7]: pairwise_accuracy(whole, pred_im, False),
        })
    return rows


# ------------------------------------------------------------------------------------------
# whole-brain grid search semantics (ridge_regression_decoding.py:111-135)
# ------------------------------------------------------------------------------------------
def kfold_bounds(n: int, n_splits: int = 5):
    """KFold(n_splits, shuffle=False): contiguous folds, the first n % n_splits one longer."""
    sizes = np.full(n_splits, n // n_splits, dtype=np.int64)
    sizes[: n % n_splits] += 1
    ends = np.cumsum(sizes)
    return list(zip(ends - sizes, ends))


def grid_search_fold_scores(x, y, alphas, n_splits: int = 5, dtype=np.float64):
    """cv_results_["split<f>_test_score"] of GridSearchCV(Ridge(fit_intercept=False), alpha grid,
    scoring=make_scorer(pairwise_accuracy), cv=n_splits) (ridge_regression_decoding.py:111-117): (n_alphas x n_splits);
    fold score = pairwise accuracy of the un-standardised validation predictions (eval.py:116-123)."""
    x = np.asarray(x, dtype=dtype)
    y = np.asarray(y, dtype=dtype)
    n = len(x)
    scores = np.zeros((len(alphas), n_splits))
    for f, (lo, hi) in enumerate(kfold_bounds(n, n_splits)):
        tr = np.r_[0:lo, hi:n]
        for a, alpha in enumerate(alphas):
            w = ridge_fit(x[tr], y[tr], alpha)
            scores[a, f] = pairwise_accuracy(y[lo:hi], np.asarray(x[lo:hi], dtype=np.float64) @ w, False)
    return scores


def grid_search_ridge(x, y, alphas, n_splits: int = 5, dtype=np.float64):
    """GridSearchCV(..., refit): mean over folds; first alpha in grid order wins ties.
    Returns (best_alpha, mean_scores, W_refit)."""
    scores = grid_search_fold_scores(x, y, alphas, n_splits, dtype)
    mean = scores.mean(1)
    best = int(np.argmax(mean))                 # argmax returns the first maximum
    return alphas[best], mean, ridge_fit(np.asarray(x, dtype=dtype), np.asarray(y, dtype=dtype), alphas[best])
