"""Drop-in for the hot path of the reference's `analyses/decoding/searchlight/searchlight.py`.

Same function names, argument order and return types as the reference (file:line cited per function); the work is
done by the sm_100a library behind `multimodal_decoding_b200.engine.Engine` (C ABI in include/mmd_b200.h).  There is
no CPU path here: without the library or a CUDA device every call raises.

What changes for a caller: `n_jobs` no longer means CPU processes.  The centres are split into contiguous ranges
exactly like the reference's `GroupIterator` (searchlight.py:130) — one range per GPU taking part
(`torch.distributed` world if initialised, else the single current device) — and the concatenated result has the
reference's vertex order.
"""
from __future__ import annotations

import argparse
import gc
import os
import pickle
import time

import numpy as np

from ....data import (TEST_STIM_TYPES, get_fmri_surface_data, SELECT_DEFAULT, LatentFeatsConfig,
                      create_null_distr_shuffled_indices, standardize_fmri_betas, standardize_latents,
                      get_latent_features, SPLIT_TRAIN, MODALITY_AGNOSTIC, SPLIT_TEST, SPLIT_IMAGERY, TRAINING_MODES,
                      FEATURE_COMBINATION_CHOICES, VISION_FEAT_COMBINATION_CHOICES, LANG_FEAT_COMBINATION_CHOICES,
                      get_data_provider, _CURRENT_MODEL)
from ....engine import Engine, score_keys
from ....eval import ACC_CAPTIONS, ACC_IMAGES, ACC_IMAGERY, ACC_IMAGERY_WHOLE_TEST
from ....utils import SUBJECTS, DATA_DIR, DEFAULT_RESOLUTION, FMRI_BETAS_SURFACE_DIR, DEFAULT_MODEL

DEFAULT_N_JOBS = 10                                   # searchlight.py:30 (kept for CLI compatibility)
SEARCHLIGHT_OUT_DIR = os.path.join(DATA_DIR, "searchlight")                                     # searchlight.py:32
SEARCHLIGHT_PERMUTATION_TESTING_RESULTS_DIR = os.path.join(SEARCHLIGHT_OUT_DIR, "permutation_testing_results")
DERANGEMENTS_THREE_DIMS = [[1, 2, 0], [2, 0, 1]]      # searchlight.py:35

_ENGINES: dict[int, Engine] = {}
NUM_CV_SPLITS = 5                       # ridge_regression_decoding.py:20
LAST_CV: dict = {}                      # best_alpha / mean_test_score of the last cross-validated call (this rank's centres)
LAST_TIMING: dict[str, float] = {}      # host milliseconds of the phases of the last custom_search_light call (diagnostics)


class _Phase:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        self.t0 = time.perf_counter()

    def __exit__(self, *exc):
        LAST_TIMING[self.name] = LAST_TIMING.get(self.name, 0.0) + (time.perf_counter() - self.t0) * 1e3


def get_engine(device: int | None = None) -> Engine:
    """One engine (CUDA stream + workspaces) per device, created on first use."""
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if device not in _ENGINES:
        _ENGINES[device] = Engine(device)
        # one process per GPU (torchrun): share the host cores between the ranks' planner threads
        ranks = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
        if ranks > 1:
            try:
                cores = len(os.sched_getaffinity(0))
            except AttributeError:
                cores = os.cpu_count() or 1
            _ENGINES[device].configure(host_threads=max(2, cores // ranks))
    return _ENGINES[device]


def _estimator_alpha(estimator) -> float:
    """The loop only relies on Ridge(alpha, fit_intercept=False) (searchlight.py:240)."""
    if getattr(estimator, "fit_intercept", False):
        raise ValueError("the searchlight path fits without intercept (searchlight.py:240); got fit_intercept=True")
    alpha = getattr(estimator, "alpha", None)
    if alpha is None or np.ndim(alpha) != 0:
        raise ValueError("estimator must carry a scalar `alpha` (sklearn.linear_model.Ridge)")
    if not float(alpha) >= 0.0:                        # sklearn: InvalidParameterError (a ValueError) at fit time
        raise ValueError(f"The 'alpha' parameter of Ridge must be a float in the range [0.0, inf). Got {alpha!r} instead.")
    return float(alpha)


def _grid_search_spec(estimator):
    """(alphas, n_splits) when `estimator` is GridSearchCV(Ridge(fit_intercept=False), param_grid=dict(alpha=[...]), cv=int)
    -- the construction of ridge_regression_decoding.py:111-117 -- else None."""
    if not (hasattr(estimator, "param_grid") and hasattr(estimator, "estimator")):
        return None
    grid = estimator.param_grid
    if not isinstance(grid, dict) or set(grid) != {"alpha"}:
        raise ValueError("the searchlight grid search varies `alpha` only (param_grid=dict(alpha=[...]))")
    if getattr(estimator.estimator, "fit_intercept", False):
        raise ValueError("the searchlight path fits without intercept (searchlight.py:240); got fit_intercept=True")
    alphas = [float(a) for a in grid["alpha"]]
    if not alphas or not all(a > 0.0 for a in alphas):
        raise ValueError("param_grid['alpha'] must hold positive values")
    cv = getattr(estimator, "cv", None)
    n_splits = 5 if cv is None else cv                 # sklearn's default
    if not isinstance(n_splits, (int, np.integer)):
        raise ValueError("cv must be an integer number of contiguous folds (KFold without shuffling)")
    return alphas, int(n_splits)


def search_light_cv(eng, A, alphas, n_splits, n_train):
    """Per-centre GridSearchCV on the uploaded problem: returns (scores (n, 8) of the refitted best alpha, flags,
    best_alpha (n,), mean_test_score (n, n_alphas)).  Fold scores come from `mmd_searchlight_cv`; the mean over the folds
    and the first-best rule (GridSearchCV: rank "min" of the mean test score, first index of rank 1) are applied here with
    numpy as sklearn does; the refit runs one fixed-alpha searchlight per selected alpha."""
    from ..ridge_regression_decoding import kfold_bounds
    A = _require_rectangular(A, "the cross-validated searchlight")
    if n_train < n_splits:
        raise ValueError(f"Cannot have number of splits n_splits={n_splits} greater than the number of samples: "
                         f"n_samples={n_train}.")
    bounds = [0] + [hi for _, hi in kfold_bounds(n_train, n_splits)]
    fold_scores, flags = eng.searchlight_cv(A, alphas, bounds)
    mean = np.average(fold_scores, axis=2)             # _search.py: array_means = np.average(array, axis=1)
    best_idx = np.argmax(mean, axis=1)
    best_alpha = np.asarray(alphas, dtype=np.float64)[best_idx]
    scores = np.empty((len(A), len(score_keys())), dtype=np.float64)
    for i in np.unique(best_idx):
        sel = np.flatnonzero(best_idx == i)
        s, f = eng.searchlight(np.asarray(A)[sel], alphas[i])
        scores[sel] = s
        flags[sel] |= f
    return scores, flags, best_alpha, mean


def _as_f16(a, name):
    """float16 view of an input of the searchlight.  The reference casts to float16 before the loop
    (searchlight.py:247-248) and this path stores float16; `custom_search_light` itself takes any float array
    (searchlight.py:117-129) and sklearn would fit it at its own precision, so a float32 / float64 array is accepted
    exactly when the cast loses nothing (every value is a float16 number, e.g. float16 data that was widened): the result
    is then the float64-exact one for those values.  Anything else raises instead of being rounded silently."""
    if hasattr(a, "data_ptr"):                        # torch tensor (possibly already on the GPU)
        import torch
        return a if a.dtype == torch.float16 else a.to(torch.float16)
    a = np.asarray(a)
    if a.dtype == np.float16:
        return a
    if a.dtype in (np.float32, np.float64) and a.ndim == 2:
        out = np.empty(a.shape, dtype=np.float16)
        step = max(1, (1 << 24) // max(1, a.shape[1]))            # ~16 Mi values per block: the temporaries stay small
        with np.errstate(over="ignore", invalid="ignore"):
            for r in range(0, a.shape[0], step):
                blk = a[r:r + step]
                out[r:r + step] = blk
                if not np.array_equal(out[r:r + step].astype(a.dtype), blk, equal_nan=True):
                    raise TypeError(f"{name} is {a.dtype} and holds values that are not float16 numbers: this path computes on "
                                    f"float16 data like the arrays searchlight.run builds (searchlight.py:247-248); cast it "
                                    f"explicitly (X.astype(np.float16)) if that rounding is intended")
        return out
    raise TypeError(f"{name} must be float16 like the arrays searchlight.run builds (searchlight.py:247-248), or float32 / "
                    f"float64 holding float16 values; got {a.dtype}")


def _is_caption_mask(n_test):
    from ....data import TEST_STIM_TYPES
    st = np.asarray(TEST_STIM_TYPES)
    if len(st) != n_test:
        raise ValueError(f"{n_test} test trials but TEST_STIM_TYPES has {len(st)} entries (data.py:182-187)")
    return st == "caption"


_PYHELPERS = None


def _pyhelpers():
    """libmmd_pyhelpers.so (csrc/pydicts.c) through ctypes.PyDLL, or False when it was not built / does not match this Python."""
    global _PYHELPERS
    if _PYHELPERS is None:
        import ctypes
        import sys
        path = os.path.join(os.path.dirname(os.path.abspath(sys.modules["multimodal_decoding_b200"].__file__)),
                            "libmmd_pyhelpers.so")
        try:
            lib = ctypes.PyDLL(path)
            lib.mmd_build_score_dicts.restype = ctypes.py_object
            lib.mmd_build_score_dicts.argtypes = [ctypes.py_object, ctypes.c_void_p, ctypes.c_ssize_t, ctypes.c_void_p]
            lib.mmd_prebuild_score_dicts.restype = ctypes.py_object
            lib.mmd_prebuild_score_dicts.argtypes = [ctypes.py_object, ctypes.c_ssize_t, ctypes.c_void_p, ctypes.c_void_p]
            lib.mmd_fill_score_floats.restype = ctypes.py_object
            lib.mmd_fill_score_floats.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_ssize_t]
            probe = np.empty(2, dtype=object)
            lib.mmd_build_score_dicts(("a", "b"), np.array([[1.0, 2.0], [3.0, 4.0]]).ctypes.data, 2, probe.ctypes.data)
            _PYHELPERS = lib if probe[1] == {"a": 3.0, "b": 4.0} else False
        except (OSError, AttributeError):
            _PYHELPERS = False
    return _PYHELPERS


def _scores_to_dicts(scores: np.ndarray) -> np.ndarray:
    """(n, 8) float64 block -> the reference's object array of 8-key dicts (searchlight.py:114, :149).  The dicts are built
    through the CPython API by csrc/pydicts.c when that helper is present (about 40 ms for a 163,842-vertex hemisphere),
    else by a dict display over the unpacked columns, the fastest way pure Python offers (about 80 ms); the cyclic garbage
    collector is paused meanwhile (the dicts hold only floats)."""
    import gc
    keys = score_keys()
    scores = np.ascontiguousarray(np.asarray(scores, dtype=np.float64))
    out = np.empty(len(scores), dtype=object)
    was_enabled = gc.isenabled()
    gc.disable()
    try:
        lib = _pyhelpers()
        if lib and scores.ndim == 2 and scores.shape[1] == len(keys) and len(scores):
            lib.mmd_build_score_dicts(tuple(keys), scores.ctypes.data, len(scores), out.ctypes.data)
        elif len(keys) == 8:
            cols = [c.tolist() for c in scores.T]
            k0, k1, k2, k3, k4, k5, k6, k7 = keys
            out[:] = [{k0: a0, k1: a1, k2: a2, k3: a3, k4: a4, k5: a5, k6: a6, k7: a7}
                      for a0, a1, a2, a3, a4, a5, a6, a7 in zip(*cols)]
        else:
            cols = [c.tolist() for c in scores.T]
            out[:] = [dict(zip(keys, row)) for row in zip(*cols)]
    finally:
        if was_enabled:
            gc.enable()
    return out


class _DictPrebuild:
    """The score dicts of one call, built in two phases so that the expensive one overlaps the GPU: `start()` (called right
    before the long ctypes call into the library, which releases the GIL) launches a helper thread that creates the `n`
    dicts with their keys and placeholder float objects (csrc/pydicts.c, in chunks so the main thread gets the GIL back
    within a millisecond whenever it needs it); `finish(scores)` joins it and writes the values into those floats
    (~2 ms for a hemisphere instead of ~55 ms for building the dicts after the scores have arrived).  Without the helper
    library, or if `start()` was never reached, `finish` builds the dicts the plain way."""
    CHUNK = 4096

    def __init__(self, n):
        self.n = int(n)
        self.thread = None
        self.error = None
        self.out = self.slots = None

    def start(self):
        lib = _pyhelpers()
        if self.thread is not None or not lib or self.n == 0:
            return
        import threading
        self.nk = len(score_keys())
        self.out = np.empty(self.n, dtype=object)
        self.slots = np.empty((self.n, self.nk), dtype=np.uintp)
        self.thread = threading.Thread(target=self._build, args=(lib,), name="mmd-score-dicts", daemon=True)
        self.thread.start()

    def _build(self, lib):
        import gc
        # give the GIL straight back: the caller is about to enter the library (Thread.start() returns once this thread
        # runs, and would otherwise wait a switch interval for the first chunk below)
        time.sleep(0.002)
        keys = tuple(score_keys())
        was_enabled = gc.isenabled()
        gc.disable()                                   # the dicts hold only str keys and floats
        try:
            o, sl, step = self.out.ctypes.data, self.slots.ctypes.data, np.dtype(np.uintp).itemsize
            for i in range(0, self.n, self.CHUNK):
                m = min(self.CHUNK, self.n - i)
                lib.mmd_prebuild_score_dicts(keys, m, o + i * step, sl + i * self.nk * step)
        except BaseException as e:                     # surfaced by finish(), which then builds the dicts the plain way
            self.error = e
        finally:
            if was_enabled:
                gc.enable()

    def cancel(self):
        if self.thread is not None:
            self.thread.join()
        self.thread = self.out = self.slots = None

    def finish(self, scores):
        scores = np.ascontiguousarray(np.asarray(scores, dtype=np.float64))
        if self.thread is None:
            return _scores_to_dicts(scores)
        self.thread.join()
        out, slots = self.out, self.slots
        self.thread = self.out = self.slots = None
        if self.error is not None or scores.shape != (self.n, self.nk):
            return _scores_to_dicts(scores)
        _pyhelpers().mmd_fill_score_floats(slots.ctypes.data, scores.ctypes.data, scores.size)
        return out


def _report_flags(scores, flags, first_centre=0):
    """A centre whose normal matrix was not positive definite has no valid fit here (sklearn would fall back from its
    Cholesky solver to an SVD solve behind searchlight.py:56; this path has no such fallback): its scores are returned as
    NaN and the centres are named in a RuntimeWarning -- never a silently wrong number."""
    bad = np.flatnonzero(np.asarray(flags) != 0)
    if len(bad):
        import warnings
        scores[bad] = np.nan
        shown = ", ".join(str(first_centre + int(b)) for b in bad[:10]) + (", ..." if len(bad) > 10 else "")
        warnings.warn(f"searchlight: the normal matrix of {len(bad)} centre(s) is not positive definite (centres {shown}); "
                      "their scores are NaN.  Use alpha > 0 or drop duplicated / constant vertices.", RuntimeWarning,
                      stacklevel=3)
    return scores


def _is_ragged(A):
    return not (isinstance(A, np.ndarray) and A.dtype != object and A.ndim == 2)


def _rectangular(A):
    """(n, k) array for k-NN adjacency; a list of index arrays for the ragged radius-mode adjacency
    (searchlight.py:164-166)."""
    if isinstance(A, np.ndarray) and A.dtype != object and A.ndim == 2:
        return A
    rows = [np.asarray(a).ravel() for a in A]
    if len({len(r) for r in rows}) > 1:
        return rows
    return np.asarray(rows) if len(rows) else np.zeros((0, 1), dtype=np.int64)


def _require_rectangular(A, what):
    if _is_ragged(A):
        raise NotImplementedError(f"{what} is only built for fixed-size (k-NN) neighbourhoods; "
                                  "radius-mode neighbourhoods are scored without it")
    return A


def _active_vertices(A, n_vertices):
    """Ascending ids of the vertices the neighbourhoods in `A` use; None when that is (nearly) all of them."""
    from .... import _lib
    mask = np.zeros(n_vertices, dtype=np.uint8)
    if _is_ragged(A):
        flat = np.ascontiguousarray(np.concatenate([np.asarray(r).ravel() for r in A]))
    else:
        flat = np.ascontiguousarray(A)
    if flat.dtype not in (np.int32, np.int64):
        flat = flat.astype(np.int64)
    rc = _lib.load().mmd_mark_vertices(flat.ctypes.data, int(flat.dtype == np.int64), flat.size, n_vertices, mask.ctypes.data)
    if rc == 2:
        raise ValueError(f"adjacency holds a vertex index outside [0, {n_vertices})")
    active = np.flatnonzero(mask)
    return None if len(active) > 0.9 * n_vertices else active


def _wants_active_set(A, n_vertices):
    """Whether a shard should restrict the upload's X^T Y rows to the vertices its neighbourhoods use.  Finding them is one
    pass of the host over the shard's adjacency (18 ms for 1/8 of a 750-neighbour hemisphere); it saves GPU time of the
    upload (15 ms for all 163,842 vertices) that runs in the shadow of the library's own pass over the same adjacency
    anyway.  At N = 2 the end-to-end step is 41 ms shorter without it (profiles/r2_bench_n2_v4.json -> v5) and the N = 8
    breakdown says the same (18 ms of marking against < 12 ms of hidden kernels), so it is taken only when the shard is a
    small part of the mesh (more than 16 ranks per hemisphere).  MMD_B200_ACTIVE=1 / 0 forces it on / off."""
    mode = os.environ.get("MMD_B200_ACTIVE", "auto").lower()
    if mode in ("1", "on", "yes"):
        return True
    if mode in ("0", "off", "no"):
        return False
    return len(A) * 16 < n_vertices


def search_light_scores(X, y, estimator, A, train_ids, test_ids, imagery_ids, *, device=None, is_caption=None,
                        upload=True, anchors=None, shard_only=False, prebuild=None):
    """(n_centres x 8) float64 score block and per-centre flags for the centres in `A` on one GPU.
    `anchors` (ragged `A` only): the vertex each centre sits on, used to group nearby centres."""
    eng = get_engine(device)
    grid = _grid_search_spec(estimator)
    alpha = None if grid else _estimator_alpha(estimator)
    A = _rectangular(A)
    if upload:
        if is_caption is None:
            is_caption = _is_caption_mask(len(test_ids))
        # a rank of a vertex-sharded call stages the cross-covariance rows of its own neighbourhoods only
        with _Phase("active_vertices"):
            active = _active_vertices(A, np.shape(X)[1]) if shard_only and len(A) and _wants_active_set(A, np.shape(X)[1]) else None
        Xd = _as_f16(X, "X")
        if shard_only and isinstance(Xd, np.ndarray):
            # every rank holds the same host X: each copies 1/N of its rows over PCIe, NVLink carries the rest
            from ....sharding import device_copy_via_peers
            with _Phase("x_via_peers"):
                Xpeer = device_copy_via_peers(Xd)
            Xd = Xd if Xpeer is None else Xpeer
        with _Phase("upload"):
            eng.upload(Xd, _as_f16(y, "y"), train_ids, test_ids, imagery_ids, is_caption, active_vertices=active)
    with _Phase("searchlight"):
        if prebuild is not None:
            prebuild.start()                           # the helper thread works while the calls below wait for the GPU
        if grid:
            scores, flags, best_alpha, mean = search_light_cv(eng, A, grid[0], grid[1], len(train_ids))
            LAST_CV.clear()
            LAST_CV.update(best_alpha=best_alpha, mean_test_score=mean, alphas=list(grid[0]), n_splits=grid[1])
            return scores, flags
        if _is_ragged(A):
            return eng.searchlight_ragged(A, alpha, anchors)
        return eng.searchlight(A, alpha)


def train_and_test(estimator, X, y=None, *, train_ids, test_ids, imagery_ids, null_distr_dir=None,
                   shuffled_indices=None, list_i=None):
    """One centre on its pre-gathered neighbourhood `X` (N_total x k); searchlight.py:38-90."""
    X = _as_f16(X, "X")
    k = X.shape[1]
    res = custom_group_iter_search_light([np.arange(k)], [list_i], estimator, X, y, train_ids, test_ids,
                                         imagery_ids, thread_id=1, null_distr_dir=null_distr_dir,
                                         shuffled_indices=shuffled_indices)
    return res[0]


def _null_branch(eng, A, list_indices, alpha, n_imagery, null_distr_dir, shuffled_indices):
    """Scores + the reference's per-centre null pickles (searchlight.py:61-79) for the centres in `A`."""
    from .null_distribution import null_scores, write_null_distribution
    if shuffled_indices is None:
        raise ValueError("null_distr_dir is set but shuffled_indices is None (searchlight.py:63 iterates over it)")
    scores, flags, null = null_scores(eng, A, alpha, shuffled_indices, n_imagery=n_imagery)
    write_null_distribution(null, list_indices, null_distr_dir)
    return scores, flags


def custom_group_iter_search_light(list_rows, list_indices, estimator, X, y, train_ids, test_ids, imagery_ids,
                                   thread_id, null_distr_dir=None, shuffled_indices=None):
    """A chunk of centres; returns a list of score dicts (searchlight.py:93-114)."""
    if null_distr_dir is None:
        scores, flags = search_light_scores(X, y, estimator, list_rows, train_ids, test_ids, imagery_ids)
        _report_flags(scores, flags)
    else:
        eng = get_engine()
        eng.upload(_as_f16(X, "X"), _as_f16(y, "y"), train_ids, test_ids, imagery_ids, _is_caption_mask(len(test_ids)))
        scores, _ = _null_branch(eng, _rectangular(list_rows), list_indices, _estimator_alpha(estimator),
                                 len(imagery_ids), null_distr_dir, shuffled_indices)
    return list(_scores_to_dicts(scores))


def custom_search_light(X, y, estimator, A, train_ids, test_ids, imagery_ids, n_jobs=-1, verbose=0,
                        null_distr_dir=None, shuffled_indices=None):
    """All centres; returns an object ndarray (V,) of 8-key dicts in vertex order (searchlight.py:117-149).

    Under `torch.distributed` (one process per GPU) rank r scores the r-th contiguous range of
    `np.array_split(arange(V), world_size)` — the reference's own chunking — and the score blocks are gathered
    with one all-gather; every rank returns the full array.  `n_jobs=1` keeps the whole call on this process's
    GPU even inside a process group (one whole searchlight per GPU, SURVEY.md §8e / BASELINE.json configs[4]).
    """
    from ....sharding import gather_scores, shard_range
    LAST_TIMING.clear()
    t_call = time.perf_counter()
    A = _rectangular(A)
    local_only = n_jobs == 1
    lo, hi = (0, len(A)) if local_only else shard_range(len(A))
    prebuild = _DictPrebuild(len(A) if not local_only else hi - lo)
    if null_distr_dir is None:
        # radius mode: centre i sits on vertex i (searchlight.py:143-148 enumerates the rows of A by vertex)
        anchors = np.arange(lo, hi) if _is_ragged(A) and len(A) == np.shape(X)[1] else None
        try:
            scores, flags = search_light_scores(X, y, estimator, A[lo:hi], train_ids, test_ids, imagery_ids,
                                                anchors=anchors, shard_only=(hi - lo) < len(A), prebuild=prebuild)
        except BaseException:
            prebuild.cancel()
            raise
    else:
        eng = get_engine()
        eng.upload(_as_f16(X, "X"), _as_f16(y, "y"), train_ids, test_ids, imagery_ids, _is_caption_mask(len(test_ids)))
        scores, flags = _null_branch(eng, A[lo:hi], np.arange(lo, hi), _estimator_alpha(estimator), len(imagery_ids),
                                     null_distr_dir, shuffled_indices)
    _report_flags(scores, flags, lo)
    if verbose:
        print(f"searchlight: centres {lo}..{hi} done on device {get_engine().device}, "
              f"{int(flags.sum())} not positive definite")
    if not local_only:
        with _Phase("gather_scores"):
            scores = gather_scores(scores, len(A))
    with _Phase("score_dicts"):
        out = prebuild.finish(scores)
    LAST_TIMING["total"] = (time.perf_counter() - t_call) * 1e3
    return out


def get_adjacency_matrix(hemi, resolution, nan_locations=None, radius=None, num_neighbors=None):
    """searchlight.py:152-178.  k-NN branch: exact float64 k-NN on the GPU (`mmd_knn`), bit-identical to sklearn's
    `NearestNeighbors().kneighbors` on tie-free coordinates.  The mesh comes from the data provider (the reference
    loads nilearn's inflated fsaverage surface; nilearn is not on this box)."""
    coords = np.asarray(get_data_provider().mesh(hemi, resolution).coords)
    if nan_locations is not None:
        coords = coords[~nan_locations]
    nearest_neighbors = None
    distances = None
    if radius is not None:
        # searchlight.py:164-171 densifies a V x V graph (`.toarray()`), which only fits low resolutions; the GPU
        # kernel emits the same per-vertex index arrays (ascending vertex id, the vertex itself included) directly
        adjacency = get_engine().radius_neighbors(coords, float(radius))
        nearest_neighbors = [len(adj) for adj in adjacency]
        print(
            f"Number of neighbors within {radius}mm radius: {np.mean(nearest_neighbors):.1f} "
            f"(max: {np.max(nearest_neighbors):.0f} | min: {np.min(nearest_neighbors):.0f})")
    elif num_neighbors is not None:
        eng = get_engine()
        distances = np.empty((len(coords), num_neighbors), dtype=np.float64)
        adjacency = np.empty((len(coords), num_neighbors), dtype=np.int64)
        for lo in range(0, len(coords), 32768):
            hi = min(lo + 32768, len(coords))
            distances[lo:hi], adjacency[lo:hi] = eng.knn(coords, num_neighbors, lo, hi)
        print(f"Max radius {num_neighbors} neighbors: {distances.max():.2f}mm")
        print(f"Mean radius: {distances.max(axis=1).mean():.2f}mm")
    else:
        raise RuntimeError("Need to set either radius or n_neighbors arg!")
    return adjacency, nearest_neighbors, distances


def run(args):
    """searchlight.py:181-296: same loop, prints and result pickle; data come from the installed provider."""
    from sklearn.linear_model import Ridge
    shuffled_indices = None
    if args.create_null_distr:
        shuffled_indices = create_null_distr_shuffled_indices(args.n_permutations_per_subject)

    for subject in args.subjects:
        for training_mode in args.training_modes:
            for hemi in args.hemis:
                feats_config = LatentFeatsConfig(
                    args.model, args.features, args.test_features, args.vision_features, args.lang_features
                )
                _CURRENT_MODEL[0] = feats_config.model          # the synthetic provider sizes Y by the model
                train_fmri, train_stim_ids, train_stim_types = get_fmri_surface_data(
                    args.betas_dir, subject, SPLIT_TRAIN, training_mode, hemi, args.resolution
                )
                test_fmri, test_stim_ids, test_stim_types = get_fmri_surface_data(
                    args.betas_dir, subject, SPLIT_TEST, hemi=hemi, resolution=args.resolution
                )
                imagery_fmri, imagery_stim_ids, imagery_stim_types = get_fmri_surface_data(
                    args.betas_dir, subject, SPLIT_IMAGERY, hemi=hemi, resolution=args.resolution
                )
                nan_locations = np.isnan(train_fmri[0])
                train_fmri, test_fmri, imagery_fmri = standardize_fmri_betas(train_fmri, test_fmri, imagery_fmri)

                print(f"\nTRAIN MODE: {training_mode} | SUBJECT: {subject} | "
                      f"MODEL: {feats_config.model} | FEATURES: {feats_config.features}")

                train_latents = get_latent_features(feats_config, subject, SPLIT_TRAIN, training_mode)
                test_latents = get_latent_features(feats_config, subject, SPLIT_TEST)
                imagery_latents = get_latent_features(feats_config, subject, SPLIT_IMAGERY)
                train_latents, test_latents, imagery_latents = standardize_latents(
                    train_latents, test_latents, imagery_latents
                )
                latents = np.concatenate((train_latents, test_latents, imagery_latents))

                results_dir = get_results_dir(
                    feats_config, hemi, subject, training_mode, args.resolution, searchlight_mode_from_args(args)
                )
                os.makedirs(results_dir, exist_ok=True)

                print("Hemisphere: ", hemi)
                print(f"train_fmri shape: {train_fmri.shape}")
                print(f"test_fmri shape: {test_fmri.shape}")
                print(f"imagery_fmri shape: {imagery_fmri.shape}")

                train_ids = list(range(len(train_fmri)))
                test_ids = list(range(len(train_fmri), len(train_fmri) + len(test_fmri)))
                imagery_ids = list(range(len(train_ids) + len(test_ids),
                                         len(train_ids) + len(test_ids) + len(imagery_fmri)))

                X = np.concatenate((train_fmri, test_fmri, imagery_fmri))

                adjacency, n_neighbors, distances = get_adjacency_matrix(
                    hemi, args.resolution, nan_locations, args.radius, args.n_neighbors
                )

                model = Ridge(alpha=args.l2_regularization_alpha, fit_intercept=False)
                if args.l2_regularization_alphas:
                    # north star: a cross-validated ridge per vertex -- the GridSearchCV construction of
                    # ridge_regression_decoding.py:111-117 applied to every neighbourhood (mmd_searchlight_cv)
                    from sklearn.model_selection import GridSearchCV
                    model = GridSearchCV(estimator=Ridge(fit_intercept=False),
                                         param_grid=dict(alpha=list(args.l2_regularization_alphas)), cv=NUM_CV_SPLITS)

                null_distr_dir = None
                if args.create_null_distr:
                    null_distr_dir = os.path.join(results_dir, "null_distr")
                    os.makedirs(null_distr_dir, exist_ok=True)

                X = X.astype(np.float16)
                latents = latents.astype(np.float16)

                start = time.time()
                scores = custom_search_light(
                    X, latents, estimator=model, A=adjacency, train_ids=train_ids, test_ids=test_ids,
                    imagery_ids=imagery_ids, n_jobs=args.n_jobs, verbose=1, null_distr_dir=null_distr_dir,
                    shuffled_indices=shuffled_indices
                )
                end = time.time()
                print(f"Searchlight time: {int(end - start)}s")
                for label, key in (("captions", ACC_CAPTIONS), ("images", ACC_IMAGES), (ACC_IMAGERY, ACC_IMAGERY),
                                   (ACC_IMAGERY_WHOLE_TEST, ACC_IMAGERY_WHOLE_TEST)):
                    vals = [score[key] for score in scores]
                    print(f"Mean score ({label}): {np.mean(vals):.2f} | Max score: {np.max(vals):.2f}")

                results_dict = {
                    "nan_locations": nan_locations,
                    "adjacency": adjacency,
                    "n_neighbors": n_neighbors,
                    "distances": distances,
                    "scores": scores,
                }
                if args.l2_regularization_alphas and len(LAST_CV.get("best_alpha", ())) == len(scores):
                    results_dict["best_alpha"] = LAST_CV["best_alpha"]
                    results_dict["mean_test_score"] = LAST_CV["mean_test_score"]
                results_file_path = get_results_file_path(
                    feats_config, hemi, subject, training_mode, args.resolution, searchlight_mode_from_args(args),
                    "cv" if args.l2_regularization_alphas else args.l2_regularization_alpha
                )
                pickle.dump(results_dict, open(results_file_path, 'wb'))

                del X, latents, train_fmri
                gc.collect()


def searchlight_mode_from_args(args):
    if args.radius is not None:
        return f"radius_{args.radius}"
    elif args.n_neighbors is not None:
        return f"n_neighbors_{args.n_neighbors}"
    else:
        raise RuntimeError("Need to set either radius or n_neighbors arg!")


def get_results_dir(feats_config, hemi, subject, training_mode, resolution, mode):
    return os.path.join(
        SEARCHLIGHT_OUT_DIR, training_mode, feats_config.model, feats_config.combined_feats,
        feats_config.vision_features, feats_config.lang_features, subject, resolution, hemi, mode
    )


def get_results_file_path(feats_config, hemi, subject, training_mode, resolution, mode, l2_regularization_alpha):
    results_dir = get_results_dir(feats_config, hemi, subject, training_mode, resolution, mode)
    return os.path.join(results_dir, f"alpha_{str(l2_regularization_alpha)}.p")


def get_args(argv=None):
    """searchlight.py:321-355: flags and defaults verbatim (argparse abbreviation keeps `--training-mode` working)."""
    parser = argparse.ArgumentParser()
    parser.add_argument("--betas-dir", type=str, default=FMRI_BETAS_SURFACE_DIR)
    parser.add_argument("--training-modes", type=str, nargs="+", default=[MODALITY_AGNOSTIC], choices=TRAINING_MODES)
    parser.add_argument("--model", type=str, default=DEFAULT_MODEL)
    parser.add_argument("--features", type=str, default=SELECT_DEFAULT, choices=FEATURE_COMBINATION_CHOICES)
    parser.add_argument("--test-features", type=str, default=SELECT_DEFAULT, choices=FEATURE_COMBINATION_CHOICES)
    parser.add_argument("--vision-features", type=str, default=SELECT_DEFAULT, choices=VISION_FEAT_COMBINATION_CHOICES)
    parser.add_argument("--lang-features", type=str, default=SELECT_DEFAULT, choices=LANG_FEAT_COMBINATION_CHOICES)
    parser.add_argument("--subjects", type=str, nargs='+', default=SUBJECTS)
    parser.add_argument("--resolution", type=str, default=DEFAULT_RESOLUTION)
    parser.add_argument("--hemis", type=str, nargs="+", default=["left", "right"])
    parser.add_argument("--l2-regularization-alpha", type=float, default=1)
    # not in the reference's searchlight CLI: an alpha grid switches on 5-fold cross-validation per vertex (the flag of
    # ridge_regression_decoding.py:36); results go to alpha_cv.p with the selected alpha per vertex beside the scores
    parser.add_argument("--l2-regularization-alphas", type=float, nargs='+', default=None)
    parser.add_argument("--radius", type=float, default=None)
    parser.add_argument("--n-neighbors", type=int, default=None)
    parser.add_argument("--n-jobs", type=int, default=DEFAULT_N_JOBS)
    parser.add_argument("--create-null-distr", default=False, action="store_true")
    parser.add_argument("--n-permutations-per-subject", type=int, default=100)
    return parser.parse_args(argv)


if __name__ == "__main__":
    args = get_args()
    os.makedirs(SEARCHLIGHT_OUT_DIR, exist_ok=True)
    run(args)
