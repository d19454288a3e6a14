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

import os
import pickle

import numpy as np

from ....engine import Engine, score_keys

DEFAULT_N_JOBS = 10                                   # searchlight.py:30 (kept for CLI compatibility)
DERANGEMENTS_THREE_DIMS = [[1, 2, 0], [2, 0, 1]]      # searchlight.py:35

_ENGINES: dict[int, Engine] = {}


def get_engine(device: int | None = None) -> Engine:
    """One engine (CUDA stream + workspaces) per device, created on first use."""
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if device not in _ENGINES:
        _ENGINES[device] = Engine(device)
    return _ENGINES[device]


def _estimator_alpha(estimator) -> float:
    """The loop only relies on Ridge(alpha, fit_intercept=False) (searchlight.py:240)."""
    if getattr(estimator, "fit_intercept", False):
        raise ValueError("the searchlight path fits without intercept (searchlight.py:240); got fit_intercept=True")
    alpha = getattr(estimator, "alpha", None)
    if alpha is None or np.ndim(alpha) != 0:
        raise ValueError("estimator must carry a scalar `alpha` (sklearn.linear_model.Ridge)")
    return float(alpha)


def _as_f16(a, name):
    if hasattr(a, "data_ptr"):                        # torch tensor (possibly already on the GPU)
        import torch
        return a if a.dtype == torch.float16 else a.to(torch.float16)
    a = np.asarray(a)
    if a.dtype == np.float16:
        return a
    # the reference casts to float16 before the loop (searchlight.py:247-248); other float inputs to
    # custom_search_light would be fitted at their own precision by sklearn, which this path does not offer
    raise TypeError(f"{name} must be float16 like the arrays searchlight.run builds (searchlight.py:247-248); "
                    f"got {a.dtype}")


def _is_caption_mask(n_test):
    from ....data import TEST_STIM_TYPES
    st = np.asarray(TEST_STIM_TYPES)
    if len(st) != n_test:
        raise ValueError(f"{n_test} test trials but TEST_STIM_TYPES has {len(st)} entries (data.py:182-187)")
    return st == "caption"


def _scores_to_dicts(scores: np.ndarray) -> np.ndarray:
    keys = score_keys()
    out = np.empty(len(scores), dtype=object)
    for i, row in enumerate(scores):
        out[i] = dict(zip(keys, row.tolist()))
    return out


def _rectangular(A):
    if isinstance(A, np.ndarray) and A.ndim == 2:
        return A
    lens = {len(a) for a in A}
    if len(lens) > 1:
        raise NotImplementedError("ragged neighbourhoods (radius mode, searchlight.py:164-171) are not supported "
                                  "by the GPU path yet; use --n-neighbors")
    return np.asarray([np.asarray(a) for a in A]) if len(A) else np.zeros((0, 1), dtype=np.int64)


def search_light_scores(X, y, estimator, A, train_ids, test_ids, imagery_ids, *, device=None, is_caption=None,
                        upload=True):
    """(n_centres x 8) float64 score block and per-centre flags for the centres in `A` on one GPU."""
    eng = get_engine(device)
    alpha = _estimator_alpha(estimator)
    A = _rectangular(A)
    if upload:
        if is_caption is None:
            is_caption = _is_caption_mask(len(test_ids))
        eng.upload(_as_f16(X, "X"), _as_f16(y, "y"), train_ids, test_ids, imagery_ids, is_caption)
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
        scores, _ = search_light_scores(X, y, estimator, list_rows, train_ids, test_ids, imagery_ids)
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
    A = _rectangular(A)
    local_only = n_jobs == 1
    lo, hi = (0, len(A)) if local_only else shard_range(len(A))
    if null_distr_dir is None:
        scores, flags = search_light_scores(X, y, estimator, A[lo:hi], train_ids, test_ids, imagery_ids)
    else:
        eng = get_engine()
        eng.upload(_as_f16(X, "X"), _as_f16(y, "y"), train_ids, test_ids, imagery_ids, _is_caption_mask(len(test_ids)))
        scores, flags = _null_branch(eng, A[lo:hi], np.arange(lo, hi), _estimator_alpha(estimator), len(imagery_ids),
                                     null_distr_dir, shuffled_indices)
    if verbose:
        print(f"searchlight: centres {lo}..{hi} done on device {get_engine().device}, "
              f"{int(flags.sum())} not positive definite")
    return _scores_to_dicts(scores if local_only else gather_scores(scores, len(A)))
