"""Null-distribution branch of the reference's `train_and_test` (searchlight.py:61-79) on the GPU.

The reference re-scores each centre's fixed predictions against `len(shuffled_indices)` permuted target sets and
pickles a list of score dicts per centre (`null_distr/{centre:010d}.p`), the format
`searchlight_permutation_testing.py:395-439` reads back.  Here the permuted scores of all centres come out of one
`mmd_searchlight_null` call as a packed (n_centres x n_perm x 6) array; `write_null_distribution` then writes the
reference's per-centre pickles from it.
"""
from __future__ import annotations

import os
import pickle

import numpy as np

DERANGEMENTS_THREE_DIMS = [[1, 2, 0], [2, 0, 1]]      # searchlight.py:35

NULL_KEYS = ("pairwise_acc_captions", "pairwise_acc_images", "pairwise_acc_imagery",
             "pairwise_acc_imagery_whole_test_set", "pairwise_acc_imagery_no_std",
             "pairwise_acc_imagery_whole_test_set_no_std")    # dict order of searchlight.py:67-76


def draw_imagery_permutations(n_permutations: int, n_imagery: int, n_centres: int | None = None, seed=None) -> np.ndarray:
    """Derangements of the imagery targets, drawn like searchlight.py:65: one per centre and permutation (the reference
    draws inside the per-centre `train_and_test`), shape (n_centres, n_permutations, 3); `n_centres=None` gives one shared
    (n_permutations, 3) table.  `seed=None` uses numpy's global generator like the reference (`np.random.seed` makes a run
    reproducible); an integer seed draws from its own `RandomState`."""
    if n_imagery != 3:
        raise ValueError(f"the reference only defines imagery derangements for 3 trials (searchlight.py:35); got {n_imagery}")
    rng = np.random if seed is None else np.random.RandomState(seed)
    shape = (n_permutations,) if n_centres is None else (n_centres, n_permutations)
    choice = rng.choice(len(DERANGEMENTS_THREE_DIMS), size=shape)
    return np.asarray(DERANGEMENTS_THREE_DIMS, dtype=np.int32)[choice]


def null_scores(eng, A, alpha, shuffled_indices, imagery_permutations=None, n_imagery=3, seed=None):
    """(scores (n,8), flags (n,), null (n, n_perm, 6)) for the centres in `A` on the engine's uploaded trials.
    Without `imagery_permutations` every centre gets its own draws (reference semantics); `seed` makes them reproducible."""
    perm_test = np.asarray(shuffled_indices, dtype=np.int32)
    if perm_test.ndim != 2:
        raise ValueError("shuffled_indices must be a list of equally long index arrays (data.py:540-551)")
    if imagery_permutations is None:
        imagery_permutations = draw_imagery_permutations(len(perm_test), n_imagery, n_centres=len(A), seed=seed)
    return eng.searchlight_null(A, alpha, perm_test, imagery_permutations)      # rectangular or ragged neighbourhoods


def null_array_to_dicts(null_row: np.ndarray) -> list:
    """(n_perm x 6) -> the list of dicts the reference pickles for one centre."""
    return [dict(zip(NULL_KEYS, r.tolist())) for r in null_row]


def write_null_distribution(null: np.ndarray, list_indices, null_distr_dir: str) -> None:
    """One pickle per centre, named by the 10-digit centre index (searchlight.py:79)."""
    os.makedirs(null_distr_dir, exist_ok=True)
    for row, list_i in zip(null, list_indices):
        with open(os.path.join(null_distr_dir, f"{int(list_i):010d}.p"), "wb") as f:
            pickle.dump(null_array_to_dicts(row), f)
