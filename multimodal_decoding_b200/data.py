"""Hot-path subset of the reference's `data.py`: constants and the host-side pre-steps the searchlight uses.

File loading (GIfTI betas, pickled embeddings; data.py:345-503) is out of scope — the dataset is not on these
boxes — and is replaced by `multimodal_decoding_b200.synthetic`, which produces arrays with the same contracts.
The standardisation helpers are one-pass host pre-steps and stay on the host in numpy float64.
"""
from __future__ import annotations

import numpy as np

MODALITY_SPECIFIC_IMAGES = "images"            # data.py:14-17
MODALITY_SPECIFIC_CAPTIONS = "captions"
MODALITY_AGNOSTIC = "agnostic"
TRAINING_MODES = [MODALITY_AGNOSTIC, MODALITY_SPECIFIC_CAPTIONS, MODALITY_SPECIFIC_IMAGES]

SPLIT_TRAIN, SPLIT_TEST, SPLIT_IMAGERY = "train", "test", "imagery"   # data.py:19-21
IMAGE, CAPTION, IMAGERY = "image", "caption", "imagery"               # data.py:23-25

NUM_TEST_STIMULI = 140                          # data.py:180: 70 test images, each also as caption
INDICES_TEST_STIM_CAPTION = list(range(NUM_TEST_STIMULI // 2))
INDICES_TEST_STIM_IMAGE = list(range(NUM_TEST_STIMULI // 2, NUM_TEST_STIMULI))
TEST_STIM_TYPES = np.array([CAPTION] * 70 + [IMAGE] * 70)             # data.py:187
TEST_BATCH_SIZE = NUM_TEST_STIMULI

SELECT_DEFAULT = "default"


def _fit_scaler(train):
    """StandardScaler().fit semantics: population std (ddof=0), zero-variance columns keep scale 1."""
    train = np.asarray(train, dtype=np.float64)
    mean = train.mean(axis=0)
    scale = train.std(axis=0)
    scale[scale == 0.0] = 1.0
    return mean, scale


def standardize_latents(train_latents, test_latents, imagery_latents=None):
    """data.py:506-515: z-score every split with the TRAIN statistics."""
    mean, scale = _fit_scaler(train_latents)
    out = [(np.asarray(a, dtype=np.float64) - mean) / scale for a in (train_latents, test_latents)]
    if imagery_latents is not None:
        out.append((np.asarray(imagery_latents, dtype=np.float64) - mean) / scale)
    return tuple(out)


def standardize_fmri_betas(train_fmri_betas, test_fmri_betas, imagery_fmri_betas=None):
    """data.py:518-537: drop the columns that are NaN in the first training trial, then z-score by TRAIN stats."""
    print('standardizing betas .. ', end='')
    keep = ~np.isnan(train_fmri_betas[0])
    splits = [train_fmri_betas[:, keep], test_fmri_betas[:, keep]]
    if imagery_fmri_betas is not None:
        splits.append(imagery_fmri_betas[:, keep])
    mean, scale = _fit_scaler(splits[0])
    out = tuple((np.asarray(a, dtype=np.float64) - mean) / scale for a in splits)
    print('done.')
    return out


def create_shuffled_indices(seed):
    """data.py:554-561 (legacy global numpy RandomState, as the reference seeds it)."""
    rs = np.random.RandomState(seed)
    half = NUM_TEST_STIMULI // 2
    first = rs.choice(range(half), size=half, replace=False)
    second = rs.choice(range(half, NUM_TEST_STIMULI), size=half, replace=False)
    return np.concatenate((first, second))


def create_null_distr_shuffled_indices(n_permutations_per_subject):
    """data.py:540-551: permutations without fixed points, captions and images shuffled separately."""
    shuffled, seed = [], 0
    for _ in range(n_permutations_per_subject):
        indices = create_shuffled_indices(seed)
        while any(indices == np.arange(NUM_TEST_STIMULI)):
            seed += 1
            indices = create_shuffled_indices(seed)
        shuffled.append(indices)
        seed += 1
    return shuffled
