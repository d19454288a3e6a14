"""Hot-path subset of the reference's `data.py`: constants and the host-side pre-steps the searchlight uses.

File loading (GIfTI betas, pickled embeddings; data.py:345-503) is out of scope — the dataset is not on these
boxes — and is replaced by `multimodal_decoding_b200.synthetic`, which produces arrays with the same contracts.
The standardisation helpers are one-pass host pre-steps and stay on the host; they restate sklearn's
StandardScaler arithmetic including its dtype rule (float32 in -> float32 out), pinned by tests/golden/data_cases.npz.
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


_FLOAT_DTYPES = (np.float64, np.float32, np.float16)


def _as_float(a):
    """sklearn's input validation for the scaler (dtype=FLOAT_DTYPES): float arrays keep their dtype, the rest -> float64."""
    a = np.asarray(a)
    return a if a.dtype in _FLOAT_DTYPES else a.astype(np.float64)


def _fit_scaler(train):
    """StandardScaler().fit (data.py:507, 526) restated: sklearn's corrected two-pass variance with float64
    accumulators whatever the input dtype (`_incremental_mean_and_var` on a first batch), population variance, and
    scale 1 for the features `_is_constant_feature` cannot tell from a constant."""
    X = _as_float(train)
    nan_mask = np.isnan(X)
    sum_op = np.nansum if nan_mask.any() else np.sum
    acc = {"dtype": np.float64} if X.dtype.itemsize < 8 else {}
    new_sum = sum_op(X, axis=0, **acc)
    count = X.shape[0] - sum_op(nan_mask.astype(X.dtype), axis=0, **acc)
    count = np.asarray(count, dtype=np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        mean = new_sum / count
        temp = X - mean                      # float64: `mean` is
        correction = sum_op(temp, axis=0)
        temp **= 2
        unnormalized = sum_op(temp, axis=0)
        unnormalized -= correction ** 2 / count
        var = unnormalized / count
    n_seen = count[0] if count.size and count.max() == count.min() else count
    eps = np.finfo(np.float64).eps
    constant = var <= n_seen * eps * var + (n_seen * mean * eps) ** 2
    scale = np.sqrt(var)
    scale[constant] = 1.0
    return mean, scale


def _transform(a, mean, scale):
    """StandardScaler.transform: a copy in the input's own float dtype, centred and scaled IN that dtype (float32
    betas / latents stay float32, as in the reference; the fp16 cast of searchlight.py:247-248 then rounds the same
    values)."""
    X = np.array(_as_float(a), copy=True)
    X -= mean.astype(X.dtype)
    X /= scale.astype(X.dtype)
    return X


def standardize_latents(train_latents, test_latents, imagery_latents=None):
    """data.py:506-515: z-score every split with the TRAIN statistics."""
    mean, scale = _fit_scaler(train_latents)
    out = [_transform(a, mean, scale) for a in (train_latents, test_latents)]
    if imagery_latents is not None:
        out.append(_transform(imagery_latents, mean, scale))
    return tuple(out)


def standardize_fmri_betas(train_fmri_betas, test_fmri_betas, imagery_fmri_betas=None):
    """data.py:518-537: drop the columns that are NaN in the first training trial, then z-score by TRAIN stats."""
    print('standardizing betas .. ', end='')
    keep = ~np.isnan(train_fmri_betas[0])
    splits = [train_fmri_betas[:, keep], test_fmri_betas[:, keep]]
    if imagery_fmri_betas is not None:
        splits.append(imagery_fmri_betas[:, keep])
    mean, scale = _fit_scaler(splits[0])
    out = tuple(_transform(a, mean, scale) for a in splits)
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


# ----------------------------------------------------------------------------------------------------------
# feature-selection vocabulary of the CLIs (data.py:190-338)
# ----------------------------------------------------------------------------------------------------------
AVG_FEATS, LANG_FEATS_ONLY, VISION_FEATS_ONLY = 'avg', 'lang', 'vision'
FUSED_FEATS_CLS, FUSED_FEATS_MEAN, MATCHED_FEATS = 'fused_cls', 'fused_mean', 'matched'
VISION_MEAN_FEAT_KEY, VISION_CLS_FEAT_KEY = "vision_features_mean", "vision_features_cls"
LANG_MEAN_FEAT_KEY, LANG_CLS_FEAT_KEY = "lang_features_mean", "lang_features_cls"
FEATURE_COMBINATION_CHOICES = [AVG_FEATS, LANG_FEATS_ONLY, VISION_FEATS_ONLY, FUSED_FEATS_CLS, FUSED_FEATS_MEAN,
                               MATCHED_FEATS, SELECT_DEFAULT]
VISION_FEAT_COMBINATION_CHOICES = [VISION_MEAN_FEAT_KEY, VISION_CLS_FEAT_KEY, SELECT_DEFAULT]
LANG_FEAT_COMBINATION_CHOICES = [LANG_MEAN_FEAT_KEY, LANG_CLS_FEAT_KEY, SELECT_DEFAULT]
FEATS_NA = "n_a"

# the multimodal models of BASELINE.json's configs; (default features, vision key, lang key, embedding width)
# widths: ImageBind 1024, CLIP ViT-L/14 768 (extract_clip_features.py:14), SigLIP so400m 1152 (extract_siglip_features.py:38)
_MODEL_DEFAULTS = {
    "imagebind": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 1024),
    "random-imagebind": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 1024),
    "clip": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 768),
    "siglip": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 1152),
    "flava": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 768),
    "blip2": (AVG_FEATS, VISION_CLS_FEAT_KEY, LANG_CLS_FEAT_KEY, 768),
    "paligemma2": (AVG_FEATS, VISION_MEAN_FEAT_KEY, LANG_MEAN_FEAT_KEY, 2304),
}
DEFAULT_FEATURES = {m: v[0] for m, v in _MODEL_DEFAULTS.items()}
DEFAULT_VISION_FEATURES = {m: v[1] for m, v in _MODEL_DEFAULTS.items()}
DEFAULT_LANG_FEATURES = {m: v[2] for m, v in _MODEL_DEFAULTS.items()}
MODEL_EMBEDDING_DIMS = {m: v[3] for m, v in _MODEL_DEFAULTS.items()}


class LatentFeatsConfig:
    """data.py:317-338: resolves the 'default' feature choices per model; `combined_feats` names the result dir."""

    def __init__(self, model, features=SELECT_DEFAULT, test_features=SELECT_DEFAULT, vision_features=SELECT_DEFAULT,
                 lang_features=SELECT_DEFAULT, logging=True):
        if model not in _MODEL_DEFAULTS:
            raise KeyError(model)
        self.model = model
        self.features = DEFAULT_FEATURES[model] if features == SELECT_DEFAULT else features
        self.test_features = DEFAULT_FEATURES[model] if test_features == SELECT_DEFAULT else test_features
        self.vision_features = DEFAULT_VISION_FEATURES[model] if vision_features == SELECT_DEFAULT else vision_features
        self.lang_features = DEFAULT_LANG_FEATURES[model] if lang_features == SELECT_DEFAULT else lang_features
        self.logging = logging
        if logging:
            print(f"Selected features for {self.model}: {self.features} {self.test_features} "
                  f"{self.vision_features} {self.lang_features}")
        self.combined_feats = f"{self.features}_test_{self.test_features}"


# ----------------------------------------------------------------------------------------------------------
# data access (data.py:401-503).  The dataset (GIfTI/NIfTI betas, pickled embeddings) is not on these boxes and its
# loaders are out of scope; the functions keep the reference's signatures and return arrays with the
# same contracts from a pluggable provider.  The default provider is the seeded synthetic generator of SURVEY §8d.
# ----------------------------------------------------------------------------------------------------------
class SyntheticProvider:
    """One synthetic subject x hemisphere: icosphere mesh (fsaverage7 -> level 7, 163,842 vertices), smooth low-rank
    betas + noise, low-rank embeddings; NaN "medial wall" columns exercise `nan_locations`."""

    LEVELS = {"fsaverage7": 7, "fsaverage6": 6, "fsaverage5": 5, "fsaverage4": 4, "fsaverage3": 3, "fsaverage": 7}

    def __init__(self, n_train=4000, nan_fraction=0.08, seed=0):
        self.n_train, self.nan_fraction, self.seed = n_train, nan_fraction, seed
        self._cache = {}


SUBJECT_FOR_MESH = "sub-01"
_PROVIDER = SyntheticProvider()
_CURRENT_MODEL = ["imagebind"]


def set_data_provider(provider):
    """Install an object with `.fmri(subject, split, mode, hemi, resolution)`, `.latents(config, subject, split, mode)`
    and `.mesh(hemi, resolution)`; a real-data provider wraps the reference's nibabel loaders."""
    global _PROVIDER
    _PROVIDER = provider


def get_data_provider():
    return _PROVIDER


def _provider_trials(self, subject, hemi, resolution, model, mode):
    from . import synthetic as syn
    key = (subject, hemi, resolution, model, mode)
    if key not in self._cache:
        level = self.LEVELS.get(resolution, 7)
        s = self.seed + 1000 * (sum(map(ord, subject)) % 97) + (0 if hemi in (None, "left") else 1)
        mesh = syn.make_mesh(level=level, seed=7 + s, nan_fraction=self.nan_fraction)
        trials = syn.make_trials(mesh, n_train=self.n_train, n_dims=MODEL_EMBEDDING_DIMS.get(model, 768), seed=s,
                                 training_mode=mode)
        self._cache = {key: (mesh, trials)}
    return self._cache[key]


def _provider_fmri(self, subject, split, mode=MODALITY_AGNOSTIC, hemi=None, resolution="fsaverage7"):
    mesh, t = _provider_trials(self, subject, hemi, resolution, _CURRENT_MODEL[0], mode if split == SPLIT_TRAIN else
                               self._last_mode)
    if split == SPLIT_TRAIN:
        n = len(t.train_fmri)
        types = np.array([CAPTION] * (n // 2) + [IMAGE] * (n - n // 2)) if mode == MODALITY_AGNOSTIC else \
            np.array([CAPTION if mode == MODALITY_SPECIFIC_CAPTIONS else IMAGE] * n)
        return t.train_fmri, np.arange(n), types
    if split == SPLIT_TEST:
        return t.test_fmri, np.arange(len(t.test_fmri)) % (NUM_TEST_STIMULI // 2), t.test_stim_types
    return t.imagery_fmri, np.arange(len(t.imagery_fmri)), np.array([IMAGERY] * len(t.imagery_fmri))


def _provider_latents(self, feats_config, subject, split, mode=MODALITY_AGNOSTIC):
    for (subj, hemi, res, model, m), (mesh, t) in self._cache.items():
        if subj == subject:
            return {SPLIT_TRAIN: t.train_latents, SPLIT_TEST: t.test_latents, SPLIT_IMAGERY: t.imagery_latents}[split]
    raise RuntimeError("synthetic provider: load the fMRI betas of a subject before its latents (as run() does)")


def _provider_mesh(self, hemi, resolution):
    for (subj, h, res, model, m), (mesh, _) in self._cache.items():
        if h == hemi and res == resolution:
            return mesh
    return _provider_trials(self, SUBJECT_FOR_MESH, hemi, resolution, _CURRENT_MODEL[0], MODALITY_AGNOSTIC)[0]


SyntheticProvider._last_mode = MODALITY_AGNOSTIC
SyntheticProvider.fmri = _provider_fmri
SyntheticProvider.latents = _provider_latents
SyntheticProvider.mesh = _provider_mesh


def get_fmri_surface_data(betas_dir, subject, split, mode=MODALITY_AGNOSTIC, hemi="left", resolution="fsaverage7"):
    """data.py:448-478 contract: (betas (N, V_full) float32 with NaN columns, stim_ids, stim_types)."""
    if split == SPLIT_TRAIN and hasattr(_PROVIDER, "_last_mode"):
        _PROVIDER._last_mode = mode
    return _PROVIDER.fmri(subject, split, mode, hemi, resolution)


def get_fmri_data(betas_dir, subject, split, mode=MODALITY_AGNOSTIC, surface=False, resolution="fsaverage7"):
    """data.py:481-490: volume data, or both hemispheres concatenated when `surface` is set."""
    if not surface:
        return get_fmri_surface_data(betas_dir, subject, split, mode, None, resolution)
    parts = [get_fmri_surface_data(betas_dir, subject, split, mode, hemi, resolution) for hemi in ("left", "right")]
    return np.concatenate([p[0] for p in parts], axis=1), parts[0][1], parts[0][2]


def get_latent_features(feats_config, subject, split, mode=MODALITY_AGNOSTIC):
    """data.py:401-445 contract: (N, D) float32 embeddings in trial order."""
    _CURRENT_MODEL[0] = feats_config.model
    return _PROVIDER.latents(feats_config, subject, split, mode)


def apply_mask(mask_name, betas_list, args=None):
    """data.py:564-574: the reference's masks are pickled voxel lists of the real dataset; only `None` is meaningful
    for synthetic data."""
    if mask_name is not None:
        raise NotImplementedError("voxel masks belong to the real dataset (data.py:564-574), which is not on this box")
    return betas_list
