"""Drop-in surface of the reference's `analyses/decoding/ridge_regression_decoding.py` (whole-brain decoder).

What the searchlight needs from this module — the names it imports (searchlight.py:20-22), `NUM_CV_SPLITS`,
`DEFAULT_ALPHAS`, `get_run_str`, the CLI — is mirrored verbatim.  `run` keeps the reference's structure
(ridge_regression_decoding.py:45-187) but replaces `GridSearchCV(Ridge(fit_intercept=False), alpha grid, cv=5,
scoring=make_scorer(pairwise_accuracy))` by `grid_search_ridge_gpu`, which reproduces sklearn's semantics
(SURVEY.md §3.3): contiguous KFold(5) folds with the first `n % 5` one longer, fold score = cosine pairwise accuracy
of the un-standardised validation predictions, unweighted mean over folds, the FIRST alpha of the grid among ties,
refit on all training rows.

Status (a "next" row): the fold scoring runs in this repository's CUDA kernels
(`mmd_pairwise_accuracy`); the linear algebra of this first cut is float64 *library* code — torch.matmul (cuBLAS) for
the Gram/kernel matrix and torch.linalg.cholesky / cholesky_solve (cuSOLVER) for the solves — in the dual form
`K = X Xᵀ, (K_tr + αI) c = Y_tr, ŷ_val = K[val, tr] c` when V > N (what sklearn does, `_ridge.py:237-296`) and the primal
normal equations otherwise.  The reference fits in float32 (`:119-120`); float64 here differs from it only by the
float32 rounding of the reference itself (tests compare against the float64 oracle exactly and the float32 reference
fixture within 1e-3).  A hand-written batched solver reusing one factorisation across the alpha grid is round-2 work.
"""
from __future__ import annotations

import argparse
import os
import pickle
import time

import numpy as np

from ...data import (LatentFeatsConfig, SELECT_DEFAULT, FEATURE_COMBINATION_CHOICES, VISION_FEAT_COMBINATION_CHOICES,  # noqa: F401
                     LANG_FEAT_COMBINATION_CHOICES, apply_mask, standardize_fmri_betas, get_latent_features,
                     standardize_latents, MODALITY_AGNOSTIC, TRAINING_MODES, SPLIT_TRAIN, SPLIT_TEST, SPLIT_IMAGERY,
                     get_fmri_data)
from ...eval import (pairwise_accuracy, calc_all_pairwise_accuracy_scores, ACC_CAPTIONS, ACC_IMAGES, ACC_IMAGERY,  # noqa: F401
                     ACC_IMAGERY_WHOLE_TEST)
from ...utils import FMRI_BETAS_DIR, SUBJECTS, RESULTS_FILE, RIDGE_DECODER_OUT_DIR, DEFAULT_MODEL, DEFAULT_RESOLUTION

NUM_CV_SPLITS = 5                                      # ridge_regression_decoding.py:20
DEFAULT_ALPHAS = [1e2, 1e3, 1e4, 1e5, 1e6, 1e7]        # ridge_regression_decoding.py:21


def get_run_str(betas_dir, feats_config, mask=None, surface=False, resolution=DEFAULT_RESOLUTION):
    """ridge_regression_decoding.py:24-42."""
    run_str = f"{feats_config.model}_{feats_config.combined_feats}"
    run_str += f"_{feats_config.vision_features}"
    run_str += f"_{feats_config.lang_features}"
    if betas_dir.endswith(os.sep):
        betas_dir = betas_dir[:-1]
    run_str += f"_{betas_dir.split(os.sep)[-1]}"
    if mask is not None:
        if mask.startswith("functional_") or mask.startswith("anatomical_"):
            run_str += f"_mask_{mask}"
        else:
            mask_name = os.path.basename(mask).replace(".p", "")
            run_str += f"_mask_{mask_name}"
    if surface:
        run_str += f"_surface_{resolution}"
    return run_str


def kfold_bounds(n, n_splits=NUM_CV_SPLITS):
    """KFold(n_splits, shuffle=False) (`sklearn/model_selection/_split.py:523-530`): contiguous folds."""
    sizes = np.full(n_splits, n // n_splits, dtype=np.int64)
    sizes[: n % n_splits] += 1
    ends = np.cumsum(sizes)
    return [(int(e - s), int(e)) for s, e in zip(sizes, ends)]


class RidgeGPU:
    """The fitted-estimator protocol the reference relies on (`.alpha`, `.fit_intercept`, `.coef_ (D x V)`, `.predict`).
    The weights stay on the GPU (mmd_ridge_fit); `predict` runs the library's float64 GEMM (mmd_ridge_predict).  The engine
    holds one fitted ridge at a time: a later grid search on the same engine replaces it."""
    fit_intercept = False

    def __init__(self, alpha, engine, token):
        self.alpha, self._eng, self._token = alpha, engine, token
        self._coef = None

    def _check_live(self):
        if getattr(self._eng, "_ridge_token", None) is not self._token:
            raise RuntimeError("this RidgeGPU's weights were replaced by a later grid search on the same engine")

    @property
    def coef_(self):
        if self._coef is None:
            self._check_live()
            self._coef = self._eng.ridge_fit(self.alpha, want_coef=True)
        return self._coef

    def predict(self, X):
        self._check_live()
        return self._eng.ridge_predict(X)


def grid_search_ridge_gpu(X, Y, alphas, n_splits=NUM_CV_SPLITS, device=None, verbose=0):
    """GridSearchCV(...).fit(X, Y) semantics; returns (best_alpha, mean_test_score per alpha, RidgeGPU refit).

    Everything numeric runs in libmmd_b200.so (csrc/dense.cu): the Gram matrix (kernel form X X^T when there are more voxels
    than trials, like sklearn's cholesky solver, else X^T X), per fold and alpha a blocked Cholesky factorisation and the
    triangular solves for all targets, the validation predictions and their pairwise accuracy; fold matrices are sub-blocks
    (or down-dates) of the one Gram matrix."""
    from ..decoding.searchlight.searchlight import get_engine
    eng = get_engine(device)
    n = len(X)
    if n < n_splits:
        raise ValueError(f"Cannot have number of splits n_splits={n_splits} greater than the number of samples: n_samples={n}.")
    alphas = [float(a) for a in alphas]
    eng.ridge_load(X, Y)
    bounds = [0] + [hi for _, hi in kfold_bounds(n, n_splits)]
    scores = eng.ridge_cv(alphas, bounds)                 # (n_alphas, n_splits)
    if verbose:
        for f in range(n_splits):
            for a, alpha in enumerate(alphas):
                print(f"[CV {f + 1}/{n_splits}] END alpha={alpha};, score={scores[a, f]:.3f}")
    mean = np.average(scores, axis=1)                     # _search.py: array_means
    best = int(np.argmax(mean))                          # first maximum = rank-"min" tie break of GridSearchCV
    alpha = alphas[best]
    eng.ridge_fit(alpha, want_coef=False)
    eng._ridge_token = token = object()
    return alpha, mean, RidgeGPU(alpha, eng, token)


def run(args):
    """ridge_regression_decoding.py:45-187: same loop, prints, `results.p` keys and resume rule."""
    for training_mode in args.training_modes:
        for subject in args.subjects:
            train_fmri_betas_full, train_stim_ids, train_stim_types = get_fmri_data(
                args.betas_dir, subject, SPLIT_TRAIN, training_mode, surface=args.surface, resolution=args.resolution)
            test_fmri_betas_full, test_stim_ids, test_stim_types = get_fmri_data(
                args.betas_dir, subject, SPLIT_TEST, surface=args.surface, resolution=args.resolution)
            imagery_fmri_betas_full, imagery_stim_ids, imagery_stim_types = get_fmri_data(
                args.betas_dir, subject, SPLIT_IMAGERY, surface=args.surface, resolution=args.resolution)
            for mask in args.masks:
                mask = None if mask in ["none", "None"] else mask
                train_fmri_betas, test_fmri_betas, imagery_fmri_betas = apply_mask(
                    mask, [train_fmri_betas_full, test_fmri_betas_full, imagery_fmri_betas_full], args)
                train_fmri_betas, test_fmri_betas, imagery_fmri_betas = standardize_fmri_betas(
                    train_fmri_betas, test_fmri_betas, imagery_fmri_betas)
                print(f"\n\ntrain fMRI betas shape: {train_fmri_betas.shape}")
                print(f"test fMRI betas shape: {test_fmri_betas.shape}")
                print(f"imagery fMRI betas shape: {imagery_fmri_betas.shape}")

                for model in args.models:
                    feats_config = LatentFeatsConfig(
                        model, args.features, args.test_features, args.vision_features, args.lang_features)
                    print(f"\nTRAIN MODE: {training_mode} | SUBJECT: {subject} | "
                          f"MODEL: {model} | FEATURES: {feats_config.features} {feats_config.vision_features} "
                          f"{feats_config.lang_features} | TEST FEATURES: {feats_config.test_features}")
                    if mask is not None:
                        print("Mask: ", os.path.basename(mask))
                    run_str = get_run_str(args.betas_dir, feats_config, mask, args.surface, args.resolution)
                    results_file_path = os.path.join(RIDGE_DECODER_OUT_DIR, training_mode, subject, run_str, RESULTS_FILE)
                    if os.path.isfile(results_file_path) and not args.overwrite:
                        print(f"Skipping decoder training as results are already present at {results_file_path}")
                        continue

                    train_latents = get_latent_features(feats_config, subject, SPLIT_TRAIN, training_mode)
                    test_latents = get_latent_features(feats_config, subject, SPLIT_TEST)
                    imagery_latents = get_latent_features(feats_config, subject, SPLIT_IMAGERY)
                    train_latents, test_latents, imagery_latents = standardize_latents(
                        train_latents, test_latents, imagery_latents)
                    print(f"train latents shape: {train_latents.shape}")

                    train_latents = train_latents.astype(np.float32)          # the reference's input precision (:119-120)
                    train_fmri_betas = train_fmri_betas.astype(np.float32)

                    start = time.time()
                    best_alpha, mean_scores, best_model = grid_search_ridge_gpu(
                        train_fmri_betas, train_latents, args.l2_regularization_alphas, NUM_CV_SPLITS, verbose=3)
                    end = time.time()
                    print(f"Elapsed time: {int(end - start)}s")

                    test_fmri_betas = test_fmri_betas.astype(np.float32)
                    test_predicted_latents = best_model.predict(test_fmri_betas)
                    imagery_fmri_betas = imagery_fmri_betas.astype(np.float32)
                    imagery_predicted_latents = best_model.predict(imagery_fmri_betas)

                    results = {
                        "alpha": best_alpha,
                        "model": model,
                        "subject": subject,
                        "features": feats_config.features,
                        "test_features": feats_config.test_features,
                        "vision_features": feats_config.vision_features,
                        "lang_features": feats_config.lang_features,
                        "training_mode": training_mode,
                        "mask": mask,
                        "num_voxels": test_fmri_betas.shape[1],
                        "stimulus_ids": test_stim_ids,
                        "stimulus_types": test_stim_types,
                        "imagery_stimulus_ids": imagery_stim_ids,
                        "predictions": test_predicted_latents,
                        "imagery_predictions": imagery_predicted_latents,
                        "latents": test_latents,
                        "imagery_latents": imagery_latents,
                        "surface": args.surface,
                        "resolution": args.resolution,
                    }
                    scores = calc_all_pairwise_accuracy_scores(
                        test_latents, test_predicted_latents, test_stim_types,
                        imagery_latents, imagery_predicted_latents, standardize_predictions=True)
                    results.update(scores)
                    print(
                        f"Best alpha: {best_alpha}\n"
                        f"Pairwise acc (mean): {np.mean((results[ACC_CAPTIONS], results[ACC_IMAGES])):.4f}"
                        f" | Pairwise acc (captions): {results[ACC_CAPTIONS]:.2f}"
                        f" | Pairwise acc (images): {results[ACC_IMAGES]:.2f}"
                        f" | Pairwise acc (imagery): {results[ACC_IMAGERY]:.2f}"
                        f" | Pairwise acc (imagery whole test set): {results[ACC_IMAGERY_WHOLE_TEST]:.2f}")
                    os.makedirs(os.path.dirname(results_file_path), exist_ok=True)
                    pickle.dump(results, open(results_file_path, 'wb'))

                    no_std = calc_all_pairwise_accuracy_scores(
                        test_latents, test_predicted_latents, test_stim_types,
                        imagery_latents, imagery_predicted_latents, standardize_predictions=False)
                    print(
                        f"Without standardization of predictions:\n"
                        f"Pairwise acc (mean): {np.mean((no_std[ACC_CAPTIONS], no_std[ACC_IMAGES])):.4f}"
                        f" | Pairwise acc (captions): {no_std[ACC_CAPTIONS]:.2f}"
                        f" | Pairwise acc (images): {no_std[ACC_IMAGES]:.2f}"
                        f" | Pairwise acc (imagery): {no_std[ACC_IMAGERY]:.2f}"
                        f" | Pairwise acc (imagery whole test set): {no_std[ACC_IMAGERY_WHOLE_TEST]:.2f}")


def get_args(argv=None):
    """ridge_regression_decoding.py:190-222: flags and defaults verbatim."""
    parser = argparse.ArgumentParser()
    parser.add_argument("--betas-dir", type=str, default=FMRI_BETAS_DIR)
    parser.add_argument("--training-modes", type=str, nargs="+", default=[MODALITY_AGNOSTIC], choices=TRAINING_MODES)
    parser.add_argument("--models", type=str, nargs='+', default=[DEFAULT_MODEL])
    parser.add_argument("--features", type=str, default=SELECT_DEFAULT, choices=FEATURE_COMBINATION_CHOICES)
    parser.add_argument("--test-features", type=str, default=SELECT_DEFAULT, choices=FEATURE_COMBINATION_CHOICES)
    parser.add_argument("--vision-features", type=str, default=SELECT_DEFAULT, choices=VISION_FEAT_COMBINATION_CHOICES)
    parser.add_argument("--lang-features", type=str, default=SELECT_DEFAULT, choices=LANG_FEAT_COMBINATION_CHOICES)
    parser.add_argument("--subjects", type=str, nargs='+', default=SUBJECTS)
    parser.add_argument("--l2-regularization-alphas", type=float, nargs='+', default=DEFAULT_ALPHAS)
    parser.add_argument("--n-jobs", type=int, default=10)
    parser.add_argument("--n-pre-dispatch", type=int, default=10)
    parser.add_argument("--surface", action='store_true', default=False)
    parser.add_argument("--resolution", default=DEFAULT_RESOLUTION)
    parser.add_argument("--masks", nargs='+', type=str, default=[None])
    parser.add_argument("--overwrite", action='store_true', default=False)
    return parser.parse_args(argv)


if __name__ == "__main__":
    args = get_args()
    os.makedirs(RIDGE_DECODER_OUT_DIR, exist_ok=True)
    run(args)
