"""Drop-in for the scoring API of the reference's `eval.py` (lines 9-44 and 103-181), GPU-backed.

Same names, argument order, defaults and return types.  The arithmetic (StandardScaler z-scoring, cosine distance
matrix, strict-`<` pairwise count) runs in float64 CUDA kernels (csrc/pairwise.cu) through the C ABI
(`mmd_pairwise_accuracy`, `mmd_cosine_distances`, `mmd_dist_mat_to_pairwise_acc`); there is no CPU path.  The
searchlight itself does not call these per centre — its eight scores are fused in csrc/score.cu — they serve callers
that use eval.py directly: the whole-brain grid search (`make_scorer(pairwise_accuracy)`,
ridge_regression_decoding.py:111) and the notebooks.  RSA / CSLS helpers (eval.py:62-100, 184-223) are not on the
path and are not provided.
"""
from __future__ import annotations

import numpy as np

from .data import CAPTION, IMAGE

ACC_MODALITY_AGNOSTIC = "pairwise_acc_modality_agnostic"            # eval.py:9-44
ACC_CAPTIONS = "pairwise_acc_captions"
ACC_IMAGES = "pairwise_acc_images"
ACC_CROSS_IMAGES_TO_CAPTIONS = "pairwise_acc_cross_images_to_captions"
ACC_CROSS_CAPTIONS_TO_IMAGES = "pairwise_acc_cross_captions_to_images"
ACC_IMAGES_MOD_AGNOSTIC = "pairwise_acc_images_mod_agnostic"
ACC_CAPTIONS_MOD_AGNOSTIC = "pairwise_acc_captions_mod_agnostic"
ACC_IMAGERY_MOD_AGNOSTIC = "pairwise_acc_imagery_mod_agnostic"
ACC_IMAGERY_WHOLE_TEST_SET_MOD_AGNOSTIC = "pairwise_acc_imagery_whole_test_set_mod_agnostic"
ACC_IMAGERY_NO_STD_MOD_AGNOSTIC = "pairwise_acc_imagery_no_std_mod_agnostic"
ACC_IMAGERY_WHOLE_TEST_SET_NO_STD_MOD_AGNOSTIC = "pairwise_acc_imagery_whole_test_set_no_std_mod_agnostic"
ACC_IMAGES_MOD_SPECIFIC_IMAGES = "pairwise_acc_images_mod_specific_images"
ACC_CAPTIONS_MOD_SPECIFIC_IMAGES = "pairwise_acc_captions_mod_specific_images"
ACC_CAPTIONS_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_captions_mod_specific_captions"
ACC_IMAGES_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_images_mod_specific_captions"
ACC_IMAGERY_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_imagery_mod_specific_captions"
ACC_IMAGERY_MOD_SPECIFIC_IMAGES = "pairwise_acc_imagery_mod_specific_images"
ACC_IMAGERY_WHOLE_TEST_SET_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_imagery_whole_test_mod_specific_captions"
ACC_IMAGERY_WHOLE_TEST_SET_MOD_SPECIFIC_IMAGES = "pairwise_acc_imagery_whole_test_mod_specific_images"
ACC_IMAGERY_NO_STD_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_imagery_no_std_mod_specific_captions"
ACC_IMAGERY_NO_STD_MOD_SPECIFIC_IMAGES = "pairwise_acc_imagery_no_std_mod_specific_images"
ACC_IMAGERY_WHOLE_TEST_SET_NO_STD_MOD_SPECIFIC_CAPTIONS = "pairwise_acc_imagery_whole_test_no_std_mod_specific_captions"
ACC_IMAGERY_WHOLE_TEST_SET_NO_STD_MOD_SPECIFIC_IMAGES = "pairwise_acc_imagery_whole_test_no_std_mod_specific_images"
ACC_IMAGERY = "pairwise_acc_imagery"
ACC_IMAGERY_WHOLE_TEST = "pairwise_acc_imagery_whole_test_set"
DISTANCE_METRIC_COSINE = "cosine"


def _engine():
    from .analyses.decoding.searchlight.searchlight import get_engine
    return get_engine()


def _check_metric(metric):
    if metric != DISTANCE_METRIC_COSINE:
        raise NotImplementedError(f"only metric='cosine' (the reference's default, eval.py:103,116) runs on the GPU; got {metric!r}")


def get_distance_matrix(predictions, originals, metric='cosine'):
    """cdist(predictions, originals, metric) (eval.py:103-104): rows = predictions, columns = originals."""
    _check_metric(metric)
    return _engine().cosine_distances(predictions, originals)


def dist_mat_to_pairwise_acc(dist_mat):
    """Share of entries a row's diagonal entry is strictly smaller than (eval.py:107-113)."""
    dist_mat = np.asarray(dist_mat, dtype=np.float64)
    r, c = dist_mat.shape
    if r > c and c != 1:
        raise ValueError(f"operands could not be broadcast together with shapes ({min(r, c)},1) ({r},{c})")
    return _engine().dist_mat_to_pairwise_acc(dist_mat)


def pairwise_accuracy(latents, predictions, metric="cosine", standardize_predictions=False, standardize_latents=False):
    """eval.py:116-123.  Usable as an sklearn scorer: `make_scorer(pairwise_accuracy)` calls it as
    (y_true, y_pred) (ridge_regression_decoding.py:111)."""
    _check_metric(metric)
    latents, predictions = np.asarray(latents), np.asarray(predictions)
    if len(predictions) > len(latents) and len(latents) != 1:
        raise ValueError(f"operands could not be broadcast together with shapes ({len(latents)},1) "
                         f"({len(predictions)},{len(latents)})")
    return _engine().pairwise_accuracy(latents, predictions, standardize_predictions, standardize_latents)


def calc_all_pairwise_accuracy_scores(latents, predictions, stim_types=None, imagery_latents=None,
                                      imagery_predictions=None, metric="cosine", standardize_predictions=True,
                                      standardize_latents=False, norm_imagery_preds_with_test_preds=False,
                                      comp_cross_decoding_scores=True):
    """eval.py:126-158: within-modality, cross-modal and imagery scores as a dict (same keys, same order)."""
    latents, predictions, stim_types = np.asarray(latents), np.asarray(predictions), np.asarray(stim_types)
    results = dict()
    for modality, acc_metric_name in zip([CAPTION, IMAGE], [ACC_CAPTIONS, ACC_IMAGES]):
        results[acc_metric_name] = pairwise_accuracy(
            latents[stim_types == modality], predictions[stim_types == modality], metric, standardize_predictions,
            standardize_latents)
    if comp_cross_decoding_scores:
        for mod_preds, mod_latents, acc_metric_name in zip([CAPTION, IMAGE], [IMAGE, CAPTION],
                                                           [ACC_CROSS_CAPTIONS_TO_IMAGES, ACC_CROSS_IMAGES_TO_CAPTIONS]):
            results[acc_metric_name] = pairwise_accuracy(
                latents[stim_types == mod_latents], predictions[stim_types == mod_preds], metric,
                standardize_predictions, standardize_latents)
    if imagery_latents is not None:
        results.update(calc_imagery_pairwise_accuracy_scores(
            imagery_latents, imagery_predictions, latents, metric, standardize_predictions, standardize_latents,
            test_set_preds=predictions if norm_imagery_preds_with_test_preds else None))
    return results


def calc_imagery_pairwise_accuracy_scores(imagery_latents, imagery_predictions, additional_latents, metric="cosine",
                                          standardize_predictions=False, standardize_latents=False, test_set_preds=None):
    """eval.py:161-181."""
    imagery_latents = np.asarray(imagery_latents, dtype=np.float64)
    imagery_predictions = np.asarray(imagery_predictions, dtype=np.float64)
    results = dict()
    if test_set_preds is not None:
        # StandardScaler fitted on [imagery; test] predictions, applied to the imagery rows (eval.py:165-170):
        # a per-feature affine map of 3 rows; host-side like the other one-pass pre-steps (data.py helpers)
        from .data import _fit_scaler
        mean, scale = _fit_scaler(np.concatenate((imagery_predictions, np.asarray(test_set_preds, dtype=np.float64))))
        imagery_predictions = (imagery_predictions - mean) / scale
        standardize_predictions = False  # Do not standardize again
    results[ACC_IMAGERY] = pairwise_accuracy(
        imagery_latents, imagery_predictions, metric, standardize_predictions, standardize_latents)
    target_latents = np.concatenate((imagery_latents, np.asarray(additional_latents, dtype=np.float64)))
    results[ACC_IMAGERY_WHOLE_TEST] = pairwise_accuracy(
        target_latents, imagery_predictions, metric, standardize_predictions, standardize_latents)
    return results
