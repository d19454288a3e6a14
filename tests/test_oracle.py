"""The float64 oracle must reproduce what the unmodified reference produced (tests/golden/*.npz,
written by tests/golden/make_golden.py) before it is trusted as the checker for the CUDA path."""
import os

import numpy as np
import pytest

from oracle import searchlight_oracle as orc


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


@pytest.mark.parametrize("name", ["searchlight_small.npz", "searchlight_smooth.npz", "searchlight_captions.npz"])
def test_searchlight_scores_match_reference(golden_dir, name):
    g = _load(golden_dir, name)
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    tr, te, im = np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)
    assert list(g["keys"]) == list(orc.SCORE_KEYS)
    centres = list(range(0, len(g["adjacency"]), 3))
    got = orc.search_light(g["X"], g["latents"], g["adjacency"], tr, te, im, alpha=float(g["alpha"]), centres=centres)
    want = g["scores"][centres]
    # identical comparison counts => identical quotients (both are count / constant in float64)
    assert np.array_equal(got, want), np.abs(got - want).max()


def test_knn_matches_sklearn_kdtree(golden_dir):
    g = _load(golden_dir, "knn_small.npz")
    for c, a, d in (("coords", "adjacency", "distances"), ("coords2", "adjacency2", "distances2")):
        dist, idx = orc.knn_adjacency(g[c], g[a].shape[1])
        assert np.array_equal(idx, g[a].astype(np.int64))
        assert np.array_equal(dist, g[d])
        assert np.array_equal(idx[:, 0], np.arange(len(idx)))


def _ragged(g):
    off = g["offsets"]
    return [g["indices"][off[i]:off[i + 1]].astype(np.int64) for i in range(len(off) - 1)]


def test_radius_mode_matches_reference(golden_dir):
    """searchlight.py:164-166 neighbourhoods and the loop on the ragged adjacency."""
    g = _load(golden_dir, "searchlight_radius.npz")
    A = _ragged(g)
    got_A = orc.radius_adjacency(g["coords"], float(g["radius"]))
    assert len(got_A) == len(A) and all(np.array_equal(a, b) for a, b in zip(got_A, A))
    assert len({len(a) for a in A}) > 1 and min(map(len, A)) < 64 < max(map(len, A))   # two padded sizes
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    tr, te, im = np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)
    centres = list(range(0, len(A), 4))
    got = orc.search_light(g["X"], g["latents"], A, tr, te, im, alpha=float(g["alpha"]), centres=centres)
    assert np.array_equal(got, g["scores"][centres]), np.abs(got - g["scores"][centres]).max()


def test_eval_known_answers(golden_dir):
    g = _load(golden_dir, "eval_cases.npz")
    lat, pred, im_lat, im_pred = g["lat"], g["pred"], g["im_lat"], g["im_pred"]
    assert orc.pairwise_accuracy(lat, pred) == float(g["pa_plain"])
    assert orc.pairwise_accuracy(lat, pred, True) == float(g["pa_std"])
    assert orc.pairwise_accuracy(np.concatenate([im_lat, lat]), im_pred) == float(g["pa_nonsquare"])
    c, n = orc.pairwise_count(g["dist_mat"])
    assert c / n == float(g["dist_mat_acc"])
    assert np.allclose(orc.cosine_distance_matrix(pred[:7], lat[:5]), g["distance_matrix"], rtol=0, atol=1e-14)
    is_cap = g["test_stim_types"] == "caption"
    assert np.array_equal(is_cap, orc.default_is_caption(140))
    sc = orc.score_centre(lat, pred, im_lat, im_pred, is_cap)
    ref = dict(zip(g["all_keys"], g["all_vals"]))
    for k, v in ref.items():
        assert sc[k] == v, k
    ref_ns = dict(zip(g["im_nostd_keys"], g["im_nostd_vals"]))
    assert sc["pairwise_acc_imagery_no_std"] == ref_ns["pairwise_acc_imagery"]
    assert sc["pairwise_acc_imagery_whole_test_set_no_std"] == ref_ns["pairwise_acc_imagery_whole_test_set"]


def test_analytic_known_answers():
    assert orc.pairwise_count(1.0 - np.eye(5)) == (20, 20)
    assert orc.pairwise_count(np.full((4, 4), 0.3))[0] == 0          # strict '<'
    rng = np.random.default_rng(0)
    y = rng.standard_normal((70, 16))
    assert orc.pairwise_accuracy(y, y.copy()) == 1.0
    assert orc.pairwise_count(np.zeros((70, 70)))[1] == 4830
    assert orc.pairwise_count(np.zeros((3, 3)))[1] == 6
    assert orc.pairwise_count(np.zeros((3, 143)))[1] == 426
    p = np.array([[1.0, 5.0], [1.0, 7.0], [1.0, 9.0]])
    z = orc.zscore_rows_population(p)
    assert np.all(z[:, 0] == 0.0)                                     # zero-variance feature keeps scale 1
    assert np.allclose(z[:, 1], (p[:, 1] - 7.0) / np.sqrt(8.0 / 3.0))


def test_shuffled_indices_match_reference(golden_dir):
    g = _load(golden_dir, "eval_cases.npz")
    want = g["shuffled_indices"]
    got = np.array(orc.shuffled_test_indices(len(want)))
    assert np.array_equal(got, want)
    assert not np.any(got == np.arange(140)[None])
    assert np.all(got[:, :70] < 70) and np.all(got[:, 70:] >= 70)


def test_null_distribution_matches_reference(golden_dir):
    g = _load(golden_dir, "null_distr_small.npz")
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    X, Y, A = g["X"], g["latents"].astype(np.float64), g["adjacency"]
    is_cap = orc.default_is_caption(nte)
    tr, te, im = np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)
    keys = list(g["null_keys"])
    for ci, c in enumerate(g["centres"]):
        _, pt, pi = orc.train_and_test(X[:, A[c]], Y, tr, te, im, 1.0, is_cap, return_preds=True)
        imperms = [g["derangements"][j] for j in g["derangement_choice"][ci]]
        rows = orc.null_scores_centre(Y[te], pt, Y[im], pi, is_cap, list(g["perms"]), imperms)
        got = np.array([[r[k] for k in keys] for r in rows])
        assert np.array_equal(got, g["null_scores"][ci])
    assert list(g["file_names"]) == [f"{int(c):010d}.p" for c in g["centres"]]


@pytest.mark.parametrize("tag", ["dual", "primal"])
def test_gridsearch_semantics(golden_dir, tag):
    g = _load(golden_dir, "gridsearch_small.npz")
    x, y, n = g[f"{tag}_x"], g[f"{tag}_y"], int(g[f"{tag}_n_train"])
    alphas = list(g[f"{tag}_alphas"])
    best, mean, w = orc.grid_search_ridge(x[:n], y[:n], alphas)
    # the reference ran sklearn in float32; the oracle in float64: scores are counts / constants, so
    # they agree unless a comparison sits at float32 rounding level
    assert best == float(g[f"{tag}_best_alpha"])
    assert np.allclose(mean, g[f"{tag}_mean_test_score"], atol=2e-3)
    pred = x[n:].astype(np.float64) @ w
    assert np.allclose(pred, g[f"{tag}_pred_test"], rtol=2e-3, atol=2e-3)
    assert orc.kfold_bounds(103, 5) == [(0, 21), (21, 42), (42, 63), (63, 83), (83, 103)]


# ---- group-level statistics --------------------------------------------------------------
def test_group_stats_oracle_matches_reference(golden_dir):
    from oracle import stats_oracle as so
    g = _load(golden_dir, "group_stats.npz")
    assert np.array_equal(so.ttest_1samp_no_p(g["ttest_x"]), g["ttest_t"])
    assert np.array_equal(so.ttest_1samp_no_p(g["ttest_x"], sigma=1e-2), g["ttest_t_rel"])
    assert np.array_equal(so.ttest_1samp_no_p(g["ttest_x"], sigma=1e-2, method="absolute"), g["ttest_t_abs"])
    for hemi in ("left", "right"):
        t = so.calc_image_t_values(g[f"{hemi}_data"], 0.5)
        assert np.array_equal(t, g[f"{hemi}_t"], equal_nan=True)
        assert np.isnan(t[5]) and np.isnan(t[77]) and np.isnan(t[300:310]).all()
        assert np.array_equal(so.calc_image_t_values(g[f"{hemi}_data"], 0.5, sigma=1e-3), g[f"{hemi}_t_sigma"],
                              equal_nan=True)
        edges = g[f"{hemi}_edges"]
        labels, sizes = so.clusters_at(t, edges, float(g[f"{hemi}_cluster_threshold"]))
        want = g[f"{hemi}_cluster_labels"]
        assert np.array_equal(labels >= 0, want >= 0)
        # same partition (label names differ): vertices share a label here iff they share one there
        pairs = set(zip(labels[labels >= 0].tolist(), want[want >= 0].tolist()))
        assert len(pairs) == len(sizes) == want.max() + 1
        assert np.array_equal(so.tfce(t, edges), g[f"{hemi}_tfce_default"])
        assert np.array_equal(so.tfce(t, edges, h=1, e=0.5, dh=0.2), g[f"{hemi}_tfce_h1_e05_dh02"])
        assert np.array_equal(so.tfce(t, edges, dh=0), g[f"{hemi}_tfce_auto"])
    assert not so.tfce(-np.abs(g["left_t"]), g["left_edges"]).any() and not g["left_tfce_nonpositive"].any()


def test_oracle_per_centre_grid_search_matches_gridsearchcv(golden_dir):
    """searchlight_cv.npz: sklearn GridSearchCV per neighbourhood (make_golden_r2.py); the oracle's fold scores give the
    same mean test scores and the same selected alpha."""
    g = np.load(os.path.join(golden_dir, "searchlight_cv.npz"))
    ntr = int(g["n_train"])
    X, Y = g["X"].astype(np.float64), g["latents"].astype(np.float64)
    alphas = [float(a) for a in g["alphas"]]
    for i in range(0, len(g["centres"]), 8):
        nb = g["adjacency"][g["centres"][i]]
        fs = orc.grid_search_fold_scores(X[:ntr][:, nb], Y[:ntr], alphas, int(g["n_folds"]))
        mean = np.average(fs, axis=1)
        assert np.array_equal(mean, g["mean_test_score"][i])
        assert alphas[int(np.argmax(mean))] == g["best_alpha"][i]
