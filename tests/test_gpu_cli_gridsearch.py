"""GPU tests of the drop-in CLIs: `searchlight.run(args)` on a small synthetic subject (result pickle layout + scores
against the oracle) and the whole-brain grid search against the float32 sklearn fixture and the float64 oracle."""
import os
import pickle

import numpy as np
import pytest

from oracle import searchlight_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag", ["dual", "primal"])
def test_gridsearch_matches_reference(golden_dir, tag):
    from multimodal_decoding_b200.analyses.decoding import ridge_regression_decoding as rr
    g = np.load(os.path.join(golden_dir, "gridsearch_small.npz"))
    x, y, n = g[f"{tag}_x"], g[f"{tag}_y"], int(g[f"{tag}_n_train"])
    alphas = list(g[f"{tag}_alphas"])
    best, mean, model = rr.grid_search_ridge_gpu(x[:n], y[:n], alphas)
    o_best, o_mean, o_w = orc.grid_search_ridge(x[:n], y[:n], alphas)
    assert best == o_best == float(g[f"{tag}_best_alpha"])
    assert np.array_equal(mean, o_mean)                                    # float64 vs float64: identical counts
    assert np.allclose(mean, g[f"{tag}_mean_test_score"], atol=2e-3)      # vs the reference's float32 fit
    pred = model.predict(x[n:])
    np.testing.assert_allclose(pred, x[n:].astype(np.float64) @ o_w, rtol=1e-9, atol=1e-9)
    assert np.allclose(pred, g[f"{tag}_pred_test"], rtol=2e-3, atol=2e-3)
    assert model.coef_.shape == (y.shape[1], x.shape[1]) and model.fit_intercept is False
    assert rr.kfold_bounds(103, 5) == [(0, 21), (21, 42), (42, 63), (63, 83), (83, 103)]


def test_searchlight_cli_run(tmp_path, monkeypatch):
    from multimodal_decoding_b200 import data as d
    from multimodal_decoding_b200.analyses.searchlight import searchlight as sl      # the README's module path
    monkeypatch.setattr(sl, "SEARCHLIGHT_OUT_DIR", str(tmp_path))
    import multimodal_decoding_b200.analyses.decoding.searchlight.searchlight as real
    monkeypatch.setattr(real, "SEARCHLIGHT_OUT_DIR", str(tmp_path))
    d.set_data_provider(d.SyntheticProvider(n_train=300, nan_fraction=0.08, seed=3))
    try:
        args = sl.get_args(["--model", "clip", "--n-neighbors", "40", "--training-mode", "agnostic", "--subjects",
                            "sub-01", "--hemis", "left", "--resolution", "fsaverage3", "--create-null-distr",
                            "--n-permutations-per-subject", "4"])
        np.random.seed(0)
        sl.run(args)
        cfg = d.LatentFeatsConfig("clip", logging=False)
        path = real.get_results_file_path(cfg, "left", "sub-01", "agnostic", "fsaverage3", "n_neighbors_40", 1)
        res = pickle.load(open(path, "rb"))
        assert set(res) == {"nan_locations", "adjacency", "n_neighbors", "distances", "scores"}
        V = int((~res["nan_locations"]).sum())
        assert res["adjacency"].shape == (V, 40) and res["adjacency"].dtype == np.int64
        assert res["distances"].shape == (V, 40) and res["n_neighbors"] is None
        assert res["scores"].shape == (V,) and len(res["scores"][0]) == 8
        assert np.array_equal(res["adjacency"][:, 0], np.arange(V))        # the centre itself first
        null_dir = os.path.join(os.path.dirname(path), "null_distr")
        assert sorted(os.listdir(null_dir)) == [f"{i:010d}.p" for i in range(V)]
        assert len(pickle.load(open(os.path.join(null_dir, f"{V - 1:010d}.p"), "rb"))) == 4
        # scores against the oracle on the same arrays the CLI built
        tr_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TRAIN, "agnostic", "left", "fsaverage3")
        te_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TEST, hemi="left", resolution="fsaverage3")
        im_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_IMAGERY, hemi="left", resolution="fsaverage3")
        tr_f, te_f, im_f = d.standardize_fmri_betas(tr_f, te_f, im_f)
        lat = [d.get_latent_features(cfg, "sub-01", s) for s in (d.SPLIT_TRAIN, d.SPLIT_TEST, d.SPLIT_IMAGERY)]
        lat = np.concatenate(d.standardize_latents(*lat)).astype(np.float16)
        X = np.concatenate((tr_f, te_f, im_f)).astype(np.float16)
        ntr, nte = len(tr_f), len(te_f)
        ids = (np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + len(im_f)))
        centres = list(range(0, V, 37))
        want = orc.search_light(X, lat, res["adjacency"], *ids, 1.0, orc.default_is_caption(nte), centres)
        got = np.array([[res["scores"][c][k] for k in orc.SCORE_KEYS] for c in centres])
        assert np.array_equal(got, want)
        _, a_ref = orc.knn_adjacency(d.get_data_provider().mesh("left", "fsaverage3").coords[~res["nan_locations"]], 40)
        assert np.array_equal(res["adjacency"], a_ref)
    finally:
        d.set_data_provider(d.SyntheticProvider())


def test_searchlight_cli_run_radius(tmp_path, monkeypatch):
    """`--radius` (searchlight.py:164-171): ragged adjacency in the pickle, per-vertex neighbour counts, scores
    against the oracle."""
    from multimodal_decoding_b200 import data as d
    import multimodal_decoding_b200.analyses.decoding.searchlight.searchlight as sl
    monkeypatch.setattr(sl, "SEARCHLIGHT_OUT_DIR", str(tmp_path))
    d.set_data_provider(d.SyntheticProvider(n_train=300, nan_fraction=0.08, seed=5))
    try:
        args = sl.get_args(["--model", "clip", "--radius", "50", "--training-mode", "agnostic", "--subjects", "sub-01",
                            "--hemis", "left", "--resolution", "fsaverage3"])
        sl.run(args)
        cfg = d.LatentFeatsConfig("clip", logging=False)
        path = sl.get_results_file_path(cfg, "left", "sub-01", "agnostic", "fsaverage3", "radius_50.0", 1)
        res = pickle.load(open(path, "rb"))
        V = int((~res["nan_locations"]).sum())
        coords = d.get_data_provider().mesh("left", "fsaverage3").coords[~res["nan_locations"]]
        want_A = orc.radius_adjacency(coords, 50.0)
        assert len(res["adjacency"]) == V and res["distances"] is None
        assert all(np.array_equal(a, b) for a, b in zip(res["adjacency"], want_A))
        assert res["n_neighbors"] == [len(a) for a in want_A] and len(set(res["n_neighbors"])) > 1
        tr_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TRAIN, "agnostic", "left", "fsaverage3")
        te_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TEST, hemi="left", resolution="fsaverage3")
        im_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_IMAGERY, hemi="left", resolution="fsaverage3")
        tr_f, te_f, im_f = d.standardize_fmri_betas(tr_f, te_f, im_f)
        lat = [d.get_latent_features(cfg, "sub-01", s) for s in (d.SPLIT_TRAIN, d.SPLIT_TEST, d.SPLIT_IMAGERY)]
        lat = np.concatenate(d.standardize_latents(*lat)).astype(np.float16)
        X = np.concatenate((tr_f, te_f, im_f)).astype(np.float16)
        ntr, nte = len(tr_f), len(te_f)
        ids = (np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + len(im_f)))
        centres = list(range(0, V, 29))
        want = orc.search_light(X, lat, want_A, *ids, 1.0, orc.default_is_caption(nte), centres)
        got = np.array([[res["scores"][c][k] for k in orc.SCORE_KEYS] for c in centres])
        assert np.array_equal(got, want)
    finally:
        d.set_data_provider(d.SyntheticProvider())
