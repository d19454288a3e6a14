"""Alpha grid x 5-fold cross-validation per searchlight centre (mmd_searchlight_cv, BASELINE.json north_star (3)) against
the fixture written by sklearn's GridSearchCV + the unmodified reference's pairwise_accuracy / train_and_test
(tests/golden/make_golden_r2.py -> searchlight_cv.npz).  Needs a B200: run with `-m gpu`."""
import os

import numpy as np
import pytest

from oracle import searchlight_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from multimodal_decoding_b200.engine import Engine
    eng = Engine(0)
    yield eng
    eng.close()


def _load(golden_dir):
    g = np.load(os.path.join(golden_dir, "searchlight_cv.npz"))
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    return g, np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)


def test_fold_scores_selected_alpha_and_refit_match_gridsearchcv(engine, golden_dir):
    from multimodal_decoding_b200.analyses.decoding.searchlight.searchlight import search_light_cv
    g, tr, te, im = _load(golden_dir)
    engine.upload(g["X"], g["latents"], tr, te, im, orc.default_is_caption(len(te)))
    A = g["adjacency"][g["centres"]]
    alphas = [float(a) for a in g["alphas"]]
    scores, flags, best_alpha, mean = search_light_cv(engine, A, alphas, int(g["n_folds"]), len(tr))
    assert flags.sum() == 0
    # mean fold scores are sums of five count ratios: the counts must agree exactly
    assert np.array_equal(mean, g["mean_test_score"])
    assert np.array_equal(best_alpha, g["best_alpha"])
    assert np.array_equal(scores, g["scores"])


def test_fold_scores_are_row_separable_and_fold_sizes_may_differ(engine, golden_dir):
    """Folds longer than one chunk of 144 validation trials (two chunks per fold here) and unequal folds: the fold scores
    equal the oracle's GridSearchCV restatement."""
    from multimodal_decoding_b200 import synthetic as syn
    from multimodal_decoding_b200.analyses.decoding.ridge_regression_decoding import kfold_bounds
    mesh = syn.make_mesh(level=None, n_points=300, seed=5)
    trials = syn.make_trials(mesh, n_train=803, n_dims=40, n_factors=6, iid=False, seed=2, signal=0.6)
    inp = syn.make_searchlight_inputs(mesh, trials)
    _, A = orc.knn_adjacency(inp.coords, 30)
    centres = np.arange(0, len(A), 13)
    X, Y = inp.X.astype(np.float16), inp.latents.astype(np.float16)
    engine.upload(X, Y, inp.train_ids, inp.test_ids, inp.imagery_ids, orc.default_is_caption(len(inp.test_ids)))
    alphas = [1.0, 1e3]
    bounds = [0] + [hi for _, hi in kfold_bounds(len(inp.train_ids), 5)]
    assert bounds[1] - bounds[0] == 161 and bounds[-1] - bounds[-2] == 160
    got, flags = engine.searchlight_cv(A[centres], alphas, bounds)
    assert flags.sum() == 0
    tr = np.asarray(inp.train_ids)
    for ci, c in enumerate(centres):
        want = orc.grid_search_fold_scores(X[tr][:, A[c]].astype(np.float64), Y[tr].astype(np.float64), alphas, 5)
        assert np.array_equal(got[ci], want), (ci, got[ci], want)


def test_cv_errors(engine, golden_dir):
    g, tr, te, im = _load(golden_dir)
    engine.upload(g["X"], g["latents"], tr, te, im, orc.default_is_caption(len(te)))
    A = g["adjacency"][:4]
    with pytest.raises(RuntimeError, match="n_splits"):
        engine.searchlight_cv(A, [1.0], [0, len(tr)])
    with pytest.raises(RuntimeError, match="fold bounds"):
        engine.searchlight_cv(A, [1.0], [0, 100, 200])
    with pytest.raises(RuntimeError, match="alpha must be positive"):
        engine.searchlight_cv(A, [0.0], [0, 150, len(tr)])


def test_searchlight_cli_with_an_alpha_grid(tmp_path, monkeypatch):
    """`--l2-regularization-alphas`: a cross-validated ridge per vertex; the pickle carries the selected alpha per vertex and
    the scores are those of the oracle's fit with that alpha."""
    import pickle
    from multimodal_decoding_b200 import data as d
    import multimodal_decoding_b200.analyses.decoding.searchlight.searchlight as sl
    monkeypatch.setattr(sl, "SEARCHLIGHT_OUT_DIR", str(tmp_path))
    d.set_data_provider(d.SyntheticProvider(n_train=300, nan_fraction=0.08, seed=3))
    try:
        args = sl.get_args(["--model", "clip", "--n-neighbors", "40", "--training-mode", "agnostic", "--subjects", "sub-01",
                            "--hemis", "left", "--resolution", "fsaverage3", "--l2-regularization-alphas", "1", "1000", "100000"])
        sl.run(args)
        cfg = d.LatentFeatsConfig("clip", logging=False)
        path = sl.get_results_file_path(cfg, "left", "sub-01", "agnostic", "fsaverage3", "n_neighbors_40", "cv")
        res = pickle.load(open(path, "rb"))
        V = int((~res["nan_locations"]).sum())
        assert res["best_alpha"].shape == (V,) and res["mean_test_score"].shape == (V, 3)
        assert set(np.unique(res["best_alpha"])) <= {1.0, 1000.0, 100000.0}
        tr_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TRAIN, "agnostic", "left", "fsaverage3")
        te_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_TEST, hemi="left", resolution="fsaverage3")
        im_f, _, _ = d.get_fmri_surface_data("", "sub-01", d.SPLIT_IMAGERY, hemi="left", resolution="fsaverage3")
        tr_f, te_f, im_f = d.standardize_fmri_betas(tr_f, te_f, im_f)
        lat = [d.get_latent_features(cfg, "sub-01", s) for s in (d.SPLIT_TRAIN, d.SPLIT_TEST, d.SPLIT_IMAGERY)]
        lat = np.concatenate(d.standardize_latents(*lat)).astype(np.float16)
        X = np.concatenate((tr_f, te_f, im_f)).astype(np.float16)
        ntr, nte = len(tr_f), len(te_f)
        ids = (np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + len(im_f)))
        for c in range(0, V, 61):
            nb = res["adjacency"][c]
            fs = orc.grid_search_fold_scores(X[:ntr][:, nb].astype(np.float64), lat[:ntr].astype(np.float64), [1.0, 1000.0, 100000.0], 5)
            mean = np.average(fs, axis=1)
            assert np.array_equal(mean, res["mean_test_score"][c])
            alpha = [1.0, 1000.0, 100000.0][int(np.argmax(mean))]
            assert alpha == res["best_alpha"][c]
            want = orc.search_light(X, lat, res["adjacency"], *ids, alpha, orc.default_is_caption(nte), [c])
            assert np.array_equal(np.array([res["scores"][c][k] for k in orc.SCORE_KEYS]), want[0])
    finally:
        d.set_data_provider(d.SyntheticProvider())
