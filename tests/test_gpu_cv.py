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
