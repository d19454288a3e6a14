"""GPU parity of the eval.py drop-in and of the null-distribution branch against fixtures written by the
unmodified reference (tests/golden/make_golden.py).  Needs a B200: run with `-m gpu`."""
import os
import pickle

import numpy as np
import pytest

from oracle import searchlight_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ev():
    from multimodal_decoding_b200 import eval as ev_mod
    return ev_mod


def test_eval_api_matches_reference(ev, golden_dir):
    g = np.load(os.path.join(golden_dir, "eval_cases.npz"))
    lat, pred, im_lat, im_pred = g["lat"], g["pred"], g["im_lat"], g["im_pred"]
    stim = g["test_stim_types"]
    assert ev.pairwise_accuracy(lat, pred) == float(g["pa_plain"])
    assert ev.pairwise_accuracy(lat, pred, standardize_predictions=True) == float(g["pa_std"])
    assert ev.pairwise_accuracy(np.concatenate([im_lat, lat]), im_pred) == float(g["pa_nonsquare"])
    allsc = ev.calc_all_pairwise_accuracy_scores(lat, pred, stim, im_lat, im_pred, standardize_predictions=True)
    assert list(allsc.keys()) == list(g["all_keys"])
    assert np.array_equal(np.array(list(allsc.values())), g["all_vals"])
    nostd = ev.calc_imagery_pairwise_accuracy_scores(im_lat, im_pred, lat, standardize_predictions=False)
    assert list(nostd.keys()) == list(g["im_nostd_keys"])
    assert np.array_equal(np.array(list(nostd.values())), g["im_nostd_vals"])
    nocross = ev.calc_all_pairwise_accuracy_scores(lat, pred, stim, im_lat, im_pred, standardize_predictions=False,
                                                   comp_cross_decoding_scores=False)
    assert list(nocross.keys()) == list(g["nocross_keys"])
    assert np.array_equal(np.array(list(nocross.values())), g["nocross_vals"])
    normed = ev.calc_all_pairwise_accuracy_scores(lat, pred, stim, im_lat, im_pred,
                                                  norm_imagery_preds_with_test_preds=True)
    assert np.array_equal(np.array(list(normed.values())), g["normed_vals"])
    assert ev.dist_mat_to_pairwise_acc(g["dist_mat"]) == float(g["dist_mat_acc"])
    np.testing.assert_allclose(ev.get_distance_matrix(pred[:7], lat[:5]), g["distance_matrix"], rtol=0, atol=1e-13)


def test_eval_known_answers_and_errors(ev):
    rng = np.random.default_rng(0)
    lat = rng.standard_normal((37, 19))
    assert ev.pairwise_accuracy(lat, lat) == 1.0                       # distinct rows predicted exactly
    assert ev.dist_mat_to_pairwise_acc(1.0 - np.eye(6)) == 1.0
    assert ev.dist_mat_to_pairwise_acc(np.full((4, 7), 0.25)) == 0.0   # strict '<'
    const = np.ones((5, 8))
    const[:, 0] = np.arange(5)                                          # zero-variance columns keep scale 1
    cand = rng.standard_normal((5, 8))
    assert ev.pairwise_accuracy(cand, const, standardize_predictions=True) == orc.pairwise_accuracy(cand, const, True)
    with pytest.raises(ValueError):
        ev.pairwise_accuracy(lat[:3], lat[:5])                          # more predictions than candidates
    with pytest.raises(NotImplementedError):
        ev.pairwise_accuracy(lat, lat, metric="euclidean")
    # random shapes against the oracle
    for n_lat, n_pred, d in ((143, 3, 64), (70, 70, 33), (200, 180, 129), (801, 800, 1024)):
        a, b = rng.standard_normal((n_lat, d)), rng.standard_normal((n_pred, d))
        for std in (False, True):
            assert ev.pairwise_accuracy(a, b, standardize_predictions=std) == orc.pairwise_accuracy(a, b, std)


def test_null_distribution_matches_reference(golden_dir, tmp_path):
    from multimodal_decoding_b200.analyses.decoding.searchlight import null_distribution as nd
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    from sklearn.linear_model import Ridge
    g = np.load(os.path.join(golden_dir, "null_distr_small.npz"))
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    tr, te, im = np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)
    eng = sl.get_engine(0)
    eng.upload(g["X"], g["latents"], tr, te, im, orc.default_is_caption(nte))
    assert list(nd.NULL_KEYS) == list(g["null_keys"])
    for ci, c in enumerate(g["centres"]):
        im_perm = g["derangements"][g["derangement_choice"][ci]]       # the draws the reference made for this centre
        scores, flags, null = eng.searchlight_null(g["adjacency"][[c]], 1.0, g["perms"], im_perm)
        assert flags.sum() == 0
        assert np.array_equal(null[0], g["null_scores"][ci])
    # all centres in one call, every centre with the draws the reference made for it (mmd_searchlight_null_pc)
    per_centre = np.stack([g["derangements"][g["derangement_choice"][ci]] for ci in range(len(g["centres"]))])
    scores, flags, null = eng.searchlight_null(g["adjacency"][g["centres"]], 1.0, g["perms"], per_centre)
    assert flags.sum() == 0 and np.array_equal(null, g["null_scores"])
    # ragged (radius-mode) neighbourhoods go through the same branch (mmd_searchlight_null_ragged)
    rows = [g["adjacency"][c] for c in g["centres"]]
    ragged = [rows[0], rows[1][:-5], rows[2]]
    s_r, f_r, null_r = eng.searchlight_null(ragged, 1.0, g["perms"], per_centre)
    assert f_r.sum() == 0 and np.array_equal(null_r[[0, 2]], g["null_scores"][[0, 2]])
    s_1, _, null_1 = eng.searchlight_null(rows[1][None, :-5], 1.0, g["perms"], per_centre[[1]])
    assert np.array_equal(null_r[1], null_1[0]) and np.array_equal(s_r[1], s_1[0])
    a = nd.draw_imagery_permutations(5, 3, n_centres=4, seed=7)
    assert a.shape == (4, 5, 3) and np.array_equal(a, nd.draw_imagery_permutations(5, 3, n_centres=4, seed=7))
    assert {tuple(r) for r in a.reshape(-1, 3)} <= {(1, 2, 0), (2, 0, 1)} and nd.draw_imagery_permutations(5, 3).shape == (5, 3)
    # the drop-in writes the reference's per-centre pickles
    np.random.seed(1234)
    out = sl.custom_group_iter_search_light([g["adjacency"][c] for c in g["centres"]], list(g["centres"]),
                                            Ridge(alpha=1.0, fit_intercept=False), g["X"], g["latents"], list(tr),
                                            list(te), list(im), thread_id=1, null_distr_dir=str(tmp_path),
                                            shuffled_indices=list(g["perms"]))
    assert len(out) == 3 and len(out[0]) == 8
    assert sorted(os.listdir(tmp_path)) == list(g["file_names"])
    first = pickle.load(open(tmp_path / g["file_names"][0], "rb"))
    assert len(first) == len(g["perms"]) and list(first[0].keys()) == list(g["null_keys"])
    # columns that do not depend on the imagery draw are identical to the reference's file
    assert np.array_equal(np.array([[p[k] for k in nd.NULL_KEYS[:2]] for p in first]), g["null_scores"][0][:, :2])
