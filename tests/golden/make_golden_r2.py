"""Round-2 golden fixtures, written by running the UNMODIFIED reference (and the sklearn calls it makes) here.

    python tests/golden/make_golden_r2.py [data] [config2] [cv]

data_cases.npz          standardize_fmri_betas / standardize_latents (data.py:506-537) on float32 and float64 inputs with
                        NaN columns and a constant column; create_null_distr_shuffled_indices (data.py:540-551)
searchlight_config2.npz custom_search_light (searchlight.py:117-149) at the headline shapes -- k = 750, N_train = 4000,
                        D = 1024, smooth betas -- on 48 centres of a 10,242-vertex mesh.  Only seeds, the centres' adjacency
                        rows and the reference's scores are stored: the test regenerates X / latents from the seeds
                        (multimodal_decoding_b200.synthetic is deterministic) and checks their checksum first.
searchlight_cv.npz      per-centre GridSearchCV(Ridge(fit_intercept=False), alphas, cv=5, scoring=pairwise_accuracy), the
                        semantics of ridge_regression_decoding.py:111-117 applied to every neighbourhood, on 96 centres:
                        mean fold scores, selected alpha, and the 8 scores of the refitted estimator through train_and_test
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from multimodal_decoding_b200 import synthetic as syn  # noqa: E402
from _reference_loader import load_reference  # noqa: E402

ref_sl, ref_eval, ref_data, ref_ridge = load_reference()
from sklearn.linear_model import Ridge  # noqa: E402
from sklearn.metrics import make_scorer  # noqa: E402
from sklearn.model_selection import GridSearchCV  # noqa: E402
from sklearn.neighbors import NearestNeighbors  # noqa: E402

from make_golden import KEYS, scores_to_array  # noqa: E402


def checksum(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def data_cases():
    rng = np.random.default_rng(21)
    out = {}
    for tag, dt in (("f32", np.float32), ("f64", np.float64)):
        tr = (rng.standard_normal((120, 90)) * rng.uniform(0.1, 30.0, 90) + rng.uniform(-5, 5, 90)).astype(dt)
        te = (rng.standard_normal((30, 90)) * 3.0).astype(dt)
        im = (rng.standard_normal((3, 90)) * 3.0).astype(dt)
        tr[0, [4, 50, 51]] = np.nan                       # dropped columns (nan_locations, searchlight.py:198)
        tr[:, 7] = 2.5                                    # constant feature: scale 1
        tr[:, 8] = 1e-3 * tr[:, 8] + 1e4                  # large offset, small spread
        lat_tr = (rng.standard_normal((120, 24)) * rng.uniform(0.5, 2.0, 24)).astype(dt)
        lat_te = rng.standard_normal((30, 24)).astype(dt)
        lat_im = rng.standard_normal((3, 24)).astype(dt)
        a = ref_data.standardize_fmri_betas(tr.copy(), te.copy(), im.copy())
        b = ref_data.standardize_latents(lat_tr.copy(), lat_te.copy(), lat_im.copy())
        c = ref_data.standardize_latents(lat_tr.copy(), lat_te.copy())
        for name, arr in zip(("tr", "te", "im", "lat_tr", "lat_te", "lat_im"), (tr, te, im, lat_tr, lat_te, lat_im)):
            out[f"{tag}_{name}"] = arr
        for name, arr in zip(("tr", "te", "im"), a):
            out[f"{tag}_betas_{name}_out"] = arr
        for name, arr in zip(("tr", "te", "im"), b):
            out[f"{tag}_latents_{name}_out"] = arr
        assert len(c) == 2 and np.array_equal(c[0], b[0])
    out["shuffled_indices_6"] = np.array(ref_data.create_null_distr_shuffled_indices(6))
    np.savez_compressed(os.path.join(HERE, "data_cases.npz"), **out)
    print("data_cases", {k: (v.shape, v.dtype) for k, v in out.items() if k.endswith("_out")})


CONFIG2 = dict(level=5, mesh_seed=7, trial_seed=0, n_train=4000, n_dims=1024, k=750, alpha=1.0)


def config2_inputs():
    mesh = syn.make_mesh(level=CONFIG2["level"], seed=CONFIG2["mesh_seed"])
    trials = syn.make_trials(mesh, n_train=CONFIG2["n_train"], n_dims=CONFIG2["n_dims"], seed=CONFIG2["trial_seed"])
    return syn.make_searchlight_inputs(mesh, trials)


def config2_case():
    inp = config2_inputs()
    k = CONFIG2["k"]
    dist, adj = NearestNeighbors().fit(inp.coords).kneighbors(inp.coords, n_neighbors=k)       # searchlight.py:173
    centres = np.concatenate([np.arange(0, len(adj), 331)[:32], np.arange(5000, 5016)])          # spread + one contiguous patch
    scores = ref_sl.custom_search_light(
        inp.X, inp.latents, estimator=Ridge(alpha=CONFIG2["alpha"], fit_intercept=False), A=adj[centres],
        train_ids=inp.train_ids, test_ids=inp.test_ids, imagery_ids=inp.imagery_ids, n_jobs=8)
    out = scores_to_array(scores)
    np.savez_compressed(os.path.join(HERE, "searchlight_config2.npz"), centres=centres, adjacency=adj[centres].astype(np.int32),
                        scores=out, keys=np.array(KEYS), x_sha256=checksum(inp.X), latents_sha256=checksum(inp.latents),
                        coords_sha256=checksum(inp.coords), **{f"cfg_{k_}": v for k_, v in CONFIG2.items()})
    print("config2", out.shape, "mean", out.mean(0).round(3))


def cv_case():
    """GridSearchCV per neighbourhood.  n_train = 300 gives validation folds of 60 rows."""
    mesh = syn.make_mesh(level=3, seed=19)
    trials = syn.make_trials(mesh, n_train=300, n_dims=48, n_factors=8, iid=False, seed=4, signal=0.6)
    inp = syn.make_searchlight_inputs(mesh, trials)
    k = 40
    dist, adj = NearestNeighbors().fit(inp.coords).kneighbors(inp.coords, n_neighbors=k)
    centres = np.arange(0, len(adj), 6)[:96]
    alphas = [1e-2, 1.0, 1e2, 1e3, 1e4, 1e6]
    X = inp.X.astype(np.float16)
    Y = inp.latents.astype(np.float16)
    tr = np.asarray(inp.train_ids)
    mean_scores, best, final = [], [], []
    for c in centres:
        Xc = X[:, adj[c]]
        clf = GridSearchCV(estimator=Ridge(fit_intercept=False), param_grid=dict(alpha=alphas),
                           scoring=make_scorer(ref_eval.pairwise_accuracy, greater_is_better=True),
                           cv=ref_ridge.NUM_CV_SPLITS, n_jobs=1, refit=True)          # ridge_regression_decoding.py:111-117
        clf.fit(Xc[tr], Y[tr])
        mean_scores.append(clf.cv_results_["mean_test_score"])
        best.append(clf.best_params_["alpha"])
        # the 8 searchlight scores of the selected alpha, through the reference's own train_and_test (searchlight.py:38-90)
        s = ref_sl.train_and_test(Ridge(alpha=clf.best_params_["alpha"], fit_intercept=False), Xc, Y, train_ids=inp.train_ids,
                                  test_ids=inp.test_ids, imagery_ids=inp.imagery_ids)
        final.append([s[k_] for k_ in KEYS])
    np.savez_compressed(os.path.join(HERE, "searchlight_cv.npz"), X=X, latents=Y, adjacency=adj.astype(np.int32), centres=centres,
                        n_train=len(inp.train_ids), n_test=len(inp.test_ids), n_imagery=len(inp.imagery_ids),
                        alphas=np.array(alphas), mean_test_score=np.array(mean_scores), best_alpha=np.array(best),
                        scores=np.array(final), keys=np.array(KEYS), n_folds=ref_ridge.NUM_CV_SPLITS)
    print("cv", np.array(mean_scores).shape, "selected", {a: int((np.array(best) == a).sum()) for a in alphas})


if __name__ == "__main__":
    which = sys.argv[1:] or ["data", "config2", "cv"]
    if "data" in which:
        data_cases()
    if "config2" in which:
        config2_case()
    if "cv" in which:
        cv_case()
