"""Generate the golden fixtures in this directory by running the UNMODIFIED reference.

Run from the repo root inside the build container (needs /root/reference):

    python tests/golden/make_golden.py

Outputs (committed):  searchlight_small.npz, searchlight_smooth.npz, searchlight_radius.npz, group_stats.npz, knn_small.npz,
eval_cases.npz, null_distr_small.npz, gridsearch_small.npz.
Each file stores the exact inputs next to the reference outputs so the tests never depend on
random-stream stability.
"""
import os
import pickle
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from multimodal_decoding_b200 import synthetic as syn  # noqa: E402
from _reference_loader import load_reference  # noqa: E402

ref_sl, ref_eval, ref_data, ref_ridge = load_reference()
from sklearn.linear_model import Ridge  # noqa: E402
from sklearn.neighbors import NearestNeighbors  # noqa: E402
from sklearn.metrics import make_scorer  # noqa: E402
from sklearn.model_selection import GridSearchCV  # noqa: E402

KEYS = [
    "pairwise_acc_captions", "pairwise_acc_images", "pairwise_acc_cross_captions_to_images",
    "pairwise_acc_cross_images_to_captions", "pairwise_acc_imagery", "pairwise_acc_imagery_whole_test_set",
    "pairwise_acc_imagery_no_std", "pairwise_acc_imagery_whole_test_set_no_std"]


def scores_to_array(scores):
    return np.array([[s[k] for k in KEYS] for s in scores], dtype=np.float64)


def searchlight_case(name, level, k, n_train, n_dims, alpha, iid, seed, nan_fraction=0.0, training_mode="agnostic"):
    mesh = syn.make_mesh(level=level, nan_fraction=nan_fraction, seed=7 + seed)
    trials = syn.make_trials(mesh, n_train=n_train, n_dims=n_dims, n_factors=8, iid=iid, seed=seed,
                             training_mode=training_mode, signal=0.6)
    inp = syn.make_searchlight_inputs(mesh, trials)
    dist, adj = NearestNeighbors().fit(inp.coords).kneighbors(inp.coords, n_neighbors=k)   # searchlight.py:173
    scores = ref_sl.custom_search_light(
        inp.X, inp.latents, estimator=Ridge(alpha=alpha, fit_intercept=False), A=adj,
        train_ids=inp.train_ids, test_ids=inp.test_ids, imagery_ids=inp.imagery_ids, n_jobs=1)
    out = scores_to_array(scores)
    np.savez_compressed(os.path.join(HERE, name), X=inp.X, latents=inp.latents, adjacency=adj, distances=dist,
                        coords=inp.coords, n_train=len(inp.train_ids), n_test=len(inp.test_ids),
                        n_imagery=len(inp.imagery_ids), alpha=alpha, scores=out, keys=np.array(KEYS))
    print(name, out.shape, "mean", out.mean(0).round(3))


def knn_case():
    mesh = syn.make_mesh(level=None, n_points=1500, seed=11)
    dist, adj = NearestNeighbors().fit(mesh.coords).kneighbors(mesh.coords, n_neighbors=48)
    mesh2 = syn.make_mesh(level=4, seed=3)     # 2562-vertex icosphere, near-regular spacing (many near-ties)
    dist2, adj2 = NearestNeighbors().fit(mesh2.coords).kneighbors(mesh2.coords, n_neighbors=40)
    np.savez_compressed(os.path.join(HERE, "knn_small.npz"), coords=mesh.coords, adjacency=adj.astype(np.int32),
                        distances=dist, coords2=mesh2.coords, adjacency2=adj2.astype(np.int32), distances2=dist2)
    print("knn", adj.shape, adj2.shape)


def radius_case():
    """Radius-mode neighbourhoods exactly as searchlight.py:160-166 builds them, then the unmodified loop on the
    ragged adjacency."""
    mesh = syn.make_mesh(level=3, seed=13)
    trials = syn.make_trials(mesh, n_train=200, n_dims=48, n_factors=8, iid=False, seed=5, signal=0.6)
    inp = syn.make_searchlight_inputs(mesh, trials)
    radius = 63.0
    nn = NearestNeighbors(radius=radius)
    adjacency = [np.argwhere(arr == 1)[:, 0] for arr in
                 nn.fit(inp.coords).radius_neighbors_graph(inp.coords).toarray()]      # searchlight.py:164-165
    lens = np.array([len(a) for a in adjacency])
    scores = ref_sl.custom_search_light(
        inp.X, inp.latents, estimator=Ridge(alpha=1.0, fit_intercept=False), A=adjacency,
        train_ids=inp.train_ids, test_ids=inp.test_ids, imagery_ids=inp.imagery_ids, n_jobs=1)
    out = scores_to_array(scores)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "searchlight_radius.npz"), X=inp.X, latents=inp.latents,
                        indices=np.concatenate(adjacency).astype(np.int32), offsets=offsets, coords=inp.coords,
                        radius=radius, n_train=len(inp.train_ids), n_test=len(inp.test_ids),
                        n_imagery=len(inp.imagery_ids), alpha=1.0, scores=out, keys=np.array(KEYS))
    print("radius", out.shape, "neighbours min/mean/max", lens.min(), lens.mean().round(1), lens.max(),
          "mean", out.mean(0).round(3))


def stats_case():
    """Group-level statistics: the unmodified t-map and TFCE functions on synthetic maps."""
    import importlib
    import types
    sys.modules.setdefault("h5py", types.ModuleType("h5py"))          # only used for file I/O in that module
    ref_perm = importlib.import_module("analyses.decoding.searchlight.searchlight_permutation_testing")
    ref_clu = importlib.import_module("analyses.cluster_analysis")
    rng = np.random.default_rng(17)
    out = {}
    meshes = {"left": syn.make_mesh(level=4, seed=31), "right": syn.make_mesh(level=4, seed=32)}
    hemis = list(ref_clu.HEMIS)
    t_values, edge_dicts = {}, {}
    for hemi in hemis:
        mesh = meshes[hemi]
        edges = syn.mesh_edges(mesh.faces)
        edges = np.array([(a, b) if a < b else (b, a) for a, b in edges])
        lengths = np.sqrt(((mesh.coords[edges[:, 0]] - mesh.coords[edges[:, 1]]) ** 2).sum(1))
        edge_dicts[hemi] = {(int(a), int(b)): float(l) for (a, b), l in zip(edges, lengths)}
        v = len(mesh.coords)
        # per-subject accuracy maps: chance 0.5 plus smooth blobs of signal; a few constant / NaN vertices
        centres = mesh.coords[rng.integers(0, v, 6)]
        d2 = ((mesh.coords[:, None, :] - centres[None]) ** 2).sum(-1)
        signal = (0.08 * np.exp(-0.5 * d2 / 25.0 ** 2)).sum(1)
        data = 0.5 + signal[None] + 0.03 * rng.standard_normal((6, v))
        data[:, 5] = 0.5
        data[:, 77] = data[0, 77]
        data[2, 300:310] = np.nan
        t = ref_perm.calc_image_t_values(data, 0.5)
        t_sigma = ref_perm.calc_image_t_values(data, 0.5, sigma=1e-3)
        t_values[hemi] = {"m": t}
        out[f"{hemi}_edges"], out[f"{hemi}_data"], out[f"{hemi}_t"], out[f"{hemi}_t_sigma"] = edges, data, t, t_sigma
        thr = 2.5
        cl = ref_clu.calc_clusters(t, thr, edge_dicts[hemi], return_clusters=True)["clusters"]
        lab = np.full(v, -1, dtype=np.int64)
        for i, c in enumerate(sorted(cl, key=min)):
            lab[list(c)] = i
        out[f"{hemi}_cluster_threshold"], out[f"{hemi}_cluster_labels"] = thr, lab
    for tag, kw in {"default": dict(), "h1_e05_dh02": dict(h=1, e=0.5, dh=0.2), "auto": dict(dh=0)}.items():
        tf = ref_clu.calc_tfce_values(t_values, edge_dicts, "m", **kw)
        for hemi in hemis:
            out[f"{hemi}_tfce_{tag}"] = tf[hemi]["m"]
    neg = {hemi: {"m": -np.abs(t_values[hemi]["m"])} for hemi in hemis}
    out["left_tfce_nonpositive"] = ref_clu.calc_tfce_values(neg, edge_dicts, "m")["left"]["m"]
    x = rng.standard_normal((6, 40)) + 0.3
    out["ttest_x"], out["ttest_t"] = x, ref_perm.ttest_1samp_no_p(x)
    out["ttest_t_rel"] = ref_perm.ttest_1samp_no_p(x, sigma=1e-2)
    out["ttest_t_abs"] = ref_perm.ttest_1samp_no_p(x, sigma=1e-2, method="absolute")
    np.savez_compressed(os.path.join(HERE, "group_stats.npz"), **out)
    print("stats", {k: (np.asarray(v).shape, str(np.asarray(v).dtype)) for k, v in out.items() if "tfce" in k},
          "max tfce", float(np.nanmax(out["left_tfce_default"])))


def eval_cases():
    rng = np.random.default_rng(5)
    out = {}
    lat = rng.standard_normal((140, 48))
    pred = lat * 0.3 + rng.standard_normal((140, 48))
    im_lat = rng.standard_normal((3, 48))
    im_pred = im_lat * 0.5 + rng.standard_normal((3, 48))
    out["lat"], out["pred"], out["im_lat"], out["im_pred"] = lat, pred, im_lat, im_pred
    out["pa_plain"] = ref_eval.pairwise_accuracy(lat, pred)
    out["pa_std"] = ref_eval.pairwise_accuracy(lat, pred, standardize_predictions=True)
    out["pa_nonsquare"] = ref_eval.pairwise_accuracy(np.concatenate([im_lat, lat]), im_pred)
    allsc = ref_eval.calc_all_pairwise_accuracy_scores(lat, pred, ref_data.TEST_STIM_TYPES, im_lat, im_pred,
                                                       standardize_predictions=True)
    out["all_keys"] = np.array(list(allsc.keys()))
    out["all_vals"] = np.array(list(allsc.values()))
    nostd = ref_eval.calc_imagery_pairwise_accuracy_scores(im_lat, im_pred, lat, standardize_predictions=False)
    out["im_nostd_keys"] = np.array(list(nostd.keys()))
    out["im_nostd_vals"] = np.array(list(nostd.values()))
    nocross = ref_eval.calc_all_pairwise_accuracy_scores(lat, pred, ref_data.TEST_STIM_TYPES, im_lat, im_pred,
                                                         standardize_predictions=False,
                                                         comp_cross_decoding_scores=False)
    out["nocross_keys"] = np.array(list(nocross.keys()))
    out["nocross_vals"] = np.array(list(nocross.values()))
    normed = ref_eval.calc_all_pairwise_accuracy_scores(lat, pred, ref_data.TEST_STIM_TYPES, im_lat, im_pred,
                                                        norm_imagery_preds_with_test_preds=True)
    out["normed_keys"] = np.array(list(normed.keys()))
    out["normed_vals"] = np.array(list(normed.values()))
    dm = rng.standard_normal((5, 9))
    out["dist_mat"] = dm
    out["dist_mat_acc"] = ref_eval.dist_mat_to_pairwise_acc(dm)
    out["distance_matrix"] = ref_eval.get_distance_matrix(pred[:7], lat[:5])
    out["test_stim_types"] = ref_data.TEST_STIM_TYPES
    perms = ref_data.create_null_distr_shuffled_indices(6)
    out["shuffled_indices"] = np.array(perms)
    np.savez_compressed(os.path.join(HERE, "eval_cases.npz"), **out)
    print("eval", {k: np.asarray(v).shape for k, v in out.items() if k.startswith("pa")})


def null_distr_case():
    mesh = syn.make_mesh(level=2, seed=21)            # 162 vertices
    trials = syn.make_trials(mesh, n_train=120, n_dims=24, n_factors=6, seed=4, signal=0.6)
    inp = syn.make_searchlight_inputs(mesh, trials)
    dist, adj = NearestNeighbors().fit(inp.coords).kneighbors(inp.coords, n_neighbors=12)
    perms = ref_data.create_null_distr_shuffled_indices(5)
    centres = [0, 5, 77]
    with tempfile.TemporaryDirectory() as d:
        np.random.seed(1234)                          # makes the unseeded draw of searchlight.py:65 reproducible
        ref_sl.custom_group_iter_search_light(
            [adj[c] for c in centres], centres, Ridge(alpha=1.0, fit_intercept=False), inp.X, inp.latents,
            inp.train_ids, inp.test_ids, inp.imagery_ids, thread_id=1, null_distr_dir=d, shuffled_indices=perms)
        files = sorted(os.listdir(d))
        null = [pickle.load(open(os.path.join(d, f), "rb")) for f in files]
    np.random.seed(1234)
    drawn = [int(np.random.choice(2)) for _ in range(len(centres) * len(perms))]
    null_keys = list(null[0][0].keys())
    arr = np.array([[[p[k] for k in null_keys] for p in per_centre] for per_centre in null])
    np.savez_compressed(os.path.join(HERE, "null_distr_small.npz"), X=inp.X, latents=inp.latents, adjacency=adj,
                        n_train=len(inp.train_ids), n_test=len(inp.test_ids), n_imagery=len(inp.imagery_ids),
                        perms=np.array(perms), centres=np.array(centres), file_names=np.array(files),
                        derangement_choice=np.array(drawn).reshape(len(centres), len(perms)),
                        derangements=np.array(ref_sl.DERANGEMENTS_THREE_DIMS),
                        null_keys=np.array(null_keys), null_scores=arr)
    print("null", arr.shape, files)


def gridsearch_case():
    rng = np.random.default_rng(9)
    out = {}
    for tag, (n, v, d) in {"dual": (103, 260, 20), "primal": (203, 60, 20)}.items():
        z = rng.standard_normal((n + 140, 6))
        x = (z @ rng.standard_normal((6, v)) + rng.standard_normal((n + 140, v))).astype(np.float32)
        y = (0.4 * z @ rng.standard_normal((6, d)) + rng.standard_normal((n + 140, d))).astype(np.float32)
        alphas = [1e-1, 1e1, 1e2, 1e3, 1e5]
        clf = GridSearchCV(estimator=Ridge(fit_intercept=False), param_grid=dict(alpha=alphas),
                           scoring=make_scorer(ref_eval.pairwise_accuracy, greater_is_better=True),
                           cv=ref_ridge.NUM_CV_SPLITS, n_jobs=1, refit=True)      # ridge_regression_decoding.py:111-117
        clf.fit(x[:n], y[:n])
        out[f"{tag}_x"], out[f"{tag}_y"], out[f"{tag}_n_train"] = x, y, n
        out[f"{tag}_alphas"] = np.array(alphas)
        out[f"{tag}_best_alpha"] = clf.best_params_["alpha"]
        out[f"{tag}_mean_test_score"] = clf.cv_results_["mean_test_score"]
        out[f"{tag}_pred_test"] = clf.best_estimator_.predict(x[n:])
    np.savez_compressed(os.path.join(HERE, "gridsearch_small.npz"), **out)
    print("gridsearch", out["dual_best_alpha"], out["primal_best_alpha"], out["dual_mean_test_score"].round(3),
          out["primal_mean_test_score"].round(3))


if __name__ == "__main__":
    # 642-vertex icosphere, iid betas (well conditioned) and smooth betas with NaN medial wall
    searchlight_case("searchlight_small.npz", level=3, k=40, n_train=300, n_dims=64, alpha=1.0, iid=True, seed=0)
    searchlight_case("searchlight_smooth.npz", level=3, k=48, n_train=260, n_dims=96, alpha=1.0, iid=False, seed=1,
                     nan_fraction=0.08)
    searchlight_case("searchlight_captions.npz", level=3, k=33, n_train=400, n_dims=32, alpha=10.0, iid=False,
                     seed=2, training_mode="captions")
    knn_case()
    radius_case()
    stats_case()
    eval_cases()
    null_distr_case()
    gridsearch_case()
