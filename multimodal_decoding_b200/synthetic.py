"""Synthetic stand-ins for the data the reference loads from disk.

The reference reads per-trial GIfTI betas and pickled embedding dicts
(`/root/reference/data.py:448-503`, `:401-445`) and the fsaverage inflated mesh
(`analyses/decoding/searchlight/searchlight.py:152-158`).  None of that exists
on the build or GPU boxes, so the generators below produce arrays with the same
contracts (shapes, dtypes, ordering) from a seed:

* `icosphere(level)`          -> mesh with 10*4**level + 2 vertices (level 7 = 163,842 = fsaverage7)
* `make_mesh(...)`            -> jittered, tie-free coordinates (+ optional medial-wall NaN mask)
* `make_trials(...)`          -> betas (N x V) and latents (N x D) with shared low-rank structure
* `make_searchlight_inputs`   -> exactly the arrays `custom_search_light` receives
                                 (`searchlight.py:229-248`): fp16 X, fp16 latents, id lists

Everything is numpy on the host; sizes used by the tests finish in seconds.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import scipy.sparse as sp

N_TEST_PER_MODALITY = 70       # data.py:182-187  (70 captions then 70 images)
N_IMAGERY = 3                  # data.py:27-40    (three imagery stimuli)


# ----------------------------------------------------------------------------------------------
# mesh
# ----------------------------------------------------------------------------------------------
def _icosahedron():
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0],
                  [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11],
                  [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9],
                  [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    return v, f


def icosphere(level: int):
    """Unit icosphere; vertices of coarser levels come first (the fsaverage ordering class)."""
    v, f = _icosahedron()
    for _ in range(level):
        nv = len(v)
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        e.sort(axis=1)
        key = e[:, 0] * nv + e[:, 1]
        ukey, inv = np.unique(key, return_inverse=True)
        a, b = ukey // nv, ukey % nv
        mid = v[a] + v[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        v = np.concatenate([v, mid])
        m = nv + inv.reshape(3, -1)           # midpoints of edges 01, 12, 20 per face
        m01, m12, m20 = m[0], m[1], m[2]
        f = np.concatenate([
            np.stack([f[:, 0], m01, m20], 1),
            np.stack([f[:, 1], m12, m01], 1),
            np.stack([f[:, 2], m20, m12], 1),
            np.stack([m01, m12, m20], 1)])
    return v, f


def mesh_edges(faces: np.ndarray) -> np.ndarray:
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]])
    e.sort(axis=1)
    return np.unique(e, axis=0)


def random_sphere_points(n: int, rng) -> np.ndarray:
    p = rng.standard_normal((n, 3))
    return p / np.linalg.norm(p, axis=1, keepdims=True)


@dataclass
class Mesh:
    coords: np.ndarray          # (V_full, 3) float32, like nilearn's load_surf_mesh output
    faces: np.ndarray | None    # (F, 3) int64 or None for point clouds
    nan_locations: np.ndarray   # (V_full,) bool: vertices without data ("medial wall")


def make_mesh(level: int | None = 7, n_points: int | None = None, radius_mm: float = 100.0,
              jitter_mm: float = 0.5, nan_fraction: float = 0.0, seed: int = 7) -> Mesh:
    """Sphere-like cortical surface stand-in.

    `level` gives an icosphere (level 7 -> 163,842 vertices); `n_points` gives a random point
    cloud of any size instead.  A smooth low-order radial deformation plus iid jitter removes
    exact distance ties so k-NN indices are well defined (SURVEY.md §7 "bit-exact neighbour lists").
    """
    rng = np.random.default_rng(seed)
    if n_points is not None:
        unit, faces = random_sphere_points(n_points, rng), None
    else:
        unit, faces = icosphere(level)
    # smooth radial bumps (a few random low-frequency lobes)
    lobes = random_sphere_points(6, rng)
    amp = rng.uniform(-0.08, 0.08, size=6)
    r = 1.0 + (np.exp(4.0 * (unit @ lobes.T - 1.0)) * amp).sum(1)
    coords = unit * (radius_mm * r)[:, None]
    coords = coords + rng.normal(scale=jitter_mm, size=coords.shape)
    coords = coords.astype(np.float32)
    nan_locations = np.zeros(len(coords), dtype=bool)
    if nan_fraction > 0:
        pole = random_sphere_points(1, rng)[0]
        proj = unit @ pole
        nan_locations = proj > np.quantile(proj, 1.0 - nan_fraction)
    return Mesh(coords=coords, faces=faces, nan_locations=nan_locations)


# ----------------------------------------------------------------------------------------------
# trials
# ----------------------------------------------------------------------------------------------
@dataclass
class Trials:
    train_fmri: np.ndarray      # (N_train, V) float32, may contain NaN columns
    test_fmri: np.ndarray       # (140, V)
    imagery_fmri: np.ndarray    # (3, V)
    train_latents: np.ndarray   # (N_train, D) float32
    test_latents: np.ndarray    # (140, D)
    imagery_latents: np.ndarray  # (3, D)
    test_stim_types: np.ndarray  # (140,) "caption"/"image"


def _smooth_operator(n_vertices: int, edges: np.ndarray, lam: float = 0.5):
    a = sp.coo_matrix((np.ones(len(edges)), (edges[:, 0], edges[:, 1])), shape=(n_vertices, n_vertices))
    a = (a + a.T).tocsr()
    deg = np.asarray(a.sum(1)).ravel()
    deg[deg == 0] = 1
    w = sp.diags(1.0 / deg) @ a
    return ((1 - lam) * sp.identity(n_vertices) + lam * w).tocsr()


def make_trials(mesh: Mesh, n_train: int = 4000, n_dims: int = 1024, n_factors: int = 32,
                smooth_steps: int = 4, signal: float = 0.35, bump_sigma_mm: float = 20.0,
                iid: bool = False, seed: int = 0, training_mode: str = "agnostic",
                n_test_per_mod: int = N_TEST_PER_MODALITY, n_imagery: int = N_IMAGERY) -> Trials:
    """Betas = stimulus factors x smooth spatial loadings + spatially correlated noise.

    `iid=True` gives the well-conditioned control (white N(0,1) betas).  Test caption i and test
    image i share a stimulus (hence the same latent), as with the reference's `avg` features
    (`data.py:217-224`).  Train rows are captions first, then images (`data.py:353-355` glob order);
    `training_mode` "captions"/"images" keeps only that half (`data.py:14-17`).
    """
    rng = np.random.default_rng(seed)
    v_full = len(mesh.coords)
    n_test = 2 * n_test_per_mod
    n_stim_train = n_train
    z_train = rng.standard_normal((n_stim_train, n_factors))
    z_test_half = rng.standard_normal((n_test_per_mod, n_factors))
    z_test = np.concatenate([z_test_half, z_test_half])
    z_im = rng.standard_normal((n_imagery, n_factors))
    mix = rng.standard_normal((n_factors, n_dims))

    def latents(z):
        return (z @ mix + 0.5 * rng.standard_normal((len(z), n_dims))).astype(np.float32)

    y_train, y_im = latents(z_train), latents(z_im)
    y_test_half = latents(z_test_half)
    y_test = np.concatenate([y_test_half, y_test_half])

    n_all = n_stim_train + n_test + n_imagery
    z_all = np.concatenate([z_train, z_test, z_im])
    noise = rng.standard_normal((v_full, n_all)).astype(np.float32)   # vertex-major for the smoother
    if iid:
        x = noise.T
    else:
        centres = mesh.coords[rng.integers(0, v_full, size=n_factors)].astype(np.float64)
        d2 = ((mesh.coords[:, None, :].astype(np.float64) - centres[None]) ** 2).sum(-1)
        load = np.exp(-0.5 * d2 / bump_sigma_mm ** 2).astype(np.float32)       # (V, r)
        if mesh.faces is not None and smooth_steps > 0:
            op = _smooth_operator(v_full, mesh_edges(mesh.faces)).astype(np.float32)
            for _ in range(smooth_steps):
                noise = op @ noise
            noise /= noise.std(axis=1, keepdims=True)
        x = (signal * (load @ z_all.T.astype(np.float32)) + noise).T
    x = np.ascontiguousarray(x, dtype=np.float32)
    x[:, mesh.nan_locations] = np.nan
    tr, te, im = x[:n_stim_train], x[n_stim_train:n_stim_train + n_test], x[n_stim_train + n_test:]
    if training_mode in ("captions", "images"):
        half = n_stim_train // 2
        sl = slice(0, half) if training_mode == "captions" else slice(half, None)
        tr, y_train = tr[sl], y_train[sl]
    stim_types = np.array(["caption"] * n_test_per_mod + ["image"] * n_test_per_mod)
    return Trials(tr, te, im, y_train, y_test, y_im, stim_types)


# ----------------------------------------------------------------------------------------------
# the arrays custom_search_light receives
# ----------------------------------------------------------------------------------------------
@dataclass
class SearchlightInputs:
    X: np.ndarray               # (N_total, V) float16
    latents: np.ndarray         # (N_total, D) float16
    train_ids: list
    test_ids: list
    imagery_ids: list
    nan_locations: np.ndarray
    coords: np.ndarray          # (V, 3) coordinates of the kept vertices (float32)


def _zscore_by_train(train, *others):
    """StandardScaler semantics (ddof=0, zero-variance columns keep scale 1); data.py:506-537."""
    train = train.astype(np.float64)
    mu = train.mean(0)
    sd = train.std(0)
    sd[sd == 0] = 1.0
    return [(a.astype(np.float64) - mu) / sd for a in (train,) + others]


def make_searchlight_inputs(mesh: Mesh, trials: Trials) -> SearchlightInputs:
    """The pre-steps of `searchlight.run` (`searchlight.py:198-248`) on synthetic trials."""
    nan_locations = np.isnan(trials.train_fmri[0])
    keep = ~nan_locations
    tr, te, im = _zscore_by_train(trials.train_fmri[:, keep], trials.test_fmri[:, keep],
                                  trials.imagery_fmri[:, keep])
    ytr, yte, yim = _zscore_by_train(trials.train_latents, trials.test_latents, trials.imagery_latents)
    n_tr, n_te, n_im = len(tr), len(te), len(im)
    X = np.concatenate([tr, te, im]).astype(np.float16)
    latents = np.concatenate([ytr, yte, yim]).astype(np.float16)
    return SearchlightInputs(
        X=X, latents=latents,
        train_ids=list(range(n_tr)),
        test_ids=list(range(n_tr, n_tr + n_te)),
        imagery_ids=list(range(n_tr + n_te, n_tr + n_te + n_im)),
        nan_locations=nan_locations, coords=mesh.coords[keep])
