"""Host-side handle around the C ABI: one `Engine` per GPU.

Inputs may be numpy arrays (host buffers, copied by the library) or CUDA torch tensors on the engine's
device (borrowed).  Nothing here computes on the CPU: every method forwards to libmmd_b200.so and raises
`RuntimeError` with the library's message on failure.
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np

from . import _lib

SCORE_KEYS = None


def score_keys():
    global SCORE_KEYS
    if SCORE_KEYS is None:
        lib = _lib.load()
        SCORE_KEYS = tuple(lib.mmd_score_name(i).decode() for i in range(_lib.N_SCORES))
    return SCORE_KEYS


def _ptr(a):
    """(address, keep-alive object) of a numpy array or a torch tensor."""
    if a is None:
        return None, None
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            a = np.ascontiguousarray(a)
        return a.ctypes.data, a
    if hasattr(a, "data_ptr"):                      # torch tensor
        if not a.is_contiguous():
            a = a.contiguous()
        return a.data_ptr(), a
    raise TypeError(f"unsupported buffer type {type(a)}")


class Engine:
    live = weakref.WeakSet()         # every open engine (the guard-mode check of tests/conftest.py walks it)

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        rc = self._lib.mmd_create(int(device), C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"mmd_create(device={device}) failed with code {rc}: no usable CUDA device "
                               "(the searchlight engine has no CPU fallback)")
        self.device = int(device)
        self.n_dims = None
        self.n_eval = None
        Engine.live.add(self)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.mmd_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            msg = self._lib.mmd_last_error(self._h).decode()
            # sklearn's input validation (Ridge.fit / predict -> check_array) raises ValueError for NaN / infinity
            raise ValueError(msg) if msg.startswith("Input ") else RuntimeError(msg)

    def configure(self, cluster_size=0, batch_centres=0, host_threads=0, group_size=0):
        self._check(self._lib.mmd_configure(self._h, cluster_size, batch_centres, host_threads))
        if group_size:
            self._check(self._lib.mmd_set_group_size(self._h, group_size))

    # -- adjacency ---------------------------------------------------------------------------
    def knn(self, coords, k: int, q_begin: int = 0, q_end: int | None = None):
        coords = np.ascontiguousarray(np.asarray(coords, dtype=np.float64))
        if coords.ndim != 2 or coords.shape[1] != 3:
            raise ValueError("coords must be (n_vertices, 3)")
        v = len(coords)
        q_end = v if q_end is None else q_end
        idx = np.empty((q_end - q_begin, k), dtype=np.int64)
        dist = np.empty((q_end - q_begin, k), dtype=np.float64)
        self._check(self._lib.mmd_knn(self._h, coords.ctypes.data, v, int(k), int(q_begin), int(q_end),
                                      idx.ctypes.data, dist.ctypes.data))
        return dist, idx

    def radius_neighbors(self, coords, radius: float):
        """The radius branch of get_adjacency_matrix (searchlight.py:164-166): a list with one int64 index array per
        vertex (ascending vertex id, the vertex itself included)."""
        coords = np.ascontiguousarray(np.asarray(coords, dtype=np.float64))
        if coords.ndim != 2 or coords.shape[1] != 3:
            raise ValueError("coords must be (n_vertices, 3)")
        v = len(coords)
        counts = np.zeros(v, dtype=np.int32)
        self._check(self._lib.mmd_radius_neighbors(self._h, coords.ctypes.data, v, float(radius), 0, v, None,
                                                   counts.ctypes.data, None))
        offsets = np.zeros(v + 1, dtype=np.int64)
        np.cumsum(counts, out=offsets[1:])
        indices = np.empty(max(int(offsets[-1]), 1), dtype=np.int32)
        self._check(self._lib.mmd_radius_neighbors(self._h, coords.ctypes.data, v, float(radius), 0, v,
                                                   offsets.ctypes.data, None, indices.ctypes.data))
        idx64 = indices[:offsets[-1]].astype(np.int64)
        return [idx64[offsets[i]:offsets[i + 1]] for i in range(v)]

    # -- trials ------------------------------------------------------------------------------
    def upload(self, X, Y, train_ids, test_ids, imagery_ids, test_is_caption, active_vertices=None):
        """`active_vertices` (optional, ascending vertex ids): the vertices the coming searchlight calls will use
        (mmd_upload_active: X^T Y rows and digit planes are built for them only).

        Ordering contract for borrowed CUDA tensors: the library reads them on its own non-blocking stream, so the
        producing torch stream is synchronised here first (see include/mmd_b200.h, mmd_upload)."""
        n_total, n_vertices = X.shape
        if Y.shape[0] != n_total:
            raise ValueError(f"X has {n_total} trials but y has {Y.shape[0]}")
        for name, a in (("X", X), ("y", Y)):
            dt = str(a.dtype).replace("torch.", "")
            if dt != "float16":
                raise TypeError(f"{name} must be float16 (got {dt}); cast like searchlight.py:247-248 does")
        tr = np.ascontiguousarray(np.asarray(train_ids, dtype=np.int32))
        te = np.ascontiguousarray(np.asarray(test_ids, dtype=np.int32))
        im = np.ascontiguousarray(np.asarray(imagery_ids, dtype=np.int32))
        cap = np.ascontiguousarray(np.asarray(test_is_caption, dtype=np.uint8))
        if len(cap) != len(te):
            raise ValueError("test_is_caption must have one entry per test trial")
        xp, xk = _ptr(X)
        yp, yk = _ptr(Y)
        if hasattr(xk, "data_ptr") or hasattr(yk, "data_ptr"):
            import torch
            for t in (xk, yk):
                if hasattr(t, "data_ptr") and t.is_cuda:
                    torch.cuda.current_stream(t.device).synchronize()     # whatever produced the tensor has finished
        act = None if active_vertices is None else np.ascontiguousarray(np.asarray(active_vertices, dtype=np.int32))
        self._check(self._lib.mmd_upload_active(self._h, xp, n_total, n_vertices, yp, int(Y.shape[1]),
                                                tr.ctypes.data, len(tr), te.ctypes.data, len(te), cap.ctypes.data,
                                                im.ctypes.data, len(im), None if act is None else act.ctypes.data,
                                                0 if act is None else len(act)))
        self.n_dims = int(Y.shape[1])
        self.n_eval = len(te) + len(im)
        del xk, yk

    # -- the hot path ------------------------------------------------------------------------
    @staticmethod
    def _is_ragged(adjacency):
        if isinstance(adjacency, np.ndarray) and adjacency.dtype != object:
            return adjacency.ndim != 2
        lens = {len(r) for r in adjacency}
        return len(lens) > 1

    def searchlight_ragged(self, adjacency, alpha: float = 1.0, anchors=None):
        """Radius-mode neighbourhoods (searchlight.py:164-171): `adjacency` is a sequence of index arrays of
        different lengths.  `anchors[i]` = the vertex centre i sits on (default: i when there is one centre per
        vertex, else the first entry of each list); it only steers the grouping of centres, never the result."""
        rows = [np.asarray(r, dtype=np.int32).ravel() for r in adjacency]
        n = len(rows)
        offsets = np.zeros(n + 1, dtype=np.int64)
        np.cumsum([len(r) for r in rows], out=offsets[1:])
        indices = np.ascontiguousarray(np.concatenate(rows)) if n else np.zeros(0, dtype=np.int32)
        anc = None if anchors is None else np.ascontiguousarray(np.asarray(anchors, dtype=np.int32))
        if anc is not None and len(anc) != n:
            raise ValueError("anchors must have one entry per centre")
        scores = np.empty((n, _lib.N_SCORES), dtype=np.float64)
        flags = np.zeros(n, dtype=np.int32)
        self._check(self._lib.mmd_searchlight_ragged(self._h, indices.ctypes.data, offsets.ctypes.data,
                                                     None if anc is None else anc.ctypes.data, n, float(alpha),
                                                     scores.ctypes.data, flags.ctypes.data))
        return scores, flags

    def searchlight(self, adjacency, alpha: float = 1.0):
        if self._is_ragged(adjacency):
            return self.searchlight_ragged(adjacency, alpha)
        adj = np.asarray(adjacency)
        if adj.ndim != 2:
            raise ValueError("adjacency must be a sequence of index arrays")
        # sklearn's int64 indices go to the library as they are (converted there by host threads)
        wide = adj.dtype == np.int64
        adj = np.ascontiguousarray(adj if wide else adj.astype(np.int32, copy=False))
        n, k = adj.shape
        scores = np.empty((n, _lib.N_SCORES), dtype=np.float64)
        flags = np.zeros(n, dtype=np.int32)
        fn = self._lib.mmd_searchlight_i64 if wide else self._lib.mmd_searchlight
        self._check(fn(self._h, adj.ctypes.data, n, k, float(alpha), scores.ctypes.data, flags.ctypes.data))
        return scores, flags

    def searchlight_cv(self, adjacency, alphas, fold_bounds):
        """Alpha grid x K-fold cross-validation per centre (mmd_searchlight_cv): returns (fold_scores (n, n_alphas,
        n_folds), flags (n,)).  `fold_bounds`: n_folds + 1 positions into train_ids (contiguous folds, KFold order)."""
        adj = np.ascontiguousarray(np.asarray(adjacency, dtype=np.int32))
        if adj.ndim != 2:
            raise ValueError("adjacency must be rectangular (n_centres, k)")
        al = np.ascontiguousarray(np.asarray(alphas, dtype=np.float64))
        fb = np.ascontiguousarray(np.asarray(fold_bounds, dtype=np.int32))
        n, k = adj.shape
        out = np.empty((n, len(al), len(fb) - 1), dtype=np.float64)
        flags = np.zeros(n, dtype=np.int32)
        self._check(self._lib.mmd_searchlight_cv(self._h, adj.ctypes.data, n, k, al.ctypes.data, len(al), fb.ctypes.data,
                                                 len(fb) - 1, out.ctypes.data, flags.ctypes.data))
        return out, flags

    # -- whole-brain ridge (ridge_regression_decoding.py:111-135) on the library's dense float64 kernels ------------
    @staticmethod
    def _dense(a):
        """float32 / float64 C-contiguous numpy array (other dtypes are widened to float64) and its is_f32 flag."""
        a = np.asarray(a)
        if a.dtype not in (np.float32, np.float64):
            a = a.astype(np.float64)
        a = np.ascontiguousarray(a)
        return a, int(a.dtype == np.float32)

    def ridge_load(self, X, Y):
        x, xf = self._dense(X)
        y, yf = self._dense(Y)
        if x.ndim != 2 or y.ndim != 2 or len(x) != len(y):
            raise ValueError(f"X {x.shape} and Y {y.shape} must be 2-D with one row per trial")
        self._check(self._lib.mmd_ridge_load(self._h, x.ctypes.data, xf, x.shape[0], x.shape[1], y.ctypes.data, yf, y.shape[1]))
        self._ridge_shape = (x.shape[0], x.shape[1], y.shape[1])

    def ridge_cv(self, alphas, fold_bounds):
        """(n_alphas, n_folds) fold scores of the loaded problem (mmd_ridge_cv)."""
        al = np.ascontiguousarray(np.asarray(alphas, dtype=np.float64))
        fb = np.ascontiguousarray(np.asarray(fold_bounds, dtype=np.int32))
        out = np.empty((len(al), len(fb) - 1), dtype=np.float64)
        self._check(self._lib.mmd_ridge_cv(self._h, al.ctypes.data, len(al), fb.ctypes.data, len(fb) - 1, out.ctypes.data))
        return out

    def ridge_fit(self, alpha, want_coef=True):
        """Refit on all trials; returns sklearn's coef_ (d x v) when asked."""
        n, v, d = self._ridge_shape
        coef = np.empty((d, v), dtype=np.float64) if want_coef else None
        self._check(self._lib.mmd_ridge_fit(self._h, float(alpha), None if coef is None else coef.ctypes.data))
        return coef

    def ridge_predict(self, X):
        x, xf = self._dense(X)
        n, v, d = self._ridge_shape
        if x.ndim != 2 or x.shape[1] != v:
            raise ValueError(f"X has {x.shape[1] if x.ndim == 2 else '?'} features, but the ridge was fitted with {v}")
        out = np.empty((x.shape[0], d), dtype=np.float64)
        self._check(self._lib.mmd_ridge_predict(self._h, x.ctypes.data, xf, x.shape[0], out.ctypes.data))
        return out

    def searchlight_null(self, adjacency, alpha, perm_test, perm_imagery, anchors=None):
        """Scores plus the null-distribution scores (searchlight.py:61-79) for target permutations given as index
        arrays into the test / imagery trials; returns (scores (n,8), flags (n,), null (n, n_perm, 6)).
        `perm_imagery`: (n_perm, n_imagery) shared by all centres, or (n_centres, n_perm, n_imagery) -- one draw per centre
        and permutation like the reference (mmd_searchlight_null_pc).  `adjacency` may be ragged (radius mode)."""
        ragged = self._is_ragged(adjacency)
        if ragged:
            rows = [np.asarray(r, dtype=np.int32).ravel() for r in adjacency]
            n = len(rows)
            offsets = np.zeros(n + 1, dtype=np.int64)
            np.cumsum([len(r) for r in rows], out=offsets[1:])
            indices = np.ascontiguousarray(np.concatenate(rows)) if n else np.zeros(0, dtype=np.int32)
            anc = None if anchors is None else np.ascontiguousarray(np.asarray(anchors, dtype=np.int32))
        else:
            adj = np.ascontiguousarray(np.asarray(adjacency, dtype=np.int32))
            if adj.ndim != 2:
                raise ValueError("adjacency must be (n_centres, k) or a sequence of index arrays")
            n, k = adj.shape
        pt = np.ascontiguousarray(np.asarray(perm_test, dtype=np.int32))
        pi = np.ascontiguousarray(np.asarray(perm_imagery, dtype=np.int32))
        per_centre = pi.ndim == 3
        if pt.ndim != 2 or pi.ndim not in (2, 3) or pi.shape[-2] != len(pt) or (per_centre and len(pi) != n):
            raise ValueError("perm_test (n_perm, n_test) and perm_imagery (n_perm, n_imagery) or (n_centres, n_perm, n_imagery) "
                             "must have one row per permutation")
        n_perm = len(pt)
        scores = np.empty((n, _lib.N_SCORES), dtype=np.float64)
        flags = np.zeros(n, dtype=np.int32)
        null = np.empty((n, n_perm, _lib.N_NULL_SCORES), dtype=np.float64)
        shared = np.ascontiguousarray(pi[0]) if per_centre else pi
        pc = pi.ctypes.data if per_centre else None
        if ragged:
            self._check(self._lib.mmd_searchlight_null_ragged(
                self._h, indices.ctypes.data, offsets.ctypes.data, None if anc is None else anc.ctypes.data, n, float(alpha),
                pt.ctypes.data, shared.ctypes.data, pc, n_perm, scores.ctypes.data, flags.ctypes.data, null.ctypes.data))
        else:
            self._check(self._lib.mmd_searchlight_null_pc(self._h, adj.ctypes.data, n, k, float(alpha), pt.ctypes.data,
                                                          shared.ctypes.data, pc, n_perm, scores.ctypes.data,
                                                          flags.ctypes.data, null.ctypes.data))
        return scores, flags, null

    # -- eval.py primitives (arbitrary shapes, float64) -----------------------------------------
    @staticmethod
    def _f64(a, name):
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
        if a.ndim != 2:
            raise ValueError(f"{name} must be 2-D (rows x features)")
        return a

    def cosine_distances(self, predictions, originals):
        p, o = self._f64(predictions, "predictions"), self._f64(originals, "originals")
        if p.shape[1] != o.shape[1]:
            raise ValueError(f"XA and XB must have the same number of columns ({p.shape[1]} != {o.shape[1]})")
        out = np.empty((len(p), len(o)), dtype=np.float64)
        self._check(self._lib.mmd_cosine_distances(self._h, p.ctypes.data, len(p), o.ctypes.data, len(o), p.shape[1],
                                                   out.ctypes.data))
        return out

    def dist_mat_to_pairwise_acc(self, dist_mat):
        d = self._f64(dist_mat, "dist_mat")
        acc = C.c_double()
        self._check(self._lib.mmd_dist_mat_to_pairwise_acc(self._h, d.ctypes.data, d.shape[0], d.shape[1], C.byref(acc)))
        return acc.value

    def pairwise_accuracy(self, latents, predictions, standardize_predictions=False, standardize_latents=False):
        lat, pred = self._f64(latents, "latents"), self._f64(predictions, "predictions")
        if lat.shape[1] != pred.shape[1]:
            raise ValueError(f"XA and XB must have the same number of columns ({pred.shape[1]} != {lat.shape[1]})")
        acc = C.c_double()
        self._check(self._lib.mmd_pairwise_accuracy(self._h, lat.ctypes.data, len(lat), pred.ctypes.data, len(pred),
                                                    lat.shape[1], int(bool(standardize_predictions)),
                                                    int(bool(standardize_latents)), C.byref(acc)))
        return acc.value

    # -- group-level statistics on score maps -----------------------------------
    @staticmethod
    def _edges(edges):
        e = np.ascontiguousarray(np.asarray(edges, dtype=np.int32))
        if e.ndim != 2 or e.shape[1] != 2:
            raise ValueError("edges must be (n_edges, 2) vertex pairs")
        return e

    def ttest_1samp(self, data, popmean: float, sigma: float = 0.0, relative: bool = True, constant_is_nan: bool = True):
        d = self._f64(data, "data")
        out = np.empty(d.shape[1], dtype=np.float64)
        self._check(self._lib.mmd_ttest_1samp(self._h, d.ctypes.data, d.shape[0], d.shape[1], float(popmean),
                                              float(sigma), int(bool(relative)) | (int(bool(constant_is_nan)) << 1),
                                              out.ctypes.data))
        return out

    def tfce(self, values, edges, h: float = 2.0, e: float = 1.0, dh: float = 0.1):
        """values: (V,) or (n_maps, V); returns float32 of the same shape."""
        v = np.asarray(values, dtype=np.float64)
        maps = np.ascontiguousarray(np.atleast_2d(v))
        ed = self._edges(edges)
        out = np.empty(maps.shape, dtype=np.float32)
        self._check(self._lib.mmd_tfce(self._h, maps.ctypes.data, maps.shape[0], maps.shape[1], ed.ctypes.data, len(ed),
                                       float(h), float(e), float(dh), out.ctypes.data))
        return out.reshape(v.shape)

    def cluster_labels(self, values, edges, threshold: float):
        v = np.ascontiguousarray(np.asarray(values, dtype=np.float64).ravel())
        ed = self._edges(edges)
        out = np.empty(len(v), dtype=np.int32)
        self._check(self._lib.mmd_cluster_labels(self._h, v.ctypes.data, len(v), ed.ctypes.data, len(ed),
                                                 float(threshold), out.ctypes.data))
        return out

    def predict_centre(self, neighbours, alpha: float = 1.0):
        nb = np.ascontiguousarray(np.asarray(neighbours, dtype=np.int32))
        out = np.empty((self.n_eval, self.n_dims), dtype=np.float64)
        self._check(self._lib.mmd_predict_centre(self._h, nb.ctypes.data, len(nb), float(alpha), out.ctypes.data))
        return out

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t of the handle (wrap with torch.cuda.ExternalStream to record events on it)."""
        return int(self._lib.mmd_stream(self._h) or 0)

    def check_guards(self):
        """Guard mode (MMD_B200_GUARD=1): raise if a kernel wrote outside one of the library's device buffers."""
        if getattr(self, "_h", None) and self._lib.mmd_check_guards(self._h) != 0:
            raise RuntimeError(self._lib.mmd_last_error(self._h).decode())

    def guard_selftest(self) -> int:
        """0 = guard mode is on and detects a one-byte overrun on either side; -1 = guard mode off."""
        return int(self._lib.mmd_guard_selftest(self._h))

    def set_predict_mode(self, mode: int):
        """1 (default): predictions on the tcgen05 tensor cores (int8 digit GEMM, guarded); 0: float64 DMMA kernel.
        Call before upload()."""
        self._check(self._lib.mmd_set_predict_mode(self._h, int(mode)))

    def set_syrk_mode(self, mode: int):
        """1 = exact integer cluster Gram on tcgen05 (default), 0 = float64 DMMA kernel."""
        self._check(self._lib.mmd_set_syrk_mode(self._h, int(mode)))

    def enable_margins(self, on=True):
        self._check(self._lib.mmd_enable_margins(self._h, 1 if on else 0))

    def margins(self, n_centres: int) -> np.ndarray:
        """Per centre: smallest |d(own candidate) - d(other candidate)| over the comparisons of the last searchlight call."""
        out = np.empty(n_centres, dtype=np.float64)
        self._check(self._lib.mmd_get_margins(self._h, out.ctypes.data_as(C.c_void_p), int(n_centres)))
        return out

    # -- statistics --------------------------------------------------------------------------
    def reset_stats(self):
        self._lib.mmd_reset_stats(self._h)

    def stats(self):
        ms = (C.c_float * _lib.N_STAGES)()
        launches, clusters, flops = C.c_int64(), C.c_int64(), C.c_double()
        self._check(self._lib.mmd_get_stats(self._h, ms, C.byref(launches), C.byref(clusters), C.byref(flops)))
        names = [self._lib.mmd_stage_name(i).decode() for i in range(_lib.N_STAGES)]
        cnt, solve = (C.c_int64 * _lib.N_COUNTERS)(), C.c_double()
        self._check(self._lib.mmd_get_counters(self._h, cnt, C.byref(solve)))
        return {"stage_ms": dict(zip(names, [float(x) for x in ms])), "launches": launches.value,
                "clusters": clusters.value, "syrk_flops": flops.value, "batches": cnt[0], "core_tasks": cnt[1],
                "grouped_centres": cnt[2], "sum_shared_prefix": cnt[3], "host_plan_ms": cnt[4] / 1e3,
                "host_adjacency_h2d_ms": cnt[5] / 1e3, "rerouted_centres": cnt[6], "plan_cache_hits": cnt[7], "solve_flops": solve.value}
