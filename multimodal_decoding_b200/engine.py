"""Host-side handle around the C ABI: one `Engine` per GPU.

Inputs may be numpy arrays (host buffers, copied by the library) or CUDA torch tensors on the engine's
device (borrowed).  Nothing here computes on the CPU: every method forwards to libmmd_b200.so and raises
`RuntimeError` with the library's message on failure.
"""
from __future__ import annotations

import ctypes as C

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
            raise RuntimeError(self._lib.mmd_last_error(self._h).decode())

    def configure(self, cluster_size=0, batch_centres=0, host_threads=0):
        self._check(self._lib.mmd_configure(self._h, cluster_size, batch_centres, host_threads))

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

    # -- trials ------------------------------------------------------------------------------
    def upload(self, X, Y, train_ids, test_ids, imagery_ids, test_is_caption):
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
        self._check(self._lib.mmd_upload(self._h, xp, n_total, n_vertices, yp, int(Y.shape[1]),
                                         tr.ctypes.data, len(tr), te.ctypes.data, len(te), cap.ctypes.data,
                                         im.ctypes.data, len(im)))
        self.n_dims = int(Y.shape[1])
        self.n_eval = len(te) + len(im)
        del xk, yk

    # -- the hot path ------------------------------------------------------------------------
    def searchlight(self, adjacency, alpha: float = 1.0):
        adj = np.ascontiguousarray(np.asarray(adjacency, dtype=np.int32))
        if adj.ndim != 2:
            raise ValueError("adjacency must be rectangular (n_centres, k); ragged radius mode is not supported")
        n, k = adj.shape
        scores = np.empty((n, _lib.N_SCORES), dtype=np.float64)
        flags = np.zeros(n, dtype=np.int32)
        self._check(self._lib.mmd_searchlight(self._h, adj.ctypes.data, n, k, float(alpha), scores.ctypes.data,
                                              flags.ctypes.data))
        return scores, flags

    def predict_centre(self, neighbours, alpha: float = 1.0):
        nb = np.ascontiguousarray(np.asarray(neighbours, dtype=np.int32))
        out = np.empty((self.n_eval, self.n_dims), dtype=np.float64)
        self._check(self._lib.mmd_predict_centre(self._h, nb.ctypes.data, len(nb), float(alpha), out.ctypes.data))
        return out

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t of the handle (wrap with torch.cuda.ExternalStream to record events on it)."""
        return int(self._lib.mmd_stream(self._h) or 0)

    # -- statistics --------------------------------------------------------------------------
    def reset_stats(self):
        self._lib.mmd_reset_stats(self._h)

    def stats(self):
        ms = (C.c_float * _lib.N_STAGES)()
        launches, clusters, flops = C.c_int64(), C.c_int64(), C.c_double()
        self._check(self._lib.mmd_get_stats(self._h, ms, C.byref(launches), C.byref(clusters), C.byref(flops)))
        names = [self._lib.mmd_stage_name(i).decode() for i in range(_lib.N_STAGES)]
        return {"stage_ms": dict(zip(names, [float(x) for x in ms])), "launches": launches.value,
                "clusters": clusters.value, "syrk_flops": flops.value}
