"""CPU-side checks: the C-ABI library loads and exports every symbol include/mmd_b200.h declares (no compute calls),
and the vertex sharding + score gather work across two processes (gloo)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from multimodal_decoding_b200 import _lib, build
    lib_path = build.build(force=False)
    header = open(os.path.join(ROOT, "include", "mmd_b200.h")).read()
    declared = set(re.findall(r"\b(mmd_[a-z_0-9]+)\s*\(", header))
    declared.discard("mmd_score_name()")
    cdll = ctypes.CDLL(lib_path)
    missing = [s for s in sorted(declared) if not hasattr(cdll, s)]
    assert not missing, missing
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert cdll.mmd_version is not None


def test_engine_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    from multimodal_decoding_b200.engine import Engine
    try:
        Engine(0)
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("Engine(0) must raise without a CUDA device")


def test_estimator_protocol_and_score_dicts():
    """Host logic of the drop-in: the loop only accepts what searchlight.py:240 builds (Ridge(alpha,
    fit_intercept=False)) and hands back the reference's object array of 8-key dicts (searchlight.py:114, :149)."""
    import pytest
    from sklearn.linear_model import Ridge
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    from multimodal_decoding_b200.engine import score_keys
    assert sl._estimator_alpha(Ridge(alpha=3, fit_intercept=False)) == 3.0
    assert sl._estimator_alpha(Ridge(alpha=0.0, fit_intercept=False)) == 0.0
    for bad in (Ridge(alpha=1.0), Ridge(alpha=-1.0, fit_intercept=False),
                Ridge(alpha=np.ones(4), fit_intercept=False), object()):
        with pytest.raises(ValueError):
            sl._estimator_alpha(bad)
    # float32 / float64 inputs are taken when the cast to float16 loses nothing, refused otherwise (never rounded silently)
    exact = np.random.default_rng(2).standard_normal((37, 50)).astype(np.float16)
    exact[3, 4] = np.nan
    for wide in (np.float32, np.float64):
        got = sl._as_f16(exact.astype(wide), "X")
        assert got.dtype == np.float16 and np.array_equal(got, exact, equal_nan=True)
        lossy = exact.astype(wide)
        lossy[30, 7] = 0.1                                         # not a float16 number
        with pytest.raises(TypeError, match="float16"):
            sl._as_f16(lossy, "X")
        with pytest.raises(TypeError, match="float16"):
            sl._as_f16(np.full((2, 2), 1e6, dtype=wide), "X")      # overflows float16
    with pytest.raises(TypeError, match="float16"):
        sl._as_f16(np.zeros((3, 3), dtype=np.int32), "X")
    assert sl._as_f16(exact, "X") is exact
    block = np.random.default_rng(0).random((5, 8))
    out = sl._scores_to_dicts(block)
    assert out.shape == (5,) and out.dtype == object
    assert list(out[2].keys()) == list(score_keys()) and [out[2][k] for k in score_keys()] == block[2].tolist()
    assert all(type(v) is float for v in out[0].values())
    assert sl._scores_to_dicts(np.zeros((0, 8))).shape == (0,)
    # ragged (radius-mode) adjacency is told apart from the rectangular k-NN matrix
    assert sl._is_ragged([np.arange(3), np.arange(5)]) and not sl._is_ragged(np.zeros((4, 3), dtype=np.int64))
    assert isinstance(sl._rectangular([np.arange(3), np.arange(3)]), np.ndarray)
    assert isinstance(sl._rectangular([np.arange(3), np.arange(5)]), list)


def test_score_dicts_helper_matches_python_construction():
    """csrc/pydicts.c (CPython API) and the pure-Python construction give the same object array of dicts."""
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    from multimodal_decoding_b200.engine import score_keys
    rng = np.random.default_rng(3)
    block = rng.random((1000, 8))
    block[7, 3] = np.nan
    got = sl._scores_to_dicts(block)
    keys = score_keys()
    assert got.dtype == object and got.shape == (1000,)
    for i in (0, 7, 999):
        assert type(got[i]) is dict and list(got[i].keys()) == list(keys)
        want = dict(zip(keys, block[i].tolist()))
        assert all((got[i][k] == want[k]) or (np.isnan(got[i][k]) and np.isnan(want[k])) for k in keys)
        assert all(type(v) is float for v in got[i].values())
    if sl._pyhelpers():                                        # built wherever Python.h is present
        saved = sl._PYHELPERS
        try:
            sl._PYHELPERS = False
            slow = sl._scores_to_dicts(block)
        finally:
            sl._PYHELPERS = saved
        assert all(list(a.items())[:3] == list(b.items())[:3] for a, b in zip(got[:7], slow[:7]))


def test_score_dicts_built_ahead_of_the_scores():
    """_DictPrebuild: the dicts are created on a helper thread while the library call runs, the values are written when the
    scores arrive; the result equals the plain construction, also when the scores arrive first or the call never starts."""
    import ctypes
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    rng = np.random.default_rng(11)
    n = 20000
    block = rng.random((n, 8))
    block[3, 5] = np.nan
    want = sl._scores_to_dicts(block)

    def same(a, b):
        return len(a) == len(b) and all(
            list(x.keys()) == list(y.keys()) and all((u == v) or (u != u and v != v) for u, v in zip(x.values(), y.values()))
            for x, y in zip(a, b))

    pb = sl._DictPrebuild(n)
    pb.start()
    ctypes.CDLL(None).usleep(100000)                  # a call that releases the GIL, like the searchlight call
    got = pb.finish(block)
    assert got.dtype == object and got.shape == (n,) and same(got, want)
    assert all(type(v) is float for v in got[3].values())
    assert len({id(v) for d in got[:500] for v in d.values()}) == 4000          # no float object is shared between dicts
    # the scores arrive before the helper thread is done
    pb = sl._DictPrebuild(n)
    pb.start()
    assert same(pb.finish(block), want)
    # start() never reached (an error before the searchlight call), a second finish, an unexpected block shape, no centres
    pb = sl._DictPrebuild(n)
    assert same(pb.finish(block), want)
    pb.start()
    pb.cancel()
    assert pb.thread is None and same(pb.finish(block[:10]), want[:10])
    pb = sl._DictPrebuild(n)
    pb.start()
    assert same(pb.finish(block[:7]), want[:7])
    assert sl._DictPrebuild(0).finish(np.zeros((0, 8))).shape == (0,)
    # the values of a finished array are not touched by a later call
    first = sl._DictPrebuild(4)
    first.start()
    a = first.finish(block[:4])
    second = sl._DictPrebuild(4)
    second.start()
    b = second.finish(block[4:8])
    assert same(a, want[:4]) and same(b, want[4:8])


def test_mark_vertices_host_helper():
    """mmd_mark_vertices (host threads, no GPU): the vertices a shard's neighbourhoods use, as the sharded upload needs them."""
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    rng = np.random.default_rng(5)
    A = rng.integers(0, 5000, size=(3000, 700))                      # int64, > 1 Mi entries: several host threads
    got = sl._active_vertices(A[:3], 5000)
    assert np.array_equal(got, np.unique(A[:3]))
    assert sl._active_vertices(A, 5000) is None                       # (nearly) every vertex: no subset
    ragged = [A[0, :5].astype(np.int32), A[1, :9].astype(np.int32)]
    assert np.array_equal(sl._active_vertices(ragged, 5000), np.unique(np.concatenate(ragged)))
    with pytest.raises(ValueError, match="outside"):
        sl._active_vertices(np.array([[1, 2, 5000]]), 5000)


def test_active_set_is_taken_for_small_shards_only(monkeypatch):
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    monkeypatch.delenv("MMD_B200_ACTIVE", raising=False)
    V = 163842
    assert not sl._wants_active_set(np.empty((V // 2, 3)), V) and not sl._wants_active_set(np.empty((V // 8, 3)), V)
    assert sl._wants_active_set(np.empty((V // 32, 3)), V)
    monkeypatch.setenv("MMD_B200_ACTIVE", "1")
    assert sl._wants_active_set(np.empty((V // 2, 3)), V)
    monkeypatch.setenv("MMD_B200_ACTIVE", "0")
    assert not sl._wants_active_set(np.empty((V // 32, 3)), V)


def test_split_bounds_equals_array_split():
    from multimodal_decoding_b200.sharding import split_bounds, spatial_order
    for n, parts in ((163842, 8), (10, 3), (7, 7), (5, 8)):
        want = [(int(a[0]), int(a[-1]) + 1) if len(a) else None for a in np.array_split(np.arange(n), parts)]
        got = split_bounds(n, parts)
        for w, g in zip(want, got):
            if w is None:
                assert g[0] == g[1]
            else:
                assert w == g
    c = np.random.default_rng(1).standard_normal((500, 3))
    assert sorted(spatial_order(c)) == list(range(500))


_WORKER = r"""
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from multimodal_decoding_b200.sharding import gather_scores, shard_range, split_bounds
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[2]}", rank=int(sys.argv[3]), world_size=2)
n = 1001
lo, hi = shard_range(n)
assert (lo, hi) == split_bounds(n, 2)[dist.get_rank()]
full = np.arange(n * 8, dtype=np.float64).reshape(n, 8) * 0.5
out = gather_scores(full[lo:hi], n)
assert out.shape == (n, 8) and np.array_equal(out, full), "gathered score map differs"
dist.destroy_process_group()
print("ok", dist.get_rank() if False else lo, hi)
"""


def test_two_rank_gather_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


_WORKER_CSL = r"""
import sys, time, warnings
import numpy as np
import torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from multimodal_decoding_b200.sharding import split_bounds
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
from multimodal_decoding_b200.engine import score_keys
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[2]}", rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
V, k = 1003, 9
A = np.random.default_rng(0).integers(0, V, size=(V, k))
expect = (A[:, :8] + 0.25 * np.arange(8)).astype(np.float64)          # a score block any rank can recompute
calls = []

def fake_scores(X, y, estimator, rows, train_ids, test_ids, imagery_ids, *, anchors=None, shard_only=False, prebuild=None):
    # stands in for the GPU part of search_light_scores: same signature, same moment of prebuild.start()
    calls.append((len(rows), shard_only))
    prebuild.start()
    time.sleep(0.05)                                                   # the library call releases the GIL
    flags = np.zeros(len(rows), dtype=np.int32)
    if rank == 1:
        flags[2] = 1                                                   # one centre of rank 1 is not positive definite
    return (np.asarray(rows)[:, :8] + 0.25 * np.arange(8)).astype(np.float64), flags

sl.search_light_scores = fake_scores
lo, hi = split_bounds(V, 2)[rank]
with warnings.catch_warnings(record=True) as w:
    warnings.simplefilter("always")
    out = sl.custom_search_light(None, None, None, A, [0], [1], [2], n_jobs=2)
assert calls == [(hi - lo, True)], calls
assert (len(w) == 1 and str(split_bounds(V, 2)[1][0] + 2) in str(w[0].message)) if rank == 1 else len(w) == 0
bad = split_bounds(V, 2)[1][0] + 2
keys = list(score_keys())
assert out.dtype == object and out.shape == (V,)
for i in range(V):
    assert list(out[i].keys()) == keys
    vals = list(out[i].values())
    if i == bad:
        assert all(v != v for v in vals)                               # NaN on every rank, not only the one that saw the flag
    else:
        assert vals == list(expect[i]), (i, vals, expect[i])
# n_jobs=1: the whole call stays on this process (no collective), every centre scored locally
calls.clear()
out1 = sl.custom_search_light(None, None, None, A[:50], [0], [1], [2], n_jobs=1)
assert calls == [(50, False)] and len(out1) == 50
dist.destroy_process_group()
print("ok", rank)
"""


def test_custom_search_light_host_logic_two_ranks_gloo(tmp_path):
    """The host side of the vertex-sharded call on two gloo ranks, the GPU part replaced by a stub: each rank scores its
    range of `np.array_split`, the blocks are gathered, every rank returns all dicts (prepared by the helper thread during
    the 'GPU' call), a flagged centre is NaN everywhere."""
    script = tmp_path / "worker_csl.py"
    script.write_text(_WORKER_CSL)
    port = str(31500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o
