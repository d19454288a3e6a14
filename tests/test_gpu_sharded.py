"""Vertex-sharded searchlight (SURVEY.md §8e / north_star (5)), the plan cache and the active-vertex upload.  Needs a B200
(the 2-rank NCCL case needs two): run with `-m gpu`."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import searchlight_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def engine():
    from multimodal_decoding_b200.engine import Engine
    eng = Engine(0)
    yield eng
    eng.close()


def _case(golden_dir):
    g = np.load(os.path.join(golden_dir, "searchlight_smooth.npz"))
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    return g, np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)


def test_plan_cache_reuses_the_plan_of_an_unchanged_adjacency(engine, golden_dir):
    g, tr, te, im = _case(golden_dir)
    engine.upload(g["X"], g["latents"], tr, te, im, orc.default_is_caption(len(te)))
    A = g["adjacency"]
    engine.reset_stats()
    first, _ = engine.searchlight(A, float(g["alpha"]))
    assert engine.stats()["plan_cache_hits"] == 0
    again, _ = engine.searchlight(A.astype(np.int64), 2.0 * float(g["alpha"]))     # same rows, other alpha and index width
    assert engine.stats()["plan_cache_hits"] == 1
    third, _ = engine.searchlight(A, float(g["alpha"]))
    assert engine.stats()["plan_cache_hits"] == 2
    assert np.array_equal(first, g["scores"]) and np.array_equal(third, g["scores"])
    want = orc.search_light(g["X"], g["latents"], A, tr, te, im, alpha=2.0 * float(g["alpha"]), centres=[0, 5, len(A) - 1])
    assert np.array_equal(again[[0, 5, len(A) - 1]], want)
    # one changed entry is another adjacency
    B = A.copy()
    B[3, 1], B[3, 2] = B[3, 2], B[3, 1]
    swapped, _ = engine.searchlight(B, float(g["alpha"]))
    assert engine.stats()["plan_cache_hits"] == 2
    assert np.array_equal(swapped, g["scores"])             # the order inside a neighbourhood does not matter


def test_active_vertex_upload_matches_full_upload(engine, golden_dir):
    g, tr, te, im = _case(golden_dir)
    is_cap = orc.default_is_caption(len(te))
    A = g["adjacency"]
    centres = np.sort(A[0][:24])                            # a compact patch: the 24 vertices nearest to vertex 0
    shard = A[centres]
    active = np.unique(shard)
    assert len(active) < g["X"].shape[1]
    engine.upload(g["X"], g["latents"], tr, te, im, is_cap, active_vertices=active)
    got, flags = engine.searchlight(shard, float(g["alpha"]))
    assert flags.sum() == 0 and np.array_equal(got, g["scores"][centres])
    outside = np.setdiff1d(np.arange(g["X"].shape[1]), active)
    bad = shard[:2].copy()
    bad[1, 0] = outside[0]
    with pytest.raises(RuntimeError, match="mmd_upload_active"):
        engine.searchlight(bad, float(g["alpha"]))
    with pytest.raises(RuntimeError, match="ascending"):
        engine.upload(g["X"], g["latents"], tr, te, im, is_cap, active_vertices=active[::-1])
    engine.upload(g["X"], g["latents"], tr, te, im, is_cap)
    got, _ = engine.searchlight(A, float(g["alpha"]))
    assert np.array_equal(got, g["scores"])


_WORKER = r"""
import os, sys
import numpy as np
import torch, torch.distributed as dist
root, port, rank, fixture = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
sys.path.insert(0, root)
os.environ.update(LOCAL_RANK=str(rank), LOCAL_WORLD_SIZE="2")
torch.cuda.set_device(rank)
dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=2,
                        device_id=torch.device("cuda", rank))
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
from multimodal_decoding_b200 import data
from sklearn.linear_model import Ridge
g = np.load(fixture)
ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
data.TEST_STIM_TYPES = np.array(["caption"] * (nte // 2) + ["image"] * (nte - nte // 2))
sl.TEST_STIM_TYPES = data.TEST_STIM_TYPES
ids = (list(range(ntr)), list(range(ntr, ntr + nte)), list(range(ntr + nte, ntr + nte + nim)))
out = sl.custom_search_light(g["X"], g["latents"], Ridge(alpha=float(g["alpha"]), fit_intercept=False), g["adjacency"],
                             ids[0], ids[1], ids[2], n_jobs=2)
keys = [str(k) for k in g["keys"]]
got = np.array([[d[k] for k in keys] for d in out])
assert got.shape == g["scores"].shape and np.array_equal(got, g["scores"]), "sharded map differs from the reference's"
assert sl.get_engine().device == rank
dist.barrier()
dist.destroy_process_group()
print("ok", rank)
"""


def test_custom_search_light_two_ranks_nccl(tmp_path, golden_dir):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    fixture = os.path.join(golden_dir, "searchlight_smooth.npz")
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r), fixture], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o


def test_singular_neighbourhood_is_reported_not_scored(golden_dir):
    """alpha = 0 with an all-zero vertex: the Gram has a zero pivot.  sklearn falls back to an SVD solve; this path has no
    fallback, so the centre is flagged, its scores are NaN and a RuntimeWarning names it -- the other centres are unaffected."""
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    from sklearn.linear_model import Ridge
    g, tr, te, im = _case(golden_dir)
    X = g["X"].copy()
    A = g["adjacency"][:6].copy()
    dead = [v for v in A[2] if v not in A[0]][0]
    X[:, dead] = 0                                         # a vertex without signal in centre 2's neighbourhood
    with pytest.warns(RuntimeWarning, match="not positive definite"):
        out = sl.custom_search_light(X, g["latents"], Ridge(alpha=0.0, fit_intercept=False), A, tr, te, im, n_jobs=1)
    assert all(np.isnan(v) for v in out[2].values())
    assert all(np.isfinite(v) for v in out[0].values())
    want = orc.search_light(X, g["latents"], A, tr, te, im, alpha=0.0, centres=[0])
    assert np.array_equal(np.array([out[0][str(k)] for k in g["keys"]]), want[0])
