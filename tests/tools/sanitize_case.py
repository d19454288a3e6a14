"""One small pass through every kernel of the library with the results checked against the oracle; meant to run in
guard mode (`MMD_B200_GUARD=1`, see include/mmd_b200.h: guard bands around every device buffer, buffers pre-filled
with 0xA5), where `check_guards()` after every stage names a kernel that wrote out of bounds.  compute-sanitizer is
closed on the GPU pool; `tools/gpu_guard.sh` runs this case and the whole GPU test suite in guard mode instead.

    MMD_B200_GUARD=1 python tests/tools/sanitize_case.py [--quick]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from multimodal_decoding_b200 import synthetic as syn          # noqa: E402
from multimodal_decoding_b200.engine import Engine             # noqa: E402
from oracle import searchlight_oracle as orc                   # noqa: E402  (checker only)
from oracle import stats_oracle as so                          # noqa: E402


def main():
    quick = "--quick" in sys.argv
    n_points, k, n_train, n_dims = (360, 200, 256, 64) if quick else (900, 200, 320, 96)
    mesh = syn.make_mesh(level=None, n_points=n_points, seed=11)
    trials = syn.make_trials(mesh, n_train=n_train, n_dims=n_dims, n_factors=8, iid=True, seed=2)
    inp = syn.make_searchlight_inputs(mesh, trials)
    is_cap = orc.default_is_caption(len(inp.test_ids))
    eng = Engine(0)
    guard = eng.guard_selftest()
    assert guard in (0, -1), "guard mode is on but does not detect an overrun"
    print("guard mode:", "on, detector self-test passed" if guard == 0 else "off (set MMD_B200_GUARD=1)")

    # neighbourhoods: k-NN (bit-exact) and radius (CSR)
    _, A = eng.knn(inp.coords, k)
    _, A_ref = orc.knn_adjacency(inp.coords, k)
    assert np.array_equal(A, A_ref)
    radius = float(np.sort(np.linalg.norm(inp.coords - inp.coords[0], axis=1))[40])
    rows = eng.radius_neighbors(inp.coords, radius)
    rows_ref = orc.radius_adjacency(inp.coords, radius)
    assert all(np.array_equal(a, b) for a, b in zip(rows, rows_ref))
    eng.check_guards()

    # searchlight with prefix sharing (k = 200 > 128 -> shared cores), without, and ragged
    eng.upload(inp.X, inp.latents, inp.train_ids, inp.test_ids, inp.imagery_ids, is_cap)
    sample = list(range(0, len(A), 45))
    want = orc.search_light(inp.X, inp.latents, A, inp.train_ids, inp.test_ids, inp.imagery_ids, 1.0, is_cap, sample)
    for g in (8, 1):
        eng.configure(group_size=g)
        eng.reset_stats()
        scores, flags = eng.searchlight(A, 1.0)
        st = eng.stats()
        assert flags.sum() == 0 and np.abs(scores[sample] - want).max() <= 1e-3
        assert (st["core_tasks"] > 0) == (g > 1)
        eng.check_guards()
    eng.configure(group_size=8)
    ragged, rflags = eng.searchlight_ragged(rows, 1.0)
    rs = sample[:3]
    rwant = orc.search_light(inp.X, inp.latents, rows, inp.train_ids, inp.test_ids, inp.imagery_ids, 1.0, is_cap, rs)
    assert rflags.sum() == 0 and np.abs(ragged[rs] - rwant).max() <= 1e-3
    eng.check_guards()

    # null-distribution scoring of the same predictions
    rng = np.random.default_rng(5)
    n_test = len(inp.test_ids)
    half = n_test // 2
    perms = np.stack([np.concatenate([rng.permutation(half), half + rng.permutation(n_test - half)])
                      for _ in range(7)]).astype(np.int64)
    im_perm = np.tile(np.array([1, 2, 0]), (len(perms), 1))
    nscores, nflags, null = eng.searchlight_null(A[:64], 1.0, perms, im_perm)
    assert nflags.sum() == 0 and null.shape[:2] == (64, 7) and np.abs(nscores - scores[:64]).max() <= 1e-3
    eng.check_guards()

    # group-level statistics: t-map, clusters, TFCE on the k-NN graph of the same mesh
    data = 0.5 + 0.1 * rng.standard_normal((6, n_points))
    t = eng.ttest_1samp(data, 0.5)
    assert np.array_equal(t, so.calc_image_t_values(data, 0.5), equal_nan=True)
    edges = np.array([(i, int(j)) for i in range(n_points) for j in A[i, 1:4] if i < j], dtype=np.int32)
    out = eng.tfce(t, edges, 2.0, 1.0, 0.1)
    assert np.array_equal(out, so.tfce(t, edges, 2.0, 1.0, 0.1))
    labels = eng.cluster_labels(t, edges, 0.5)
    ref_labels, _ = so.clusters_at(t, edges, 0.5)
    assert np.array_equal(labels >= 0, ref_labels >= 0)
    eng.check_guards()
    eng.close()
    print(f"sanitize case ok ({'quick' if quick else 'full'}): {n_points} centres, k={k}, {st['launches']} launches "
          "in the last searchlight")


if __name__ == "__main__":
    main()
