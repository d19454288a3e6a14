"""GPU parity of the group-level statistics against the golden fixture written by the unmodified
reference (`tests/golden/group_stats.npz`) and the numpy oracle, through the C ABI and the drop-in modules."""
import os

import numpy as np
import pytest

from oracle import stats_oracle as so

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g(golden_dir):
    return np.load(os.path.join(golden_dir, "group_stats.npz"))


@pytest.fixture(scope="module")
def engine():
    from multimodal_decoding_b200.analyses.decoding.searchlight.searchlight import get_engine
    return get_engine()


def test_t_values_bit_exact(engine, g):
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight_permutation_testing as pt
    for hemi in ("left", "right"):
        data = g[f"{hemi}_data"]
        assert np.array_equal(pt.calc_image_t_values(data, 0.5), g[f"{hemi}_t"], equal_nan=True)
        assert np.array_equal(pt.calc_image_t_values(data, 0.5, sigma=1e-3), g[f"{hemi}_t_sigma"], equal_nan=True)
        assert np.array_equal(pt.calc_t_value(data[:, 11], 0.5), g[f"{hemi}_t"][11])
        assert np.isnan(pt.calc_t_value(data[:, 5], 0.5))
    x = g["ttest_x"]
    assert np.array_equal(pt.ttest_1samp_no_p(x), g["ttest_t"])
    assert np.array_equal(pt.ttest_1samp_no_p(x, sigma=1e-2, method="absolute"), g["ttest_t_abs"])
    np.testing.assert_allclose(pt.ttest_1samp_no_p(x, sigma=1e-2), g["ttest_t_rel"], rtol=1e-12)
    # more subjects than numpy's 8-way unrolled summation block
    rng = np.random.default_rng(3)
    big = 0.5 + 0.1 * rng.standard_normal((37, 300))
    assert np.array_equal(engine.ttest_1samp(big, 0.5), so.calc_image_t_values(big, 0.5), equal_nan=True)


def test_cluster_labels_match_reference(engine, g):
    from multimodal_decoding_b200.analyses import cluster_analysis as ca
    for hemi in ("left", "right"):
        t, edges, thr = g[f"{hemi}_t"], g[f"{hemi}_edges"], float(g[f"{hemi}_cluster_threshold"])
        want = g[f"{hemi}_cluster_labels"]
        got = engine.cluster_labels(t, edges, thr)
        assert np.array_equal(got >= 0, want >= 0)
        pairs = set(zip(got[got >= 0].tolist(), want[want >= 0].tolist()))
        assert len(pairs) == want.max() + 1 == len(set(got[got >= 0].tolist()))
        res = ca.calc_clusters(t, thr, edges, return_clusters=True, return_agg_t_values=True, return_cluster_map=True)
        ref_sets = sorted((sorted(np.flatnonzero(want == i).tolist()) for i in range(want.max() + 1)))
        assert sorted(sorted(c) for c in res["clusters"]) == ref_sets
        assert len(res["agg_t_values"]) == len(ref_sets) and (res["cluster_map"][want < 0] == 0).all()


def test_tfce_matches_reference(engine, g):
    from multimodal_decoding_b200.analyses import cluster_analysis as ca
    t_values = {h: {"m": g[f"{h}_t"]} for h in ("left", "right")}
    edges = {h: g[f"{h}_edges"] for h in ("left", "right")}
    out = ca.calc_tfce_values(t_values, edges, "m")
    for hemi in ("left", "right"):
        assert out[hemi]["m"].dtype == np.float32
        assert np.array_equal(out[hemi]["m"], g[f"{hemi}_tfce_default"])                # h=2, e=1: bit-exact
    auto = ca.calc_tfce_values(t_values, edges, "m", dh=0)
    frac = ca.calc_tfce_values(t_values, edges, "m", h=1, e=0.5, dh=0.2)
    for hemi in ("left", "right"):
        assert np.array_equal(auto[hemi]["m"], g[f"{hemi}_tfce_auto"])
        np.testing.assert_allclose(frac[hemi]["m"], g[f"{hemi}_tfce_h1_e05_dh02"], rtol=2e-6)  # pow(): <= 2 ulp in fp64
    neg = ca.calc_tfce_values({h: {"m": -np.abs(g[f"{h}_t"])} for h in ("left", "right")}, edges, "m")
    assert not neg["left"]["m"].any() and neg["left"]["m"].dtype == np.float64


def test_tfce_batch_of_maps_and_cutoff(engine, g):
    """A null distribution is a batch of maps: one launch, every map equal to its single-map result / the oracle."""
    from multimodal_decoding_b200.analyses import cluster_analysis as ca
    rng = np.random.default_rng(5)
    edges = g["left_edges"]
    base = g["left_t"]
    maps = np.stack([np.where(np.isnan(base), np.nan, 0.2 * base * rng.standard_normal()) + rng.standard_normal(len(base))
                     for _ in range(6)])
    maps[3] = -np.abs(maps[3])                                    # non-positive map -> zeros
    maps[4, :] = np.nan                                           # NaN-only map -> zeros
    got = engine.tfce(maps, edges)
    assert got.shape == maps.shape and got.dtype == np.float32
    for i in range(len(maps)):
        assert np.array_equal(got[i], so.tfce(maps[i], edges)), i
    assert not got[3].any() and not got[4].any()
    null = [{"left": {"m": got[i]}, "right": {"m": got[i] * 0.5}} for i in range(len(maps))]
    cutoff, distr = ca.calc_significance_cutoff(null, "m", p_value_threshold=0.2)
    assert np.array_equal(distr, np.sort([np.nanmax(got[i]) for i in range(len(maps))]))
    assert cutoff == np.quantile(distr, 0.8, method="closest_observation")


def test_stats_errors(engine, g):
    with pytest.raises(RuntimeError, match="edge endpoint"):
        engine.tfce(g["left_t"], np.array([[0, 10 ** 6]]))
    with pytest.raises(RuntimeError, match="subjects"):
        engine.ttest_1samp(np.zeros((1, 10)), 0.5)
