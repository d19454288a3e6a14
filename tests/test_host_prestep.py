"""Host-side pre-steps of the PRODUCT (multimodal_decoding_b200.data) against fixtures the unmodified reference wrote
(tests/golden/make_golden_r2.py -> data_cases.npz): SURVEY.md §8 a2.  float32 in -> float32 out, bit for bit, so that the
fp16 cast of searchlight.py:247-248 rounds the same numbers the reference's run() would round."""
import os

import numpy as np
import pytest

from multimodal_decoding_b200 import data as pdata


@pytest.mark.parametrize("tag,dtype", [("f32", np.float32), ("f64", np.float64)])
def test_standardize_matches_reference(golden_dir, tag, dtype):
    g = np.load(os.path.join(golden_dir, "data_cases.npz"))
    tr, te, im = g[f"{tag}_tr"], g[f"{tag}_te"], g[f"{tag}_im"]
    keep = tr.copy()
    out = pdata.standardize_fmri_betas(tr, te, im)
    assert np.array_equal(tr, keep, equal_nan=True), "inputs must not be modified"
    for name, got in zip(("tr", "te", "im"), out):
        want = g[f"{tag}_betas_{name}_out"]
        assert got.dtype == want.dtype == dtype
        assert np.array_equal(got, want, equal_nan=True)
    # what the searchlight feeds the kernels: the fp16 cast of the standardised arrays
    with np.errstate(over="ignore"):         # the test split of the large-offset column leaves the fp16 range
        assert np.array_equal(np.concatenate(out).astype(np.float16),
                              np.concatenate([g[f"{tag}_betas_{n}_out"] for n in ("tr", "te", "im")]).astype(np.float16),
                              equal_nan=True)
    lat = pdata.standardize_latents(g[f"{tag}_lat_tr"], g[f"{tag}_lat_te"], g[f"{tag}_lat_im"])
    for name, got in zip(("tr", "te", "im"), lat):
        want = g[f"{tag}_latents_{name}_out"]
        assert got.dtype == want.dtype == dtype and np.array_equal(got, want)
    two = pdata.standardize_latents(g[f"{tag}_lat_tr"], g[f"{tag}_lat_te"])
    assert len(two) == 2 and np.array_equal(two[1], g[f"{tag}_latents_te_out"])


def test_standardize_other_dtypes():
    """sklearn's validation keeps float16, turns everything else into float64."""
    rng = np.random.default_rng(0)
    a = rng.integers(-5, 5, size=(20, 6))
    out = pdata.standardize_latents(a, a[:3])
    assert out[0].dtype == np.float64 and abs(out[0].mean()) < 1e-12
    h = rng.standard_normal((20, 6)).astype(np.float16)
    assert pdata.standardize_latents(h, h[:3])[1].dtype == np.float16


def test_null_shuffled_indices_match_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "data_cases.npz"))
    got = np.array(pdata.create_null_distr_shuffled_indices(6))
    assert np.array_equal(got, g["shuffled_indices_6"])
