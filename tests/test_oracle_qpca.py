"""oracle/qpca.py against fixtures produced by the reference's qPCA class."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import qpca as oq


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "qpca_reference.npz"))


def test_fit_full_matches_reference(g):
    out = oq.fit_full(g["X"], 5)
    for name in ("mean_", "components_", "singular_values_", "explained_variance_", "explained_variance_ratio_",
                 "left_sv"):
        np.testing.assert_allclose(out[name], g[name], rtol=1e-9, atol=1e-12)
    assert out["muA"] == pytest.approx(float(g["muA"]), rel=1e-12)
    assert out["norm_muA"] == str(g["norm_muA"][0])
    assert out["frob_norm"] == pytest.approx(float(g["frob_norm"]), rel=1e-12)


def test_prescribed_singular_values(g):
    # sklearn/decomposition/tests/test_pca.py:159-189
    out = oq.fit_full(g["Xs_hat"], 3)
    np.testing.assert_allclose(out["singular_values_"], [3.142, 2.718, 1.0], rtol=1e-12)
    np.testing.assert_allclose(g["Xs_sv"], [3.142, 2.718, 1.0], rtol=1e-12)


def test_singular_value_estimates_match_reference(g):
    # (3 phase-estimation outcome tables of 2^21 entries each: the reference's own O(M) cost)
    rng = np.random.default_rng(0)
    est = oq.estimate_singular_values(g["singular_values_"][:3], float(g["muA"]), 0.5, 12, rng)
    ref = g["est_sv"][0][:3]
    np.testing.assert_allclose(est, ref, rtol=1e-9)
