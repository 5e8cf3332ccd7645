"""qPCA on the GPU (Gram route K4 + K5 signs + K6 + K7 + K1) against the reference's outputs."""
import os
import warnings

import numpy as np
import pytest
from scipy import stats

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
P_MIN = 1e-3


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "qpca_reference.npz"))


def test_classical_part_matches_reference(cuda, g):
    from sqlearn.decomposition import qPCA
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = qPCA(n_components=5, svd_solver='full').fit(g["X"])
    # singular values within 1e-10 relative (north_star), vectors to the accuracy the Gram route allows
    np.testing.assert_allclose(m.singular_values_, g["singular_values_"], rtol=1e-10)
    np.testing.assert_allclose(m.mean_, g["mean_"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(m.components_, g["components_"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(m.left_sv, g["left_sv"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(m.explained_variance_, g["explained_variance_"], rtol=1e-10)
    np.testing.assert_allclose(m.explained_variance_ratio_, g["explained_variance_ratio_"], rtol=1e-10)
    assert m.noise_variance_ == pytest.approx(float(g["noise_variance_"]), rel=1e-9)
    assert m.spectral_norm == pytest.approx(float(g["spectral_norm"]), rel=1e-10)
    assert m.frob_norm == pytest.approx(float(g["frob_norm"]), rel=1e-10)
    assert m.muA == pytest.approx(float(g["muA"]), rel=1e-10) and m.norm_muA == str(g["norm_muA"][0])
    np.testing.assert_allclose(m.transform(g["X"][:40]), g["transform"], rtol=0, atol=1e-7)
    back = m.inverse_transform(m.transform(g["X"][:40]))
    assert back.shape == (40, 12)


def test_prescribed_singular_values(cuda, g):
    # the stock test's anchor (sklearn/decomposition/tests/test_pca.py:159-189): sigma = [3.142, 2.718, 1.0]
    from sqlearn.decomposition import qPCA
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = qPCA(n_components=3, svd_solver='full').fit(g["Xs_hat"])
    np.testing.assert_allclose(m.singular_values_, [3.142, 2.718, 1.0], rtol=1e-10)
    Xt = m.transform(g["Xs_hat"])
    assert np.sum(m.singular_values_ ** 2) == pytest.approx(np.linalg.norm(Xt, "fro") ** 2, rel=1e-10)


@pytest.mark.parametrize("n,d,k", [(50, 7, 3), (500, 33, 8), (2000, 128, 20), (1500, 300, 40), (60, 100, 10)])
def test_shapes_against_oracle(cuda, n, d, k):
    from oracle import qpca as oq
    from sqlearn.decomposition import qPCA
    rs = np.random.RandomState(n + d)
    r = min(n, d)
    X = rs.randn(n, r) @ np.diag(np.linspace(5, 1, r)) @ rs.randn(r, d) + rs.randn(d)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = qPCA(n_components=k, svd_solver='full').fit(X)
    want = oq.fit_full(X, k)
    np.testing.assert_allclose(m.singular_values_, want["singular_values_"], rtol=1e-10)
    np.testing.assert_allclose(m.components_, want["components_"], rtol=0, atol=1e-8)
    np.testing.assert_allclose(m.left_sv[:, :k], want["left_sv"][:, :k], rtol=0, atol=1e-8)
    assert m.muA == pytest.approx(want["muA"], rel=1e-10)


def test_quantum_estimates_match_reference_law(cuda, g):
    from sqlearn import QuantumUtility as Q
    from sqlearn.decomposition import qPCA
    theta = float(g["theta"][0])
    Q.seed(7)
    right_err, left_err, rep_err, topk, hits = [], [], [], [], []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(40):
            m = qPCA(n_components=5, svd_solver='full').fit(g["X"], estimate_all=True, eps=0.5, delta=0.25,
                                                             theta_major=theta, true_tomography=True)
            topk.append(m.topk)
            # consistent phase estimation returns the modal cell "almost always" (P ~ 0.998 per value here,
            # computed from the closed-form pmf); the heavy 1/x^2 tails put the rest far away
            hits.append(np.isclose(m.estimate_s_values, g["est_sv"][0][:m.topk], rtol=1e-8))
            assert np.abs(m.estimate_s_values / g["est_sv"][0][:m.topk] - 1).max() < 0.2
            right_err.extend(np.linalg.norm(m.estimate_right_sv - m.topk_right_singular_vectors, axis=1))
            lv = m.topk_left_singular_vectors
            left_err.extend(np.linalg.norm(m.estimate_left_sv - lv / np.linalg.norm(lv, axis=1, keepdims=True), axis=1))
            r = m.transform(g["X"][:30], classic_transform=False, epsilon_delta=0.3, quantum_representation=True,
                            norm='est_representation')
            A, e, f = r['quantum_representation_results']
            Y = (g["X"][:30] - m.mean_) @ m.components_.T
            rep_err.extend(np.linalg.norm(A - Y / np.linalg.norm(Y, axis=1, keepdims=True), axis=1))
            assert e == 0.3 and f == pytest.approx(np.linalg.norm(Y - A), rel=1e-9)
        assert set(topk) == set(g["topk"].tolist())
        assert np.concatenate(hits).mean() >= 0.97
        assert m.estimate_fs_ratio == pytest.approx(m.estimate_s_values ** 2 / m.frob_norm ** 2, rel=1e-12)
        assert stats.ks_2samp(right_err, g["right_err"]).pvalue > P_MIN
        assert stats.ks_2samp(left_err, g["left_err"]).pvalue > P_MIN
        assert stats.ks_2samp(rep_err, g["rep_err"]).pvalue > P_MIN
        assert max(rep_err) <= 0.3
        p_est = [m.quantum_factor_score_ratio_sum(0.05, theta / m.muA, 0.05) for _ in range(200)]
        assert stats.ks_2samp(p_est, g["p_est"]).pvalue > P_MIN
        # the remaining estimators run and land near the truth
        m2 = qPCA(n_components=5, svd_solver='full').fit(g["X"], theta_estimate=True, eps_theta=2.0, eta=0.2, p=0.7,
                                                        quantum_retained_variance=True, eps=0.5, theta_major=theta)
        assert 0 <= m2.est_theta <= m2.muA and 0 <= m2.p <= 1
        sn = m.spectral_norm_estimation(epsilon=5.0, delta=0.05)
        assert sn > 0
        q, c = m.runtime_comparison(1000, 100, None)
        assert q.shape == c.shape == (100, 100)
        other = m.transform(g["X"][:10], classic_transform=False, psi=0.2, epsilon_delta=0.2,
                            quantum_representation=True, norm='f_norm')['quantum_representation_results']
        assert np.linalg.norm(other) == pytest.approx(1.0, rel=1e-9)
        qs = m.transform(g["X"][:10], classic_transform=False, psi=0.2, epsilon_delta=0.2,
                         quantum_representation=True, norm='q_state')['quantum_representation_results']
        assert abs(sum(qs.probabilities) - 1) < 1e-9
        with pytest.raises(ValueError, match="eps must be > 0"):
            qPCA(n_components=2).fit(g["X"], quantum_retained_variance=True)
