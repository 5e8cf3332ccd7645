"""QLSSVC on the GPU against the reference's own outputs (tests/golden/qsvm_reference.npz) and the CPU oracle."""
import os

import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
P_MIN = 1e-3

CONFIGS = {
    "lin": dict(kernel='linear', penalty=2.0),
    "rbf": dict(kernel='rbf', penalty=1.0),
    "poly": dict(kernel='poly', penalty=0.5, degree=3, coef0=1.0),
    "sig": dict(kernel='sigmoid', penalty=1.5, coef0=0.5, gamma='auto'),
    "linlr": dict(kernel='linear', penalty=2.0, low_rank=True, var=0.9),
    "rbflr": dict(kernel='rbf', penalty=1.0, low_rank=True, var=5),
    "rbfg": dict(kernel='rbf', penalty=1.0, gamma=0.3),
}


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(HERE, "golden", "qsvm_reference.npz"))


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_fit_matches_the_reference(cuda, g, name):
    from sqlearn.svm import QLSSVC
    X, y, Xt = g["X"], g["y"], g["Xt"]
    m = QLSSVC(**CONFIGS[name]).fit(X, y)
    tol = dict(rtol=1e-8, atol=1e-10)      # fp64 throughout; the pseudo-inverse amplifies rounding by cond <= ~1e3
    np.testing.assert_allclose(m.b, g[name + "_b"], **tol)
    np.testing.assert_allclose(m.alpha, g[name + "_alpha"], **tol)
    np.testing.assert_allclose(m.cond, g[name + "_cond"], rtol=1e-9)
    np.testing.assert_allclose(m.normF, g[name + "_normF"], rtol=1e-12)
    np.testing.assert_allclose(m.singular_values_F_, g[name + "_sv"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(m.Nu, g[name + "_Nu"], rtol=1e-8)
    np.testing.assert_allclose(m.alpha_F, g[name + "_alpha_F"], rtol=1e-12)
    if CONFIGS[name]["kernel"] == "linear":
        np.testing.assert_allclose(m.coef_, g[name + "_coef"], **tol)
    np.testing.assert_allclose(m.get_h(Xt), g[name + "_h"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(m.get_betas(Xt), g[name + "_betas"], rtol=1e-8)
    np.testing.assert_allclose(m.get_P(Xt), g[name + "_P"], rtol=1e-8)
    assert np.array_equal(m.classical_predict(Xt), g[name + "_cpred"])
    np.testing.assert_allclose(m.get_training_complexity(), g[name + "_train_cx"], rtol=1e-8)
    np.testing.assert_allclose(m.get_classification_complexity(Xt), g[name + "_class_cx"], rtol=1e-8)
    np.testing.assert_allclose(m.get_classification_complexity(Xt, relative_error=True), g[name + "_class_cx_rel"], rtol=1e-7)
    att = m.get_all_attributes(Xt)
    np.testing.assert_allclose(att[4], g[name + "_rel_comp"], rtol=1e-7)
    np.testing.assert_allclose(att[5], g[name + "_abs_comp"], rtol=1e-8)
    # with the default error of 0.01 the noisy prediction agrees with the classical one away from the boundary
    far = np.abs(g[name + "_h"]) > 0.05
    assert np.array_equal(m.predict(Xt)[far], g[name + "_cpred"][far])
    assert m.score(Xt[far], g[name + "_cpred"][far]) == 1.0


def test_kernel_matrix_against_oracle(cuda, g):
    from oracle import qsvm as oq
    from sqlearn.svm import QLSSVC
    rs = np.random.RandomState(1)
    X, Y = rs.randn(300, 37), rs.randn(129, 37)
    for kind, kw in [("linear", {}), ("poly", dict(degree=2, coef0=0.3)), ("rbf", {}), ("sigmoid", dict(coef0=-0.2))]:
        m = QLSSVC(kernel=kind, **kw)
        m.n_features_in_ = 37
        gam = oq.gamma_of(X, 'scale', 37)
        np.testing.assert_allclose(m.get_kernel(X, Y), oq.kernel(X, Y, kind, gam, kw.get("coef0", 0.0), kw.get("degree", 3)),
                                   rtol=1e-11, atol=1e-12)
        K = m.get_kernel(X)
        np.testing.assert_allclose(K, oq.kernel(X, None, kind, gam, kw.get("coef0", 0.0), kw.get("degree", 3)),
                                   rtol=1e-11, atol=1e-12)
        if kind == "rbf":
            assert (np.diag(K) == 1.0).all()


def test_noisy_paths_follow_the_reference_law(cuda, g):
    from sqlearn import QuantumUtility as Q
    from sqlearn.svm import QLSSVC
    X, y, Xt = g["X"], g["y"], g["Xt"]
    Q.seed(9)
    for name, kw in {"absn": dict(kernel='rbf', penalty=1.0, absolute_error=40.0),
                     "reln": dict(kernel='rbf', penalty=1.0, error_type='relative', relative_error=0.9)}.items():
        m = QLSSVC(**kw).fit(X, y)
        reps = 300
        mine = np.stack([m.predict(Xt) for _ in range(reps)])
        ref = g[name + "_pred"]
        assert set(np.unique(mine)) <= {-1.0, 1.0}
        f_me, f_ref = (mine > 0).mean(axis=0), (ref > 0).mean(axis=0)
        assert np.abs(f_me - f_ref).max() < 0.15
        tab = np.array([[(mine > 0).sum(), (mine < 0).sum()], [(ref > 0).sum(), (ref < 0).sum()]])
        assert stats.chi2_contingency(tab)[1] > P_MIN
        Pa = np.stack([m.get_P(Xt[:4], approx=True) for _ in range(reps)])
        ha = np.stack([m.get_h(Xt[:4], approx=True) for _ in range(reps)])
        for j in range(4):
            assert stats.ks_2samp(Pa[:, j], g[name + "_Papprox"][:, j]).pvalue > P_MIN
            assert stats.ks_2samp(ha[:, j], g[name + "_happrox"][:, j]).pvalue > P_MIN
    m = QLSSVC(kernel='linear', penalty=2.0, relative_error=0.8).fit(X, y)
    hp = [m.get_approximated_hyperplane(Xt[:1]) for _ in range(300)]
    assert stats.ks_2samp(np.array([h[0] for h in hp]), g["hyp_b"]).pvalue > P_MIN
    assert stats.ks_2samp(np.stack([h[1] for h in hp])[:, 0], g["hyp_coef"][:, 0]).pvalue > P_MIN
    rr = np.array([m.relative_error_routine(0.1, 7.0, 1.3) for _ in range(1000)])
    assert stats.ks_2samp(rr[:, 0], g["rer"][:, 0]).pvalue > P_MIN
    vals = np.union1d(np.round(rr[:, 2], 12), np.round(g["rer"][:, 2], 12))
    tab = np.array([[np.sum(np.isclose(rr[:, 2], v)) for v in vals], [np.sum(np.isclose(g["rer"][:, 2], v)) for v in vals]])
    tab = tab[:, tab.sum(axis=0) >= 10]
    assert tab.shape[1] < 2 or stats.chi2_contingency(tab)[1] > P_MIN


def test_larger_problem_and_errors(cuda):
    from oracle import qsvm as oq
    from sqlearn.svm import QLSSVC
    rs = np.random.RandomState(2)
    N, d = 1500, 40
    X = rs.randn(N, d)
    y = np.where(X[:, 0] * X[:, 1] + 0.2 * rs.randn(N) > 0, 1.0, -1.0)
    m = QLSSVC(kernel='rbf', penalty=10.0).fit(X, y)
    fit = oq.classical_fit(X, y, kind='rbf', penalty=10.0)
    np.testing.assert_allclose(m.alpha, fit["alpha"], rtol=1e-7, atol=1e-9)
    np.testing.assert_allclose(m.b, fit["b"], rtol=1e-7, atol=1e-9)
    assert (m.classical_predict(X) == y).mean() > 0.9
    with pytest.raises(Exception, match="absolute"):
        QLSSVC(error_type='other')
    with pytest.raises(ValueError):
        QLSSVC().fit(X, y[:-1])
    with pytest.raises(ValueError):
        m.predict(X[:, :5])
