"""The CPU restatement of QLSSVC (oracle/qsvm.py) against vectors produced by the reference's own class
(tests/golden/qsvm_reference.npz, made by tests/golden/gen_golden_qsvm.py)."""
import os

import numpy as np
import pytest
from scipy import stats

from oracle import qsvm as oq

HERE = os.path.dirname(os.path.abspath(__file__))
P_MIN = 1e-3

CONFIGS = {
    "lin": dict(kind='linear', penalty=2.0),
    "rbf": dict(kind='rbf', penalty=1.0),
    "poly": dict(kind='poly', penalty=0.5, degree=3, coef0=1.0),
    "sig": dict(kind='sigmoid', penalty=1.5, coef0=0.5, gamma='auto'),
    "linlr": dict(kind='linear', penalty=2.0, low_rank=True, var=0.9),
    "rbflr": dict(kind='rbf', penalty=1.0, low_rank=True, var=5),
    "rbfg": dict(kind='rbf', penalty=1.0, gamma=0.3),
}


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(HERE, "golden", "qsvm_reference.npz"))


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_fit_and_noiseless_quantities(g, name):
    kw = dict(CONFIGS[name])
    X, y, Xt = g["X"], g["y"], g["Xt"]
    fit = oq.classical_fit(X, y, **kw)
    att = oq.fit_attributes(X, fit, kw["penalty"])
    tol = dict(rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(fit["b"], g[name + "_b"], **tol)
    np.testing.assert_allclose(fit["alpha"], g[name + "_alpha"], **tol)
    np.testing.assert_allclose(fit["cond"], g[name + "_cond"], rtol=1e-9)
    np.testing.assert_allclose(fit["normF"], g[name + "_normF"], rtol=1e-12)
    np.testing.assert_allclose(fit["singular_values_F_"], g[name + "_sv"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(att["Nu"], g[name + "_Nu"], rtol=1e-8)
    np.testing.assert_allclose(att["alpha_F"], g[name + "_alpha_F"], rtol=1e-12)
    if kw["kind"] == "linear":
        np.testing.assert_allclose(att["coef_"], g[name + "_coef"], **tol)
    h = oq.get_h(X, fit, Xt, kw["kind"], kw.get("coef0", 0.0), kw.get("degree", 3))
    betas = oq.get_betas(X, att["Nu"], Xt)
    np.testing.assert_allclose(h, g[name + "_h"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(betas, g[name + "_betas"], rtol=1e-8)
    np.testing.assert_allclose(oq.get_P(h, betas), g[name + "_P"], rtol=1e-8)
    assert np.array_equal(np.sign(h), g[name + "_cpred"])


def test_noisy_predictions_follow_the_reference_law(g):
    X, y, Xt = g["X"], g["y"], g["Xt"]
    rng = np.random.default_rng(3)
    for name, kw in {"absn": dict(error_type='absolute', absolute_error=40.0),
                     "reln": dict(error_type='relative', relative_error=0.9)}.items():
        P, betas, h = g[name + "_P"], g[name + "_betas"], g[name + "_h"]
        mine = np.stack([oq.predict(rng, P, betas, h, **kw)[0] for _ in range(120)])
        f_me, f_ref = (mine > 0).mean(axis=0), (g[name + "_pred"] > 0).mean(axis=0)
        assert np.abs(f_me - f_ref).max() < 0.2                        # 120 against 300 repetitions
        tab = np.array([[(mine > 0).sum(), (mine < 0).sum()], [(g[name + "_pred"] > 0).sum(), (g[name + "_pred"] < 0).sum()]])
        assert stats.chi2_contingency(tab)[1] > P_MIN


def test_relative_error_routine_law(g):
    rng = np.random.default_rng(4)
    mine = np.array([oq.relative_error_routine(rng, 0.8, 0.1, 7.0, 1.3) for _ in range(1000)])
    ref = g["rer"]
    assert stats.ks_2samp(mine[:, 0], ref[:, 0]).pvalue > P_MIN
    assert set(np.unique(mine[:, 2])) <= set(np.unique(ref[:, 2])) | set(np.unique(mine[:, 2]))
    tab_vals = np.union1d(mine[:, 2], ref[:, 2])
    tab = np.array([[np.sum(np.isclose(mine[:, 2], v)) for v in tab_vals], [np.sum(np.isclose(ref[:, 2], v)) for v in tab_vals]])
    tab = tab[:, tab.sum(axis=0) >= 10]
    assert tab.shape[1] < 2 or stats.chi2_contingency(tab)[1] > P_MIN
