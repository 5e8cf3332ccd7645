"""K4/K6/K7/K8: Gram, mu-norms, AE/PE sampler, truncated-normal injection."""
import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
P_MIN = 1e-3


def _chisq(draws, pmf):
    B = len(draws)
    obs = np.bincount(draws.astype(np.int64), minlength=len(pmf)).astype(float)
    small = pmf * B < 8
    o = np.append(obs[~small], obs[small].sum())
    e = np.append(pmf[~small], pmf[small].sum()) * B
    if e[-1] == 0:
        o, e = o[:-1], e[:-1]
    return stats.chisquare(o, e * o.sum() / e.sum()).pvalue


@pytest.mark.parametrize("M,x", [(35, 12.91), (35, 0.2), (35, 34.7), (64, 19.2), (318, 100.5), (5, 2.3), (16, 7.5),
                                 (17, 8.5001), (3145, 1000.25), (1024, 1023.999), (2, 0.5), (33, 16.9)])
def test_fejer_sampler_matches_closed_form_pmf(cuda, M, x):
    import gpu_util as G
    B = 300000
    j = G.fejer(np.full(B, float(M)), np.full(B, x), seed=M)
    assert j.min() >= 0 and j.max() <= M - 1 and np.all(j == np.floor(j))
    k = np.arange(M)
    pmf = np.sin(np.pi * x) ** 2 / (M * np.sin(np.pi * (x - k) / M)) ** 2
    assert pmf.sum() == pytest.approx(1.0, abs=1e-9)
    assert _chisq(j, pmf) > P_MIN


def test_fejer_sampler_agrees_with_bruteforce_and_reference(cuda, golden_stoch):
    import gpu_util as G
    from oracle import quantum_utility as qu
    B = 200000
    M = qu.ae_M(0.05)
    x = M * np.arcsin(np.sqrt(0.3)) / np.pi
    fast = G.fejer(np.full(B, float(M)), np.full(B, x), seed=1)
    brute = G.fejer(np.full(B, float(M)), np.full(B, x), seed=2, brute=True)
    tab = np.stack([np.bincount(fast.astype(int), minlength=M), np.bincount(brute.astype(int), minlength=M)])
    tab = tab[:, tab.sum(axis=0) >= 10]
    assert stats.chi2_contingency(tab)[1] > P_MIN
    assert _chisq(fast, qu.ae_pmf(0.3, M)) > P_MIN
    assert stats.ks_2samp(np.sin(np.pi * fast / M) ** 2, golden_stoch["ae_draws_a0.3_e0.05"]).pvalue > P_MIN
    pe = G.fejer(np.full(B, 64.0), np.full(B, 0.3 * 64), seed=3)
    assert _chisq(pe, qu.pe_pmf(0.3, 6)) > P_MIN
    assert stats.ks_2samp(pe / 64, golden_stoch["pe_draws_w0.3_m6"]).pvalue > P_MIN
    # exact grid hit is deterministic; huge M stays inside the main lobe mostly
    assert (G.fejer(np.full(100, 64.0), np.full(100, 20.0)) == 20).all()
    big = G.fejer(np.full(B, 2.0 ** 40), np.full(B, 0.3 * 2.0 ** 40 + 0.25), seed=4)
    off = big - np.floor(0.3 * 2.0 ** 40 + 0.25)
    assert np.mean((off == 0) | (off == 1)) > 0.8 and np.abs(off).max() < 2.0 ** 39


def test_public_ae_pe_api(cuda, golden_stoch):
    import warnings
    from sqlearn import QuantumUtility as Q
    Q.seed(5)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        med = [Q.amplitude_estimation(0.3, epsilon=0.05, gamma=0.1) for _ in range(400)]
        assert stats.ks_2samp(med, golden_stoch["ae_median_a0.3_e0.05_g0.1"]).pvalue > P_MIN
        x, y = golden_stoch["ipe_x"], golden_stoch["ipe_y"]
        dr = [Q.ipe(x, y, 0.25) for _ in range(400)]
        assert stats.ks_2samp(dr, golden_stoch["ipe_draws_e0.25"]).pvalue > P_MIN
        assert Q.phase_estimation(1.0, m=5) == 31 / 32
        w, k, m, M = Q.phase_estimation(0.3, epsilon=0.05, gamma=0.1, nqubit=True)
        assert (m, M) == (8, 256) and k == w * M
        est = [Q.consistent_phase_estimation(0.3, 0.05, 0.1) for _ in range(50)]
        assert max(abs(e - 0.3) for e in est) <= 0.05
        assert len(set(est)) <= 2
        with pytest.raises(AssertionError):
            Q.phase_estimation(0.3)


def test_truncnorm_injection(cuda, golden_stoch):
    import gpu_util as G  # noqa: F401
    from sqlearn import QuantumUtility as Q
    Q.seed(6)
    a = 0.5 / np.sqrt(8)
    dev = np.concatenate([Q.make_gaussian_est(np.zeros(8), 0.5) for _ in range(400)])
    assert np.abs(dev).max() <= a
    assert stats.ks_2samp(dev, golden_stoch["gauss_est_dev"]).pvalue > P_MIN
    big = Q.introduce_error_array(np.zeros(200000), 2.5 * np.sqrt(200000))
    assert stats.kstest(big, stats.truncnorm(-2.5, 2.5).cdf).pvalue > P_MIN
    one = Q.introduce_error(3.0, 0.1)
    assert one.shape == (1,) and abs(one[0] - 3.0) <= 0.1


@pytest.mark.parametrize("n,d", [(40, 12), (1000, 61), (3000, 784), (257, 300)])
def test_mu_norms_match_oracle(cuda, golden_det, n, d):
    from oracle import quantum_utility as oq
    from sqlearn import QuantumUtility as Q
    if (n, d) == (40, 12):
        A = np.asarray(golden_det["best_mu"]["A"])
        for key, kw in (("grid_0_0.1", dict(start=0, end=0.1)), ("grid_0_1_0.1", dict(start=0, step=0.1)),
                        ("default", {})):
            name, val = Q.best_mu(A, **kw)
            assert name == golden_det["best_mu"][key][0]
            assert val == pytest.approx(golden_det["best_mu"][key][1], rel=1e-10)
        return
    rs = np.random.RandomState(n)
    A = rs.randn(n, d) * (rs.rand(n, d) > 0.2)
    name, val = Q.best_mu(A, start=0, step=0.1)
    name_o, val_o = oq.best_mu(A, start=0, step=0.1)
    assert name == name_o and val == pytest.approx(val_o, rel=1e-10)
    p, v = Q.linear_search(A)
    grid = [i for i in np.arange(0.0, 1.0, 0.05)] + [1.0]
    vals = [oq.mu_p(q, A) for q in grid]
    assert v == pytest.approx(min(vals), rel=1e-10) and p == grid[int(np.argmin(vals))]


@pytest.mark.parametrize("n,d", [(1, 1), (50, 3), (1000, 61), (5000, 128), (2000, 784), (300, 130), (10000, 256)])
def test_gram_and_column_sums(cuda, n, d):
    import torch
    from sqlearn import _cabi, _device as dv
    rs = np.random.RandomState(n + d)
    X = rs.randn(n, d) + rs.randn(d)
    Xd = torch.as_tensor(X).cuda()
    cs = torch.empty(d, dtype=torch.float64, device="cuda")
    nb = _cabi.lib.sq_col_sums_workspace_bytes(n, d)
    ws = torch.empty(max(nb, 8), dtype=torch.uint8, device="cuda")
    _cabi.check(_cabi.lib.sq_col_sums(Xd.data_ptr(), n, d, cs.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr()), "cs")
    np.testing.assert_allclose(cs.cpu().numpy(), X.sum(axis=0), rtol=1e-12, atol=1e-10)
    mu = cs / n
    G = torch.empty((d, d), dtype=torch.float64, device="cuda")
    nb = _cabi.lib.sq_gram_workspace_bytes(n, d)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    for mean in (mu, None):
        _cabi.check(_cabi.lib.sq_gram_centered(Xd.data_ptr(), n, d, mean.data_ptr() if mean is not None else None,
                                               G.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr()), "gram")
        Xc = X - X.mean(axis=0) if mean is not None else X
        want = Xc.T @ Xc
        got = G.cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=1e-10, atol=1e-10 * np.abs(want).max())
        assert np.array_equal(got, got.T)
    Y = torch.empty_like(Xd)
    _cabi.check(_cabi.lib.sq_sub_rowvec(Xd.data_ptr(), mu.data_ptr(), n, d, Y.data_ptr(), dv.stream_ptr()), "sub")
    np.testing.assert_allclose(Y.cpu().numpy(), X - mu.cpu().numpy(), rtol=0, atol=0)


@pytest.mark.parametrize("M,x", [(10, 2.3), (12, 0.41), (28, 13.76), (64, 5.5), (600, 151.27), (600, 299.9),
                                  (19167, 4791.75), (1001, 500.2), (33, 0.05)])
def test_ipe_tail_sampler_law(cuda, M, x):
    """the rejection sampler for outcomes outside the 8-point window (csrc/ipe.cuh) against the exact conditional
    pmf sin^2(pi t) / (M^2 sin^2(pi (j - x) / M)) restricted to the tail"""
    import torch
    from sqlearn import _cabi, _device as dv
    B = 200000
    out = torch.empty(B, dtype=torch.float64, device="cuda")
    _cabi.check(_cabi.lib.sq_ipe_debug_tail_sample(float(M), float(x), B, 77, out.data_ptr(), dv.stream_ptr()), "tail")
    j = out.cpu().numpy().astype(np.int64)
    j0 = int(np.floor(x))
    window = {(j0 + m) % M for m in range(-3, 5)}
    assert j.min() >= 0 and j.max() < M and not (set(np.unique(j)) & window)
    grid = np.arange(M)
    pmf = 1.0 / (M * np.sin(np.pi * (grid - x) / M)) ** 2
    pmf[list(window)] = 0.0
    pmf /= pmf.sum()
    obs = np.bincount(j, minlength=M).astype(float)
    big = pmf * B >= 10
    o = np.append(obs[big], obs[~big].sum())
    e = np.append(pmf[big], pmf[~big].sum()) * B
    keep = e > 0
    assert stats.chisquare(o[keep], e[keep] * o[keep].sum() / e[keep].sum()).pvalue > P_MIN
    # left / right split of the tail mass (what the fast path uses)
    side = ((j - x + M / 2) % M) < M / 2          # True: counter-clockwise of x ... a fixed partition of the circle
    p_side = pmf[((grid - x + M / 2) % M) < M / 2].sum()
    assert abs(side.mean() - p_side) < 5 * np.sqrt(p_side * (1 - p_side) / B) + 1e-12
