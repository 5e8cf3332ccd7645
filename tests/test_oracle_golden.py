"""The CPU oracle against fixtures produced by the reference's own source
(tests/golden/gen_golden.py).  Deterministic quantities must agree exactly;
stochastic ones in distribution (two-sample KS / chi-square, fixed seeds)."""
import numpy as np
import pytest
from scipy import stats

from oracle import qmeans as oq
from oracle import quantum_utility as qu

P_MIN = 1e-3  # fixed seeds: a failure is a real mismatch, not bad luck


def test_N_schedule_Q_M_m(golden_det):
    for d, delta, norm, N in golden_det["N"]:
        assert qu.tomography_N(d, delta, norm) == N
    for N, fmi, sched in golden_det["schedule"]:
        assert [int(x) for x in qu.schedule(N, fmi)] == sched
    for g, Q in golden_det["median_Q"]:
        assert qu.median_Q(g) == Q
    for e, M in golden_det["ae_M"]:
        assert qu.ae_M(e) == M
    for e, g, m in golden_det["pe_m"]:
        assert qu.pe_m(e, g) == m


def test_consistent_bucket(golden_det):
    for eps, gamma, pe, want in golden_det["cpe_bucket"]:
        got, _ = qu.consistent_bucket(pe, eps, gamma)
        assert got == want


def test_best_mu(golden_det):
    A = np.asarray(golden_det["best_mu"]["A"])
    for key, kw in (("grid_0_0.1", dict(start=0, end=0.1)), ("grid_0_1_0.1", dict(start=0, step=0.1)), ("default", {})):
        name, val = qu.best_mu(A, **kw)
        assert name == golden_det["best_mu"][key][0]
        assert val == pytest.approx(golden_det["best_mu"][key][1], rel=1e-14)


@pytest.mark.parametrize("case", ["a", "b", "c", "toy"])
def test_noiseless_qmeans_bit_exact(golden_qmeans, case):
    g = golden_qmeans
    X, C0 = g[f"{case}_X"], g[f"{case}_C0"]
    labels, inertia, centers, n_iter = oq.kmeans_single_lloyd(
        X, C0, False, True, True, True, max_iter=10 if case == "toy" else 8, tol=0, delta=0,
        rng=np.random.default_rng(0))
    assert n_iter == int(g[f"{case}_n_iter"])
    assert np.array_equal(labels, g[f"{case}_labels"])
    np.testing.assert_allclose(centers, g[f"{case}_centers"], rtol=1e-12, atol=1e-13)
    assert inertia == pytest.approx(float(g[f"{case}_inertia"]), rel=1e-12)


def test_tomography_default_mode_law(golden_stoch):
    v = golden_stoch["tomo_v"]
    rng = np.random.default_rng(1)
    errs, stops = [], []
    for _ in range(600):
        r = qu.real_tomography(v, delta=0.3, rng=rng)
        n = list(r.keys())[-1]
        errs.append(np.linalg.norm(v - r[n]))
        stops.append(n)
    assert stats.ks_2samp(errs, golden_stoch["tomo_default_err"]).pvalue > P_MIN
    assert stats.ks_2samp(stops, golden_stoch["tomo_default_N"]).pvalue > P_MIN
    assert max(errs) <= 0.3


@pytest.mark.parametrize("norm", ["L2", "inf"])
def test_tomography_single_shot_law(golden_stoch, norm):
    v = golden_stoch["tomo_v"]
    rng = np.random.default_rng(2)
    errs, c0 = [], []
    for _ in range(300):
        est = list(qu.real_tomography(v, delta=0.3, norm=norm, incremental_measure=False, rng=rng).values())[0]
        errs.append(np.linalg.norm(v - est, ord=2 if norm == "L2" else np.inf))
        c0.append(est[0])
    assert stats.ks_2samp(errs, golden_stoch[f"tomo_single_{norm}_err"]).pvalue > P_MIN
    assert stats.ks_2samp(c0, golden_stoch[f"tomo_single_{norm}_c0"]).pvalue > P_MIN


def test_tomography_small_N_signs(golden_stoch):
    v = golden_stoch["tomo_v"]
    rng = np.random.default_rng(3)
    wrong, c3 = 0, []
    for _ in range(1500):
        est = list(qu.real_tomography(v, N=40, delta=0.3, incremental_measure=False, rng=rng).values())[0]
        wrong += int(np.sum(np.sign(est) != np.sign(v)))
        c3.append(est[3])
    ref_wrong, total = golden_stoch["tomo_N40_signerr"]
    # two binomial proportions
    p = (wrong + ref_wrong) / (2 * total)
    z = (wrong - ref_wrong) / np.sqrt(2 * total * p * (1 - p))
    assert abs(z) < 4
    assert stats.ks_2samp(c3, golden_stoch["tomo_N40_c3"]).pvalue > P_MIN


def test_tomography_returns_unit_norm(golden_stoch):
    v = golden_stoch["tomo_v"]
    out = qu.tomography(5 * v, 0.3, rng=np.random.default_rng(4))
    assert np.linalg.norm(out) == pytest.approx(1.0, abs=1e-12)
    np.testing.assert_allclose(golden_stoch["tomo_scaled_norm"], 1.0, atol=1e-12)


def _chisq_against_pmf(draws, support, pmf):
    obs = np.array([np.sum(np.isclose(draws, s)) for s in support], dtype=float)
    keep = pmf * len(draws) >= 5
    o = np.append(obs[keep], obs[~keep].sum())
    e = np.append(pmf[keep], pmf[~keep].sum()) * len(draws)
    if e[-1] == 0:
        o, e = o[:-1], e[:-1]
    return stats.chisquare(o, e * o.sum() / e.sum()).pvalue


def test_ae_pe_closed_form_pmf_matches_reference(golden_stoch):
    M = qu.ae_M(0.05)
    pmf = qu.ae_pmf(0.3, M)
    support = np.sin(np.pi * np.arange(M) / M) ** 2
    # a_tilde values j and M-j coincide: merge them
    vals = np.unique(np.round(support, 12))
    merged = np.array([pmf[np.isclose(support, v)].sum() for v in vals])
    assert _chisq_against_pmf(golden_stoch["ae_draws_a0.3_e0.05"], vals, merged) > P_MIN
    pm = qu.pe_pmf(0.3, 6)
    assert _chisq_against_pmf(golden_stoch["pe_draws_w0.3_m6"], np.arange(64) / 64, pm) > P_MIN


def test_ae_median_and_ipe_law(golden_stoch):
    rng = np.random.default_rng(5)
    med = [qu.amplitude_estimation(0.3, epsilon=0.05, gamma=0.1, rng=rng) for _ in range(1500)]
    assert stats.ks_2samp(med, golden_stoch["ae_median_a0.3_e0.05_g0.1"]).pvalue > P_MIN
    x, y = golden_stoch["ipe_x"], golden_stoch["ipe_y"]
    dr = [qu.ipe(x, y, 0.25, rng=rng) for _ in range(1500)]
    assert stats.ks_2samp(dr, golden_stoch["ipe_draws_e0.25"]).pvalue > P_MIN
    dr2 = [qu.ipe(x, golden_stoch["ipe_y2"], 0.25, rng=rng) for _ in range(1500)]
    assert len(np.unique(np.round(dr2, 9))) >= 2
    assert stats.ks_2samp(dr2, golden_stoch["ipe2_draws_e0.25"]).pvalue > P_MIN


def test_window_and_ipe_labels(golden_stoch):
    rng = np.random.default_rng(6)
    cnt = np.zeros(4)
    for _ in range(3000):
        lab, _, _ = oq.labels_estimation(golden_stoch["window_X"], golden_stoch["window_C"], 0.5, False, rng)
        cnt[lab[0]] += 1
    ref = golden_stoch["window_counts"]
    assert cnt[3] == 0 and ref[3] == 0
    assert stats.chi2_contingency(np.stack([cnt[:3], ref[:3]]))[1] > P_MIN
    X, C, ref_lab = golden_stoch["ipe_X"], golden_stoch["ipe_C"], golden_stoch["ipe_labels"]
    mine = np.array([oq.labels_estimation(X, C, 0.5, True, rng)[0] for _ in range(40)])
    # per-label totals over all rows and repetitions
    tab = np.stack([np.bincount(mine.ravel(), minlength=3), np.bincount(ref_lab.ravel(), minlength=3)])
    assert stats.chi2_contingency(tab)[1] > P_MIN


def test_gaussian_est_law(golden_stoch):
    dev = np.concatenate([qu.make_gaussian_est(np.zeros(8), 0.5, rng=np.random.default_rng(7 + i)) for i in range(400)])
    assert stats.ks_2samp(dev, golden_stoch["gauss_est_dev"]).pvalue > P_MIN
    assert np.abs(dev).max() <= 0.5 / np.sqrt(8)
