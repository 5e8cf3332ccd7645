"""Distributional parity at the sample sizes north_star names: the fixtures of tests/golden/gen_golden_r2.py hold
up to 10^4 draws of the REFERENCE's own routines (tomography at d = 16 and at the BASELINE shapes d = 128
delta = 0.25 and d = 784, amplitude / phase estimation, IPE, the delta-window pick, the IPE-mode E-step) and
runs of its noisy Lloyd loop; the CUDA path must follow the same laws (two-sample KS / chi-square at fixed
seeds, p > 1e-3) and keep the theorem's bound ||x~ - x|| <= delta."""
import os
import warnings

import numpy as np
import pytest
from scipy import stats

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
P_MIN = 1e-3


@pytest.fixture(scope="module")
def ref():
    return np.load(os.path.join(GOLDEN, "stochastic_reference_r2.npz"))


@pytest.fixture(scope="module")
def prob():
    return np.load(os.path.join(GOLDEN, "r2_problems.npz"))


def _ks(a, b):
    """two-sample KS on values rounded to 8 decimals: with few measurements the errors live on atoms, and the same
    atom summed in two different orders differs in the last bit - unrounded, its whole mass would show up as a gap"""
    return stats.ks_2samp(np.round(np.asarray(a, dtype=float), 8), np.round(np.asarray(b, dtype=float), 8)).pvalue


def _same_discrete_law(a, b):
    """chi-square test of homogeneity for two samples of a discrete variable (cells with few counts merged)"""
    vals = np.unique(np.concatenate([a, b]))
    ca = np.array([(a == v).sum() for v in vals], dtype=float)
    cb = np.array([(b == v).sum() for v in vals], dtype=float)
    small = (ca + cb) < 10
    if small.sum() > 1:
        ca = np.append(ca[~small], ca[small].sum())
        cb = np.append(cb[~small], cb[small].sum())
    keep = (ca + cb) > 0
    if keep.sum() < 2:
        return 1.0
    return stats.chi2_contingency(np.stack([ca[keep], cb[keep]]))[1]


TOMO = {  # fixture -> (vector, delta, norm, incremental, N override, GPU trials)
    "tomo16_default": ("v16", 0.3, "L2", True, None, 10000),
    "tomo16_default_inf": ("v16", 0.3, "inf", True, None, 10000),
    "tomo16_single_L2": ("v16", 0.3, "L2", False, None, 10000),
    "tomo16_single_inf": ("v16", 0.3, "inf", False, None, 10000),
    "tomo16_N40": ("v16", 0.3, "L2", False, 40, 10000),
    "tomo128_default": ("v128", 0.25, "L2", True, None, 10000),
    "tomo128_default_inf": ("v128", 0.25, "inf", True, None, 10000),
    "tomo128_single_L2": ("v128", 0.25, "L2", False, None, 10000),
    "tomo128_single_inf": ("v128", 0.25, "inf", False, None, 10000),
    "tomo784_default": ("v784", 0.5, "L2", True, None, 10000),
    "tomo784_default_inf": ("v784", 0.1, "inf", True, None, 10000),
    "tomo784_single_L2": ("v784", 0.5, "L2", False, None, 4000),
    "tomo784_single_inf": ("v784", 0.1, "inf", False, None, 10000),
}


@pytest.mark.parametrize("name", sorted(TOMO))
def test_tomography_law_against_reference_draws(cuda, ref, prob, name):
    """error, stopping N, two components and the number of wrong signs: GPU trials against the reference's"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    vname, delta, norm, inc, N, B = TOMO[name]
    v = prob[vname]
    d = len(v)
    Nfull = N if N is not None else qu.tomography_N(d, delta, norm)
    sched = qu.schedule(Nfull).astype(float) if inc else [float(Nfull)]
    est, n_used, err = G.tomography(np.repeat(v[None], B, axis=0), delta, sched, norm_inf=int(norm == "inf"),
                                    stop=int(inc), seed=2026 + d, vec_offset=7)
    r_err, r_N = ref[name + "__err"], ref[name + "__N"]
    assert len(r_err) >= 400
    assert _ks(err, r_err) > P_MIN
    if inc:
        assert (err <= delta).all()                       # the stop rule: the returned attempt met the accuracy
        assert _same_discrete_law(n_used, r_N.astype(float)) > P_MIN
    else:
        assert (n_used == Nfull).all() and (r_N == Nfull).all()
        if N is None:
            # theorem 4.1's guarantee, failure probability far inside 1/poly(d)
            assert (err > delta).mean() < 1e-3 and (r_err > delta).mean() < 1e-2
    assert _ks(est[:, 0], ref[name + "__c0"]) > P_MIN
    assert _ks(est[:, 3], ref[name + "__c3"]) > P_MIN
    wrong = (np.sign(est) != np.sign(v)[None]).sum(axis=1)
    r_wrong = ref[name + "__signerr"]
    # mean number of wrong signs per trial (z-test on the means; the counts are small integers)
    se = np.sqrt(wrong.var() / len(wrong) + r_wrong.var() / len(r_wrong)) + 1e-12
    assert abs(wrong.mean() - r_wrong.mean()) / se < 4.5


@pytest.mark.parametrize("d,delta,norm", [(128, 0.25, "L2"), (128, 0.25, "inf"), (784, 0.1, "L2"), (784, 0.1, "inf"),
                                           (1000, 0.5, "L2")])
def test_failure_rate_inside_the_theorem_bound(cuda, d, delta, norm):
    """single-shot tomography with the full N: ||x~ - x|| <= delta fails with probability <= 1/poly(d)
    (Theorem 4.1 of the tomography paper; north_star); 2 * 10^4 trials on different random vectors"""
    import gpu_util as G
    from oracle import quantum_utility as qu
    rs = np.random.RandomState(d)
    B = 20000
    V = rs.uniform(-1, 1, (B, d))
    V /= np.linalg.norm(V, axis=1, keepdims=True)
    N = qu.tomography_N(d, delta, norm)
    est, n_used, err = G.tomography(V, delta, [float(N)], norm_inf=int(norm == "inf"), seed=88)
    want = np.linalg.norm(V - est, ord=2 if norm == "L2" else np.inf, axis=1)
    np.testing.assert_allclose(err, want, rtol=1e-10)
    assert (err > delta).mean() <= 1.0 / d
    # default mode on the same vectors always returns an estimate within delta
    est2, n2, err2 = G.tomography(V[:4000], delta, qu.schedule(N).astype(float), norm_inf=int(norm == "inf"), stop=1, seed=89)
    assert (err2 <= delta).all() and (n2 <= N).all()


SCALAR = {"ae": ("amplitude_estimation", dict(a=0.3, epsilon=0.05)),
          "ae_med": ("amplitude_estimation", dict(a=0.3, epsilon=0.05, gamma=0.1)),
          "ae_small": ("amplitude_estimation", dict(a=0.013, epsilon=0.01)),
          "pe": ("phase_estimation", dict(omega=0.3, m=6)),
          "pe_eps": ("phase_estimation", dict(omega=0.7123, epsilon=0.02, gamma=0.1)),
          "cpe": ("consistent_phase_estimation", dict(omega=0.3, epsilon=0.05, gamma=0.1))}


@pytest.mark.parametrize("name", sorted(SCALAR))
def test_scalar_estimators_against_10k_reference_draws(cuda, ref, name):
    from sqlearn import QuantumUtility as Q
    fn, kw = SCALAR[name]
    Q.seed(31)
    n = 10000 if name not in ("ae_med",) else 4000
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        mine = np.array([float(getattr(Q, fn)(**kw)) for _ in range(n)])
    r = ref[name + "__x"]
    assert len(r) >= 400
    # the outcomes live on a grid: compare cell frequencies
    assert _same_discrete_law(np.round(mine, 12), np.round(r, 12)) > P_MIN


def test_ipe_against_10k_reference_draws(cuda, ref):
    from sqlearn import QuantumUtility as Q
    x = np.random.RandomState(4).randn(6)
    y = np.random.RandomState(5).randn(6)
    Q.seed(32)
    for name, (a, b, eps, n) in {"ipe": (x, y, 0.25, 6000), "ipe2": (x, y * 0.816, 0.25, 6000),
                                 "ipe3": (x * 0.2, x * 0.21 + 0.01, 0.05, 3000)}.items():
        mine = np.array([Q.ipe(a, b, eps) for _ in range(n)])
        assert _same_discrete_law(np.round(mine, 10), np.round(ref[name + "__x"], 10)) > P_MIN


def test_error_injection_and_window_pick(cuda, ref, golden_stoch):
    import gpu_util as G
    from sqlearn import QuantumUtility as Q
    Q.seed(33)
    g = np.concatenate([Q.make_gaussian_est(np.zeros(8), 0.5) for _ in range(1250)])
    assert _ks(g, ref["gauss__x"]) > P_MIN
    e = np.array([float(Q.introduce_error(1.5, 0.3)[0]) for _ in range(5000)])
    assert _ks(e, ref["intro__x"]) > P_MIN
    Xw, Cw = golden_stoch["window_X"], golden_stoch["window_C"]
    labels, _, _, _ = G.assign(np.repeat(Xw, 10000, axis=0), Cw, mode=1, delta=0.5, seed=34)
    mine, want = np.bincount(labels, minlength=4), np.bincount(ref["window__label"], minlength=4)
    assert mine[3] == 0 and want[3] == 0
    assert stats.chi2_contingency(np.stack([mine[:3], want[:3]]))[1] > P_MIN


def test_ipe_mode_labels_against_reference_frequencies(cuda, ref, prob):
    """labels_estimation(true_distance_estimate=True) on 30 rows x 3 centres: per-row label frequencies over
    1000 reference repetitions against the GPU's"""
    import gpu_util as G
    Xi, Ci = prob["ipe_X"], prob["ipe_C"]
    R = 4000
    labels, _, _, _ = G.assign(np.tile(Xi, (R, 1)), Ci, mode=2, delta=0.5, seed=35)
    mine = labels.reshape(R, len(Xi))
    rl = ref["ipe_labels__labels"]
    assert len(rl) >= 40
    pvals = []
    for row in range(len(Xi)):
        a = np.bincount(mine[:, row], minlength=3).astype(float)
        b = np.bincount(rl[:, row], minlength=3).astype(float)
        keep = (a + b) > 0
        pvals.append(1.0 if keep.sum() < 2 else stats.chi2_contingency(np.stack([a[keep], b[keep]]))[1])
    # 30 tests: Bonferroni
    assert min(pvals) > P_MIN / len(pvals)


def test_quantum_state_api(cuda, ref):
    """QuantumState.measure / estimate_wald / coupon_collect (reference: Utility.py:25-85)"""
    from sqlearn import QuantumUtility as Q
    amp = np.array([0.6, -0.2, 0.45, 0.1, -0.5, 0.35])
    q = Q.QuantumState(registers=list(range(6)), amplitudes=amp)
    p = amp ** 2 / np.sum(amp ** 2)
    np.testing.assert_allclose(q.probabilities, p, rtol=1e-12)
    assert q.get_state() == {i: q.probabilities[i] for i in range(6)}
    Q.seed(36)
    draws = q.measure(10000)
    assert draws.shape == (10000,) and set(np.unique(draws)) <= set(range(6))
    obs = np.bincount(draws, minlength=6)
    assert stats.chisquare(obs, p * 10000).pvalue > P_MIN
    # the order of the draws carries no information: first and second half have the same law
    assert stats.chi2_contingency(np.stack([np.bincount(draws[:5000], minlength=6), np.bincount(draws[5000:], minlength=6)]))[1] > P_MIN
    w = Q.estimate_wald(draws)
    assert abs(sum(w.values()) - 1) < 1e-12 and all(w[i] == obs[i] / 10000 for i in w)
    assert Q.estimate_wald(["a", "b", "a", "a"]) == {"a": 0.75, "b": 0.25}
    # relative frequencies of 50 measurements against the reference's (outcome 0 and outcome 3)
    w50 = np.array([[Q.estimate_wald(q.measure(50)).get(i, 0.0) for i in range(6)] for _ in range(1500)])
    for i in (0, 3):
        assert _same_discrete_law(np.round(w50[:, i] * 50), np.round(ref["qstate__wald50"][:, i] * 50)) > P_MIN
    cc = np.array([Q.coupon_collect(q) for _ in range(300)])
    assert cc.min() >= 6
    assert _ks(cc, ref["qstate__coupon"]) > P_MIN


# ------------------------------------------------------------------ the noisy Lloyd loop
@pytest.fixture(scope="module")
def lloyd_ref():
    return np.load(os.path.join(GOLDEN, "noisy_lloyd_reference.npz"))


@pytest.mark.parametrize("which,runs", [("w", 400), ("i", 300)])
def test_noisy_lloyd_loop_follows_the_reference_law(cuda, lloyd_ref, prob, which, runs):
    """_kmeans_single_lloyd(delta = 0.5, intermediate_err = True) with the delta-window ('w') and the IPE ('i')
    E-step: final inertia, per-iteration inertia and centre shift, cluster sizes - against runs of the
    reference's own loop (the centres come back unit-norm, so the inertia jumps after iteration 1)"""
    import torch
    from sqlearn import QuantumUtility as Q
    from sqlearn.cluster._dmeans import _kmeans_single_lloyd
    from sqlearn.cluster._engine import QMeansEngine
    g = lloyd_ref
    X, C0 = prob[f"lloyd_{which}_X"], prob[f"lloyd_{which}_C0"]
    max_iter = 8 if which == "w" else 5
    key = f"lloyd_{which}__"
    ok = g[key + "failed"] == 0
    assert ok.sum() >= 1
    Xd = torch.as_tensor(X).cuda().contiguous()
    Cd = torch.as_tensor(C0).cuda().contiguous()
    eng = QMeansEngine(Xd, C0.shape[0])
    Q.seed(41)
    from sqlearn.cluster import _dmeans as D
    fin, its, shifts, norms, sizes = [], [], [], [], []
    host = torch.empty((2,), dtype=torch.float64).pin_memory()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(runs):
            labels, inertia, centers, n_iter = _kmeans_single_lloyd(
                eng, None, Cd, True, True, True, None, which == "i", max_iter=max_iter, tol=0, delta=0.5)
            assert n_iter == max_iter
            fin.append(inertia)
            norms.append(np.linalg.norm(centers.cpu().numpy(), axis=1))
            sizes.append(np.sort(np.bincount(labels.cpu().numpy(), minlength=C0.shape[0])))
            # the loop body once more, iteration by iteration, for the per-iteration inertia and centre shift
            # (the reference prints both, _dmeans.py:643-644, :652)
            st = D.LloydState(Cd)
            row_i, row_s = [], []
            for _it in range(max_iter):
                i_, s_ = D._lloyd_iteration(eng, st, host, 0.5, True, True, True, None, which == "i")
                row_i.append(i_)
                row_s.append(s_)
            its.append(row_i)
            shifts.append(row_s)
    fin, its = np.array(fin), np.array(its)
    assert (g[key + "n_iter"] == max_iter).all()
    np.testing.assert_allclose(np.array(norms), 1.0, atol=1e-9)
    np.testing.assert_allclose(g[key + "center_norms"], 1.0, atol=1e-9)
    r_it = g[key + "it_inertia"]
    shifts, r_sh = np.array(shifts), g[key + "shifts"]
    if which == "w":
        # the window mode reports the sum of the TRUE minimum distances (_dmeans.py:751), and iteration 0 starts
        # from the same centres: one number (the reference prints it with 3 decimals)
        np.testing.assert_allclose(its[:, 0], r_it[0, 0], atol=2e-3)
        assert np.allclose(r_it[:, 0], r_it[0, 0])
    if len(r_it) >= 100:
        assert _ks(fin, g[key + "inertia"]) > P_MIN
        for it in ((1, 2, max_iter - 1) if which == "w" else (0, 1, 2, max_iter - 1)):
            assert _ks(its[:, it], r_it[:, it]) > P_MIN
        for it in (0, 1, max_iter - 1):
            assert _ks(shifts[:, it], r_sh[:, it]) > P_MIN
        r_sizes = np.sort(g[key + "label_counts"], axis=1)
        assert _ks(np.array(sizes)[:, 0], r_sizes[:, 0]) > P_MIN
        assert _ks(np.array(sizes)[:, -1], r_sizes[:, -1]) > P_MIN
    # the inflated inertia of SURVEY D5 reproduces: after the first update the inertia is far above iteration 0's
    assert np.median(its[:, 1]) > 1.5 * its[0, 0] and np.median(r_it[:, 1]) > 1.5 * r_it[0, 0]
