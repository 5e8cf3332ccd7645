"""Parity of the q-means kernels (K2 E-step, K3 M-step, Lloyd loop) with the
oracle and with fixtures produced by the reference's source.  Noiseless:
labels bit-exact, distances / centres within 1e-10 relative (north_star).
Noisy: same law (chi-square / KS at fixed seeds)."""
import warnings

import numpy as np
import pytest
from scipy import stats

pytestmark = pytest.mark.gpu
P_MIN = 1e-3


@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_exact_assign_matches_reference_first_estep(cuda, golden_qmeans, case):
    import gpu_util as G
    from oracle import qmeans as oq
    X, C0 = golden_qmeans[f"{case}_X"], golden_qmeans[f"{case}_C0"]
    labels, mind, inertia, _ = G.assign(X, C0, mode=0)
    assert np.array_equal(labels, golden_qmeans[f"{case}_labels_first"])
    D = oq.cdist_sq(X, C0)
    np.testing.assert_allclose(mind, D.min(axis=1), rtol=1e-10, atol=1e-12)
    assert inertia == pytest.approx(float(golden_qmeans[f"{case}_inertia_first"]), rel=1e-10)


@pytest.mark.parametrize("n,d,k", [(1, 1, 1), (5, 2, 1), (37, 3, 2), (1000, 7, 10), (777, 16, 16), (513, 33, 17),
                                   (2049, 64, 33), (1500, 100, 64), (900, 130, 100), (1300, 128, 129),
                                   (2000, 128, 256), (640, 61, 200), (1111, 48, 300), (700, 20, 600),
                                   (70000, 32, 10)])
def test_exact_assign_shapes_bit_exact_labels(cuda, n, d, k):
    import gpu_util as G
    from oracle import qmeans as oq
    rs = np.random.RandomState(n + d + k)
    C = rs.randn(k, d) * 2
    X = C[rs.randint(0, k, n)] + rs.randn(n, d)
    labels, mind, inertia, _ = G.assign(X, C, mode=0)
    D = oq.cdist_sq(X, C)
    assert np.array_equal(labels, D.argmin(axis=1))
    np.testing.assert_allclose(mind, D.min(axis=1), rtol=1e-10, atol=1e-11)
    assert inertia == pytest.approx(D.min(axis=1).sum(), rel=1e-10)


def test_exact_assign_near_ties_use_reference_arithmetic(cuda):
    """rows equidistant (up to an ulp) from two centres must follow cdist()**2's rounding"""
    import gpu_util as G
    from oracle import qmeans as oq
    rs = np.random.RandomState(0)
    d, k = 24, 9
    C = rs.randn(k, d)
    mids = []
    for _ in range(400):
        i, j = rs.choice(k, 2, replace=False)
        m = 0.5 * (C[i] + C[j]) + rs.randn(d) * 1e-17
        mids.append(m * (1 + rs.randint(-2, 3) * 2.0 ** -52))
    X = np.vstack([np.asarray(mids), rs.randn(300, d)])
    labels, mind, _, nfix = G.assign(X, C, mode=0)
    D = oq.cdist_sq(X, C)
    mins = D.min(axis=1)
    ties = (D == mins[:, None]).sum(axis=1)
    assert nfix >= 1
    single = ties == 1
    assert np.array_equal(labels[single], D.argmin(axis=1)[single])
    # exact ties: any of the minimisers is legal (reference: random.choice, _dmeans.py:772)
    assert all(D[r, labels[r]] == mins[r] for r in np.where(~single)[0])
    flagged_exact = np.isclose(mind, mins, rtol=1e-10, atol=1e-13)
    assert flagged_exact.all()


def test_exact_ties_are_broken_uniformly(cuda):
    import gpu_util as G
    X = np.zeros((6000, 2))
    C = np.array([[1.0, 0], [0, 1.0], [-1.0, 0], [2.0, 2.0]])
    labels, _, _, nfix = G.assign(X, C, mode=0, seed=5)
    assert nfix == 6000
    cnt = np.bincount(labels, minlength=4)
    assert cnt[3] == 0
    assert stats.chisquare(cnt[:3]).pvalue > P_MIN


def test_centroid_accumulate_matches_numpy_and_is_deterministic(cuda):
    import gpu_util as G
    rs = np.random.RandomState(1)
    for n, d, k in [(1, 1, 1), (1000, 5, 3), (5000, 128, 256), (3000, 61, 7), (40000, 64, 10), (2500, 784, 10),
                    (3000, 1100, 4)]:
        X = rs.randn(n, d)
        lab = rs.randint(0, k, n).astype(np.int32)
        if k > 2:
            lab[lab == 1] = 0  # an empty cluster
        sums, counts, bad = G.accumulate(X, lab, k)
        assert bad == 0
        want = np.zeros((k, d))
        np.add.at(want, lab, X)
        np.testing.assert_allclose(sums, want, rtol=1e-12, atol=1e-11)
        assert np.array_equal(counts, np.bincount(lab, minlength=k))
        sums2, _, _ = G.accumulate(X, lab, k)
        assert np.array_equal(sums, sums2)
    _, _, bad = G.accumulate(rs.randn(10, 2), np.array([0, 1, 5, 0, 0, 0, 0, 0, 0, 0], dtype=np.int32), 3)
    assert bad == 1


@pytest.mark.parametrize("case", ["a", "b", "c", "toy"])
def test_lloyd_noiseless_matches_reference(cuda, golden_qmeans, case):
    import torch
    from sqlearn.cluster._dmeans import _kmeans_single_lloyd
    from sqlearn.cluster._engine import QMeansEngine
    g = golden_qmeans
    X = torch.as_tensor(g[f"{case}_X"]).cuda().contiguous()
    C0 = torch.as_tensor(g[f"{case}_C0"]).cuda().contiguous()
    eng = QMeansEngine(X, C0.shape[0])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        labels, inertia, centers, n_iter = _kmeans_single_lloyd(
            eng, None, C0, False, True, True, None, True, max_iter=10 if case == "toy" else 8, tol=0, delta=0)
    assert n_iter == int(g[f"{case}_n_iter"])
    assert np.array_equal(labels.cpu().numpy(), g[f"{case}_labels"])
    np.testing.assert_allclose(centers.cpu().numpy(), g[f"{case}_centers"], rtol=1e-10, atol=1e-12)
    assert inertia == pytest.approx(float(g[f"{case}_inertia"]), rel=1e-10)


def test_window_mode_law_and_support(cuda, golden_stoch):
    import gpu_util as G
    from oracle import qmeans as oq
    Xw, Cw, ref = golden_stoch["window_X"], golden_stoch["window_C"], golden_stoch["window_counts"]
    labels, mind, inertia, _ = G.assign(np.repeat(Xw, 3000, axis=0), Cw, mode=1, delta=0.5, seed=9)
    cnt = np.bincount(labels, minlength=4)
    assert cnt[3] == 0
    assert stats.chi2_contingency(np.stack([cnt[:3], ref[:3]]))[1] > P_MIN
    np.testing.assert_allclose(mind, 1.0, rtol=1e-12)           # inertia uses the minimum, not the pick (:751)
    # random data, several tile shapes: every pick lies inside the window, picks vary with the seed
    rs = np.random.RandomState(2)
    for n, d, k, delta in [(3000, 8, 12, 3.0), (2000, 128, 256, 30.0), (1500, 33, 300, 8.0), (800, 16, 100, 4.0)]:
        X, C = rs.randn(n, d), rs.randn(k, d)
        D = oq.cdist_sq(X, C)
        mins = D.min(axis=1)
        l1, m1, in1, _ = G.assign(X, C, mode=1, delta=delta, seed=1)
        l2, _, _, _ = G.assign(X, C, mode=1, delta=delta, seed=2)
        assert (D[np.arange(n), l1] <= mins + delta + 1e-9).all()
        np.testing.assert_allclose(m1, mins, rtol=1e-10, atol=1e-11)
        assert in1 == pytest.approx(mins.sum(), rel=1e-10)
        ncand = (D <= (mins + delta)[:, None]).sum(axis=1)
        assert (l1 != l2).mean() > 0.2 * (ncand > 1).mean()
        assert np.array_equal(l1[ncand == 1], D.argmin(axis=1)[ncand == 1])
        # uniformity over candidates: rank of the pick among the row's candidates
        two = np.where(ncand == 2)[0]
        if len(two) > 200:
            first = np.array([np.where(D[r] <= mins[r] + delta)[0][0] for r in two])
            hits = (l1[two] == first).sum()
            assert stats.binomtest(int(hits), len(two), 0.5).pvalue > P_MIN


def _ipe_distance_pmf(x, y, eps):
    """exact law of |x|^2+|y|^2-2*ipe(x,y,eps): median of 13 amplitude-estimation outcomes (closed-form pmf)"""
    from oracle import quantum_utility as qu
    from scipy import stats as st
    S = x @ x + y @ y
    inner = x @ y
    a = (S - 2 * inner) / (2 * S)
    M = qu.ae_M(eps * max(1, abs(inner)) / S)
    pmf = qu.ae_pmf(a, M)
    at = np.round(np.sin(np.pi * np.arange(M) / M) ** 2, 12)
    vals = np.unique(at)
    single = np.array([pmf[at == v].sum() for v in vals])
    G = np.cumsum(single)
    med_cdf = st.binom.sf(6, 13, np.clip(G, 0, 1))        # P(at least 7 of 13 draws <= v)
    med_pmf = np.diff(np.concatenate([[0.0], med_cdf]))
    return 2 * S * vals, med_pmf


def test_ipe_mode_distance_law(cuda, golden_stoch, mode=2):
    """k = 1: the reported min distance IS the IPE-estimated distance |x|^2+|c|^2-2*ipe(x,c).  Checked against
    the exact law (closed-form amplitude-estimation pmf -> median of 13) and against the reference's draws."""
    import gpu_util as G
    n = 20000
    x = golden_stoch["ipe_x"]
    for ykey, rkey in (("ipe_y", "ipe_draws_e0.25"), ("ipe_y2", "ipe2_draws_e0.25")):
        y, ref = golden_stoch[ykey], golden_stoch[rkey]
        _, mind, _, _ = G.assign(np.repeat(x[None], n, axis=0), y[None], mode=mode, delta=0.5, seed=3)
        support, pmf = _ipe_distance_pmf(x, y, 0.25)
        idx = np.array([np.argmin(np.abs(support - v)) for v in mind])
        assert np.allclose(support[idx], mind, rtol=1e-9, atol=1e-9)        # every value is a grid value
        obs = np.bincount(idx, minlength=len(support)).astype(float)
        big = pmf * n >= 8
        o = np.append(obs[big], obs[~big].sum())
        e = np.append(pmf[big], pmf[~big].sum()) * n
        keep = e > 0
        assert stats.chisquare(o[keep], e[keep] * o[keep].sum() / e[keep].sum()).pvalue > P_MIN
        # the reference's own draws follow the same law
        ref_d = x @ x + y @ y - 2 * ref
        ridx = np.array([np.argmin(np.abs(support - v)) for v in ref_d])
        robs = np.bincount(ridx, minlength=len(support)).astype(float)
        ro = np.append(robs[big], robs[~big].sum())
        assert stats.chisquare(ro[keep], e[keep] * ro[keep].sum() / e[keep].sum()).pvalue > P_MIN


def test_ipe_mode_label_law(cuda, golden_stoch, mode=2):
    import gpu_util as G
    X, C, ref_lab = golden_stoch["ipe_X"], golden_stoch["ipe_C"], golden_stoch["ipe_labels"]
    reps = 200
    labels, _, _, _ = G.assign(np.tile(X, (reps, 1)), C, mode=mode, delta=0.5, seed=4)
    mine = labels.reshape(reps, len(X))
    tab = np.stack([np.bincount(mine.ravel(), minlength=3) / reps * len(ref_lab),
                    np.bincount(ref_lab.ravel(), minlength=3)])
    assert stats.chi2_contingency(np.round(tab))[1] > P_MIN
    # per-row agreement with the reference's empirical label frequencies
    f_me = np.stack([(mine == j).mean(axis=0) for j in range(3)])
    f_ref = np.stack([(ref_lab == j).mean(axis=0) for j in range(3)])
    assert np.abs(f_me - f_ref).max() < 0.35  # 40 reference repetitions only


def test_qmeans_fit_end_to_end(cuda, golden_qmeans):
    from oracle import qmeans as oq
    from sqlearn import QuantumUtility as Q
    from sqlearn.cluster import qMeans_
    g = golden_qmeans
    X = g["a_X"] + 3.0        # fit centres internally; reference centres are for the centred data
    C0 = g["a_C0"] + 3.0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        Q.seed(1)
        est = qMeans_(n_clusters=5, init=C0, n_init=1, max_iter=8, tol=0, delta=0, algorithm='full').fit(X)
    assert est.labels_ == [int(v) for v in g["a_labels"]]
    np.testing.assert_allclose(est.cluster_centers_, g["a_centers"] + 3.0, rtol=1e-9, atol=1e-9)
    assert est.n_iter_ == int(g["a_n_iter"])
    assert est.inertia_ == pytest.approx(float(g["a_inertia"]), rel=1e-9)
    Xc = g["a_X"]
    assert est.eta == pytest.approx(np.max(np.einsum("ij,ij->i", X, X)), rel=1e-12)
    assert est.condition_number == pytest.approx(1 / np.linalg.svd(X, compute_uv=False).min(), rel=1e-8)
    from oracle import quantum_utility as oqu
    assert est.muA == pytest.approx(oqu.best_mu(X, 0, 0.1)[1], rel=1e-10)
    assert est.predict(X[:50]) == est.labels_[:50]
    assert est.score(X) == pytest.approx(-oq.cdist_sq(X, est.cluster_centers_).min(axis=1).sum(), rel=1e-9)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with pytest.raises(TypeError, match="elkan"):
            qMeans_(n_clusters=5, delta=0.5).fit(X)
        with pytest.raises(ValueError, match="intermediate_error"):
            qMeans_(n_clusters=5, delta=0, intermediate_error=True, algorithm='full').fit(X)
        # noisy run: every mode runs, centroid tomography returns unit-norm centres (D5)
        Q.seed(2)
        e2 = qMeans_(n_clusters=5, init=C0, n_init=1, max_iter=3, tol=0, delta=0.5, algorithm='full',
                     intermediate_error=True, true_distance_estimate=False).fit(X)
        norms = np.linalg.norm(e2.cluster_centers_ - X.mean(axis=0), axis=1)
        np.testing.assert_allclose(norms, 1.0, atol=1e-9)
        e3 = qMeans_(n_clusters=5, init='k-means++', n_init=2, max_iter=3, delta=0.5, algorithm='full',
                     random_state=0).fit(X)
        assert len(e3.labels_) == len(X) and np.isfinite(e3.inertia_)


def test_kmeans_alias_reaches_the_quantum_branch(cuda, golden_qmeans):
    from sqlearn.cluster import KMeans
    g = golden_qmeans
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        est = KMeans(n_clusters=5, init=g["a_C0"], n_init=1, max_iter=8, tol=0, delta=0).fit(g["a_X"])
        assert est.labels_ == [int(v) for v in g["a_labels"]]
        noisy = KMeans(n_clusters=5, init=g["a_C0"], n_init=1, max_iter=2, delta=0.5, true_distance_estimate=False,
                       intermediate_error=True).fit(g["a_X"])
        assert len(noisy.labels_) == len(g["a_X"])
    assert KMeans(n_clusters=3, delta=0.1).get_params()["algorithm"] == "full"


def test_device_kmeans_plusplus_equals_host_version(cuda):
    """same candidates, same centres, same consumption of random_state as the reference's seeding"""
    import torch
    from sqlearn.cluster._dmeans import _kmeans_plusplus, _kmeans_plusplus_device
    from sqlearn.cluster._engine import QMeansEngine
    rs = np.random.RandomState(0)
    for n, d, k in [(500, 6, 7), (3000, 33, 20), (20000, 128, 40)]:
        X = rs.randn(n, d) + 3 * rs.randint(0, 5, n)[:, None]
        xs = np.einsum("ij,ij->i", X, X)
        r1, r2 = np.random.RandomState(42), np.random.RandomState(42)
        want, ids = _kmeans_plusplus(X, k, xs, r1)
        eng = QMeansEngine(torch.as_tensor(X).cuda().contiguous(), k)
        got = _kmeans_plusplus_device(eng, k, r2).cpu().numpy()
        np.testing.assert_array_equal(got, want)
        assert r1.random_sample() == r2.random_sample()      # both generators are in the same state


IPE_REGIMES = [
    # (x, c, delta): regimes of the amplitude-estimation grid (M points, true position x_grid = M*theta/pi)
    ((3.0, 0.0), (3.05, 0.1), 0.5),          # a ~ 0: the window wraps around grid point 0
    ((3.0, 0.0), (3.0, 0.0), 0.5),           # a = 0: deterministic
    ((30.0, 0.0), (0.0, 25.0), 0.5),         # orthogonal, large norms: M ~ 2e4, x_grid = M/4
    ((2.0, 0.0), (-2.1, 0.05), 0.5),         # a ~ 1: the window straddles M/2
    ((1.0, 0.2), (0.9, 0.5), 3.0),           # M <= 9: whole grid enumerated
    ((1.0, 0.2), (0.9, 0.5), 1.5),           # M ~ 12: one or two tail cells per side
    ((1.0, 0.2), (0.7, -0.5), 1.0),          # M ~ 20
    ((5.0, 1.0), (1.0, 0.3), 0.5),           # generic, M ~ 100
    ((0.3, 0.1), (0.2, -0.4), 0.2),          # |dot| < 1
    ((12.0, -3.0), (11.0, -2.0), 0.1),       # close pair, fine grid
]


@pytest.mark.parametrize("replay", [0, 1])
@pytest.mark.parametrize("case", range(len(IPE_REGIMES)))
def test_ipe_distance_law_regimes(cuda, case, replay):
    """k = 1, n identical rows: the reported distance is one IPE estimate per row; its law must be the exact
    median-of-13 law of the reference's amplitude-estimation pmf in every regime of the grid.  replay = 1 sends
    every pair through the draw-by-draw path that normally serves only the rare pairs whose median may leave the
    enumerated window."""
    import gpu_util as G
    from sqlearn import _cabi
    xv, cv, delta = IPE_REGIMES[case]
    x, c = np.array(xv), np.array(cv)
    n = 30000
    _cabi.lib.sq_ipe_debug_force_replay(replay)
    try:
        _, mind, _, _ = G.assign(np.repeat(x[None], n, axis=0), c[None], mode=2, delta=delta, seed=21 + case + 100 * replay)
    finally:
        _cabi.lib.sq_ipe_debug_force_replay(0)
    support, pmf = _ipe_distance_pmf(x, c, delta / 2)
    idx = np.array([np.argmin(np.abs(support - v)) for v in mind])
    assert np.allclose(support[idx], mind, rtol=1e-9, atol=1e-9)
    obs = np.bincount(idx, minlength=len(support)).astype(float)
    big = pmf * n >= 8
    o = np.append(obs[big], obs[~big].sum())
    e = np.append(pmf[big], pmf[~big].sum()) * n
    keep = e > 1e-9
    if keep.sum() == 1:
        assert o[keep][0] == n
    else:
        assert o[~keep].sum() == 0
        assert stats.chisquare(o[keep], e[keep] * o[keep].sum() / e[keep].sum()).pvalue > P_MIN


@pytest.mark.parametrize("n,d,k", [(1000, 33, 300), (777, 128, 256), (100, 7, 10), (5000, 64, 600), (64, 16, 1)])
def test_ipe_mode_tends_to_exact_labels_for_small_delta(cuda, n, d, k):
    """every tile shape, several centroid tiles (k > 256), odd d (no TMA): with delta -> 0 the estimates converge to
    the true distances, so on separated blobs the labels are the exact ones and the reported distances are close"""
    import gpu_util as G
    rs = np.random.RandomState(n + k)
    cen = rs.randn(k, d) * 6
    lab = rs.randint(0, k, n)
    X = cen[lab] + 0.3 * rs.randn(n, d)
    le, me, _, _ = G.assign(X, cen, mode=0, delta=0.0, seed=1)
    li, mi, _, _ = G.assign(X, cen, mode=2, delta=1e-3, seed=2)
    assert np.array_equal(le, lab)
    assert np.array_equal(li, le)
    S = (X * X).sum(axis=1) + (cen[lab] ** 2).sum(axis=1)
    assert np.all(np.abs(mi - me) <= 0.01 * S + 1e-9)


PRUNE_CASES = [  # n, d, k, delta, how X and C are made
    (5000, 16, 40, 0.5, "blobs"), (3000, 128, 256, 0.5, "blobs"), (2000, 8, 300, 0.3, "rand"),
    (4000, 32, 17, 1.0, "unit"), (1000, 64, 600, 0.2, "blobs"), (3000, 5, 3, 2.0, "rand"),
    (2000, 128, 256, 0.5, "datacentres"), (777, 3, 1, 0.5, "rand"), (1500, 20, 33, 0.05, "unit"),
    (1200, 16, 24, 0.5, "zeros"), (1500, 24, 40, 0.5, "huge"), (1500, 24, 40, 1e-4, "tiny"), (900, 12, 20, 0.5, "opposite"),
]


@pytest.mark.parametrize("case", range(len(PRUNE_CASES)))
def test_ipe_pairs_left_out_change_nothing(cuda, case):
    """The IPE E-step does not sample a pair whose estimate certainly exceeds one already sampled in the same
    row (csrc/ipe_lean.cuh: ipe_out_possible - a certificate from the pair's own uniform).  With the hook at 2
    every pair is sampled: labels, distances and inertia must be the same bits - structured data, structureless
    data, unit-norm centroids (the steady state of the noisy loop), centres that are data points (its first
    iteration), one centroid, several centroid tiles."""
    import gpu_util as G
    from sqlearn import _cabi
    n, d, k, delta, kind = PRUNE_CASES[case]
    rng = np.random.default_rng(100 + case)
    cen = rng.normal(size=(k, d)) * 4
    X = cen[rng.integers(0, k, n)] + rng.normal(size=(n, d))
    if kind == "blobs":
        C = cen + 0.1 * rng.normal(size=(k, d))
    elif kind == "rand":
        X, C = rng.normal(size=(n, d)), rng.normal(size=(k, d))
    elif kind == "unit":
        C = cen / np.linalg.norm(cen, axis=1, keepdims=True)
    elif kind == "zeros":                 # null rows, a null centroid, duplicated centroids
        C = cen.copy()
        X[::7] = 0.0
        C[3] = 0.0
        C[5] = C[4]
    elif kind == "huge":                  # norms of 1e6 and of 1e-6 side by side
        C = cen * 1e6
        X = X * 1e6
        X[::5] *= 1e-12
    elif kind == "tiny":
        C = cen * 1e-3
        X = X * 1e-3
    elif kind == "opposite":              # nearly opposite vectors: the general estimator's pairs
        C = np.concatenate([-X[:k // 2] * (1.0 + 1e-9 * rng.normal(size=(k // 2, 1))), cen[:k - k // 2]])
    else:
        C = X[:k].copy()
    out = []
    for hook in (0, 2):
        _cabi.lib.sq_ipe_debug_force_replay(hook)
        try:
            lab, mind, inertia, _ = G.assign(X, C, mode=2, delta=delta, seed=11 + case)
        finally:
            _cabi.lib.sq_ipe_debug_force_replay(0)
        out.append((np.asarray(lab), np.asarray(mind), float(inertia)))
    np.testing.assert_array_equal(out[0][0], out[1][0])
    np.testing.assert_array_equal(out[0][1], out[1][1])
    assert out[0][2] == out[1][2]
    assert np.isfinite(out[0][1]).all() and (out[0][0] >= 0).all() and (out[0][0] < k).all()


FILTER_CASES = [  # n, d, k, delta, mode, data
    (5000, 128, 256, 0.0, 0, "blobs"), (5000, 128, 256, 0.5, 1, "blobs"), (3000, 16, 40, 0.5, 1, "blobs"),
    (3000, 64, 17, 0.0, 0, "rand"), (4000, 128, 256, 0.5, 1, "unit"), (2000, 32, 200, 2.0, 1, "rand"),
    (4097, 128, 256, 0.0, 0, "dup"), (1000, 48, 3, 0.3, 1, "blobs"), (70, 128, 256, 0.5, 1, "blobs"),
    (3000, 128, 256, 50.0, 1, "blobs"), (3000, 128, 100, 0.0, 0, "huge"), (2500, 96, 130, 0.0, 0, "tiny"),
    (1800, 128, 256, 0.5, 1, "datacentres"), (1200, 64, 64, 0.0, 0, "nan"),
    (2000, 10, 7, 0.0, 0, "blobs"), (2500, 50, 33, 0.5, 1, "blobs"), (1500, 100, 256, 0.0, 0, "rand"), (900, 2, 5, 0.3, 1, "blobs"),
    (3100, 126, 200, 0.5, 1, "unit"),
]


@pytest.mark.parametrize("case", range(len(FILTER_CASES)))
def test_bf16_filter_changes_nothing(cuda, case):
    """Exact and delta-window E-step: narrowing every row down to a few columns on the bf16 tensor cores (with a
    rigorous error bound) and deciding on fp64 distances of those columns (csrc/qmeans_filter.cuh) must give the
    labels of the fp64 tensor-core kernel, and its distances up to the rounding of two fp64 summation orders -
    whatever the scale of the data, for any even number of features up to 128, with duplicated centroids, null rows, windows wider than the filter keeps
    columns for, centres that are data points, and non-finite input (which both paths hand to the exact fix-up)."""
    import gpu_util as G
    from sqlearn import _cabi
    n, d, k, delta, mode, kind = FILTER_CASES[case]
    rng = np.random.default_rng(200 + case)
    cen = rng.normal(size=(k, d)) * 4
    X = cen[rng.integers(0, k, n)] + rng.normal(size=(n, d))
    C = cen + 0.1 * rng.normal(size=(k, d))
    if kind == "rand":
        X, C = rng.normal(size=(n, d)), rng.normal(size=(k, d))
    elif kind == "unit":
        C = cen / np.linalg.norm(cen, axis=1, keepdims=True)
    elif kind == "dup":
        C[5] = C[4]
        C[7] = C[4]
        X[::9] = 0.0
    elif kind == "huge":
        X, C = X * 1e6, C * 1e6
    elif kind == "tiny":
        X, C = X * 1e-20, C * 1e-20
    elif kind == "datacentres":
        C = X[:k].copy()
    elif kind == "nan":
        X[::101, 3] = np.nan
        X[50::101, 7] = np.inf
    out = []
    for on in (1, 0):
        _cabi.lib.sq_qmeans_debug_filter(on)
        try:
            lab, mind, inertia, nfix = G.assign(X, C, mode=mode, delta=delta, seed=5 + case)
        finally:
            _cabi.lib.sq_qmeans_debug_filter(1)
        out.append((np.asarray(lab), np.asarray(mind), float(inertia)))
    ok = np.isfinite(X).all(axis=1)
    np.testing.assert_array_equal(out[0][0][ok], out[1][0][ok])
    np.testing.assert_allclose(out[0][1][ok], out[1][1][ok], rtol=1e-11, atol=1e-11 * float(np.abs(X[ok]).max()) ** 2)
    if ok.all():
        assert abs(out[0][2] - out[1][2]) <= 1e-11 * abs(out[1][2])
