"""Generate the golden fixtures of tests/golden/ by running the REFERENCE's own
source (/root/reference, loaded unmodified by oracle/refload.py).

Run once in the build container:  python tests/golden/gen_golden.py
The GPU box has no /root/reference; tests there read only the committed files.
Versions in effect when these were generated are recorded in deterministic.json.
"""
import contextlib
import io
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import refload  # noqa: E402

warnings.filterwarnings("ignore")
u = refload.load_utility()
dm = refload.load_dmeans(u)
warnings.filterwarnings("ignore")  # Utility.py switches UserWarning back to 'always'
refload.seed_everything(u, 20261106)


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return fn(*a, **k)


# ---------------------------------------------------------------- deterministic
def ref_N(d, delta, norm):
    return int((36 * d * np.log(d)) / (delta ** 2)) if norm == "L2" else int((36 * np.log(d)) / (delta ** 2))


det = {"versions": {"numpy": np.__version__, "scipy": __import__("scipy").__version__,
                    "python": sys.version.split()[0]}}
# N is computed inline in real_tomography (Utility.py:307-311); recover it from a single-shot run's dict key
det["N"] = []
for d, delta in [(4, 0.5), (16, 0.3), (128, 0.25), (784, 0.25), (1000, 0.5)]:
    for norm in ("L2", "inf"):
        v = np.ones(d) / np.sqrt(d)
        if ref_N(d, delta, norm) <= 4_000_000:
            key = list(quiet(u.real_tomography, v, delta=delta, norm=norm, incremental_measure=False).keys())[0]
        else:
            key = ref_N(d, delta, norm)
        det["N"].append([d, delta, norm, int(key)])
det["schedule"] = []
for N, fmi in [(357730, 0), (17744, 0), (994, 0), (50, 0), (357730, 7), (4973583800867, 0)]:
    s = u.check_measure(np.geomspace(1, N, num=100, dtype=np.int64), faster_measure_increment=fmi)
    det["schedule"].append([N, fmi, [int(x) for x in s]])
det["median_Q"] = []
for g in (0.1, 0.05, 0.01, 0.3):
    calls = []
    quiet(u.median_evaluation, lambda: calls.append(1) or 0.0, gamma=g)
    det["median_Q"].append([g, len(calls)])
det["ae_M"] = [[e, int(quiet(u.amplitude_estimation, 0.3, epsilon=e, nqubit=True)[2])] for e in (0.1, 0.01, 0.001, 0.25)]
det["pe_m"] = [[e, g, int(quiet(u.phase_estimation, 0.3, epsilon=e, gamma=g, nqubit=True)[2])]
               for e, g in [(0.05, 0.1), (0.001, 0.1), (0.01, 0.5)]]
det["check_division"] = [[v, j, u.check_division(v, j)] for v, j in [(10, 3), (7, 7), (100, 8)]]
# consistent phase estimation bucketing with the PE outcome pinned
det["cpe_bucket"] = []
orig_pe = u.phase_estimation
for eps, gamma in [(0.05, 0.1), (0.01, 0.9987), (0.003, 0.5)]:
    for pe in (0.0, 0.123456, 0.5, 0.73, 0.999):
        u.phase_estimation = lambda omega, epsilon, gamma, _pe=pe: _pe
        det["cpe_bucket"].append([eps, gamma, pe, float(u.consistent_phase_estimation(0.3, eps, gamma))])
u.phase_estimation = orig_pe
det["wrap"] = [[x, u.wrapper_phase_est_arguments(x), u.unwrap_phase_est_arguments(u.wrapper_phase_est_arguments(x) / (0.01 + np.pi), 0.01)]
               for x in (0.1, 0.5, 0.9)]
rs = np.random.RandomState(5)
A = rs.randn(40, 12) * (rs.rand(40, 12) > 0.3)
det["best_mu"] = {"A": A.tolist(),
                  "grid_0_0.1": list(u.best_mu(A, 0, 0.1)), "grid_0_1_0.1": list(u.best_mu(A, start=0, step=0.1)),
                  "default": list(u.best_mu(A))}
with open(os.path.join(HERE, "deterministic.json"), "w") as f:
    json.dump(det, f, indent=1, default=float)

# ---------------------------------------------------------------- noiseless q-means
cases = {}
rs = np.random.RandomState(11)


def blobs(n, d, k, spread):
    cen = rs.randn(k, d) * spread
    lab = rs.randint(0, k, n)
    return cen[lab] + rs.randn(n, d)


for name, (n, d, k, spread) in {"a": (400, 8, 5, 4.0), "b": (300, 33, 3, 2.0), "c": (1000, 128, 17, 1.0)}.items():
    X = blobs(n, d, k, spread)
    X = X - X.mean(axis=0)
    C0 = X[rs.permutation(n)[:k]].copy()
    labels, inertia, centers, n_iter = quiet(dm._kmeans_single_lloyd, X.copy(), None, C0.copy(), False, True, True,
                                             None, True, max_iter=8, tol=0, delta=0)
    lab1, _, in1 = quiet(dm.labels_estimation, X, C0, 0, None, True)
    cases[name] = dict(X=X, C0=C0, labels=np.asarray(labels, dtype=np.int32), inertia=inertia, centers=centers,
                       n_iter=n_iter, labels_first=np.asarray(lab1, dtype=np.int32), inertia_first=in1)
# the docstring toy (_dmeans.py:1005-1012): integer data with exact ties in later iterations
Xt = np.array([[1, 2], [1, 4], [1, 0], [10, 2], [10, 4], [10, 0]], dtype=float)
Xc = Xt - Xt.mean(axis=0)
labels, inertia, centers, n_iter = quiet(dm._kmeans_single_lloyd, Xc.copy(), None, Xc[[0, 3]].copy(), False, True,
                                         True, None, True, max_iter=10, tol=0, delta=0)
cases["toy"] = dict(X=Xc, C0=Xc[[0, 3]], labels=np.asarray(labels, dtype=np.int32), inertia=inertia,
                    centers=centers, n_iter=n_iter)
np.savez_compressed(os.path.join(HERE, "qmeans_noiseless.npz"),
                    **{f"{c}_{k}": v for c, dct in cases.items() for k, v in dct.items()})

# ---------------------------------------------------------------- stochastic reference samples
st = {}
v16 = np.random.RandomState(3).uniform(-1, 1, 16)
v16 /= np.linalg.norm(v16)
st["tomo_v"] = v16
# default mode (incremental + stop), delta 0.3: error and stopping N
errs, stops = [], []
for _ in range(600):
    r = quiet(u.real_tomography, v16, delta=0.3)
    n_used = list(r.keys())[-1]
    errs.append(np.linalg.norm(v16 - np.asarray(r[n_used])))
    stops.append(n_used)
st["tomo_default_err"], st["tomo_default_N"] = np.sort(errs), np.sort(stops)
# single shot, L2 and Linf
for norm in ("L2", "inf"):
    errs, comp0 = [], []
    for _ in range(300):
        r = quiet(u.real_tomography, v16, delta=0.3, norm=norm, incremental_measure=False)
        est = np.asarray(list(r.values())[0])
        errs.append(np.linalg.norm(v16 - est, ord=2 if norm == "L2" else np.inf))
        comp0.append(est[0])
    st[f"tomo_single_{norm}_err"], st[f"tomo_single_{norm}_c0"] = np.sort(errs), np.sort(comp0)
# small-N attempts: estimate of component 3 and sign error counts at N = 40
sgn = 0
c3 = []
for _ in range(1500):
    r = quiet(u.real_tomography, v16, N=40, delta=0.3, incremental_measure=False)
    est = np.asarray(list(r.values())[0])
    sgn += int(np.sum(np.sign(est) != np.sign(v16)))
    c3.append(est[3])
st["tomo_N40_signerr"], st["tomo_N40_c3"] = np.asarray([sgn, 1500 * 16]), np.sort(c3)
# un-normalised input is normalised (D5)
st["tomo_scaled_norm"] = np.asarray([np.linalg.norm(quiet(u.tomography, 5 * v16, 0.3)) for _ in range(20)])
# amplitude / phase estimation draws
st["ae_draws_a0.3_e0.05"] = np.sort([quiet(u.amplitude_estimation, 0.3, epsilon=0.05) for _ in range(4000)])
st["ae_median_a0.3_e0.05_g0.1"] = np.sort([quiet(u.amplitude_estimation, 0.3, epsilon=0.05, gamma=0.1) for _ in range(1500)])
st["pe_draws_w0.3_m6"] = np.sort([quiet(u.phase_estimation, 0.3, m=6) for _ in range(4000)])
x = np.random.RandomState(4).randn(6)
y = np.random.RandomState(5).randn(6)
st["ipe_x"], st["ipe_y"] = x, y
st["ipe_draws_e0.25"] = np.sort([quiet(u.ipe, x, y, 0.25) for _ in range(1500)])
# truncated normal error injection
st["gauss_est_dev"] = np.sort(np.concatenate([quiet(u.make_gaussian_est, np.zeros(8), 0.5) for _ in range(400)]))
# delta-window label choice: one row, three centres inside the window, one outside
Xw = np.zeros((1, 2))
Cw = np.array([[1.0, 0.0], [0.0, 1.05], [-1.1, 0.0], [3.0, 0.0]])
cnt = np.zeros(4, dtype=np.int64)
for _ in range(3000):
    lab, _, _ = quiet(dm.labels_estimation, Xw, Cw, 0.5, None, False)
    cnt[int(lab[0])] += 1
st["window_X"], st["window_C"], st["window_counts"] = Xw, Cw, cnt
# IPE-mode labels: fraction assigned to the true nearest centre
Xi = np.random.RandomState(6).randn(30, 4)
Ci = Xi[:3] + 0.1
hits = 0
truth = np.argmin(((Xi[:, None, :] - Ci[None]) ** 2).sum(-1), axis=1)
ipe_labels = []
for _ in range(40):
    lab, _, _ = quiet(dm.labels_estimation, Xi, Ci, 0.5, None, True)
    ipe_labels.append(np.asarray(lab))
st["ipe_X"], st["ipe_C"], st["ipe_labels"] = Xi, Ci, np.asarray(ipe_labels, dtype=np.int32)
# a pair whose amplitude-estimation phase falls between two grid points: the median of 13 varies
y2 = y * 0.816
st["ipe_y2"] = y2
st["ipe2_draws_e0.25"] = np.sort([quiet(u.ipe, x, y2, 0.25) for _ in range(1500)])
np.savez_compressed(os.path.join(HERE, "stochastic_reference.npz"), **st)
print("golden fixtures written to", HERE)

# ---------------------------------------------------------------- qPCA (reference class, AST-loaded)
QP = refload.load_qpca(u)
warnings.filterwarnings("ignore")
qp = {}
rs = np.random.RandomState(21)
Xp = rs.randn(300, 12) @ np.diag(np.linspace(6, 0.5, 12)) @ rs.randn(12, 12) + rs.randn(12) * 3
qp["X"] = Xp
ref = quiet(lambda: QP(n_components=5, svd_solver='full').fit(Xp.copy()))
for name in ("mean_", "components_", "singular_values_", "explained_variance_", "explained_variance_ratio_",
             "left_sv", "noise_variance_", "spectral_norm", "frob_norm", "muA"):
    qp[name] = np.asarray(getattr(ref, name))
qp["norm_muA"] = np.asarray([ref.norm_muA])
qp["transform"] = quiet(ref.transform, Xp[:40])
# the stock-test anchor: a matrix with prescribed singular values (sklearn/decomposition/tests/test_pca.py:159-189)
rs2 = np.random.RandomState(0)
Xs = rs2.randn(100, 110)
Us, _, Vs = np.linalg.svd(Xs - Xs.mean(axis=0), full_matrices=False)
Xs_hat = (Us[:, :3] * np.array([3.142, 2.718, 1.0])) @ Vs[:3]
qp["Xs_hat"] = Xs_hat
qp["Xs_sv"] = quiet(lambda: QP(n_components=3, svd_solver='full').fit(Xs_hat.copy())).singular_values_
# stochastic: estimated singular values, error of the estimated vectors, representation error
theta = float(ref.singular_values_[3]) * 0.98
est_sv, right_err, left_err, topk, frep = [], [], [], [], []
for _ in range(60):
    m = quiet(lambda: QP(n_components=5, svd_solver='full').fit(Xp.copy(), estimate_all=True, eps=0.5, delta=0.25,
                                                                 theta_major=theta, true_tomography=True))
    topk.append(m.topk)
    est_sv.append(np.pad(m.estimate_s_values, (0, 5 - len(m.estimate_s_values)), constant_values=np.nan))
    right_err.extend(np.linalg.norm(m.estimate_right_sv - m.topk_right_singular_vectors, axis=1))
    lv = m.topk_left_singular_vectors
    left_err.extend(np.linalg.norm(m.estimate_left_sv - lv / np.linalg.norm(lv, axis=1, keepdims=True), axis=1))
    r = quiet(m.transform, Xp[:30], classic_transform=False, epsilon_delta=0.3, quantum_representation=True,
              norm='est_representation')
    Y = m.transform(Xp[:30]) if False else (Xp[:30] - m.mean_) @ m.components_.T
    A = r['quantum_representation_results'][0]
    frep.extend(np.linalg.norm(A - Y / np.linalg.norm(Y, axis=1, keepdims=True), axis=1))
qp["theta"] = np.asarray([theta])
qp["topk"] = np.asarray(topk)
qp["est_sv"] = np.asarray(est_sv)
qp["right_err"], qp["left_err"], qp["rep_err"] = np.sort(right_err), np.sort(left_err), np.sort(frep)
p_est = [quiet(m.quantum_factor_score_ratio_sum, 0.05, theta / m.muA, 0.05) for _ in range(200)]
qp["p_est"] = np.sort(p_est)
np.savez_compressed(os.path.join(HERE, "qpca_reference.npz"), **qp)
print("qPCA fixtures written")
