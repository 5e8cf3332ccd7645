"""Second set of golden fixtures: 10^4-trial samples of the reference's stochastic routines,
draws at the BASELINE shapes (d = 128 delta = 0.25, d = 784), runs of the reference's NOISY Lloyd
loop, and the qPCA binary-search estimators.  Everything is produced by the REFERENCE's own source
(/root/reference, loaded unmodified by oracle/refload.py); nothing here is computed by this repo.

  python tests/golden/gen_golden_r2.py [--jobs 8] [--only NAME[,NAME...]]

Each task is cut into chunks; a chunk runs in a worker process with the reference's RNGs seeded
from (task name, chunk index), so a rerun reproduces the files bit for bit.  Build container only
(the GPU box has no /root/reference); tests read the committed .npz files.
"""
import argparse
import contextlib
import io
import os
import re
import sys
import time
import warnings
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

_state = {}


def _ref():
    """Reference modules, loaded once per worker process."""
    if not _state:
        from oracle import refload
        warnings.filterwarnings("ignore")
        u = refload.load_utility()
        dm = refload.load_dmeans(u)
        warnings.filterwarnings("ignore")
        _state.update(u=u, dm=dm, refload=refload)
    return _state["u"], _state["dm"], _state["refload"]


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        return fn(*a, **k)


def unit_vector(d, seed=3):
    v = np.random.RandomState(seed).uniform(-1, 1, d)
    return v / np.linalg.norm(v)


# ------------------------------------------------------------------ tasks: (count) -> dict of lists
def t_tomo(count, d, delta, norm, incremental, N=None):
    u, _, _ = _ref()
    v = unit_vector(d)
    out = {"err": [], "N": [], "c0": [], "c3": [], "signerr": []}
    for _ in range(count):
        r = quiet(u.real_tomography, v, N=N, delta=delta, norm=norm, incremental_measure=incremental)
        n_used = list(r.keys())[-1]
        est = np.asarray(r[n_used])
        out["err"].append(np.linalg.norm(v - est, ord=2 if norm == "L2" else np.inf))
        out["N"].append(n_used)
        out["c0"].append(est[0])
        out["c3"].append(est[3])
        out["signerr"].append(int(np.sum(np.sign(est) != np.sign(v))))
    return out


def t_scalar(count, which):
    u, _, _ = _ref()
    x = np.random.RandomState(4).randn(6)
    y = np.random.RandomState(5).randn(6)
    fn = {"ae": lambda: u.amplitude_estimation(0.3, epsilon=0.05),
          "ae_med": lambda: u.amplitude_estimation(0.3, epsilon=0.05, gamma=0.1),
          "ae_small": lambda: u.amplitude_estimation(0.013, epsilon=0.01),
          "pe": lambda: u.phase_estimation(0.3, m=6),
          "pe_eps": lambda: u.phase_estimation(0.7123, epsilon=0.02, gamma=0.1),
          "cpe": lambda: u.consistent_phase_estimation(0.3, 0.05, 0.1),
          "ipe": lambda: u.ipe(x, y, 0.25),
          "ipe2": lambda: u.ipe(x, y * 0.816, 0.25),
          "ipe3": lambda: u.ipe(x * 0.2, x * 0.21 + 0.01, 0.05),
          "gauss": lambda: u.make_gaussian_est(np.zeros(8), 0.5)[0],
          "intro": lambda: float(u.introduce_error(1.5, 0.3)[0])}[which]
    return {"x": [float(quiet(fn)) for _ in range(count)]}


def t_window(count):
    _, dm, _ = _ref()
    Xw = np.zeros((1, 2))
    Cw = np.array([[1.0, 0.0], [0.0, 1.05], [-1.1, 0.0], [3.0, 0.0]])
    return {"label": [int(quiet(dm.labels_estimation, Xw, Cw, 0.5, None, False)[0][0]) for _ in range(count)]}


def ipe_label_problem():
    Xi = np.random.RandomState(6).randn(30, 4)
    return Xi, Xi[:3] + 0.1


def t_ipe_labels(count):
    _, dm, _ = _ref()
    Xi, Ci = ipe_label_problem()
    return {"labels": [np.asarray(quiet(dm.labels_estimation, Xi, Ci, 0.5, None, True)[0], dtype=np.int32)
                       for _ in range(count)]}


def lloyd_problem(which):
    """Two small problems for the noisy loop: 'w' (400 x 8, k = 5: delta-window E-step) and
    'i' (60 x 4, k = 3, small norms so that one IPE call costs milliseconds)."""
    rs = np.random.RandomState(11 if which == "w" else 12)
    n, d, k, spread = (400, 8, 5, 4.0) if which == "w" else (60, 4, 3, 1.5)
    cen = rs.randn(k, d) * spread
    X = cen[rs.randint(0, k, n)] + rs.randn(n, d) * (1.0 if which == "w" else 0.5)
    X = X - X.mean(axis=0)
    C0 = X[rs.permutation(n)[:k]].copy()
    return X, C0


def t_lloyd(count, which, max_iter):
    """_kmeans_single_lloyd(delta = 0.5, intermediate_err = True) of the reference.  The per-iteration
    inertia and centre shift are read from what the function prints (_dmeans.py:643-644, :652)."""
    _, dm, _ = _ref()
    X, C0 = lloyd_problem(which)
    out = {"inertia": [], "n_iter": [], "shifts": [], "it_inertia": [], "center_norms": [], "label_counts": [],
           "failed": []}
    for _ in range(count):
        buf = io.StringIO()
        try:
            with contextlib.redirect_stdout(buf), warnings.catch_warnings():
                warnings.simplefilter("ignore")
                labels, inertia, centers, n_iter = dm._kmeans_single_lloyd(
                    X.copy(), None, C0.copy(), True, True, True, None, which == "i", max_iter=max_iter, verbose=True,
                    tol=0, delta=0.5)
        except (IndexError, ValueError):
            # an emptied cluster makes the reference's M-step fail (_dmeans.py:812-823): counted, not sampled
            out["failed"].append(1)
            continue
        out["failed"].append(0)
        its = [float(m) for m in re.findall(r"Iteration\s+\d+, inertia ([-0-9.e+]+)", buf.getvalue())]
        sh = [float(line.split()[0]) for line in buf.getvalue().splitlines()
              if re.fullmatch(r"[-0-9.e+]+ [-0-9.e+]+", line.strip())]
        assert len(its) == n_iter and len(sh) == n_iter, (len(its), len(sh), n_iter)
        out["inertia"].append(inertia)
        out["n_iter"].append(n_iter)
        out["shifts"].append(sh)
        out["it_inertia"].append(its)
        out["center_norms"].append(np.linalg.norm(centers, axis=1))
        out["label_counts"].append(np.bincount(np.asarray(labels, dtype=np.int64), minlength=len(C0)))
    return out


def qpca_problem():
    rs = np.random.RandomState(21)
    return rs.randn(300, 12) @ np.diag(np.linspace(6, 0.5, 12)) @ rs.randn(12, 12) + rs.randn(12) * 3


def _qpca_fitted():
    if "qp" not in _state:
        u, _, refload = _ref()
        QP = refload.load_qpca(u)
        warnings.filterwarnings("ignore")
        Xp = qpca_problem()
        _state["QP"] = QP
        _state["qp"] = quiet(lambda: QP(n_components=5, svd_solver='full').fit(Xp.copy()))
    return _state["QP"], _state["qp"]


def t_qpca_scalar(count, which):
    _, m = _qpca_fitted()
    theta = float(m.singular_values_[3]) * 0.98
    fn = {"p_est": lambda: m.quantum_factor_score_ratio_sum(0.05, theta / m.muA, 0.05),
          "theta": lambda: m.estimate_theta(2.0, 0.2, 0.7),
          "theta_fine": lambda: m.estimate_theta(0.5, 0.1, 0.85),
          "spectral": lambda: m.spectral_norm_estimation(5.0, 0.05),
          "spectral_fine": lambda: m.spectral_norm_estimation(1.0, 0.02)}[which]
    out = []
    for _ in range(count):
        try:
            out.append(float(quiet(fn)))
        except ValueError:          # "The binary search doesn't found any values" (_qPCA.py:1022)
            out.append(np.nan)
    return {"x": out}


def t_qpca_fit(count):
    QP, ref = _qpca_fitted()
    Xp = qpca_problem()
    theta = float(ref.singular_values_[3]) * 0.98
    out = {"topk": [], "est_sv": [], "right_err": [], "left_err": [], "rep_err": []}
    for _ in range(count):
        m = quiet(lambda: QP(n_components=5, svd_solver='full').fit(
            Xp.copy(), estimate_all=True, eps=0.5, delta=0.25, theta_major=theta, true_tomography=True))
        out["topk"].append(m.topk)
        out["est_sv"].append(np.pad(m.estimate_s_values, (0, 5 - len(m.estimate_s_values)), constant_values=np.nan))
        out["right_err"].append(np.pad(np.linalg.norm(m.estimate_right_sv - m.topk_right_singular_vectors, axis=1),
                                       (0, 5 - m.topk), constant_values=np.nan))
        lv = m.topk_left_singular_vectors
        out["left_err"].append(np.pad(
            np.linalg.norm(m.estimate_left_sv - lv / np.linalg.norm(lv, axis=1, keepdims=True), axis=1),
            (0, 5 - m.topk), constant_values=np.nan))
        r = quiet(m.transform, Xp[:30], classic_transform=False, epsilon_delta=0.3, quantum_representation=True,
                  norm='est_representation')
        Y = (Xp[:30] - m.mean_) @ m.components_.T
        A = r['quantum_representation_results'][0]
        out["rep_err"].append(np.linalg.norm(A - Y / np.linalg.norm(Y, axis=1, keepdims=True), axis=1))
    return out


def rank_deficient_problem():
    """200 x 10 of rank 7: the trailing singular values are zero to rounding, which is what
    least_k_sv_extractors looks for (np.isclose(singular_values_, 0), _qPCA.py:1078)."""
    rs = np.random.RandomState(31)
    return rs.randn(200, 7) @ np.diag(np.linspace(5, 1, 7)) @ rs.randn(7, 10) + rs.randn(10)


def t_qpca_least(count):
    QP, _ = _qpca_fitted()
    Xr = rank_deficient_problem()
    base = quiet(lambda: QP(n_components=10, svd_solver='full').fit(Xr.copy()))
    theta_minor = float(base.singular_values_[3]) * 1.03
    out = {"least_k": [], "least_sv": [], "right_err": [], "left_err": [], "least_k_p": []}
    for _ in range(count):
        m = quiet(lambda: QP(n_components=10, svd_solver='full').fit(
            Xr.copy(), estimate_least_k=True, eps=0.3, delta=0.25, theta_minor=theta_minor, true_tomography=True))
        out["least_k"].append(m.least_k)
        out["least_k_p"].append(m.least_k_p)
        ls = np.asarray(m.estimate_least_s_values, dtype=float)
        out["least_sv"].append(np.pad(ls, (0, 10 - len(ls)), constant_values=np.nan))
        out["right_err"].append(np.pad(np.linalg.norm(m.estimate_least_right_sv - m.leastk_right_singular_vectors, axis=1),
                                       (0, 10 - m.least_k), constant_values=np.nan))
        lv = m.leastk_left_singular_vectors
        out["left_err"].append(np.pad(
            np.linalg.norm(m.estimate_least_left_sv - lv / np.linalg.norm(lv, axis=1, keepdims=True), axis=1),
            (0, 10 - m.least_k), constant_values=np.nan))
    return out


def t_qstate(count):
    """QuantumState.measure / estimate_wald / coupon_collect (Utility.py:25-85) on a 6-outcome register."""
    u, _, _ = _ref()
    amp = np.array([0.6, -0.2, 0.45, 0.1, -0.5, 0.35])
    q = u.QuantumState(registers=list(range(6)), amplitudes=amp)
    out = {"coupon": [], "wald50": []}
    for _ in range(count):
        out["coupon"].append(int(u.coupon_collect(q)))
        w = u.estimate_wald(q.measure(50))
        out["wald50"].append([w.get(i, 0.0) for i in range(6)])
    return out


# name -> (function, total trials, chunk size, kwargs)
TASKS = {
    "tomo16_default":   (t_tomo, 10000, 500, dict(d=16, delta=0.3, norm="L2", incremental=True)),
    "tomo16_single_L2": (t_tomo, 10000, 250, dict(d=16, delta=0.3, norm="L2", incremental=False)),
    "tomo16_single_inf": (t_tomo, 10000, 500, dict(d=16, delta=0.3, norm="inf", incremental=False)),
    "tomo16_default_inf": (t_tomo, 10000, 500, dict(d=16, delta=0.3, norm="inf", incremental=True)),
    "tomo16_N40":       (t_tomo, 10000, 500, dict(d=16, delta=0.3, norm="L2", incremental=False, N=40)),
    "tomo128_default":  (t_tomo, 10000, 250, dict(d=128, delta=0.25, norm="L2", incremental=True)),
    "tomo128_default_inf": (t_tomo, 10000, 250, dict(d=128, delta=0.25, norm="inf", incremental=True)),
    "tomo128_single_L2": (t_tomo, 2500, 50, dict(d=128, delta=0.25, norm="L2", incremental=False)),
    "tomo128_single_inf": (t_tomo, 10000, 500, dict(d=128, delta=0.25, norm="inf", incremental=False)),
    "tomo784_default":  (t_tomo, 1000, 25, dict(d=784, delta=0.5, norm="L2", incremental=True)),
    "tomo784_single_L2": (t_tomo, 400, 10, dict(d=784, delta=0.5, norm="L2", incremental=False)),
    "tomo784_single_inf": (t_tomo, 4000, 100, dict(d=784, delta=0.1, norm="inf", incremental=False)),
    "tomo784_default_inf": (t_tomo, 2000, 50, dict(d=784, delta=0.1, norm="inf", incremental=True)),
    "ae":       (t_scalar, 10000, 2500, dict(which="ae")),
    "ae_med":   (t_scalar, 10000, 1000, dict(which="ae_med")),
    "ae_small": (t_scalar, 10000, 1000, dict(which="ae_small")),
    "pe":       (t_scalar, 10000, 2500, dict(which="pe")),
    "pe_eps":   (t_scalar, 10000, 1000, dict(which="pe_eps")),
    "cpe":      (t_scalar, 10000, 1000, dict(which="cpe")),
    "ipe":      (t_scalar, 10000, 500, dict(which="ipe")),
    "ipe2":     (t_scalar, 10000, 500, dict(which="ipe2")),
    "ipe3":     (t_scalar, 4000, 100, dict(which="ipe3")),
    "gauss":    (t_scalar, 10000, 2500, dict(which="gauss")),
    "intro":    (t_scalar, 10000, 2500, dict(which="intro")),
    "window":   (t_window, 10000, 2500, {}),
    "ipe_labels": (t_ipe_labels, 1000, 25, {}),
    "lloyd_w":  (t_lloyd, 400, 10, dict(which="w", max_iter=8)),
    "lloyd_i":  (t_lloyd, 200, 2, dict(which="i", max_iter=5)),
    "qpca_p_est": (t_qpca_scalar, 10000, 500, dict(which="p_est")),
    "qpca_theta": (t_qpca_scalar, 1000, 25, dict(which="theta")),
    "qpca_theta_fine": (t_qpca_scalar, 300, 10, dict(which="theta_fine")),
    "qpca_spectral": (t_qpca_scalar, 1000, 25, dict(which="spectral")),
    "qpca_spectral_fine": (t_qpca_scalar, 200, 5, dict(which="spectral_fine")),
    "qpca_fit": (t_qpca_fit, 400, 10, {}),
    "qpca_least": (t_qpca_least, 400, 10, {}),
    "qstate":   (t_qstate, 10000, 1000, {}),
}
FILES = {"stochastic_reference_r2.npz": [k for k in TASKS if k.startswith(("tomo", "ae", "pe", "cpe", "ipe", "gauss",
                                                                           "intro", "window", "qstate"))],
         "noisy_lloyd_reference.npz": ["lloyd_w", "lloyd_i"],
         "qpca_reference_r2.npz": [k for k in TASKS if k.startswith("qpca")]}


def run_chunk(job):
    name, idx, count = job
    fn, _, _, kw = TASKS[name]
    u, _, refload = _ref()
    refload.seed_everything(u, zlib.crc32(("%s/%d" % (name, idx)).encode()) & 0x7FFFFFFF)
    t0 = time.perf_counter()
    res = fn(count, **kw)
    return name, idx, res, time.perf_counter() - t0


def main():
    import multiprocessing as mp
    ap = argparse.ArgumentParser()
    ap.add_argument("--jobs", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--only", default="")
    ap.add_argument("--scale", type=float, default=1.0, help="multiply every trial count (smoke runs)")
    args = ap.parse_args()
    only = [s for s in args.only.split(",") if s]
    jobs = []
    for name, (_, total, chunk, _) in TASKS.items():
        if only and name not in only:
            continue
        total = max(1, int(total * args.scale))
        chunk = min(chunk, total)
        jobs += [(name, i, min(chunk, total - i * chunk)) for i in range((total + chunk - 1) // chunk)]
    # long chunks first
    jobs.sort(key=lambda j: -{"lloyd_i": 9, "qpca_spectral_fine": 8, "qpca_theta_fine": 7, "qpca_theta": 6, "qpca_spectral": 5}.get(j[0], 0))
    parts, spent = {}, {}
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(args.jobs) as pool:
        for name, idx, res, dt in pool.imap_unordered(run_chunk, jobs):
            parts.setdefault(name, {})[idx] = res
            spent[name] = spent.get(name, 0.0) + dt
            done = sum(len(v) for v in parts.values())
            if done % 20 == 0:
                print("%d / %d chunks, %.0f s" % (done, len(jobs), time.perf_counter() - t0), flush=True)
    for fname, names in FILES.items():
        names = [n for n in names if n in parts]
        if not names:
            continue
        path = os.path.join(HERE, fname)
        data = dict(np.load(path)) if (only and os.path.exists(path)) else {}
        for name in names:
            for key in [k for k in data if k.startswith(name + "__")]:
                del data[key]
            chunks = [parts[name][i] for i in sorted(parts[name])]
            for key in chunks[0]:
                rows = [r for c in chunks for r in c[key]]
                data["%s__%s" % (name, key)] = np.asarray(rows)
            data["%s__cpu_seconds" % name] = np.asarray([spent[name]])
        np.savez_compressed(path, **data)
        print("wrote", path, {n: round(spent[n], 1) for n in names})
    # problems the tests must rebuild identically
    X, C0 = lloyd_problem("w")
    Xi, C0i = lloyd_problem("i")
    np.savez_compressed(os.path.join(HERE, "r2_problems.npz"), lloyd_w_X=X, lloyd_w_C0=C0, lloyd_i_X=Xi, lloyd_i_C0=C0i,
                        v16=unit_vector(16), v128=unit_vector(128), v784=unit_vector(784), qpca_X=qpca_problem(),
                        ipe_X=ipe_label_problem()[0], ipe_C=ipe_label_problem()[1], rankdef_X=rank_deficient_problem())


if __name__ == "__main__":
    main()
