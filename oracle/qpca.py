"""NumPy restatement of the quantum part of sklearn/decomposition/_qPCA.py (reference).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py."""
import numpy as np

from . import quantum_utility as qu


def svd_flip_u(U, Vt):
    """sklearn/utils/extmath.py:552-557 (u-based decision)."""
    idx = np.argmax(np.abs(U), axis=0)
    signs = np.sign(U[idx, range(U.shape[1])])
    return U * signs, Vt * signs[:, None]


def fit_full(X, n_components):
    """_qPCA.py:578-636: centre, SVD, flip, variances, the first n_components rows of U, norms."""
    X = np.asarray(X, dtype=np.float64)
    mean = X.mean(axis=0)
    Xc = X - mean
    U, S, Vt = np.linalg.svd(Xc, full_matrices=False)
    U, Vt = svd_flip_u(U, Vt)
    ev = S ** 2 / (X.shape[0] - 1)
    out = dict(mean_=mean, components_=Vt[:n_components], singular_values_=S[:n_components],
               explained_variance_=ev[:n_components], explained_variance_ratio_=(ev / ev.sum())[:n_components],
               left_sv=U[:n_components], spectral_norm=S[0], frob_norm=np.linalg.norm(Xc),
               noise_variance_=ev[n_components:].mean() if n_components < min(X.shape) else 0.0)
    out["norm_muA"], out["muA"] = qu.best_mu(Xc, start=0, step=0.1)
    return out


def consistent_phase_estimation(omega, epsilon, gamma, rng):
    """Utility.py:773-792."""
    n = qu.pe_m(epsilon, gamma)
    delta_prime = epsilon * (gamma / n) / 2
    pe = qu.phase_estimation(omega, epsilon=delta_prime, gamma=gamma, rng=rng)
    return qu.consistent_bucket(pe, epsilon, gamma)[0]


def estimate_singular_values(sv, muA, eps, n_features, rng):
    """_qPCA.py:1031-1037."""
    e = eps / muA
    theta = [2 * np.arccos(s / muA) / (e + np.pi) for s in sv]
    est = [consistent_phase_estimation(t, e, 1 - 1 / n_features, rng) for t in theta]
    return np.array([np.cos(t * (e + np.pi) / 2) * muA for t in est])
