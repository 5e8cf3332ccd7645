"""TEST INFRASTRUCTURE - CPU restatement of the reference's QLSSVC (sklearn/svm/_qSVM.py:10-404), the third
consumer of QuantumUtility.  NumPy only; every function cites the lines it follows.  Pinned by
tests/golden/qsvm.npz, produced by running the reference's own class (oracle/refload.py: load_qsvm).  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this package; the product never does."""
import numpy as np
from scipy.stats import truncnorm


def gamma_of(X, gamma, n_features):
    """_get_gamma (_qSVM.py:392-404): 'scale' -> 1 / (n_features * X.var()), 'auto' -> 1 / n_features."""
    X = np.asarray(X)
    if gamma == 'scale':
        return 1.0 / (X.shape[1] * X.var())
    if gamma == 'auto':
        return 1.0 / n_features
    return None   # the reference returns None for a float gamma too (the pairwise kernels then use 1 / n_features)


def kernel(X, Y, kind, gamma=None, coef0=0.0, degree=3):
    """get_kernel (_qSVM.py:375-389) over scikit-learn's pairwise kernels (metrics/pairwise.py): linear X Y^T,
    poly (g X Y^T + c)^degree, rbf exp(-g ||x - y||^2), sigmoid tanh(g X Y^T + c); g defaults to 1 / n_features."""
    X = np.asarray(X, dtype=np.float64)
    Y = X if Y is None else np.asarray(Y, dtype=np.float64)
    G = X @ Y.T
    if kind == 'linear':
        return G
    g = 1.0 / X.shape[1] if gamma is None else gamma
    if kind == 'poly':
        return (g * G + coef0) ** degree
    if kind == 'sigmoid':
        return np.tanh(g * G + coef0)
    if kind == 'rbf':
        d2 = (X * X).sum(axis=1)[:, None] + (Y * Y).sum(axis=1)[None, :] - 2.0 * G
        np.maximum(d2, 0.0, out=d2)
        if Y is X:
            np.fill_diagonal(d2, 0.0)      # euclidean_distances zeroes the diagonal when Y is X
        return np.exp(-g * d2)
    raise ValueError(kind)


def classical_fit(X, y, kind='linear', penalty=0.1, gamma='scale', coef0=0.0, degree=3, low_rank=False, var=0.9):
    """_classical_fit (_qSVM.py:84-130): F = [[0, 1^T], [1, K + I / penalty]], pseudo-inverse through the
    Hermitian SVD (optionally truncated to the leading singular values), [b; alpha] = F^+ [0; y].
    Returns dict(b, alpha, cond, normF, singular_values_F_)."""
    X = np.asarray(X, dtype=np.float64)
    N = len(X)
    g = gamma_of(X, gamma, X.shape[1])
    K = kernel(X, None, kind, g, coef0, degree)
    F = np.zeros((N + 1, N + 1))
    F[0, 1:] = 1.0
    F[1:, 0] = 1.0
    F[1:, 1:] = K + np.identity(N) / penalty
    w, U = np.linalg.eigh(F)
    order = np.argsort(-np.abs(w), kind='stable')
    s, u = np.abs(w[order]), U[:, order]
    vh = (u * np.sign(w[order])).T                 # np.linalg.svd(F, hermitian=True)
    if low_rank:
        s_new = np.zeros(len(s))
        if 0 <= var < 1.0:
            p, sums = 0.0, np.sum(s ** 2)
            for index, v in enumerate(s):          # :92-99
                s_new[index] = v
                p = p + v ** 2 / sums
                if p >= var:
                    break
        elif var >= 1.0:
            for index, v in enumerate(s):          # :107-112
                s_new[index] = v
                if index == var - 1:
                    break
        else:
            raise Exception("QLSSVC.var shoud be greater than 0")
        cond, normF = s_new[0] / s_new[index], s_new[0]
        s = s_new
    else:
        cond, normF = s[0] / s[-1], s[0]
    inv = np.array([1.0 / x if x > 0 else x for x in s])
    sol = (u * inv) @ vh @ np.append(0.0, np.asarray(y, dtype=np.float64))
    return dict(b=sol[0], alpha=sol[1:], cond=cond, normF=normF, singular_values_F_=s, gamma=g)


def fit_attributes(X, fit, penalty):
    """fit (_qSVM.py:159-171): alpha_F, Nu and, for the linear kernel, coef_."""
    X = np.asarray(X, dtype=np.float64)
    N = len(X)
    alpha_F = np.sqrt(N) + 1.0 / penalty + np.linalg.norm(X, ord='fro') ** 2
    Nu = fit["b"] ** 2 + np.sum(fit["alpha"] ** 2 * (X * X).sum(axis=1))
    return dict(alpha_F=alpha_F, Nu=Nu, coef_=fit["alpha"] @ X)


def get_h(Xtrain, fit, Xtest, kind='linear', coef0=0.0, degree=3):
    """get_h without noise (_qSVM.py:263-267): alpha . K(X_train, x) + b for every test row (gamma from X_train)."""
    return fit["alpha"] @ kernel(Xtrain, Xtest, kind, fit["gamma"], coef0, degree) + fit["b"]


def get_betas(Xtrain, Nu, Xtest):
    """get_betas (_qSVM.py:278-282)."""
    Xtest = np.asarray(Xtest, dtype=np.float64)
    return np.sqrt((len(Xtrain) * (Xtest * Xtest).sum(axis=1) + 1.0) * Nu)


def get_P(h, betas):
    """get_P without noise (_qSVM.py:284-289)."""
    return 0.5 * (1.0 - h / betas)


def relative_error_routine(rng, relative_error, delta, Xmax, Xreal):
    """relative_error_routine (_qSVM.py:245-261); rng is a numpy Generator."""
    r, Xr, Xhat, delta_r, eps = 0.0, Xmax, 0.0, 0.0, 0.0
    while Xr > Xhat:
        r += 1.0
        Xr = Xmax / 2 ** r
        eps = relative_error * Xr / 2
        delta_r = 6 * delta / (np.pi ** 2 * r ** 2)
        Xhat = float(Xreal + truncnorm.rvs(-eps, eps, random_state=rng))
    return Xhat, delta_r, eps


def predict(rng, P, betas, h, error_type='absolute', absolute_error=0.01, relative_error=0.5):
    """predict (_qSVM.py:199-214): P + truncnorm(+-err / (2 beta)), then the sign rule P <= 0.5 -> +1."""
    P = np.array(P, dtype=np.float64)
    for i in range(len(P)):
        if error_type == 'absolute':
            e = absolute_error / (2 * betas[i])
        else:
            e = relative_error_routine(rng, relative_error, 0.1, betas[i], abs(h[i]))[2] / (2 * betas[i])
        P[i] = P[i] + truncnorm.rvs(-e, e, random_state=rng)
    return np.where(P <= 0.5, 1.0, -1.0), P
