"""NumPy restatement of the q-means path of sklearn/cluster/_dmeans.py (reference).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.
"""
import numpy as np

from . import quantum_utility as qu


def cdist_sq(X, C):
    """np.square(cdist(X, C, 'euclidean')) (_dmeans.py:736-737), restated with the same
    operation order as scipy's kernel: per pair, sum_t (x_t - c_t)^2 accumulated in index
    order, then sqrt, then square.  (Checked bit-for-bit against scipy in the tests.)"""
    X = np.asarray(X, dtype=np.float64)
    C = np.asarray(C, dtype=np.float64)
    s = np.zeros((X.shape[0], C.shape[0]))
    for t in range(X.shape[1]):
        df = X[:, t][:, None] - C[:, t][None, :]
        s = s + df * df
    r = np.sqrt(s)
    return r * r


def labels_estimation(X, centers, delta, true_distance_estimate, rng, scipy_cdist=False):
    """_dmeans.py:732-777.  Returns (labels, distances, inertia).  ``scipy_cdist`` calls scipy's compiled cdist as
    the reference does (:736) instead of the operation-order restatement above - same values (checked in the
    tests), an order of magnitude faster; bench.py times this one."""
    if scipy_cdist:
        from scipy.spatial.distance import cdist
        D = np.square(cdist(X, centers, "euclidean"))
    else:
        D = cdist_sq(X, centers)
    if delta > 0:
        if not true_distance_estimate:
            mins = D.min(axis=1)
            labels = np.array([rng.choice(np.where(row <= mn + delta)[0]) for row, mn in zip(D, mins)])
            return labels, D, float(np.sum(mins))
        xn = np.einsum("ij,ij->i", X, X)
        cn = np.einsum("ij,ij->i", centers, centers)
        D = np.array([[abs(xn[i]) + abs(cn[j]) - 2 * qu.ipe(X[i], centers[j], delta / 2, rng=rng)
                       for j in range(len(centers))] for i in range(len(X))])
    mins = D.min(axis=1)
    labels = np.array([rng.choice(np.where(row <= mn)[0]) for row, mn in zip(D, mins)])
    return labels, D, float(np.sum(mins))


def centers_update(X, labels, error, intermediate_error, true_tomography, stop_when_reached_accuracy, rng,
                   n_clusters=None, centers_old=None):
    """_dmeans.py:780-830.  Empty clusters keep ``centers_old`` (the reference leaves the row
    uninitialised); with every cluster populated this is exactly the reference."""
    labels = np.asarray(labels)
    k = n_clusters if n_clusters is not None else len(np.unique(labels))
    centers = np.zeros((k, X.shape[1])) if centers_old is None else np.array(centers_old, dtype=np.float64)
    for j in np.unique(labels):
        members = np.where(labels == j)[0]
        centers[int(j)] = np.true_divide(np.sum(X[members], axis=0), len(members))
    if intermediate_error:
        centers = qu.tomography(centers, error, true_tomography=true_tomography,
                                stop_when_reached_accuracy=stop_when_reached_accuracy, rng=rng)
    return centers


def kmeans_single_lloyd(X, centers_init, intermediate_err, true_tomography, stop_when_reached_accuracy,
                        true_distance_estimate, max_iter=300, tol=1e-4, delta=None, rng=None):
    """_dmeans.py:534-671.  Returns (labels, inertia, centers, n_iter)."""
    rng = rng or np.random.default_rng()
    delta = delta or 0
    centers = np.array(centers_init, dtype=np.float64)
    best = None
    shift = 0.0
    for i in range(max_iter):
        old = centers.copy()
        labels, _, inertia = labels_estimation(X, centers, delta, true_distance_estimate, rng)
        centers = centers_update(X, labels, delta / 2, intermediate_err, true_tomography,
                                 stop_when_reached_accuracy, rng, n_clusters=len(old), centers_old=old)
        if best is None or inertia < best[1]:
            best = (labels.copy(), inertia, centers.copy())
        shift = float(np.sum((old - centers) ** 2))
        if shift <= tol:
            break
    labels, inertia, centers = best
    if shift > 0:
        labels, _, inertia = labels_estimation(X, centers, delta, true_distance_estimate, rng)
    return labels, inertia, centers, i + 1


def tolerance(X, tol):
    """_dmeans.py:253-261."""
    return 0 if tol == 0 else float(np.mean(np.var(X, axis=0)) * tol)
