"""q-means (delta-means) clustering on a B200.

Host-side mirror of the quantum part of the reference's
``sklearn/cluster/_dmeans.py``: ``qMeans_`` with its constructor arguments,
``fit / predict / transform / fit_predict / fit_transform / score`` and the
helpers ``_kmeans_single_lloyd``, ``_labels_inertia``, ``labels_estimation``,
``_centers_update``, ``select_labels``, ``_tolerance`` keep their names and
meaning.  The arithmetic runs in CUDA kernels (see ``_engine.py`` and
``include/sqb200.h``); arrays stay on the GPU between the E-step, the M-step
and the centroid tomography.

Deliberate differences from the reference, all listed in DESIGN.md:
  * helpers take the device engine instead of host arrays, and return the
    per-row minimum distance instead of the n x k distance matrix (which is
    never materialised);
  * the unconditional ``print(center_shift_total, tol)`` (_dmeans.py:652) is
    dropped;
  * an empty cluster keeps its previous centre and warns (the reference
    indexes an uninitialised array, _dmeans.py:816-823);
  * ``predict`` / ``score`` implement their evident intent (the reference
    calls ``_labels_inertia`` with the wrong arity, _dmeans.py:1387, :1401);
  * ``sample_weight`` is accepted and ignored, as in the reference (:812).
"""
import numbers
import random
import warnings

import numpy as np
import torch

from .. import _device as dv
from ..QuantumUtility import Utility as qu
from ._engine import MODE_EXACT, MODE_IPE, MODE_WINDOW, QMeansEngine

# the reference seeds Python's global generator when the module is imported (_dmeans.py:22)
random.seed(a=1234)


class ConvergenceWarning(UserWarning):
    """Same role as sklearn.exceptions.ConvergenceWarning."""


def check_random_state(seed):
    """np.random.RandomState from None / int / instance (sklearn.utils.check_random_state)."""
    if seed is None or seed is np.random:
        return np.random.mtrand._rand
    if isinstance(seed, numbers.Integral):
        return np.random.RandomState(seed)
    if isinstance(seed, np.random.RandomState):
        return seed
    raise ValueError('%r cannot be used to seed a numpy.random.RandomState instance' % seed)


def select_labels(labels):
    """Uniform choice among candidate centroid indices (reference: _dmeans.py:2252-2257).
    On the device the same choice is made inside K2; this host version serves callers
    that hold a candidate list."""
    if len(labels) == 1:
        return labels[0]
    return random.choice(labels)


def _tolerance(engine, tol):
    """tol scaled by the mean per-feature variance (reference: _dmeans.py:253-261).
    Uses the centred shard on the device: variance is shift invariant."""
    if tol == 0:
        return 0
    # sum of squares per column through the mu-norm column pass (exponent 2)
    X = engine.X
    _, d = X.shape
    from .. import _cabi
    rowmax, colsum, fr = dv.empty((1,)), dv.empty((1, d)), dv.empty((1,))
    nb = _cabi.lib.sq_mu_norms_workspace_bytes(engine.n, d, 1)
    ws = dv.workspace(nb)
    e2 = np.asarray([2.0])
    _cabi.check(_cabi.lib.sq_mu_norms(dv.ptr(X), engine.n, d, e2.ctypes.data, e2.ctypes.data, 1, dv.ptr(rowmax),
                                      dv.ptr(colsum), dv.ptr(fr), dv.ptr(ws), nb, dv.stream_ptr()), "sq_mu_norms")
    csum = engine.col_sums()
    cnt = torch.tensor([float(engine.n)], dtype=torch.float64, device=X.device)
    dv.all_reduce_sum(colsum)
    dv.all_reduce_sum(csum)
    dv.all_reduce_sum(cnt)
    n_glob = float(cnt.item())
    var = colsum[0].cpu().numpy() / n_glob - (csum.cpu().numpy() / n_glob) ** 2
    return float(np.mean(var) * tol)


def _kmeans_plusplus(X, n_clusters, x_squared_norms, random_state, n_local_trials=None):
    """k-means++ seeding on the host (reference: _dmeans.py:153-247, stock
    scikit-learn): first centre uniform, then each next centre is the best of
    2+log(k) candidates drawn with probability proportional to the squared
    distance to the closest centre chosen so far.  Consumes ``random_state``
    in the same order as the reference (randint, then random_sample per step)."""
    n_samples, n_features = X.shape
    centers = np.empty((n_clusters, n_features), dtype=X.dtype)
    if n_local_trials is None:
        n_local_trials = 2 + int(np.log(n_clusters))
    center_id = random_state.randint(n_samples)
    indices = np.full(n_clusters, -1, dtype=int)
    centers[0] = X[center_id]
    indices[0] = center_id

    def sqdist(rows):
        d2 = x_squared_norms[rows][:, None] - 2.0 * (X[rows] @ X.T) + x_squared_norms[None, :]
        np.maximum(d2, 0, out=d2)
        return d2

    closest = sqdist(np.asarray([center_id]))[0]
    current_pot = closest.sum()
    for c in range(1, n_clusters):
        rand_vals = random_state.random_sample(n_local_trials) * current_pot
        cand = np.searchsorted(np.cumsum(closest, dtype=np.float64), rand_vals)
        np.clip(cand, None, closest.size - 1, out=cand)
        d2 = sqdist(cand)
        np.minimum(closest, d2, out=d2)
        pots = d2.sum(axis=1)
        best = np.argmin(pots)
        current_pot = pots[best]
        closest = d2[best]
        centers[c] = X[cand[best]]
        indices[c] = cand[best]
    return centers, indices


def _kmeans_plusplus_device(engine, n_clusters, random_state, n_local_trials=None):
    """k-means++ seeding with X on the device (same algorithm and same consumption of
    ``random_state`` as :func:`_kmeans_plusplus` / the reference's _dmeans.py:153-247).  Per step one
    pass over X on the tensor cores gives the dot products with all candidates (``sq_project``) and one
    elementwise kernel the new closest distances (``sq_kpp_update``); the prefix sum, the binary search and
    the reductions over n run as device tensor ops.  Returns device centres."""
    from .. import _cabi
    X, n, d = engine.X, engine.n, engine.d
    if n_local_trials is None:
        n_local_trials = 2 + int(np.log(n_clusters))
    T = n_local_trials
    if T > 16:
        raise ValueError("n_local_trials=%d exceeds the 16 candidates one pass handles" % T)
    dev = X.device
    centers = torch.empty((n_clusters, d), dtype=torch.float64, device=dev)
    dots = torch.empty((n, T), dtype=torch.float64, device=dev)
    d2buf = torch.empty((T, n), dtype=torch.float64, device=dev)
    zeros = torch.zeros((T,), dtype=torch.float64, device=dev)

    def sqdist_min(ids, closest):
        """min(closest, squared distance to X[ids[t]]) for every candidate t: the dot products on the tensor
        cores (one pass over X), then one elementwise kernel"""
        t = len(ids)
        idx = torch.as_tensor(np.asarray(ids, dtype=np.int64), device=dev)
        cand = X[idx].contiguous()
        cn = engine.xnorm[idx].contiguous()
        _cabi.check(_cabi.lib.sq_project(dv.ptr(X), n, d, dv.ptr(cand), t, dv.ptr(zeros), dv.ptr(dots),
                                         dv.stream_ptr()), "sq_project")
        _cabi.check(_cabi.lib.sq_kpp_update(dv.ptr(dots), dv.ptr(engine.xnorm), dv.ptr(cn), dv.ptr(closest), n, t,
                                            dv.ptr(d2buf), dv.stream_ptr()), "sq_kpp_update")
        return d2buf[:t]

    center_id = int(random_state.randint(n))
    centers[0] = X[center_id]
    closest = sqdist_min([center_id], None)[0].clone()
    current_pot = float(closest.sum().item())
    for c in range(1, n_clusters):
        rand_vals = random_state.random_sample(T) * current_pot
        cum = torch.cumsum(closest, dim=0)
        cand = torch.searchsorted(cum, torch.as_tensor(rand_vals, device=dev))
        cand = torch.clamp(cand, max=n - 1).cpu().numpy()
        d2 = sqdist_min(cand, closest)
        pots = d2.sum(dim=1)
        best = int(torch.argmin(pots).item())
        current_pot = float(pots[best].item())
        closest = d2[best].clone()
        centers[c] = X[int(cand[best])]
    return centers


# kernel launches of the centroid tomography per Lloyd iteration (attempt-lane search + replay in one kernel)
TOMO_LAUNCHES_PER_ITER = 1


def _mode_of(delta, true_distance_estimate):
    if not delta or delta <= 0:
        return MODE_EXACT
    return MODE_IPE if true_distance_estimate else MODE_WINDOW


def labels_estimation(engine, centers, delta, pool, true_distance_estimate):
    """E-step on the device (reference: _dmeans.py:732-777).

    Returns (labels int32 device tensor, per-row min distance device tensor,
    inertia device scalar).  ``pool`` is accepted and unused: the IPE draws of
    every (sample, centroid) pair run inside one kernel."""
    mode = _mode_of(delta, true_distance_estimate)
    row0, total = engine.row0, engine.n_total
    seed, off = qu._take_streams(total)
    engine.assign(centers, mode, delta if delta else 0.0, seed, off + row0)
    return engine.labels, engine.mindist, engine.inertia


def _labels_inertia(engine, centers, distances=None, delta=0, pool=None, true_distance_estimate=True,
                    n_threads=None):
    """E step of the q-means EM algorithm (reference: _dmeans.py:674-724)."""
    return labels_estimation(engine, centers, delta, pool, true_distance_estimate)


def _centers_update(engine, labels, error, intermediate_error, true_tomography, stop_when_reached_accuracy,
                    centers_old=None, centers_out=None):
    """M step (reference: _dmeans.py:780-830): per-cluster means, then optional
    tomography of the centres with error ``error`` (= delta/2)."""
    engine.accumulate()
    engine.all_reduce()
    if centers_out is None:
        centers_out = dv.empty((engine.k, engine.d))
    engine.finalize(centers_old, centers_out)
    if intermediate_error:
        centers_out = _centroid_tomography(engine, centers_out, error, true_tomography, stop_when_reached_accuracy)
    return centers_out


def _centroid_tomography(engine, centers, error, true_tomography, stop_when_reached_accuracy):
    """tomography(centers, delta/2) (reference: _dmeans.py:825-828).  With several
    ranks each one estimates a contiguous block of centroids and the blocks are
    all-gathered; the Philox streams are keyed by centroid id, so the result
    is the same as on one GPU."""
    k, world, rank = engine.k, engine.world, engine.rank
    if world == 1 or not true_tomography or k % world != 0:
        # every rank draws from the same (synchronised) stream range, so all get the same centres
        return qu.tomography(centers, error, true_tomography=true_tomography,
                             stop_when_reached_accuracy=stop_when_reached_accuracy)
    import torch.distributed as dist
    per = k // world
    seed, off = qu._take_streams(k)
    mine = centers[rank * per:(rank + 1) * per].contiguous()
    est = qu._tomography_device(mine, error, None, 'L2', stop_when_reached_accuracy, True, 0,
                                vec_offset=off + rank * per, seed_=seed)[0]
    out = torch.empty_like(centers)
    dist.all_gather_into_tensor(out, est)
    return out


class LloydState:
    """The centres of a run on the device: ``centers`` is the current [k, d] tensor (a new object after every
    iteration - nothing is copied), ``spare`` a buffer of the same shape for the means of the next M-step."""

    def __init__(self, centers_init):
        self.centers = centers_init.clone()
        self.spare = torch.empty_like(self.centers)


def _lloyd_iteration(engine, state, host, delta, intermediate_err, true_tomography, stop_when_reached_accuracy, pool,
                     true_distance_estimate):
    """One pass of the loop body of _kmeans_single_lloyd (reference: _dmeans.py:619-651): E-step, M-step
    (+ centroid tomography), centre shift; ``state.centers`` becomes the new centres and (inertia, shift) come
    back through the pinned 2-double buffer ``host``."""
    old = state.centers
    labels, _, _ = _labels_inertia(engine, old, delta=delta, pool=pool, true_distance_estimate=true_distance_estimate)
    new = _centers_update(engine, labels, delta / 2, intermediate_error=intermediate_err,
                          true_tomography=true_tomography, stop_when_reached_accuracy=stop_when_reached_accuracy,
                          centers_old=old, centers_out=state.spare)
    shift_dev = engine.shift(old, new)
    if new is state.spare:                       # no tomography: the means are the centres; the old ones become spare
        state.spare = old
    state.centers = new
    host[0:1].copy_(engine.inertia, non_blocking=True)   # already summed over ranks by the M-step
    host[1:2].copy_(shift_dev, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return float(host[0]), float(host[1])


def _kmeans_single_lloyd(engine, sample_weight, centers_init, intermediate_err, true_tomography,
                         stop_when_reached_accuracy, pool, true_distance_estimate, max_iter=300, verbose=False,
                         tol=1e-4, delta=None):
    """One run of q-means Lloyd iterations (reference: _dmeans.py:534-671).

    ``centers_init`` is a device tensor [k, d].  Returns (labels, inertia,
    centers, n_iter) with labels / centers still on the device.  Per iteration
    the host reads back two doubles (inertia, centre shift) to take the same
    decisions as the reference: best-by-inertia tracking (:646-649) and the
    tolerance break (:651-658)."""
    if delta == None:  # noqa: E711
        delta = 0
    k, d = engine.k, engine.d
    state = LloydState(centers_init)
    best_labels = torch.empty_like(engine.labels)
    best_centers = torch.empty_like(state.centers)
    best_inertia = None
    host = torch.empty((2,), dtype=torch.float64).pin_memory()
    stats = {"fixed_rows": 0, "empty_clusters": 0}
    center_shift_total = 0.0
    i = 0
    for i in range(max_iter):
        labels = engine.labels
        inertia, center_shift_total = _lloyd_iteration(
            engine, state, host, delta, intermediate_err, true_tomography, stop_when_reached_accuracy, pool,
            true_distance_estimate)
        if verbose:
            print("Iteration %2d, inertia %.3f" % (i, inertia))
        if best_inertia is None or inertia < best_inertia:
            best_labels.copy_(labels)
            best_centers.copy_(state.centers)
            best_inertia = inertia
        if center_shift_total <= tol:
            if verbose:
                print("Converged at iteration %d: center shift %e within tolerance %e"
                      % (i, center_shift_total, tol))
            break
    if center_shift_total > 0:
        # rerun the E-step so that the labels match the returned centres (:660-666)
        labels, _, inertia_dev = _labels_inertia(engine, best_centers, delta=delta, pool=pool,
                                                 true_distance_estimate=true_distance_estimate)
        red = inertia_dev.clone()
        dv.all_reduce_sum(red)
        best_inertia = float(red.item())
        best_labels.copy_(labels)
    if verbose and i + 1 == max_iter:
        print("Reached {} iterations of q-means".format(i + 1))
    empty = int((engine.counts == 0).sum().item())
    if empty:
        warnings.warn("%d cluster(s) received no sample; their previous centre was kept" % empty,
                      ConvergenceWarning)
    return best_labels, best_inertia, best_centers, i + 1


from .._base import BaseEstimator, check_is_fitted  # noqa: E402
from .._spectral import refined_eigh  # noqa: E402


def _validate_X(X, copy=True):
    """C-ordered float64 2-D array (reference: _validate_data, _dmeans.py:1237-1240;
    float32 input is accepted and promoted: all kernels compute in fp64)."""
    if hasattr(X, "toarray") and not isinstance(X, np.ndarray):
        raise TypeError("sparse input has no quantum path (the reference falls back to classical k-means, "
                        "_dmeans.py:631-635); densify it first")
    X = np.array(X, dtype=np.float64, order='C', copy=copy) if copy else np.ascontiguousarray(X, dtype=np.float64)
    if X.ndim != 2:
        raise ValueError("Expected 2D array, got array with shape %r" % (X.shape,))
    return X


class qMeans_(BaseEstimator):
    """q-means clustering (reference: sklearn/cluster/_dmeans.py:833-1469).

    Same constructor as the reference (:1016-1021).  ``delta`` is the error of
    the distance estimates and (halved) of the centroid tomography;
    ``true_distance_estimate`` selects inner-product estimation (True) or the
    delta-window rule (False); ``intermediate_error`` adds tomography noise to
    the centres after every M-step.  Only ``algorithm='full'`` reaches the
    quantum path, exactly as in the reference (:1274-1278).

    Multi-GPU: under ``torch.distributed`` every rank passes its own rows;
    centres, inertia and ``n_iter_`` are global, ``labels_`` are this rank's.
    """

    def __init__(self, n_clusters=8, *, init='k-means++', n_init=10, max_iter=300, tol=1e-4,
                 precompute_distances='deprecated', verbose=0, random_state=None, copy_x=True,
                 n_jobs='deprecated', algorithm='auto', delta=None, intermediate_error=False,
                 true_tomography=True, stop_when_reached_accuracy=True, multiprocess=False,
                 true_distance_estimate=True):
        self.n_clusters = n_clusters
        self.init = init
        self.max_iter = max_iter
        self.tol = tol
        self.precompute_distances = precompute_distances
        self.n_init = n_init
        self.verbose = verbose
        self.random_state = random_state
        self.copy_x = copy_x
        self.n_jobs = n_jobs
        self.algorithm = algorithm
        self.delta = delta
        self.intermediate_error = intermediate_error
        self.true_tomography = true_tomography
        self.stop_when_reached_accuracy = stop_when_reached_accuracy
        self.multiprocess = multiprocess
        self.true_distance_estimate = true_distance_estimate

    # ---- parameter checks (reference: _dmeans.py:1041-1102) -----------------
    def _check_params(self, X):
        if self.precompute_distances != 'deprecated':
            warnings.warn("'precompute_distances' was deprecated in version 0.23 and will be removed in 1.0 "
                          "(renaming of 0.25). It has no effect", FutureWarning)
        if self.n_jobs != 'deprecated':
            warnings.warn("'n_jobs' was deprecated in version 0.23 and will be removed in 1.0 "
                          "(renaming of 0.25).", FutureWarning)
        if self.n_init <= 0:
            raise ValueError(f"n_init should be > 0, got {self.n_init} instead.")
        self._n_init = self.n_init
        if self.max_iter <= 0:
            raise ValueError(f"max_iter should be > 0, got {self.max_iter} instead.")
        if X.shape[0] < self.n_clusters:
            raise ValueError(f"n_samples={X.shape[0]} should be >= n_clusters={self.n_clusters}.")
        if self.algorithm not in ("auto", "full", "elkan"):
            raise ValueError(f"Algorithm must be 'auto', 'full' or 'elkan', got {self.algorithm} instead.")
        self._algorithm = self.algorithm
        if self._algorithm == "auto":
            self._algorithm = "full" if self.n_clusters == 1 else "elkan"
        if self._algorithm == "elkan" and self.n_clusters == 1:
            warnings.warn("algorithm='elkan' doesn't make sense for a single cluster. Using 'full' instead.",
                          RuntimeWarning)
            self._algorithm = "full"
        if not (hasattr(self.init, '__array__') or callable(self.init)
                or (isinstance(self.init, str) and self.init in ["k-means++", "random"])):
            raise ValueError(f"init should be either 'k-means++', 'random', a ndarray or a callable, "
                             f"got '{self.init}' instead.")
        if hasattr(self.init, '__array__') and self._n_init != 1:
            warnings.warn(f"Explicit initial center position passed: performing only one init in "
                          f"{self.__class__.__name__} instead of n_init={self._n_init}.", RuntimeWarning,
                          stacklevel=2)
            self._n_init = 1

    def _validate_center_shape(self, X, centers):
        if centers.shape[0] != self.n_clusters:
            raise ValueError(f"The shape of the initial centers {centers.shape} does not match the number of "
                             f"clusters {self.n_clusters}.")
        if centers.shape[1] != X.shape[1]:
            raise ValueError(f"The shape of the initial centers {centers.shape} does not match the number of "
                             f"features of the data {X.shape[1]}.")

    def _init_centroids(self, X, x_squared_norms, init, random_state, init_size=None):
        """Initial centres on the host (reference: _dmeans.py:1153-1209)."""
        n_samples = X.shape[0]
        n_clusters = self.n_clusters
        if init_size is not None and init_size < n_samples:
            idx = random_state.randint(0, n_samples, init_size)
            X, x_squared_norms, n_samples = X[idx], x_squared_norms[idx], len(idx)
        if isinstance(init, str) and init == 'k-means++':
            centers, _ = _kmeans_plusplus(X, n_clusters, random_state=random_state,
                                          x_squared_norms=x_squared_norms)
        elif isinstance(init, str) and init == 'random':
            seeds = random_state.permutation(n_samples)[:n_clusters]
            centers = X[seeds]
        elif hasattr(init, '__array__'):
            centers = init
        elif callable(init):
            centers = np.ascontiguousarray(init(X, n_clusters, random_state=random_state), dtype=X.dtype)
            self._validate_center_shape(X, centers)
        return np.ascontiguousarray(centers, dtype=np.float64)

    # ---- fit (reference: _dmeans.py:1211-1325) --------------------------------
    # best_mu(X, 0, 0.1): p = 0, 0.05, 0.1 (Utility.py:196-231), plus one row of exponents 2 for eta / Frobenius
    _MU_GRID = [i for i in np.arange(0, 0.1, 0.05)] + [0.1]

    def _preamble_chunk(self, engine, acc, r0, r1):
        """The part of the preamble (:1242-1245, :1262-1266) that is a sum or a maximum over rows, for rows
        [r0, r1) of the uncentred data: row norms, mu-norm partials, Gram partial, column sums."""
        from .. import _cabi
        d = engine.d
        n = r1 - r0
        Xc = engine.X[r0:r1]
        er = np.asarray([2 * p for p in self._MU_GRID] + [2.0])
        ec = np.asarray([2 * (1 - p) for p in self._MU_GRID] + [2.0])
        npk = len(er)
        _cabi.check(_cabi.lib.sq_row_sqnorms(dv.ptr(Xc), n, d, dv.ptr(engine.xnorm[r0:r1]), dv.stream_ptr()),
                    "sq_row_sqnorms")
        rowmax, colsum, fr = dv.empty((npk,)), dv.empty((npk, d)), dv.empty((1,))
        nb = _cabi.lib.sq_mu_norms_workspace_bytes(n, d, npk)
        ws = dv.workspace(nb)
        _cabi.check(_cabi.lib.sq_mu_norms(dv.ptr(Xc), n, d, er.ctypes.data, ec.ctypes.data, npk,
                                          dv.ptr(rowmax), dv.ptr(colsum), dv.ptr(fr), dv.ptr(ws), nb,
                                          dv.stream_ptr()), "sq_mu_norms")
        G = dv.empty((d, d))
        nb = _cabi.lib.sq_gram_workspace_bytes(n, d)
        ws = dv.workspace(nb)
        _cabi.check(_cabi.lib.sq_gram_centered(dv.ptr(Xc), n, d, None, dv.ptr(G), dv.ptr(ws), nb,
                                               dv.stream_ptr()), "sq_gram_centered")
        csum = dv.empty((d,))
        nb = _cabi.lib.sq_col_sums_workspace_bytes(n, d)
        ws = dv.workspace(nb)
        _cabi.check(_cabi.lib.sq_col_sums(dv.ptr(Xc), n, d, dv.ptr(csum), dv.ptr(ws), nb, dv.stream_ptr()),
                    "sq_col_sums")
        engine.launches += 9
        if not acc:
            acc.update(rowmax=rowmax, colsum=colsum, fr=fr, G=G, csum=csum)
        else:                                   # chunks are combined in row order: deterministic
            acc["rowmax"] = torch.maximum(acc["rowmax"], rowmax)
            acc["colsum"] += colsum
            acc["fr"] += fr
            acc["G"] += G
            acc["csum"] += csum

    def _preamble_finish(self, engine, acc):
        """eta, muA and condition_number of the uncentred data (:1242-1245).  The reference takes a full LAPACK
        SVD for min(S); here min(S) = sqrt(lambda_min(X^T X)) from the tensor-core Gram kernel and a d x d
        cuSOLVER eigensolve.  Returns the column sums of this rank's rows (for the mean, :1264)."""
        d = engine.d
        rowmax, colsum, fr, G = acc["rowmax"], acc["colsum"], acc["fr"], acc["G"]
        if engine.world > 1:
            import torch.distributed as dist
            dist.all_reduce(rowmax, op=dist.ReduceOp.MAX)
            dv.all_reduce_sum(colsum)
            dv.all_reduce_sum(fr)
            dv.all_reduce_sum(G)
        s1 = rowmax.cpu().numpy()
        s2 = colsum.max(dim=1).values.cpu().numpy()
        self.eta = float(s1[-1])                                        # max_i ||x_i||^2
        mu = list(np.sqrt(s1[:-1] * s2[:-1]))
        self.muA = min(min(mu), float(np.sqrt(fr.item())))
        # cuSOLVER, d x d; eigenvalues below 1e-6 of the largest are re-solved in their own subspace
        # (sqlearn/_spectral.py), so that min(S) is as accurate as the reference's LAPACK SVD
        lam, _ = refined_eigh(engine.X, G)
        r = min(engine.n_total, d)
        smin = float(torch.sqrt(torch.clamp(lam[d - r], min=0.0)).item())
        self.condition_number = 1 / smin if smin > 0 else float('inf')
        return acc["csum"]

    def fit(self, X, y=None, sample_weight=None):
        """Compute q-means clustering (reference: _dmeans.py:1211-1325)."""
        # copy_x protects the caller's array from the in-place centring (:1262-1269).  The device path never writes
        # the host array, so the private copy is taken only where the host rows are centred (host-side init below).
        X = _validate_X(X, copy=False)
        dv.require_cuda()
        rank, world = dv.dist_info()
        if world > 1:
            qu._sync_state()
        if X.shape[0] == 0:
            raise ValueError("Found array with 0 sample(s)")
        # The rows go up in chunks on a copy stream; everything the preamble needs from the uncentred data is a
        # sum or a maximum over rows, so it is computed chunk by chunk underneath the upload.
        state = {}

        def on_chunk(out, r0, r1):
            if "engine" not in state:
                state["engine"] = QMeansEngine(out, self.n_clusters, norms=False)
                state["acc"] = {}
            self._preamble_chunk(state["engine"], state["acc"], r0, r1)
        dv.stream_to_device(X, on_chunk)
        engine = state["engine"]
        self._engine_launches = 0
        # _assert_all_finite of the reference's check_array, evaluated on the device: ||X||_F^2 is finite
        if not bool(torch.isfinite(engine.xnorm.sum()).item()):
            raise ValueError("Input contains NaN, infinity or a value too large for dtype('float64').")
        csum = self._preamble_finish(engine, state["acc"])

        if self.delta == 0:
            warnings.warn("Attention! You are running classic version of kmeans!")
            if self.intermediate_error:
                raise ValueError("intermediate_error value cannot be True if delta is zero.")
        self._check_params(X)
        random_state = check_random_state(self.random_state)

        init = self.init
        if hasattr(init, '__array__'):
            init = np.array(init, dtype=np.float64, order='C', copy=True)
            self._validate_center_shape(X, init)

        # subtract the mean of X for more accurate distance computations (:1262-1269)
        dv.all_reduce_sum(csum)
        X_mean_dev = csum / float(engine.n_total)
        engine.subtract_rowvec(X_mean_dev)
        engine.refresh_norms()
        X_mean = X_mean_dev.cpu().numpy()
        if hasattr(init, '__array__'):
            init -= X_mean
        self._tol = _tolerance(engine, self.tol)

        if self._algorithm != "full":
            # the reference dispatches to _kmeans_single_elkan, which rejects the quantum keywords (:1294-1299)
            raise TypeError("_kmeans_single_elkan() got an unexpected keyword argument 'delta' "
                            "(use algorithm='full' for q-means, as in the reference)")

        # k-means++ seeds on the device; with several ranks every rank seeds from its own rows and rank 0's
        # centres are broadcast below (as with the host-side inits)
        device_init = isinstance(init, str) and init == 'k-means++'
        need_host = not hasattr(init, '__array__') and not device_init
        if need_host:
            if self.copy_x:
                X = X.copy()                 # a private copy when copy_x=True, as in the reference
            X -= X_mean
            x_sq_host = np.einsum('ij,ij->i', X, X)
        best_inertia = None
        for _ in range(self._n_init):
            if device_init:
                c_dev = _kmeans_plusplus_device(engine, self.n_clusters, random_state)
            else:
                centers_init = self._init_centroids(X, x_sq_host if need_host else None, init, random_state)
                c_dev = dv.to_device(centers_init)
            if world > 1:
                import torch.distributed as dist
                dist.broadcast(c_dev, src=0)
            if self.verbose:
                print("Initialization complete")
            labels, inertia, centers, n_iter_ = _kmeans_single_lloyd(
                engine, sample_weight, c_dev, max_iter=self.max_iter, verbose=self.verbose, tol=self._tol,
                delta=self.delta, intermediate_err=self.intermediate_error, true_tomography=self.true_tomography,
                stop_when_reached_accuracy=self.stop_when_reached_accuracy, pool=None,
                true_distance_estimate=self.true_distance_estimate)
            if best_inertia is None or inertia < best_inertia:
                best_labels = labels.clone()
                best_centers = centers.clone()
                best_inertia = inertia
                best_n_iter = n_iter_
        if need_host and not self.copy_x:
            X += X_mean
        best_centers = best_centers + X_mean_dev
        self._labels = best_labels.cpu().numpy()
        distinct = int((torch.bincount(best_labels.to(torch.int64), minlength=self.n_clusters) > 0).sum().item())
        if world == 1 and distinct < self.n_clusters:
            warnings.warn("Number of distinct clusters ({}) found smaller than n_clusters ({}). Possibly due to "
                          "duplicate points in X.".format(distinct, self.n_clusters), ConvergenceWarning,
                          stacklevel=2)
        self.cluster_centers_ = best_centers.cpu().numpy()
        self.inertia_ = best_inertia
        self.n_iter_ = best_n_iter
        self.n_features_in_ = X.shape[1]
        self._engine_launches = engine.launches
        return self

    @property
    def labels_(self):
        """Labels of the training rows as a list of int, as in the reference (:628, :1322)."""
        check_is_fitted(self, "_labels")
        return self._labels.tolist()

    def fit_predict(self, X, y=None, sample_weight=None):
        warnings.warn('Attention! This computational path is purely classic.')
        return self.fit(X, sample_weight=sample_weight).labels_

    def fit_transform(self, X, y=None, sample_weight=None):
        warnings.warn("Attention! This computational part is purely classical.")
        return self.fit(X, sample_weight=sample_weight)._transform(_validate_X(X))

    def transform(self, X):
        warnings.warn("Attention! This computational part is purely classical.")
        check_is_fitted(self, "cluster_centers_")
        return self._transform(self._check_test_data(X))

    def _check_test_data(self, X):
        X = _validate_X(X, copy=False)
        if X.shape[1] != self.n_features_in_:
            raise ValueError("X has %d features, but %s is expecting %d features as input."
                             % (X.shape[1], type(self).__name__, self.n_features_in_))
        return X

    def _transform(self, X):
        """Distances to the centres, n x k (reference: euclidean_distances, :1349-1351).
        Classical and off the hot path: a library call."""
        Xd = dv.stream_to_device(X)
        C = dv.to_device(self.cluster_centers_)
        return torch.cdist(Xd, C, compute_mode='donot_use_mm_for_euclid_dist').cpu().numpy()

    def _predict_engine(self, X, delta):
        from .. import _cabi
        X = self._check_test_data(X)
        if X.shape[0] == 0:
            raise ValueError("Found array with 0 sample(s)")
        state = {}

        def on_chunk(out, r0, r1):          # row norms of each chunk while the next one is on its way
            if "engine" not in state:
                state["engine"] = QMeansEngine(out, self.n_clusters, norms=False)
            e = state["engine"]
            _cabi.check(_cabi.lib.sq_row_sqnorms(dv.ptr(e.X[r0:r1]), r1 - r0, e.d, dv.ptr(e.xnorm[r0:r1]),
                                                 dv.stream_ptr()), "sq_row_sqnorms")
        dv.stream_to_device(X, on_chunk)
        engine = state["engine"]
        C = dv.to_device(self.cluster_centers_)
        labels, mind, inertia = _labels_inertia(engine, C, delta=delta, pool=None,
                                                true_distance_estimate=self.true_distance_estimate)
        return labels, inertia

    def predict(self, X, sample_weight=None, delta=None):
        """Closest (or delta-noisy closest) centre of every row of X, list of int
        (intent of the reference's predict, :1353-1391)."""
        check_is_fitted(self, "cluster_centers_")
        if delta == None:  # noqa: E711
            delta = 0
        labels, _ = self._predict_engine(X, delta)
        return labels.cpu().numpy().tolist()

    def score(self, X, y=None, sample_weight=None):
        """Opposite of the k-means objective on X (intent of :1393-1402)."""
        check_is_fitted(self, "cluster_centers_")
        _, inertia = self._predict_engine(X, 0)
        red = inertia.clone()
        dv.all_reduce_sum(red)
        return -float(red.item())

    def runtime_comparison(self, n_samples, n_features, saveas, well_clusterable=False, plot=False):
        """Classical vs quantum cost model (reference: _dmeans.py:1412-1469).  The MATLAB
        plot of the reference is left out (no MATLAB engine); the two cost grids are returned."""
        if plot:
            warnings.warn("plot=True needs the MATLAB engine the reference uses; ignored")
        n, m = np.meshgrid(np.linspace(0, n_samples, dtype=np.int64, num=100),
                           np.linspace(0, n_features, dtype=np.int64, num=100))
        c_runtime = n * m * self.n_clusters * self.n_init
        k, eta, kappa, mu, dl = self.n_clusters, self.eta, self.condition_number, self.muA, self.delta
        if not well_clusterable:
            q_runtime = (k * m * eta * kappa * (mu + k * eta / dl) / (dl ** 2)) + \
                        ((k ** 2) * (eta ** 1.5) * kappa * mu) / (dl ** 2)
        else:
            q_runtime = ((k ** 2) * m * (eta ** 2.5) / (dl ** 3)) + (k ** 2.5) * (eta ** 2) / (dl ** 3)
        return q_runtime, c_runtime
