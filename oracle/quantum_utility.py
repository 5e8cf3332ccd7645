"""NumPy restatement of sklearn/QuantumUtility/Utility.py (reference).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.  Each function cites the
reference lines it follows.  Sampling uses a numpy Generator: the reference's
own streams are not seedable end to end (Utility.py:53 uses OS entropy), so
stochastic functions agree with it in distribution; the deterministic helpers
(N, schedules, Q, M, m, pmfs, bucketing, mu-norms) agree exactly and are
pinned by tests/golden/deterministic.json.
"""
import math

import numpy as np


# ---- tomography ------------------------------------------------------------
def tomography_N(d, delta, norm="L2"):
    """Utility.py:307-311."""
    if norm == "L2":
        return int((36 * d * np.log(d)) / (delta ** 2))
    return int((36 * np.log(d)) / (delta ** 2))


def check_measure(arr, faster_measure_increment=0):
    """Utility.py:414-422: strictly increasing schedule, minimum step 5 + increment."""
    arr = np.array(arr, dtype=np.int64)
    step = 5 + faster_measure_increment
    for i in range(len(arr) - 1):
        if arr[i + 1] <= arr[i]:
            # (an equal entry is first bumped by `step`, which already makes it larger)
            arr[i + 1] = arr[i] + step
    return arr


def schedule(N, faster_measure_increment=0):
    """Utility.py:317-319."""
    return check_measure(np.geomspace(1, N, num=100, dtype=np.int64), faster_measure_increment)


def prepare_state(V):
    """Utility.py:301-304 (normalise unless already within rtol 1e-2 of unit norm)."""
    V = np.asarray(V, dtype=np.float64)
    nrm = np.linalg.norm(V)
    if not np.isclose(nrm, 1, rtol=1e-2):
        V = V / nrm
    return V


def tomography_attempt(V, n, rng):
    """One attempt with n measurements per stage (Utility.py:366-396 / :323-352).

    Stage 1: counts ~ Multinomial(n, V_i^2/sum V^2) (QuantumState, :43-55; estimate_wald, :61-64),
    P = sqrt(counts/n).  Stage 2: 2d outcomes with amplitudes (V+P)/2, (V-P)/2, renormalised;
    only the '+' counts are read; sign rule m_i > 0.4 P_i^2 n."""
    d = len(V)
    p = V ** 2 / np.sum(V ** 2)
    counts = rng.multinomial(n, p)
    P = np.sqrt(counts / n)
    amp = 0.5 * np.concatenate([V + P, V - P])
    q = amp ** 2 / np.sum(amp ** 2)
    m = rng.multinomial(n, q)[:d]
    return np.where(m > 0.4 * P ** 2 * n, P, -P)


def real_tomography(V, N=None, delta=None, stop_when_reached_accuracy=True, norm="L2", incremental_measure=True,
                    faster_measure_increment=0, rng=None):
    """Utility.py:259-402.  Returns {n_measurements: estimate} for every attempt made."""
    rng = rng or np.random.default_rng()
    V = prepare_state(V)
    d = len(V)
    if N is None:
        N = tomography_N(d, delta, norm)
    res = {}
    if incremental_measure:
        for n in schedule(N, faster_measure_increment):
            est = tomography_attempt(V, int(n), rng)
            res[int(n)] = est
            if stop_when_reached_accuracy:
                err = np.linalg.norm(V - est, ord=2 if norm == "L2" else np.inf)
                if not err > delta:
                    break
    else:
        res[N] = tomography_attempt(V, int(N), rng)
    return res


def tomography_error(V, est, norm="L2"):
    return np.linalg.norm(prepare_state(V) - est, ord=2 if norm == "L2" else np.inf)


def truncnorm_halfwidth(noise, length):
    """make_gaussian_est / introduce_error_array: +-noise/sqrt(len) (Utility.py:97, :73)."""
    return noise / np.sqrt(length)


def make_gaussian_est(vec, noise, rng=None):
    """Utility.py:88-104: vec + TruncNormal(0,1; +-noise/sqrt(len))."""
    from scipy.stats import truncnorm
    a = truncnorm_halfwidth(noise, len(vec))
    return np.asarray(vec) + truncnorm.rvs(-a, a, size=len(vec), random_state=rng)


def tomography(A, noise, true_tomography=True, stop_when_reached_accuracy=True, N=None, norm="L2",
               incremental_measure=True, faster_measure_increment=0, rng=None):
    """Utility.py:107-180."""
    assert noise >= 0
    if noise == 0:
        return A
    A = np.asarray(A, dtype=np.float64)
    if not true_tomography:
        return make_gaussian_est(A.reshape(-1), noise, rng).reshape(A.shape)
    rows = A if A.ndim == 2 else A[None]
    out = np.array([list(real_tomography(r, N=N, delta=noise, stop_when_reached_accuracy=stop_when_reached_accuracy,
                                         norm=norm, incremental_measure=incremental_measure,
                                         faster_measure_increment=faster_measure_increment, rng=rng).values())[-1]
                    for r in rows])
    return out if A.ndim == 2 else out[0]


# ---- amplitude / phase estimation -----------------------------------------------
def median_Q(gamma):
    """Utility.py:564-568."""
    Q = np.ceil(np.log(1 / gamma) / (2 * (8 / np.pi ** 2 - 0.5) ** 2))
    return int(Q + 1 if Q % 2 == 0 else Q)


def ae_M(epsilon):
    """Utility.py:484."""
    return math.ceil((np.pi / (2 * epsilon)) * (1 + np.sqrt(1 + 4 * epsilon)))


def pe_m(epsilon, gamma):
    """Utility.py:635."""
    return int(np.ceil(np.log2(1 / epsilon)) + np.ceil(np.log2(2 + 1 / (2 * gamma))))


def ae_pmf(a, M):
    """Normalised outcome law of amplitude estimation (Utility.py:494-508): weight of grid
    point j is |sin(M D pi)/(M sin(D pi))|^2 with D the circular distance between j/M and
    theta_a/pi (amplitude_est_dist, :435-439); weight 1 at distance 0."""
    w = math.asin(np.sqrt(a)) / np.pi
    j = np.arange(M)
    diff = w - j / M
    D = np.minimum(np.abs(diff - np.ceil(diff)), np.abs(diff - np.floor(diff)))
    with np.errstate(divide="ignore", invalid="ignore"):
        p = np.abs(np.sin(M * D * np.pi) / (M * np.sin(D * np.pi))) ** 2
    p[D == 0] = 1.0
    return p / p.sum()


def pe_pmf(omega, m):
    """Normalised outcome law of phase estimation (Utility.py:637-653)."""
    M = 2 ** m
    k = np.arange(M)
    with np.errstate(divide="ignore", invalid="ignore"):
        p = np.abs(np.sin((M * omega - k) * np.pi) / (M * np.sin((omega - k / M) * np.pi))) ** 2
    p[~np.isfinite(p)] = 1.0
    return p / p.sum()


def amplitude_estimation(a, epsilon=0.01, gamma=None, M=None, rng=None):
    """Utility.py:442-531 (a_tilde only)."""
    rng = rng or np.random.default_rng()
    if M is None:
        M = ae_M(epsilon)
    p = ae_pmf(a, M)
    Q = median_Q(gamma) if gamma else 1
    j = rng.choice(M, size=Q, p=p)
    return float(np.median(np.sin(np.pi * j / M) ** 2))


def phase_estimation(omega, m=None, epsilon=None, gamma=0.1, rng=None):
    """Utility.py:591-694 (omega_tilde only)."""
    rng = rng or np.random.default_rng()
    if epsilon is not None:
        m = pe_m(epsilon, gamma)
    M = 2 ** m
    if omega == 1 or np.isclose(omega, 1):
        return (M - 1) / M
    return rng.choice(M, p=pe_pmf(omega, m)) / M


def consistent_bucket(pe_estimate, epsilon, gamma, n=None, shift=None):
    """Deterministic tail of consistent_phase_estimation (Utility.py:773-792):
    the grid, the bisect and the cell midpoint, given the phase-estimation outcome."""
    from bisect import bisect
    if n is None:
        n = pe_m(epsilon, gamma)
    C = gamma / n
    delta_prime = (epsilon * C) / 2
    if shift is None:
        shift = int(np.floor(2 / C) / 2) + 1
    grid = np.arange(-1 - shift * delta_prime, 1 + epsilon - shift * delta_prime, epsilon)
    grid = np.append(grid, 1 + epsilon - shift * delta_prime)
    i = bisect(grid, pe_estimate)
    est = np.mean((grid[i - 1], grid[i]))
    return 0 if est < 0 else float(est), delta_prime


def ipe(x, y, epsilon, gamma=0.1, rng=None):
    """Utility.py:697-737."""
    S = np.linalg.norm(x) ** 2 + np.linalg.norm(y) ** 2
    inner = np.inner(x, y)
    a = (S - 2 * inner) / (2 * S)
    eps_a = epsilon * max(1, abs(inner)) / S
    if math.isclose(a, 0.0, abs_tol=1e-15):
        a = 0
    return S * (1 - 2 * amplitude_estimation(a, epsilon=eps_a, gamma=gamma, rng=rng)) / 2


# ---- mu norms ---------------------------------------------------------------------
def mu_p(p, A):
    """Utility.py:196-212: sqrt(max row sum |a|^(2p) * max column sum |a|^(2(1-p))),
    an exponent of 0 meaning count_nonzero."""
    def s(e, M):
        if e == 0:
            return np.max(np.count_nonzero(M, axis=1))
        return np.max(np.sum(np.power(np.abs(M), e), axis=1))
    return np.sqrt(s(2 * p, A) * s(2 * (1 - p), A.T))


def best_mu(A, start=0.0, end=1.0, step=0.05):
    """Utility.py:215-231."""
    grid = [i for i in np.arange(start, end, step)] + [end]
    vals = [mu_p(p, A) for p in grid]
    v = min(vals)
    p = grid[vals.index(v)]
    fro = np.linalg.norm(A)
    return (f"p={p}", v) if v <= fro else ("Frobenius", fro)
