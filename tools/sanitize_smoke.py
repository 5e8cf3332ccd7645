"""Small invocation of every kernel, meant to run under compute-sanitizer:

  compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
"""
import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "sq-learn_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import gpu_util as G  # noqa: E402
from sqlearn import QuantumUtility as Q  # noqa: E402
from sqlearn.cluster import qMeans_  # noqa: E402
from sqlearn.decomposition import qPCA  # noqa: E402

rs = np.random.RandomState(0)
for n, d, k in [(300, 16, 5), (257, 33, 40), (200, 8, 130), (150, 12, 300), (500, 128, 256)]:
    X, C = rs.randn(n, d), rs.randn(k, d)
    for mode, delta in [(0, 0.0), (1, 3.0), (2, 0.5)]:
        if mode == 2 and n * k > 30000:
            continue
        G.assign(X, C, mode=mode, delta=delta)
    G.accumulate(X, rs.randint(0, k, n).astype(np.int32), k)
v = rs.uniform(-1, 1, (5, 37))
G.tomography(v, 0.3, [50.0, 400.0, 100000.0], stop=0)
G.tomography(v, 0.3, [100000.0], norm_inf=1)
G.binomial(np.full(64, 1000.0), np.full(64, 0.3), np.full(64, 0.7))
G.fejer(np.full(64, 35.0), np.full(64, 12.3))
G.fejer(np.full(64, 7.0), np.full(64, 2.3), brute=True)
Q.seed(1)
Q.tomography(rs.uniform(-1, 1, (2, 5000)), 0.4)
Q.make_gaussian_est(np.zeros(100), 0.5)
Q.best_mu(rs.randn(100, 20))
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    X = rs.randn(400, 24) + 2
    qMeans_(n_clusters=4, algorithm='full', delta=0.5, n_init=1, max_iter=2, intermediate_error=True,
            true_distance_estimate=False, random_state=0).fit(X)
    m = qPCA(n_components=4, svd_solver='full').fit(X, estimate_all=True, eps=0.5, delta=0.3, theta_major=1.0)
    m.transform(X[:20], classic_transform=False, epsilon_delta=0.3, quantum_representation=True,
                norm='est_representation')
print("sanitize_smoke done")
