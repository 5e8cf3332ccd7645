import sys
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/sq-learn_b200")
import numpy as np
from scipy import stats
import gpu_util as G
from oracle import quantum_utility as qu
rng = np.random.default_rng(3)
for d, N in [(784, 23991), (784, 40000), (784, 5000), (300, 10000), (200, 10000), (1000, 30000), (600, 600), (600, 3000)]:
    v = np.random.RandomState(3).uniform(-1, 1, d); v /= np.linalg.norm(v)
    B = 6000
    est, n_used, err = G.tomography(np.repeat(v[None], B, axis=0), 0.1, [float(N)], norm_inf=1, seed=5)
    o = np.array([qu.tomography_attempt(v, N, rng) for _ in range(B)])
    oe = np.abs(v - o).max(axis=1)
    L = 1 << int(np.ceil(np.log2(d)))
    mode = "direct" if N <= 64 * min(max(L, 32), 512) else "tree"
    big = np.argmax(np.abs(v))
    c_me = np.round(est[:, big] ** 2 * N); c_o = np.round(o[:, big] ** 2 * N)
    small = np.argsort(np.abs(v))[d // 4]
    print("d=%d N=%d L=%d %s | err mean gpu %.5f oracle %.5f KS p=%.3g | zeros/trial %.2f vs %.2f | sign err %.2f vs %.2f | stage-1 count of largest comp mean %.2f vs %.2f (exp %.2f) var %.1f vs %.1f | sum est^2 %.6f" % (
        d, N, L, mode, err.mean(), oe.mean(), stats.ks_2samp(err, oe).pvalue, (est == 0).sum(1).mean(), (o == 0).sum(1).mean(),
        (np.sign(est) != np.sign(v)).sum(1).mean(), (np.sign(o) != np.sign(v)).sum(1).mean(), c_me.mean(), c_o.mean(), N * v[big] ** 2, c_me.var(), c_o.var(),
        (est ** 2).sum(1).mean()), flush=True)
    # per-component mean of stage-1 frequency vs v^2
    f = (est ** 2).mean(axis=0)
    z = (f - v ** 2) / np.sqrt(v ** 2 * (1 - v ** 2) / N / B + 1e-30)
    print("    stage-1 frequency z-scores: max |z| %.2f at comp %d; mean z %.3f; comps with |z|>4: %d" % (np.abs(z).max(), np.abs(z).argmax(), z.mean(), (np.abs(z) > 4).sum()))
