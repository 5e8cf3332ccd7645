"""cProfile of qPCA.fit(estimate_all=True) + quantum transform at BASELINE config 2 (70 000 x 784, 50 components)."""
import cProfile, os, pstats, sys, time, warnings
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn.decomposition import qPCA
from sqlearn.QuantumUtility import Utility as U
rs = np.random.RandomState(0)
X = rs.randn(70000, 784) + rs.randn(10, 784)[rs.randint(0, 10, 70000)] * 3
warnings.simplefilter("ignore")
U.seed(1)
def run():
    p = qPCA(n_components=50, svd_solver='full')
    s50 = None
    p.fit(X)
    s50 = p.singular_values_[49]
    t = time.perf_counter()
    p = qPCA(n_components=50, svd_solver='full').fit(X, estimate_all=True, eps=0.05, delta=0.1, theta_major=0.5 * s50,
                                                    true_tomography=True, norm='L2')
    torch.cuda.synchronize(); t1 = time.perf_counter()
    Y = p.transform(X, classic_transform=False, epsilon_delta=0.1, quantum_representation=True, norm='L2') if False else None
    return t1 - t
run()
pr = cProfile.Profile(); pr.enable(); dt = run(); pr.disable()
print("fit(estimate_all) %.1f ms" % (dt * 1e3))
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
