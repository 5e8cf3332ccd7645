import sys, time, warnings
sys.path.insert(0, "sq-learn_b200"); sys.path.insert(0, ".")
import numpy as np, torch
import bench
from sqlearn.cluster import qMeans_
from sqlearn.cluster import _dmeans as D
from sqlearn.QuantumUtility import Utility as Q
dev = torch.device("cuda", 0)
ROWS = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
X = bench.make_shard(ROWS, 0, 1, dev)
Xh = torch.empty(X.shape, dtype=torch.float64, pin_memory=True); Xh.copy_(X)
init = X[:256].cpu().numpy(); del X
torch.cuda.synchronize()
orig = {}
def wrap(obj, name):
    f = getattr(obj, name)
    def g(*a, **k):
        torch.cuda.synchronize(); t = time.perf_counter(); r = f(*a, **k); torch.cuda.synchronize()
        print(f"  {name}: {(time.perf_counter()-t)*1e3:.1f} ms"); return r
    setattr(obj, name, g)
wrap(D.dv, "stream_to_device"); wrap(qMeans_, "_preamble_finish"); wrap(D, "_kmeans_single_lloyd"); wrap(D, "_tolerance"); wrap(D, "_validate_X")
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    for rep in range(2):
        t = time.perf_counter()
        est = qMeans_(n_clusters=256, init=init, n_init=1, max_iter=10, tol=0, delta=0.5, algorithm='full', intermediate_error=True, true_distance_estimate=False, copy_x=False).fit(Xh.numpy())
        torch.cuda.synchronize(); print("fit total", (time.perf_counter()-t)*1e3, "ms")
