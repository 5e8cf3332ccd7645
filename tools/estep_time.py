"""delta-window E-step time for a given number of centroids (which tile shape runs depends on k)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200")); sys.path.insert(0, ROOT)
import bench
from sqlearn.cluster._engine import QMeansEngine
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4000000
X = bench.make_shard(rows, 0, 1, torch.device("cuda", 0))
X -= X.mean(dim=0)
for k in [int(v) for v in sys.argv[2:]] or [128]:
    eng = QMeansEngine(X, k)
    C = X[:k].contiguous()
    eng.assign(C, 1, 0.5, 1, 0); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(5): eng.assign(C, 1, 0.5, 2 + r, 0)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("k=%d: %.3f ms per %d rows, %.1f TFLOP/s" % (k, ms, rows, 2.0 * rows * k * X.shape[1] / ms / 1e9))
