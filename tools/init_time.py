"""time of the device k-means++ seeding at the headline shape"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200")); sys.path.insert(0, ROOT)
import bench
from sqlearn.cluster._engine import QMeansEngine
from sqlearn.cluster import _dmeans as D
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 256
X = bench.make_shard(rows, 0, 1, torch.device("cuda", 0))
X -= X.mean(dim=0)
eng = QMeansEngine(X, k)
for rep in range(2):
    torch.cuda.synchronize(); t = time.perf_counter()
    C = D._kmeans_plusplus_device(eng, k, np.random.RandomState(0))
    torch.cuda.synchronize(); print("k-means++ on %d x %d, k=%d: %.1f ms" % (rows, X.shape[1], k, (time.perf_counter() - t) * 1e3))
