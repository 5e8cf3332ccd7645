"""IPE E-step time with centroids that are plain means (not unit vectors): the nearest pairs sit at grid
position x ~ 1, where the window wraps around 0 and the general (unsorted) branch of ipe.cuh runs."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200")); sys.path.insert(0, ROOT)
import bench
from sqlearn.cluster._engine import QMeansEngine
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
dev = torch.device("cuda", 0)
X = bench.make_shard(rows, 0, 1, dev)
X -= X.mean(dim=0)
eng = QMeansEngine(X, 256)
eng.assign(X[:256].contiguous(), 0, 0.0, 1, 0)
torch.cuda.synchronize()
cent = torch.stack([X[eng.labels == j].mean(dim=0) if bool((eng.labels == j).any()) else X[j] for j in range(256)])
for name, C in (("plain means", cent), ("unit vectors", cent / cent.norm(dim=1, keepdim=True))):
    for mode in (1, 2):
        eng.assign(C, mode, 0.5, 1, 0); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(3): eng.assign(C, mode, 0.5, 2 + r, 0)
        e1.record(); torch.cuda.synchronize()
        print("%-12s mode %d: %.2f ms per %d rows" % (name, mode, e0.elapsed_time(e1) / 3, rows))
