"""How much of a Lloyd iteration at the 8-GPU shard size (1.25M rows per GPU) is the host waiting for the two doubles
it reads back?  Same iteration enqueued K times with and without the per-iteration synchronize."""
import importlib.util
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn.cluster import _dmeans as D  # noqa: E402
from sqlearn.cluster._engine import QMeansEngine  # noqa: E402


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1250000
    K = 40
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1)
    X = torch.randn((rows, 128), dtype=torch.float64, device=dev, generator=g)
    X += 3.0 * torch.randn((256, 128), dtype=torch.float64, device=dev, generator=g)[torch.randint(0, 256, (rows,), device=dev, generator=g)]
    eng = QMeansEngine(X, 256)
    c0 = X[:256].clone()
    host = torch.empty((2,), dtype=torch.float64).pin_memory()

    def body(st, sync):
        old = st.centers
        labels, _, _ = D._labels_inertia(eng, old, delta=0.5, pool=None, true_distance_estimate=False)
        new = D._centers_update(eng, labels, 0.25, intermediate_error=True, true_tomography=True,
                                stop_when_reached_accuracy=True, centers_old=old, centers_out=st.spare)
        sh = eng.shift(old, new)
        if new is st.spare:
            st.spare = old
        st.centers = new
        host[0:1].copy_(eng.inertia, non_blocking=True)
        host[1:2].copy_(sh, non_blocking=True)
        if sync:
            torch.cuda.current_stream().synchronize()

    for sync in (True, False, True, False):
        st = D.LloydState(c0)
        for _ in range(5):
            body(st, True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(K):
            body(st, sync)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        print("rows %d sync=%s: %.3f ms/iter (events), %.3f ms/iter (wall)" % (rows, sync, e0.elapsed_time(e1) / K, (t1 - t0) * 1e3 / K))


main()
