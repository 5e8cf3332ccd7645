"""Long-vector tomography (d > 4096): block-group kernel on / off (sq_tomography_tune 8)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn import _cabi
from sqlearn.QuantumUtility import Utility as U
g = torch.Generator(device="cuda")
for B, d, delta, inc, norm in [(4096, 10000, 0.1, False, "L2"), (64, 100000, 0.1, False, "L2"), (64, 1000000, 0.1, False, "L2"),
                               (1, 1000000, 0.1, False, "L2"), (4096, 10000, 0.1, True, "L2"), (64, 100000, 0.1, True, "L2"),
                               (1, 1000000, 0.1, True, "L2"), (64, 100000, 0.05, False, "inf"), (512, 5000, 0.2, True, "L2")]:
    g.manual_seed(7)
    V = torch.rand((B, d), generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    V /= V.norm(dim=1, keepdim=True)
    for on in (0, 1):
        _cabi.lib.sq_tomography_tune(8, float(on))
        def run():
            U.seed(3)
            return U._tomography_device(V, delta, None, norm, True, inc, 0)
        r = run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(2): r = run()
        e1.record(); torch.cuda.synchronize()
        err = r[2]
        print("B=%d d=%d delta=%g %s %s blockgroup=%d: %.3f ms  err mean %.5f max %.5f  N mean %.4g" % (
            B, d, delta, "incr" if inc else "single", norm, on, e0.elapsed_time(e1) / 2, float(err.mean()), float(err.max()), float(r[1].mean())), flush=True)
_cabi.lib.sq_tomography_tune(8, 1.0)
