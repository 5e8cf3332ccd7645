"""Time (and, under ncu, profile) one batched-tomography configuration.
usage: python tools/tomo_profile.py B d delta [incremental=0] [norm=L2] [reps=5]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn.QuantumUtility import Utility as U  # noqa: E402

B, d, delta = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3])
inc = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
norm = sys.argv[5] if len(sys.argv) > 5 else "L2"
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 5
g = torch.Generator(device="cuda")
g.manual_seed(1234)
V = torch.rand((B, d), generator=g, device="cuda", dtype=torch.float64) * 2 - 1
V /= V.norm(dim=1, keepdim=True)
U.seed(1)
U._tomography_device(V, delta, None, norm, True, inc, 0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    est, n_used, err, _, _, _ = U._tomography_device(V, delta, None, norm, True, inc, 0)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print("B=%d d=%d delta=%g inc=%d norm=%s: %.3f ms, %.3g vectors/s, %.1f GB/s algorithmic, max err %.3g, N %.3g" % (
    B, d, delta, inc, norm, ms, B / ms * 1e3, 16.0 * B * d / ms / 1e6, float(err.max()), float(n_used.max())))
