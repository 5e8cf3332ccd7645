"""IPE E-step with and without leaving pairs out (hook 2 samples every pair), on data with and without cluster
structure: the certificate must never cost much where it cannot help.
usage: python tools/ipe_prune_bench.py [rows]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn import _cabi  # noqa: E402
from sqlearn.cluster._engine import QMeansEngine, MODE_IPE  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
d, k = 128, 256
cen = torch.randn((k, d), dtype=torch.float64, device=dev, generator=g) * 4.0
lab = torch.randint(0, k, (rows,), device=dev, generator=g)
noise = torch.randn((rows, d), dtype=torch.float64, device=dev, generator=g)
cases = {
    "blobs, centres = blob centres": (cen[lab] + noise, cen.clone()),
    "blobs, unit-norm centroids (after tomography)": (cen[lab] + noise, cen / cen.norm(dim=1, keepdim=True)),
    "blobs, centres = data points": (cen[lab] + noise, (cen[lab] + noise)[:k].clone()),
    "no structure (all N(0,1))": (noise, torch.randn((k, d), dtype=torch.float64, device=dev, generator=g)),
    "no structure, unit-norm centroids": (noise, torch.nn.functional.normalize(torch.randn((k, d), dtype=torch.float64, device=dev, generator=g), dim=1)),
}
for name, (X, C) in cases.items():
    eng = QMeansEngine(X.contiguous(), k)
    out = []
    for hook in (0, 2):
        _cabi.lib.sq_ipe_debug_force_replay(hook)
        try:
            for _ in range(2):
                eng.assign(C, MODE_IPE, 0.5, 7, 0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                eng.assign(C, MODE_IPE, 0.5, 7, 0)
            e1.record()
            torch.cuda.synchronize()
            out.append((e0.elapsed_time(e1) / 3, eng.labels.clone(), float(eng.inertia.item())))
        finally:
            _cabi.lib.sq_ipe_debug_force_replay(0)
    same = bool((out[0][1] == out[1][1]).all().item()) and out[0][2] == out[1][2]
    print("%-48s %8.2f ms leaving pairs out, %8.2f ms sampling every pair, identical results: %s" % (name, out[0][0], out[1][0], same))
    del eng
