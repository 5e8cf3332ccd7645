"""E-step through the bf16 filter (csrc/qmeans_filter.cuh) against the fp64 tensor-core kernel: same labels, distances to
the rounding of two summation orders; then timing.  usage: python tools/filter_check.py [rows_for_timing]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from sqlearn import _cabi  # noqa: E402
from sqlearn.cluster._engine import QMeansEngine  # noqa: E402
import gpu_util as G  # noqa: E402

rng = np.random.default_rng(1)
bad = 0
for (n, d, k, delta, mode, kind) in [(5000, 128, 256, 0.0, 0, "blobs"), (5000, 128, 256, 0.5, 1, "blobs"), (3000, 16, 40, 0.5, 1, "blobs"),
                                     (3000, 64, 17, 0.0, 0, "rand"), (4000, 128, 256, 0.5, 1, "unit"), (2000, 32, 200, 2.0, 1, "rand"),
                                     (4097, 128, 256, 0.0, 0, "dup"), (1000, 48, 3, 0.3, 1, "blobs"), (70, 128, 256, 0.5, 1, "blobs"),
                                     (3000, 128, 256, 50.0, 1, "blobs"), (3000, 128, 100, 0.0, 0, "huge")]:
    cen = rng.normal(size=(k, d)) * 4
    X = cen[rng.integers(0, k, n)] + rng.normal(size=(n, d))
    C = cen + 0.1 * rng.normal(size=(k, d))
    if kind == "rand":
        X, C = rng.normal(size=(n, d)), rng.normal(size=(k, d))
    elif kind == "unit":
        C = cen / np.linalg.norm(cen, axis=1, keepdims=True)
    elif kind == "dup":
        C[5] = C[4]; C[7] = C[4]; X[::9] = 0.0
    elif kind == "huge":
        X *= 1e6; C *= 1e6
    out = []
    for on in (1, 0):
        _cabi.lib.sq_qmeans_debug_filter(on)
        lab, mind, inertia, nfix = G.assign(X, C, mode=mode, delta=delta, seed=3)
        out.append((np.asarray(lab), np.asarray(mind), float(inertia), int(nfix)))
    _cabi.lib.sq_qmeans_debug_filter(1)
    same_lab = (out[0][0] != out[1][0]).sum()
    rel = np.abs(out[0][1] - out[1][1]) / np.maximum(np.abs(out[1][1]), 1e-300)
    print("n=%d d=%d k=%d delta=%g mode=%d %-6s labels differing %d, max rel distance diff %.2e, inertia %.12g / %.12g, fixed rows %d / %d"
          % (n, d, k, delta, mode, kind, same_lab, rel.max(), out[0][2], out[1][2], out[0][3], out[1][3]))
    bad += int(same_lab != 0 and mode == 0) + int(rel.max() > 1e-9)
print("BAD" if bad else "OK")

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
cen = torch.randn((256, 128), dtype=torch.float64, device=dev, generator=g) * 4.0
X = cen[torch.randint(0, 256, (rows,), device=dev, generator=g)] + torch.randn((rows, 128), dtype=torch.float64, device=dev, generator=g)
eng = QMeansEngine(X.contiguous(), 256)
for name, C, mode, delta in [("exact", cen, 0, 0.0), ("window", cen, 1, 0.5), ("window, unit centroids", cen / cen.norm(dim=1, keepdim=True), 1, 0.5),
                             ("window, data points as centres", X[:256].clone(), 1, 0.5)]:
    res = []
    for on in (1, 0):
        _cabi.lib.sq_qmeans_debug_filter(on)
        for _ in range(2):
            eng.assign(C.contiguous(), mode, delta, 7, 0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            eng.assign(C.contiguous(), mode, delta, 7, 0)
        e1.record()
        torch.cuda.synchronize()
        res.append((e0.elapsed_time(e1) / 5, eng.labels.clone(), float(eng.inertia.item()), int(eng.n_fixed.item())))
    _cabi.lib.sq_qmeans_debug_filter(1)
    print("%-34s rows %d: filter %.3f ms (%.0f GB/s of X), fp64 kernel %.3f ms; labels equal %s; inertia %.15g / %.15g; fixed %d / %d"
          % (name, rows, res[0][0], rows * 128 * 8 / res[0][0] / 1e6, res[1][0], bool((res[0][1] == res[1][1]).all().item()), res[0][2], res[1][2], res[0][3], res[1][3]))
