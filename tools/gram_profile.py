"""Gram matrix X^T X at BASELINE config 2 (70 000 x 784) and at 4M x 128: time, TFLOP/s (also the ncu target)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn import _cabi, _device as dv
for n, d in ((70000, 784), (4000000, 128)):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    X = torch.randn((n, d), generator=g, device="cuda", dtype=torch.float64)
    G = dv.empty((d, d)); nb = _cabi.lib.sq_gram_workspace_bytes(n, d); ws = dv.workspace(nb)
    run = lambda: _cabi.check(_cabi.lib.sq_gram_centered(dv.ptr(X), n, d, None, dv.ptr(G), dv.ptr(ws), nb, dv.stream_ptr()), "gram")
    run(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [run() for _ in range(5)]; e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("gram %d x %d: %.3f ms, %.1f TFLOP/s (2 n d^2 flop of the full product; only the upper block triangle is computed), max |G - X^T X| %.2e" % (
        n, d, ms, 2.0 * n * d * d / ms / 1e9, float((G - X.T @ X).abs().max()) if n <= 100000 else float("nan")))
