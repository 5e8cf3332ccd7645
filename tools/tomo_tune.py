"""Sweep the threads-per-vector tunables of the tomography kernel (sq_tomography_tune 1, 2) on a few shapes.
usage: python tools/tomo_tune.py"""
import os
import sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from sqlearn import _cabi  # noqa: E402
from sqlearn import _device as dv  # noqa: E402
from sqlearn.QuantumUtility import Utility as U  # noqa: E402


def run(V, delta, inc, norm, seed=7):
    B, d = V.shape
    N = U.tomography_num_measurements(d, delta, norm)
    sched = U.tomography_schedule(int(N), 0).astype(np.float64) if inc else np.asarray([float(int(N))])
    sched = np.ascontiguousarray(sched)
    est = torch.empty_like(V)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    rc = _cabi.lib.sq_tomography_batched(V.data_ptr(), B, d, float(delta), 1 if norm == "inf" else 0, 1 if inc else 0,
                                         sched.ctypes.data, len(sched), seed, 0, est.data_ptr(), n_used.data_ptr(),
                                         err.data_ptr(), None, None, dv.stream_ptr())
    assert rc == 0, _cabi.lib.sq_last_error()
    return est


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


g = torch.Generator(device="cuda")
for B, d, delta, inc, norm in [(4096, 1000, 0.1, False, "L2"), (65536, 128, 0.25, False, "L2"), (65536, 128, 0.25, True, "L2"),
                               (4096, 1000, 0.1, True, "L2"), (16384, 300, 0.1, False, "L2"), (4096, 4096, 0.1, False, "L2")]:
    g.manual_seed(1234)
    V = torch.rand((B, d), generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    V /= V.norm(dim=1, keepdim=True)
    ref = None
    for div in (1, 2, 4, 8, 16, 32):
        for cap in (32, 64, 128):
            _cabi.lib.sq_tomography_tune(1, float(div))
            _cabi.lib.sq_tomography_tune(2, float(cap))
            ms = timed(lambda: run(V, delta, inc, norm))
            out = run(V, delta, inc, norm)
            if ref is None:
                ref = out
            print("B=%d d=%d inc=%d  L/%d cap %d: %.3f ms  same=%s" % (B, d, inc, div, cap, ms, torch.equal(out, ref)), flush=True)
