"""Group-kernel tunables: threads per CTA (256 / 512 / 1024) on a few single-shot shapes."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn import _cabi
from sqlearn import _device as dv
from sqlearn.QuantumUtility import Utility as U


def run_ws(V, delta, norm, seed=7):
    B, d = V.shape
    N = U.tomography_num_measurements(d, delta, norm)
    sched = np.asarray([float(int(N))])
    est = torch.empty_like(V)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    nb = _cabi.lib.sq_tomography_workspace_bytes(B, d)
    ws = dv.workspace(nb)
    rc = _cabi.lib.sq_tomography_batched_ws(V.data_ptr(), B, d, float(delta), 1 if norm == "inf" else 0, 0,
                                            sched.ctypes.data, 1, seed, 0, est.data_ptr(), n_used.data_ptr(),
                                            err.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr())
    assert rc == 0, _cabi.lib.sq_last_error()
    return est


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


g = torch.Generator(device="cuda")
for B, d, delta, norm in [(4096, 1000, 0.1, "L2"), (65536, 128, 0.25, "L2"), (16384, 300, 0.05, "L2"), (4096, 4096, 0.1, "L2"),
                          (1024, 1000, 0.1, "L2"), (512, 4000, 0.01, "L2"), (4096, 1000, 0.01, "inf"), (300000, 16, 0.1, "L2")]:
    g.manual_seed(1234)
    V = torch.rand((B, d), generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    V /= V.norm(dim=1, keepdim=True)
    ref = None
    for th in (1024, 512, 256):
        _cabi.lib.sq_tomography_tune(5, float(th))
        ms = timed(lambda: run_ws(V, delta, norm))
        out = run_ws(V, delta, norm)
        ref = out if ref is None else ref
        print("B=%d d=%d delta=%g %s threads %d: %.3f ms  %.1f GB/s  same=%s" % (B, d, delta, norm, th, ms, 16.0 * B * d / ms / 1e6,
                                                                          torch.equal(out, ref)), flush=True)
