"""Time batched-tomography configurations under several settings of the kernel's tunables
(sq_tomography_tune), and optionally check the estimates bit for bit against another build of the
library (e.g. the previous round's, tools/_r1/libsqb200_r1.so).

usage: python tools/tomo_bench.py [--old PATH] [--quick]
Prints one line per (shape, setting): ms, vectors/s, algorithmic GB/s.
"""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sq-learn_b200"))
from sqlearn import _cabi  # noqa: E402
from sqlearn import _device as dv  # noqa: E402
from sqlearn.QuantumUtility import Utility as U  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--old", default=None)
ap.add_argument("--quick", action="store_true")
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

SHAPES = [  # B, d, delta, incremental, norm
    (4096, 1000, 0.1, False, "L2"),
    (4096, 1000, 0.1, True, "L2"),
    (65536, 128, 0.25, True, "L2"),
    (65536, 128, 0.25, False, "L2"),
    (70000, 61, 0.1, True, "L2"),
    (64, 4096, 0.05, False, "L2"),
    (4096, 1000, 0.1, False, "inf"),
    (4096, 1000, 0.1, True, "inf"),
    (256, 128, 0.25, True, "L2"),
    (4096, 4096, 0.1, False, "L2"),
    (16384, 300, 0.05, False, "L2"),
    (1000, 1000, 0.5, False, "inf"),
    (300, 2, 0.5, False, "L2"),
    (5000, 3, 0.5, False, "L2"),
]
if args.quick:
    SHAPES = SHAPES[:3]
SETTINGS = [("default", {})]


def run_ws(V, delta, inc, norm, seed=7):
    """through sq_tomography_batched_ws (group kernel for single attempts of many vectors)"""
    B, d = V.shape
    N = U.tomography_num_measurements(d, delta, norm)
    sched = U.tomography_schedule(int(N), 0).astype(np.float64) if inc else np.asarray([float(int(N))])
    sched = np.ascontiguousarray(sched)
    est = torch.empty_like(V)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    nb = _cabi.lib.sq_tomography_workspace_bytes(B, d)
    ws = dv.workspace(nb)
    rc = _cabi.lib.sq_tomography_batched_ws(V.data_ptr(), B, d, float(delta), 1 if norm == "inf" else 0, 1 if inc else 0,
                                            sched.ctypes.data, len(sched), seed, 0, est.data_ptr(), n_used.data_ptr(),
                                            err.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr())
    assert rc == 0, _cabi.lib.sq_last_error()
    return est, n_used, err


def run(lib, V, delta, inc, norm, seed=7, spec_ok=True):
    B, d = V.shape
    N = U.tomography_num_measurements(d, delta, norm)
    sched = U.tomography_schedule(int(N), 0).astype(np.float64) if inc else np.asarray([float(int(N))])
    sched = np.ascontiguousarray(sched)
    est = torch.empty_like(V)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    rc = lib.sq_tomography_batched(V.data_ptr(), B, d, float(delta), 1 if norm == "inf" else 0, 1 if inc else 0,
                                   sched.ctypes.data, len(sched), seed, 0, est.data_ptr(), n_used.data_ptr(),
                                   err.data_ptr(), None, None, dv.stream_ptr())
    assert rc == 0, rc
    return est, n_used, err


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


old = None
if args.old:
    old = ctypes.CDLL(args.old)
    old.sq_tomography_batched.restype = ctypes.c_int
    old.sq_tomography_batched.argtypes = _cabi.SIGNATURES["sq_tomography_batched"][1]

g = torch.Generator(device="cuda")
for B, d, delta, inc, norm in SHAPES:
    g.manual_seed(1234)
    V = torch.rand((B, d), generator=g, device="cuda", dtype=torch.float64) * 2 - 1
    V /= V.norm(dim=1, keepdim=True)
    if B >= 300:
        V[3] *= 2.5          # a vector that is normalised first
        V[5] = 0.0           # and a zero vector
    tag = "B=%d d=%d delta=%g %s %s" % (B, d, delta, "incr+stop" if inc else "single", norm)
    if old is not None:
        ms_old = timed(lambda: run(old, V, delta, inc, norm), args.reps)
        ref = run(old, V, delta, inc, norm)
        print("%-44s %-10s %8.3f ms" % (tag, "OLD", ms_old), flush=True)
    ms = timed(lambda: run(_cabi.lib, V, delta, inc, norm), args.reps)
    fused = run(_cabi.lib, V, delta, inc, norm)
    same = ""
    if old is not None:
        same = "bit-identical to OLD: %s" % all(torch.equal(a, b) for a, b in zip(fused, ref))
    print("%-44s %-10s %8.3f ms  %.3g vec/s  %.1f GB/s  err max %.3g  N mean %.4g  %s" % (
        tag, "fused", ms, B / ms * 1e3, 16.0 * B * d / ms / 1e6, float(fused[2].max()), float(fused[1].mean()), same),
        flush=True)
    if not inc:
        ms = timed(lambda: run_ws(V, delta, inc, norm), args.reps)
        out = run_ws(V, delta, inc, norm)
        print("%-44s %-10s %8.3f ms  %.3g vec/s  %.1f GB/s  bit-identical to fused: %s" % (
            tag, "GROUP", ms, B / ms * 1e3, 16.0 * B * d / ms / 1e6, all(torch.equal(a, b) for a, b in zip(out, fused))),
            flush=True)
