"""Thin helpers for the -m gpu tests: call the C ABI on torch tensors."""
import numpy as np
import torch

from sqlearn import _cabi
from sqlearn import _device as dv

L = _cabi.lib


def dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a)).to(dtype).cuda()


def assign(X, C, mode=0, delta=0.0, seed=1, offset=0):
    """-> labels (np.int32), mindist, inertia, n_fixed"""
    Xd, Cd = dev(X), dev(C)
    n, d = Xd.shape
    k = Cd.shape[0]
    xn = torch.empty(n, dtype=torch.float64, device="cuda")
    _cabi.check(L.sq_row_sqnorms(Xd.data_ptr(), n, d, xn.data_ptr(), dv.stream_ptr()), "norms")
    labels = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    mind = torch.empty(n, dtype=torch.float64, device="cuda")
    inertia = torch.empty(1, dtype=torch.float64, device="cuda")
    nfix = torch.zeros(1, dtype=torch.int32, device="cuda")
    nb = L.sq_qmeans_assign_workspace_bytes(n, d, k)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    _cabi.check(L.sq_qmeans_assign(Xd.data_ptr(), xn.data_ptr(), n, d, Cd.data_ptr(), k, mode, float(delta), seed,
                                   offset, labels.data_ptr(), mind.data_ptr(), inertia.data_ptr(), nfix.data_ptr(),
                                   ws.data_ptr(), nb, dv.stream_ptr()), "assign")
    torch.cuda.synchronize()
    return labels.cpu().numpy(), mind.cpu().numpy(), float(inertia.item()), int(nfix.item())


def accumulate(X, labels, k):
    Xd = dev(X)
    n, d = Xd.shape
    lab = dev(labels, torch.int32)
    sums = torch.empty((k, d), dtype=torch.float64, device="cuda")
    counts = torch.empty(k, dtype=torch.float64, device="cuda")
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    nb = L.sq_centroid_workspace_bytes(n, d, k)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    _cabi.check(L.sq_centroid_accumulate(Xd.data_ptr(), lab.data_ptr(), n, d, k, sums.data_ptr(), counts.data_ptr(),
                                         bad.data_ptr(), ws.data_ptr(), nb, dv.stream_ptr()), "accumulate")
    torch.cuda.synchronize()
    return sums.cpu().numpy(), counts.cpu().numpy(), int(bad.item())


def tomography(V, delta, sched, norm_inf=0, stop=0, seed=1, vec_offset=0, trace=False):
    Vd = dev(np.atleast_2d(V))
    B, d = Vd.shape
    sched = np.ascontiguousarray(sched, dtype=np.float64)
    est = torch.empty_like(Vd)
    n_used = torch.empty(B, dtype=torch.float64, device="cuda")
    err = torch.empty(B, dtype=torch.float64, device="cuda")
    tr = torch.zeros((B, len(sched), d), dtype=torch.float64, device="cuda") if trace else None
    att = torch.zeros(B, dtype=torch.int32, device="cuda") if trace else None
    _cabi.check(L.sq_tomography_batched(Vd.data_ptr(), B, d, float(delta), norm_inf, stop, sched.ctypes.data,
                                        len(sched), seed, vec_offset, est.data_ptr(), n_used.data_ptr(),
                                        err.data_ptr(), tr.data_ptr() if trace else None,
                                        att.data_ptr() if trace else None, dv.stream_ptr()), "tomography")
    torch.cuda.synchronize()
    out = [est.cpu().numpy(), n_used.cpu().numpy(), err.cpu().numpy()]
    if trace:
        out += [tr.cpu().numpy(), att.cpu().numpy()]
    return out


def binomial(n, pl, pr, seed=1, offset=0):
    nd, a, b = dev(n), dev(pl), dev(pr)
    out = torch.empty_like(nd)
    _cabi.check(L.sq_binomial_sample(nd.data_ptr(), a.data_ptr(), b.data_ptr(), nd.numel(), seed, offset,
                                     out.data_ptr(), dv.stream_ptr()), "binomial")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def fejer(M, x, seed=1, offset=0, brute=False):
    Md, xd = dev(M), dev(x)
    out = torch.empty_like(xd)
    fn = L.sq_fejer_sample_bruteforce if brute else L.sq_fejer_sample
    _cabi.check(fn(Md.data_ptr(), xd.data_ptr(), xd.numel(), seed, offset, out.data_ptr(), dv.stream_ptr()), "fejer")
    torch.cuda.synchronize()
    return out.cpu().numpy()
