"""Device plumbing: torch owns memory, streams and the process group; the
kernels are reached through :mod:`sqlearn._cabi` only."""
import os

import numpy as np
import torch

from . import _cabi


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("sqlearn needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


def to_device(a, dtype=torch.float64, pin=False):
    """Host array -> device tensor (C order)."""
    dev = require_cuda()
    if isinstance(a, torch.Tensor):
        return a.to(device=dev, dtype=dtype).contiguous()
    a = np.ascontiguousarray(a)
    t = torch.from_numpy(a)
    if t.dtype != dtype:
        t = t.to(dtype)
    if pin:
        t = t.pin_memory()
    return t.to(dev, non_blocking=pin)


def empty(shape, dtype=torch.float64):
    return torch.empty(shape, dtype=dtype, device=require_cuda())


def workspace(nbytes):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=require_cuda())


# ---- distributed (one process per GPU; rows sharded) -----------------------
def dist_info():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def all_reduce_sum(t):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t
