"""Device plumbing: torch owns memory, streams and the process group; the
kernels are reached through :mod:`sqlearn._cabi` only."""
import os

import numpy as np
import torch

from . import _cabi  # noqa: F401  (fails loudly when the library is missing)


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


_STAGE_THREADS = max(1, int(os.environ.get("SQ_STAGE_THREADS", "4")))   # host threads per rank for the staging copy
_stage_pool = None


def _stage_threads():
    """4 threads move 19 GB/s on the B200 box's host (one: 6 GB/s); with several ranks on one node the cores are
    shared, so each rank takes its share."""
    world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")) or 1)
    return max(1, min(_STAGE_THREADS, (os.cpu_count() or 1) // max(world, 1)))


def _staged_copy(dst, src):
    """pageable -> pinned copy of one chunk, cut into row slices copied by a few host threads (numpy's copy
    releases the GIL and does not depend on the OpenMP thread count, which torchrun pins to one per rank)"""
    global _stage_pool
    rows = src.shape[0]
    nthr = _stage_threads()
    if nthr == 1 or rows < 4 * nthr or src.numel() < (8 << 20):
        dst.copy_(src)
        return
    if _stage_pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _stage_pool = ThreadPoolExecutor(max_workers=_STAGE_THREADS)
    dn, sn = dst.numpy(), src.numpy()
    step = (rows + nthr - 1) // nthr
    futs = [_stage_pool.submit(np.copyto, dn[a:a + step], sn[a:a + step]) for a in range(0, rows, step)]
    for f in futs:
        f.result()


def stream_to_device(a, on_chunk=None, chunk_bytes=256 << 20):
    """Host matrix [n, d] (float64) -> device tensor, copied in row chunks on a side stream; ``on_chunk(out, r0,
    r1)`` is called once per chunk with the current stream already waiting for that chunk, so the caller's
    kernels on rows [r0, r1) overlap the copy of the next chunk.  Pageable input is staged through two pinned
    chunk buffers (filled by a few host threads) instead of pinning - and copying - the whole array."""
    dev = require_cuda()
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
    n, d = t.shape
    out = torch.empty((n, d), dtype=torch.float64, device=dev)
    if n == 0:
        return out
    rows = max(1, int(chunk_bytes) // (8 * d))
    main = torch.cuda.current_stream()
    copy = torch.cuda.Stream(device=dev)
    copy.wait_stream(main)                      # the allocation above belongs to the current stream
    out.record_stream(copy)
    pinned = t.is_pinned()
    ring, ring_ev = [], [None, None]
    if not pinned:
        ring = [torch.empty((min(rows, n), d), dtype=torch.float64).pin_memory() for _ in range(2 if n > rows else 1)]
    for c, r0 in enumerate(range(0, n, rows)):
        r1 = min(n, r0 + rows)
        with torch.cuda.stream(copy):
            if pinned:
                out[r0:r1].copy_(t[r0:r1], non_blocking=True)
            else:
                slot = c % len(ring)
                if ring_ev[slot] is not None:
                    ring_ev[slot].synchronize()       # the buffer's previous upload has left the host
                _staged_copy(ring[slot][:r1 - r0], t[r0:r1])
                out[r0:r1].copy_(ring[slot][:r1 - r0], non_blocking=True)
                ring_ev[slot] = torch.cuda.Event()
                ring_ev[slot].record(copy)
            ev = torch.cuda.Event()
            ev.record(copy)
        main.wait_event(ev)
        if on_chunk is not None:
            on_chunk(out, r0, r1)
    return out


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
