"""Device-resident q-means state: buffers + the per-iteration kernel calls.

Not part of the reference's surface; ``_dmeans.py`` (the mirror of
sklearn/cluster/_dmeans.py) drives it.  One instance holds this rank's shard
of X on the GPU together with every scratch buffer the E-step (K2), the
M-step (K3) and the centroid tomography (K1) need, so a Lloyd iteration makes
no allocation.  With torch.distributed initialised the rows are sharded over
ranks and the only exchange per iteration is one all-reduce of
``[k*d sums | k counts | inertia]``.
"""
import torch

from .. import _cabi
from .. import _device as dv

MODE_EXACT, MODE_WINDOW, MODE_IPE = 0, 1, 2


class QMeansEngine:
    def __init__(self, X_dev, n_clusters, norms=True):
        assert X_dev.is_cuda and X_dev.dtype == torch.float64 and X_dev.is_contiguous()
        self.X = X_dev
        self.n, self.d = X_dev.shape
        self.k = int(n_clusters)
        n, d, k = self.n, self.d, self.k
        self.xnorm = dv.empty((n,))
        self.labels = dv.empty((n,), dtype=torch.int32)
        self.mindist = dv.empty((n,))
        # one buffer for the all-reduce: sums | counts | inertia
        self.red = dv.empty((k * d + k + 1,))
        self.sums = self.red[:k * d].view(k, d)
        self.counts = self.red[k * d:k * d + k]
        self.inertia = self.red[k * d + k:]
        self.scal = dv.empty((4,))          # [shift, ...]
        self.n_fixed = dv.empty((1,), dtype=torch.int32)
        self.bad = torch.zeros((1,), dtype=torch.int32, device=X_dev.device)   # sticky flag: a label outside [0, k) was seen
        self.ws_a_bytes = _cabi.lib.sq_qmeans_assign_workspace_bytes(n, d, k)
        self.ws_a = dv.workspace(self.ws_a_bytes)
        self.ws_c_bytes = _cabi.lib.sq_centroid_workspace_bytes(n, d, k)
        self.ws_c = dv.workspace(self.ws_c_bytes)
        self.rank, self.world = dv.dist_info()
        # position of this shard in the global row order (keys the Philox streams)
        self.row0, self.n_total = 0, n
        if self.world > 1:
            import torch.distributed as dist
            mine = torch.tensor([float(n)], dtype=torch.float64, device=X_dev.device)
            sizes = [torch.zeros_like(mine) for _ in range(self.world)]
            dist.all_gather(sizes, mine)
            sizes = [int(s.item()) for s in sizes]
            self.row0, self.n_total = sum(sizes[:self.rank]), sum(sizes)
        self.launches = 0
        if norms:                   # (fit fills xnorm chunk by chunk while the rows are still arriving)
            self.refresh_norms()

    # ---- helpers ---------------------------------------------------------
    def refresh_norms(self):
        _cabi.check(_cabi.lib.sq_row_sqnorms(dv.ptr(self.X), self.n, self.d, dv.ptr(self.xnorm), dv.stream_ptr()),
                    "sq_row_sqnorms")
        self.launches += 1

    def col_sums(self):
        out = dv.empty((self.d,))
        nb = _cabi.lib.sq_col_sums_workspace_bytes(self.n, self.d)
        ws = dv.workspace(nb)
        _cabi.check(_cabi.lib.sq_col_sums(dv.ptr(self.X), self.n, self.d, dv.ptr(out), dv.ptr(ws), nb,
                                          dv.stream_ptr()), "sq_col_sums")
        self.launches += 2
        return out

    def subtract_rowvec(self, mu):
        _cabi.check(_cabi.lib.sq_sub_rowvec(dv.ptr(self.X), dv.ptr(mu), self.n, self.d, dv.ptr(self.X),
                                            dv.stream_ptr()), "sq_sub_rowvec")
        self.launches += 1

    # ---- E-step (K2) -------------------------------------------------------
    def assign(self, centers, mode, delta, seed, offset):
        """labels / mindist / inertia of this shard for the given centres (device [k, d])."""
        _cabi.check(_cabi.lib.sq_qmeans_assign(
            dv.ptr(self.X), dv.ptr(self.xnorm), self.n, self.d, dv.ptr(centers), self.k, mode, float(delta),
            seed, offset, dv.ptr(self.labels), dv.ptr(self.mindist), dv.ptr(self.inertia), dv.ptr(self.n_fixed),
            dv.ptr(self.ws_a), self.ws_a_bytes, dv.stream_ptr()), "sq_qmeans_assign")
        # centre norms (+ counter reset), assign, (fix-up), 2 x sum
        self.launches += 4 if mode == MODE_IPE else 5

    # ---- M-step (K3) -------------------------------------------------------
    def accumulate(self):
        _cabi.check(_cabi.lib.sq_centroid_accumulate(
            dv.ptr(self.X), dv.ptr(self.labels), self.n, self.d, self.k, dv.ptr(self.sums), dv.ptr(self.counts),
            dv.ptr(self.bad), dv.ptr(self.ws_c), self.ws_c_bytes, dv.stream_ptr()), "sq_centroid_accumulate")
        self.launches += 5          # histogram, scan (+ bases), scatter, segment sums, combine

    def finalize(self, centers_old, centers_new):
        _cabi.check(_cabi.lib.sq_centroid_finalize(dv.ptr(self.sums), dv.ptr(self.counts), dv.ptr(centers_old),
                                                   self.k, self.d, dv.ptr(centers_new), dv.stream_ptr()),
                    "sq_centroid_finalize")
        self.launches += 1

    def shift(self, a, b):
        _cabi.check(_cabi.lib.sq_sqdiff_sum(dv.ptr(a), dv.ptr(b), a.numel(), dv.ptr(self.scal), dv.stream_ptr()),
                    "sq_sqdiff_sum")
        self.launches += 1
        return self.scal[:1]

    def all_reduce(self):
        """Global sums / counts / inertia (no-op on one GPU)."""
        dv.all_reduce_sum(self.red)
