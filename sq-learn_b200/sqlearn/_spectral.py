"""Eigen-decomposition of a Gram matrix with the small end of the spectrum refined.

The reference takes LAPACK SVDs of the data itself (scipy.linalg.svd in qPCA._fit_full,
sklearn/decomposition/_qPCA.py:581; np.linalg.svd in qMeans_.fit, sklearn/cluster/_dmeans.py:1244), whose
singular values are accurate to about eps * sigma_max in absolute terms.  sigma = sqrt(eig(X^T X)) alone is
not: the eigenvalues carry an absolute error of about eps * sigma_max^2, so every sigma below
sqrt(eps) * sigma_max (1.5e-8 relative) is noise - and code downstream looks at exactly those values
(np.isclose(singular_values_, 0) in least_k_sv_extractors, _qPCA.py:1078; 1 / min(S), _dmeans.py:1245).

Refinement: the eigenvectors of the small eigenvalues span the right subspace to about eps, whatever their
eigenvalues say.  Project the rows on that subspace, Y = X V2 (sq_project, fp64 tensor cores), and solve
the small problem Y^T Y, whose entries are now of the size of the values looked for; repeat on what is
still small relative to that block.  Each level multiplies the reachable range by ~1e-6 in sigma^2 ... the
floor is the eps * sigma_max of Y itself, the accuracy LAPACK's SVD has.  Nothing happens (and nothing is
paid) for well-conditioned data.
"""
import torch

from . import _cabi
from . import _device as dv

TAU = 1e-6          # eigenvalues below TAU * (largest of the block) are re-solved in their own subspace
MAX_LEVELS = 8


def refined_eigh(Xd, G):
    """Eigen-pairs of G = sum over ranks of Xd^T Xd (already all-reduced), ascending, with the small
    eigenvalues and their vectors refined as described above.  Xd: this rank's rows [n, d] on the device.
    Returns (lam [d], V [d, d] with eigenvectors in columns)."""
    lam, V = torch.linalg.eigh(G)                      # cuSOLVER, d x d
    n, d = Xd.shape
    top = float(lam[-1].item())
    if not top > 0.0:
        return lam, V
    hi = int((lam < TAU * top).sum().item())
    for _ in range(MAX_LEVELS):
        if hi == 0:
            break
        Vt2 = V[:, :hi].T.contiguous()                 # rows = vectors of the small block
        Y = dv.empty((n, hi))
        zeros = torch.zeros((hi,), dtype=torch.float64, device=Xd.device)
        if n > 0:
            _cabi.check(_cabi.lib.sq_project(dv.ptr(Xd), n, d, dv.ptr(Vt2), hi, dv.ptr(zeros), dv.ptr(Y),
                                             dv.stream_ptr()), "sq_project")
        G2 = torch.zeros((hi, hi), dtype=torch.float64, device=Xd.device)
        if n > 0:
            nb = _cabi.lib.sq_gram_workspace_bytes(n, hi)
            ws = dv.workspace(nb)
            _cabi.check(_cabi.lib.sq_gram_centered(dv.ptr(Y), n, hi, None, dv.ptr(G2), dv.ptr(ws), nb,
                                                   dv.stream_ptr()), "sq_gram_centered")
        dv.all_reduce_sum(G2)
        lam2, W = torch.linalg.eigh(G2)
        V[:, :hi] = V[:, :hi] @ W
        lam[:hi] = lam2
        top2 = float(lam2[-1].item())
        nxt = int((lam2 < TAU * top2).sum().item()) if top2 > 0.0 else 0
        if nxt == hi:                                  # the whole block is numerically zero
            break
        hi = nxt
    return lam, V
