"""QLSSVC - least-squares SVM classifier with the quantum error model (reference: sklearn/svm/_qSVM.py:10-404).

Same constructor, attributes and methods as the reference class.  The kernel matrices run on the fp64 tensor
cores (``sq_kernel_matrix``), the (N+1) x (N+1) Hermitian decomposition goes through cuSOLVER
(``torch.linalg.eigh``; the reference calls ``np.linalg.svd(F, hermitian=True)``, which is the same
factorisation), and the truncated-normal error of ``predict`` / ``get_P`` / ``get_h`` is drawn for all samples
in one launch (``sq_truncnorm_add_each``) instead of one ``scipy.stats.truncnorm.rvs`` call per sample.

Quirks kept: a float ``gamma`` is ignored (``_get_gamma`` returns None for it, :392-404, and scikit-learn's
pairwise kernels then use 1 / n_features); ``gamma='scale'`` is computed from the first argument of
``get_kernel`` (always the training set on the prediction paths); ``get_approximated_hyperplane`` adds noise of
half-width ``relative_error / beta`` whatever ``error_type`` says for the 'absolute' case (:316-317).
"""
import numpy as np
import torch

from .. import _cabi
from .. import _device as dv
from .._base import BaseEstimator, check_is_fitted
from ..QuantumUtility import Utility as qu

_KIND = {'linear': 0, 'poly': 1, 'rbf': 2, 'sigmoid': 3}


def _truncnorm_each(x, halfwidth):
    """x_i + TruncNormal(0, 1; [-a_i, a_i]) on device tensors (reference: introduce_error, Utility.py:68-69)."""
    x = x.contiguous()
    a = halfwidth.to(torch.float64).contiguous()
    out = torch.empty_like(x)
    s, off = qu._take_streams(x.numel())
    _cabi.check(_cabi.lib.sq_truncnorm_add_each(dv.ptr(x), x.numel(), dv.ptr(a), s, off, dv.ptr(out),
                                                dv.stream_ptr()), "sq_truncnorm_add_each")
    return out


class QLSSVC(BaseEstimator):
    """LSSVC ("Least squares Support Vector Machine Classifier", Suykens & Vandewalle) whose predictions carry
    the error of the quantum algorithm (reference: _qSVM.py:10-82 for the parameters)."""

    def __init__(self, kernel='linear', penalty=0.1, degree=3, gamma='scale', coef0=0.0,
                 verbose=False, algorithm='classic', low_rank=False, var=0.9, error_type='absolute',
                 relative_error=0.5, absolute_error=0.01, train_error=0.01) -> None:
        self.kernel = kernel
        self.penalty = penalty
        self.degree = degree
        self.gamma = gamma
        self.coef0 = coef0
        self.verbose = verbose
        self.algorithm = algorithm
        self.low_rank = low_rank
        self.var = var
        self.relative_error = relative_error
        self.absolute_error = absolute_error
        if error_type == 'absolute' or error_type == 'relative':
            self.error_type = error_type
        else:
            raise Exception(r"The error should be either 'absolute' or 'relative'")
        self.train_error = train_error

        self.alpha = None
        self.b = None
        self.approx_b = None
        self.approx_alpha = None
        self.is_fitted_ = False
        self.X = None
        self.normX = None
        self.coef_ = None
        self.n_features_in_ = None
        self.Nu = None
        self.cond = None
        self.alpha_F = None
        self.normF = None
        self.singular_values_F_ = None

    # ------------------------------------------------------------------ kernels
    def _get_gamma(self, X):
        """reference: _qSVM.py:392-404 (None for a float gamma: the pairwise kernels fall back to 1/n_features)."""
        if self.gamma == 'scale':
            Xd = X if isinstance(X, torch.Tensor) else dv.to_device(np.asarray(X, dtype=np.float64))
            gamma = 1 / (Xd.shape[1] * float(Xd.var(unbiased=False)))
            if self.verbose:
                print(f"Gamma: {gamma}, shape: {Xd.shape[1]}, Variance: {float(Xd.var(unbiased=False))}")
            return gamma
        if self.gamma == 'auto':
            gamma = 1 / self.n_features_in_
            if self.verbose:
                print(f"Gamma: {gamma}")
            return gamma
        return None

    def _kernel_device(self, X, Y=None):
        """k(X, Y) as a device tensor [n, m]; X, Y device tensors or arrays."""
        if self.kernel not in _KIND:
            return None                                   # the reference's get_kernel falls through to None
        Xd = X if isinstance(X, torch.Tensor) else dv.to_device(np.asarray(X, dtype=np.float64))
        Yd = Xd if Y is None else (Y if isinstance(Y, torch.Tensor) else dv.to_device(np.asarray(Y, dtype=np.float64)))
        Xd, Yd = Xd.contiguous(), Yd.contiguous()
        if Xd.dim() != 2 or Yd.dim() != 2 or Xd.shape[1] != Yd.shape[1]:
            raise ValueError("Incompatible dimension for X and Y matrices: X.shape[1] == %d while Y.shape[1] == %d"
                             % (Xd.shape[-1], Yd.shape[-1]))
        n, d = Xd.shape
        m = Yd.shape[0]
        gamma = None if self.kernel == 'linear' else self._get_gamma(Xd)
        if gamma is None:
            gamma = 1.0 / d
        K = dv.empty((n, m))
        if n == 0 or m == 0:
            return K
        nb = _cabi.lib.sq_kernel_matrix_workspace_bytes(n, m)
        ws = dv.workspace(nb)
        _cabi.check(_cabi.lib.sq_kernel_matrix(dv.ptr(Xd), n, dv.ptr(Yd), m, d, _KIND[self.kernel], float(gamma),
                                               float(self.coef0), float(self.degree), dv.ptr(K), dv.ptr(ws), nb,
                                               dv.stream_ptr()), "sq_kernel_matrix")
        return K

    def get_kernel(self, X, Y=None):
        """reference: _qSVM.py:375-389; returns a NumPy array."""
        K = self._kernel_device(X, Y)
        return None if K is None else K.cpu().numpy()

    # ------------------------------------------------------------------ fit
    def _classical_fit(self, y):
        """[b; alpha] = F^+ [0; y], F = [[0, 1^T], [1, K + I / penalty]] (reference: _qSVM.py:84-130)."""
        N = self._Xd.shape[0]
        dev = self._Xd.device
        F = torch.zeros((N + 1, N + 1), dtype=torch.float64, device=dev)
        F[0, 1:] = 1.0
        F[1:, 0] = 1.0
        F[1:, 1:] = self._kernel_device(self._Xd)
        F[1:, 1:].diagonal().add_(self.penalty ** -1)
        # np.linalg.svd(F, hermitian=True): eigendecomposition, singular values |w| in decreasing order
        w, U = torch.linalg.eigh(F)
        order = torch.argsort(-w.abs(), stable=True)
        s, u, sign = w.abs()[order], U[:, order], torch.sign(w[order])
        s_host = s.cpu().numpy()
        if self.low_rank:
            s_new = np.zeros(len(s_host))
            index = 0
            if 0 <= self.var < 1.:
                p = 0
                sums = np.sum(s_host ** 2)
                for index, i in enumerate(s_host):
                    s_new[index] = i
                    p = p + i ** 2 / sums
                    if p >= self.var:
                        break
            elif self.var >= 1.:
                for index, i in enumerate(s_host):
                    s_new[index] = i
                    if index == self.var - 1:
                        break
            else:
                raise Exception("QLSSVC.var shoud be greater than 0")
            self.cond = s_new[0] / s_new[index]
            self.normF = s_new[0]
            s_host = s_new
        else:
            self.cond = s_host[0] / s_host[-1]
            self.normF = s_host[0]
        self.singular_values_F_ = s_host
        inv = torch.as_tensor(np.array([1 / x if x > 0 else x for x in s_host]), device=dev)
        rhs = torch.cat([torch.zeros(1, dtype=torch.float64, device=dev), y])
        sol = (u * inv) @ ((u * sign).T @ rhs)
        sol = sol.cpu().numpy()
        return sol[0], sol[1:]

    def fit(self, X, y):
        """Fit the model (reference: _qSVM.py:133-175)."""
        X = np.asarray(X, dtype=np.float64)
        y = np.asarray(y)
        if X.ndim != 2:
            raise ValueError("Expected 2D array, got %dD array instead" % X.ndim)
        if y.shape[0] != X.shape[0]:
            raise ValueError("Found input variables with inconsistent numbers of samples: [%d, %d]"
                             % (X.shape[0], y.shape[0]))
        self.n_features_in_ = X.shape[1]
        self.X = X
        self._Xd = dv.to_device(X)
        N, d = X.shape
        if self.algorithm == 'classic':
            self.b, self.alpha = self._classical_fit(dv.to_device(y.astype(np.float64)))
        xn = (self._Xd * self._Xd).sum(dim=1)
        self.alpha_F = np.sqrt(N) + self.penalty ** -1 + float(xn.sum())
        alpha_d = dv.to_device(self.alpha)
        self.Nu = self.b ** 2 + float((alpha_d * alpha_d * xn).sum())
        if self.kernel == 'linear':
            self.coef_ = (alpha_d @ self._Xd).cpu().numpy()
        self.is_fitted_ = True
        return self

    # ------------------------------------------------------------------ noiseless quantities (device)
    def _check(self, X):
        check_is_fitted(self, "is_fitted_")
        if not self.is_fitted_:
            check_is_fitted(self, "_not_fitted_")
        Xd = dv.to_device(np.asarray(X, dtype=np.float64))
        if Xd.dim() != 2:
            raise ValueError("Expected 2D array, got %dD array instead" % Xd.dim())
        return Xd

    def _h_device(self, Xd):
        """alpha . k(X_train, x) + b for every row x (reference: _qSVM.py:267)."""
        K = self._kernel_device(self._Xd, Xd)                    # [N, n_test]
        return dv.to_device(self.alpha) @ K + self.b

    def _betas_device(self, Xd):
        N = self._Xd.shape[0]
        return torch.sqrt((N * (Xd * Xd).sum(dim=1) + 1) * self.Nu)

    def _relative_error_batch(self, delta, Xmax, Xreal):
        """relative_error_routine for a batch (device tensors Xmax, Xreal): every sample halves its bound until it
        drops to the noisy estimate of |h| (reference: _qSVM.py:245-261).  Returns (Xhat, delta_r, epsilon_abs)."""
        n = Xmax.numel()
        Xhat = torch.zeros_like(Xmax)
        eps = torch.zeros_like(Xmax)
        delta_r = torch.zeros_like(Xmax)
        Xr = Xmax.clone()
        active = Xr > Xhat
        r = 0.0
        while bool(active.any()):
            r += 1.0
            Xr_new = Xmax / (2 ** r)
            eps_new = (self.relative_error * Xr_new) / 2
            draw = _truncnorm_each(Xreal, torch.where(active, eps_new, torch.zeros_like(eps_new)))
            Xr = torch.where(active, Xr_new, Xr)
            eps = torch.where(active, eps_new, eps)
            delta_r = torch.where(active, torch.full_like(delta_r, (6 * delta) / (np.pi ** 2 * r ** 2)), delta_r)
            Xhat = torch.where(active, draw, Xhat)
            active = active & (Xr > Xhat)
            if r > 1100:          # Xmax / 2^r underflows to 0 long before: cannot loop forever
                break
        return Xhat, delta_r, eps

    def relative_error_routine(self, delta, Xmax, Xreal):
        """reference: _qSVM.py:245-261 (one sample)."""
        dev = dv.require_cuda()
        a = torch.tensor([float(Xmax)], dtype=torch.float64, device=dev)
        b = torch.tensor([float(Xreal)], dtype=torch.float64, device=dev)
        Xhat, delta_r, eps = self._relative_error_batch(delta, a, b)
        return float(Xhat[0]), float(delta_r[0]), float(eps[0])

    # ------------------------------------------------------------------ public API
    def predict(self, X):
        """Class labels with the quantum error on P (reference: _qSVM.py:178-214)."""
        Xd = self._check(X)
        h = self._h_device(Xd)
        betas = self._betas_device(Xd)
        P = 0.5 * (1 - h / betas)
        if self.error_type == 'absolute':
            P = _truncnorm_each(P, self.absolute_error / (2 * betas))
        else:
            _, _, abs_err = self._relative_error_batch(0.1, betas, h.abs())
            P = _truncnorm_each(P, abs_err / (2 * betas))
        return torch.where(P <= 0.5, torch.ones_like(P), -torch.ones_like(P)).cpu().numpy()

    def classical_predict(self, X):
        """sign(alpha . k(X_train, x) + b) (reference: _qSVM.py:217-243)."""
        return torch.sign(self._h_device(self._check(X))).cpu().numpy()

    def get_h(self, X, approx=False):
        """reference: _qSVM.py:263-276."""
        Xd = self._check(X)
        hs = self._h_device(Xd)
        if approx:
            if self.error_type == 'absolute':
                hs = _truncnorm_each(hs, torch.full_like(hs, float(self.absolute_error)))
            else:
                _, _, abs_err = self._relative_error_batch(0.1, self._betas_device(Xd), hs.abs())
                hs = _truncnorm_each(hs, abs_err)
        return hs.cpu().numpy()

    def get_betas(self, X):
        """reference: _qSVM.py:278-282."""
        return self._betas_device(self._check(X)).cpu().numpy()

    def get_P(self, X, approx=False):
        """reference: _qSVM.py:284-298."""
        Xd = self._check(X)
        hs = self._h_device(Xd)
        beta = self._betas_device(Xd)
        P = 0.5 * (1 - hs / beta)
        if approx:
            if self.error_type == 'absolute':
                P = _truncnorm_each(P, self.absolute_error / (2 * beta))
            else:
                _, _, abs_err = self._relative_error_batch(0.1, beta, hs.abs())
                P = _truncnorm_each(P, abs_err / (2 * beta))
        return P.cpu().numpy()

    def get_training_complexity(self):
        return self.cond * self.alpha_F

    def get_classification_complexity(self, X, relative_error=False):
        """reference: _qSVM.py:303-310."""
        betas = self.get_betas(X)
        nba = np.linalg.norm(np.append(self.b, self.alpha), ord=2)
        if relative_error:
            hs = np.abs(self.get_h(X))
            return (self.cond * betas * self.alpha_F) / (self.relative_error * np.abs(hs) * self.normF ** 2 * nba)
        return (self.cond * betas * self.alpha_F) / (self.absolute_error * self.normF ** 2 * nba)

    def get_approximated_hyperplane(self, x):
        """Noisy (b, w) for the sample(s) x (reference: _qSVM.py:313-332)."""
        beta = self.get_betas(x)
        ba = np.append(self.b, self.alpha)
        size = ba.shape[0]
        if self.error_type == 'absolute':
            norm_error = self.relative_error / beta
        else:
            norm_error = self.relative_error * np.abs(self.get_h(x)) / beta
        half = np.broadcast_to(np.asarray(norm_error, dtype=np.float64) / np.sqrt(size), (size,))
        approx_ba = _truncnorm_each(dv.to_device(ba), dv.to_device(np.ascontiguousarray(half)))
        b = float(approx_ba[0])
        approx_coef = (approx_ba[1:] @ self._Xd).cpu().numpy()
        return b, approx_coef

    def get_all_attributes(self, X):
        """reference: _qSVM.py:334-343."""
        betas = self.get_betas(X)
        hs = self.get_h(X)
        Ps = self.get_P(X)
        cond = self.cond
        rel_comp = (cond * (betas - np.abs(hs)) * self.alpha_F) / (np.abs(hs) * np.sqrt(Ps))
        abs_comp = (cond * betas * self.alpha_F)
        return (betas, hs, Ps, cond, rel_comp, abs_comp)

    def score(self, X, y):
        """Mean accuracy of predict(X) (reference: _qSVM.py:345-370)."""
        check_is_fitted(self, "is_fitted_")
        y_pred = self.predict(X=X)
        return float(np.mean(np.asarray(y) == y_pred))
