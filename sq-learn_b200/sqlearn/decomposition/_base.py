"""Shared PCA plumbing (reference: sklearn/decomposition/_base.py:97-164, with the
fork's ``use_classical_components`` switch)."""
import numpy as np
import torch

from .. import _device as dv
from .._base import BaseEstimator, check_is_fitted


class _BasePCA(BaseEstimator):
    def get_covariance(self):
        """components_.T * S**2 * components_ + sigma2 * I (stock scikit-learn)."""
        components_ = self.components_
        exp_var = self.explained_variance_
        if self.whiten:
            components_ = components_ * np.sqrt(exp_var[:, np.newaxis])
        exp_var_diff = np.maximum(exp_var - self.noise_variance_, 0.)
        cov = np.dot(components_.T * exp_var_diff, components_)
        cov.flat[::len(cov) + 1] += self.noise_variance_
        return cov

    def get_precision(self):
        return np.linalg.inv(self.get_covariance())

    def _project(self, X, comp):
        """(X - mean_) @ comp.T on the device (reference: _base.py:117-128), as X comp^T - mean.comp^T on the
        fp64 tensor cores (``sq_project``)."""
        from .. import _cabi
        Xd = dv.stream_to_device(np.asarray(X, dtype=np.float64))
        Vt = dv.to_device(np.ascontiguousarray(comp, dtype=np.float64))
        n, d = Xd.shape
        r = Vt.shape[0]
        if n == 0 or r == 0:
            return dv.empty((n, r))
        off = torch.zeros((r,), dtype=torch.float64, device=Xd.device)
        if self.mean_ is not None:
            off = Vt @ dv.to_device(self.mean_)
        out = dv.empty((n, r))
        _cabi.check(_cabi.lib.sq_project(dv.ptr(Xd), n, d, dv.ptr(Vt), r, dv.ptr(off), dv.ptr(out), dv.stream_ptr()),
                    "sq_project")
        return out

    def _transform_device(self, X, use_classical_components=True):
        """transform() with the result left on the device (the quantum representation feeds it
        straight into the batched tomography kernel)."""
        check_is_fitted(self, "components_")
        X = np.asarray(X, dtype=np.float64)
        if X.ndim != 2 or X.shape[1] != self.n_features_:
            raise ValueError("X has %d features, but %s is expecting %d features as input."
                             % (X.shape[-1], type(self).__name__, self.n_features_))
        if use_classical_components:
            Xt = self._project(X, self.components_)
            if self.whiten:
                Xt = Xt / dv.to_device(np.sqrt(self.explained_variance_))
        else:
            Xt = self._project(X, self.estimate_right_sv)
            if self.whiten:
                Xt = Xt / dv.to_device(np.sqrt(self.factor_score_estimation))
        return Xt.contiguous()

    def transform(self, X, use_classical_components=True):
        """Project X on the (classical or estimated) components (reference: _base.py:97-128)."""
        return self._transform_device(X, use_classical_components).cpu().numpy()

    def inverse_transform(self, X, use_classical_components=True):
        """Back to the original space (reference: _base.py:130-164)."""
        if use_classical_components:
            if self.whiten:
                return np.dot(X, np.sqrt(self.explained_variance_[:, np.newaxis]) * self.components_) + self.mean_
            return np.dot(X, self.components_) + self.mean_
        if self.whiten:
            return np.dot(X, np.sqrt(self.factor_score_estimation[:, np.newaxis]) * self.estimate_right_sv) + self.mean_
        return np.dot(X, self.estimate_right_sv) + self.mean_
