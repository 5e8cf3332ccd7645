"""Minimal estimator base: get_params / set_params as in sklearn/base.py:179-260
so that clone(), pickling and parameter grids keep working without importing
scikit-learn in the product path."""
import inspect


class BaseEstimator:
    @classmethod
    def _get_param_names(cls):
        init = cls.__init__
        if init is object.__init__:
            return []
        sig = inspect.signature(init)
        return sorted(p.name for p in sig.parameters.values()
                      if p.name != 'self' and p.kind not in (p.VAR_KEYWORD, p.VAR_POSITIONAL))

    def get_params(self, deep=True):
        return {name: getattr(self, name) for name in self._get_param_names()}

    def set_params(self, **params):
        valid = self._get_param_names()
        for key, value in params.items():
            if key not in valid:
                raise ValueError('Invalid parameter %s for estimator %s.' % (key, type(self).__name__))
            setattr(self, key, value)
        return self

    def __repr__(self):
        args = ", ".join("%s=%r" % kv for kv in self.get_params().items())
        return "%s(%s)" % (type(self).__name__, args)


class NotFittedError(ValueError, AttributeError):
    pass


def check_is_fitted(est, attr):
    if not hasattr(est, attr):
        raise NotFittedError("This %s instance is not fitted yet. Call 'fit' with appropriate arguments "
                             "before using this estimator." % type(est).__name__)
