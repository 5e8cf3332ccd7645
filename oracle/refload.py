"""Load the reference's hot-path source UNMODIFIED from /root/reference.

Only usable inside the build container (the GPU box has no /root/reference);
used by tests/golden/gen_golden.py to produce the committed fixtures and by the
``-m "not gpu"`` tests that re-validate the oracle when the mount is present.

The fork cannot be imported as a package here (it needs a Cython build of
scikit-learn 1.0.dev0, matlab.engine and matplotlib), so:
  * QuantumUtility/Utility.py is exec'd with empty stand-ins for matplotlib;
  * the seven hot functions of cluster/_dmeans.py are extracted by AST and
    exec'd in a namespace holding the loaded Utility symbols.
Nothing is copied into this repository.
"""
import ast
import importlib.util
import itertools
import os
import random
import sys
import types
import warnings

import numpy as np

REF = os.environ.get("SQ_REFERENCE", "/root/reference")


def available():
    return os.path.exists(os.path.join(REF, "sklearn", "QuantumUtility", "Utility.py"))


def load_utility():
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    path = os.path.join(REF, "sklearn", "QuantumUtility", "Utility.py")
    spec = importlib.util.spec_from_file_location("ref_quantum_utility", path)
    mod = importlib.util.module_from_spec(spec)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(mod)
    return mod


DMEANS_FUNCS = ("_kmeans_single_lloyd", "_labels_inertia", "wrapper", "labels_estimation", "_centers_update",
                "select_labels", "_tolerance", "_kmeans_plusplus")


def load_dmeans(util=None):
    """Namespace with the reference's q-means functions (cluster/_dmeans.py)."""
    import scipy as sc
    import scipy.sparse as sp
    import scipy.spatial  # noqa: F401
    from sklearn.utils.extmath import row_norms, squared_norm, stable_cumsum
    from sklearn.metrics.pairwise import euclidean_distances
    from sklearn.metrics.pairwise import _euclidean_distances
    from threadpoolctl import threadpool_limits
    util = util or load_utility()
    path = os.path.join(REF, "sklearn", "cluster", "_dmeans.py")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in DMEANS_FUNCS]
    ns = {k: getattr(util, k) for k in dir(util) if not k.startswith("__")}
    ns.update(np=np, sc=sc, sp=sp, itertools=itertools, random=random, warnings=warnings,
              threadpool_limits=threadpool_limits, row_norms=row_norms, squared_norm=squared_norm,
              stable_cumsum=stable_cumsum, euclidean_distances=euclidean_distances,
              _euclidean_distances=_euclidean_distances,
              _openmp_effective_n_threads=lambda n=None: 1,
              lloyd_iter_chunked_dense=None, lloyd_iter_chunked_sparse=None, _inertia_dense=None,
              _inertia_sparse=None, _k_means=None, check_random_state=None)
    mod = ast.Module(body=keep, type_ignores=[])
    exec(compile(mod, path, "exec"), ns)
    return types.SimpleNamespace(**{k: ns[k] for k in DMEANS_FUNCS}, util=util, ns=ns)


def seed_everything(util, s):
    """The reference draws from OS entropy (Utility.py:53); make it reproducible."""
    random.seed(s)
    np.random.seed(s)
    rs = np.random.RandomState(s)
    util.np = types.SimpleNamespace(**{k: getattr(np, k) for k in dir(np) if not k.startswith("__")})
    util.np.random = types.SimpleNamespace(**{k: getattr(np.random, k) for k in dir(np.random)
                                              if not k.startswith("__")})
    util.np.random.RandomState = lambda *a, **k: rs
    return rs


def load_qpca(util=None):
    """Class ``qPCA`` of the reference (decomposition/_qPCA.py), AST-extracted and exec'd over a
    minimal stand-in for ``_BasePCA`` (only _validate_data and the transform pair, restated
    from decomposition/_base.py:97-164) and the installed scikit-learn's svd_flip/stable_cumsum."""
    import numbers
    from math import log, sqrt
    from scipy import linalg
    from scipy.sparse import issparse
    from sklearn.utils.extmath import stable_cumsum, svd_flip
    util = util or load_utility()
    path = os.path.join(REF, "sklearn", "decomposition", "_qPCA.py")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "qPCA"]
    for node in ast.walk(keep[0]):      # the class decorates __init__ with a sklearn-internal helper
        if isinstance(node, ast.FunctionDef):
            node.decorator_list = []

    class _BasePCA:
        def _validate_data(self, X, dtype=None, ensure_2d=True, copy=False, reset=True):
            return np.array(X, dtype=np.float64, copy=True) if copy else np.asarray(X, dtype=np.float64)

        def transform(self, X, use_classical_components=True):
            X = np.asarray(X, dtype=np.float64)
            if self.mean_ is not None:
                X = X - self.mean_
            comp = self.components_ if use_classical_components else self.estimate_right_sv
            Xt = np.dot(X, comp.T)
            if self.whiten:
                Xt /= np.sqrt(self.explained_variance_ if use_classical_components else self.factor_score_estimation)
            return Xt

        def inverse_transform(self, X, use_classical_components=True):
            comp = self.components_ if use_classical_components else self.estimate_right_sv
            return np.dot(X, comp) + self.mean_

    ns = {k: getattr(util, k) for k in dir(util) if not k.startswith("__")}
    ns.update(np=np, linalg=linalg, issparse=issparse, svd_flip=svd_flip, stable_cumsum=stable_cumsum, log=log,
              sqrt=sqrt, numbers=numbers, warnings=warnings, random=random, _BasePCA=_BasePCA,
              check_is_fitted=lambda *a, **k: None, _infer_dimension=None)
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["qPCA"]


def load_qsvm(util=None):
    """Class ``QLSSVC`` of the reference (svm/_qSVM.py), AST-extracted and exec'd with the installed
    scikit-learn's pairwise kernels / accuracy_score and a minimal BaseEstimator (only _validate_data)."""
    from sklearn.metrics import accuracy_score
    from sklearn.metrics.pairwise import linear_kernel, polynomial_kernel, rbf_kernel, sigmoid_kernel
    util = util or load_utility()
    path = os.path.join(REF, "sklearn", "svm", "_qSVM.py")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tree = ast.parse(open(path).read())
    keep = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "QLSSVC"]

    class BaseEstimator:
        def _validate_data(self, X, y=None, **kw):
            X = np.asarray(X, dtype=np.float64)
            self.n_features_in_ = X.shape[1]
            return (X, np.asarray(y)) if y is not None else X

    ns = dict(np=np, BaseEstimator=BaseEstimator, check_array=lambda X, **k: np.asarray(X),
              check_is_fitted=lambda *a, **k: None, linear_kernel=linear_kernel, rbf_kernel=rbf_kernel,
              sigmoid_kernel=sigmoid_kernel, polynomial_kernel=polynomial_kernel, accuracy_score=accuracy_score,
              introduce_error=util.introduce_error, introduce_error_array=util.introduce_error_array)
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["QLSSVC"]
