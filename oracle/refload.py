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
