"""sqlearn — the quantum-simulation path of sq-learn on NVIDIA B200.

Mirrors the reference's public names for that path only:
``sqlearn.QuantumUtility`` (sklearn/QuantumUtility/Utility.py),
``sqlearn.cluster.qMeans_`` (sklearn/cluster/_dmeans.py),
``sqlearn.decomposition.qPCA`` (sklearn/decomposition/_qPCA.py) and
``sqlearn.svm.QLSSVC`` (sklearn/svm/_qSVM.py).  Compute runs
in hand-written sm_100a CUDA kernels behind the C ABI of ``include/sqb200.h``;
importing this package fails if that library has not been built.
"""
from . import _cabi  # noqa: F401  (fails loudly when libsqb200.so is missing)

__version__ = "0.1.0"
