"""Quantum decomposition (reference: sklearn/decomposition/__init__.py:22).  ``PCA`` is an alias of
``qPCA``: the quantum knobs are keyword arguments of ``fit`` / ``transform``."""
from ._qPCA import qPCA  # noqa: F401

PCA = qPCA

__all__ = ['qPCA', 'PCA']
