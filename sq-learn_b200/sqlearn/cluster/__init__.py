"""Quantum clustering (reference: sklearn/cluster/__init__.py:19)."""
from ._aliases import KMeans  # noqa: F401
from ._dmeans import qMeans_, ConvergenceWarning  # noqa: F401

__all__ = ['qMeans_', 'KMeans']
