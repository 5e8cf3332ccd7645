"""Quantum clustering (reference: sklearn/cluster/__init__.py:19)."""
from ._dmeans import qMeans_, ConvergenceWarning  # noqa: F401

__all__ = ['qMeans_']
