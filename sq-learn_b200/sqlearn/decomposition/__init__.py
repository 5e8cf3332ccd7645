"""Quantum decomposition (reference: sklearn/decomposition/__init__.py:22)."""
from ._qPCA import qPCA  # noqa: F401

__all__ = ['qPCA']
