"""Quantum least-squares SVM (reference: sklearn/svm/__init__.py exports QLSSVC from _qSVM.py)."""
from ._qSVM import QLSSVC  # noqa: F401

__all__ = ['QLSSVC']
