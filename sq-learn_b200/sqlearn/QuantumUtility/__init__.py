"""Quantum routines (reference: sklearn/QuantumUtility/__init__.py:5-6)."""
from .Utility import *  # noqa: F401,F403
from .Utility import seed  # noqa: F401
