"""Minimal stand-in for GPflow 1.5.1 (requirements.txt:8 of the reference), enough to execute gpsig/kernels.py eagerly
(test infrastructure, see oracle/refrun/__init__.py)."""
from . import settings, transforms, params, decors, kernels, conditionals, kullback_leiblers, dispatch, features  # noqa: F401
