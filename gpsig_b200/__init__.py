"""gpsig_b200 -- B200-native signature-kernel Gram computation.

Drop-in for the hot path of maudl3116/GPSig: ``gpsig_b200.kernels`` mirrors
``gpsig.kernels`` (truncated signature kernels) and ``gpsig_b200.kernels_pde``
mirrors ``gpsig.kernels_pde`` (untruncated / Goursat PDE kernel).  All numbers
come from hand-written sm_100a CUDA kernels behind the C ABI declared in
``include/gpsig_b200.h``; importing the package fails if that library has not
been built (``python gpsig_b200/build.py``).
"""
from . import settings  # noqa: F401
from . import _lib  # noqa: F401  (raises if libgpsig_b200.so is missing)
from . import ops, kernels, kernels_pde  # noqa: F401
from . import inducing_variables, conditionals, preprocessing  # noqa: F401

__version__ = '0.1.0'
