"""TEST INFRASTRUCTURE ONLY (like everything under ``oracle/``): runs the reference's OWN Python code for the truncated
path -- ``gpsig/signature_algs.py``, ``gpsig/kernels.py``, ``gpsig/lags.py``, imported where they lie under
``/root/reference`` -- in the build container, where TensorFlow 1.15 and GPflow 1.5.1 cannot be installed.

The two dependencies are replaced by the minimal eager stand-ins of ``stubs/``: ``tensorflow`` = the ~30 graph ops the
reference's truncated path calls, evaluated at once with numpy (same semantics: exclusive / reverse cumsums, batched
matmul with transposes, gather along an axis ...), ``gpflow`` = ``Parameter`` as a plain array, the ``params_as_tensors``
/ ``autoflow`` decorators as the identity, ``Kernel`` with its ``active_dims`` slicing, ``settings.float_type`` /
``jitter``.  No line of the reference is copied or modified: its modules are executed as they are, so what comes out
is the reference's algorithm, order of operations included, in float64 numpy arithmetic.  ``tests/golden/make_golden.py``
uses it to write ``tests/golden/trunc_ref.npz``; nothing here is imported at test or bench time on the GPU box.
"""
import importlib
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference(root='/root/reference'):
    """Returns the reference's ``gpsig.kernels`` module, executed on the numpy stand-ins."""
    sys.dont_write_bytecode = True               # nothing is written next to the reference's sources
    pkg_dir = os.path.join(root, 'gpsig')
    if not os.path.exists(os.path.join(pkg_dir, 'kernels.py')):
        raise RuntimeError('reference not found under %s' % root)
    stubs = os.path.join(HERE, 'stubs')
    if stubs not in sys.path:
        sys.path.insert(0, stubs)
    for name in ('tensorflow', 'gpflow'):
        mod = sys.modules.get(name)
        if mod is not None and not getattr(mod, '__file__', '').startswith(stubs):
            raise RuntimeError('a real %s is already imported; run the generator in a fresh process' % name)
    # the package namespace without running gpsig/__init__.py (which pulls in the models, training loops, ...)
    pkg = types.ModuleType('gpsig')
    pkg.__path__ = [pkg_dir]
    sys.modules['gpsig'] = pkg
    return importlib.import_module('gpsig.kernels')


def load_reference_pde(root='/root/reference'):
    """``gpsig.kernels_pde`` of the reference with ``implementation='cython'``: its ``.sigKer_fast`` import resolves to
    the reference's own Cython solver compiled by oracle/build_ref.py into oracle/_ref/ (built from the sources where
    they lie)."""
    load_reference(root)
    ref_dir = os.path.join(os.path.dirname(HERE), '_ref')
    pkg = sys.modules['gpsig']
    if ref_dir not in pkg.__path__:
        pkg.__path__.append(ref_dir)
    gp = os.path.join(root, 'gpsig')
    if gp not in sys.path:
        sys.path.append(gp)                       # kernels_pde.py:24-25 does ``from covariance_op import ...``
    return importlib.import_module('gpsig.kernels_pde')


def load_reference_inducing(root='/root/reference'):
    """``gpsig.inducing_variables`` of the reference (InducingTensors / InducingSequences, Kuu, Kuf, Kuu_Kuf_Kff)."""
    load_reference(root)
    return importlib.import_module('gpsig.inducing_variables')
