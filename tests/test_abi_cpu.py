"""CPU tests of the boundary: the C-ABI library loads, exports every symbol the
header declares, validates arguments without touching a GPU, and the Python
mirror of the reference classes behaves like the reference's constructors."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib():
    from gpsig_b200 import _lib
    return _lib


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, 'include', 'gpsig_b200.h')).read()
    declared = set(re.findall(r'\b(gpsig_[a-z0-9_]+)\s*\(', hdr))
    assert len(declared) >= 14
    lib = ctypes.CDLL(_lib().LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "libgpsig_b200.so does not export %s" % name


def test_version_and_workspace_queries():
    L = _lib()
    assert L.lib.gpsig_version() == 100
    # packed block: L*DP + padded L elements per sequence, 128-byte aligned regions
    assert L.lib.gpsig_trunc_diag_workspace_bytes(10, 200, 8, 8) == 10 * (200 * 8 + 200) * 8
    # the Gram workspace uses the padded row stride of the systolic kernel (an odd number of 16-byte units per row;
    # float32 rows with every coordinate stored twice for the packed-pair contraction: 2 x 4 floats = 2 units -> 3)
    assert L.lib.gpsig_trunc_gram_workspace_bytes(3, 5, 0, 0, 3, 4) == ((3 * (5 * 12 + 8) * 4 + 127) // 128) * 128
    assert L.lib.gpsig_trunc_gram_workspace_bytes(4, 50, 6, 70, 3, 8) == \
        ((4 * (50 * 6 + 50) * 8 + 127) // 128) * 128 + ((6 * (70 * 6 + 70) * 8 + 127) // 128) * 128


def test_argument_validation_without_gpu():
    L = _lib()
    desc = L.make_desc('rbf', 4, 1, True)
    rc = L.lib.gpsig_trunc_diag_f64(desc, None, 4, 10, 3, None, None, None, 0, None)
    assert rc == -1 and b'NULL' in L.lib.gpsig_last_error_string()
    bad = L.make_desc('rbf', 9, 1, True)
    buf = (ctypes.c_double * 16)()
    rc = L.lib.gpsig_trunc_diag_f64(bad, buf, 1, 4, 2, None, buf, buf, 128, None)
    assert rc == -2 and b'num_levels' in L.lib.gpsig_last_error_string()
    ho = L.make_desc('linear', 4, 5, True)
    rc = L.lib.gpsig_trunc_diag_f64(ho, buf, 1, 4, 2, None, buf, buf, 128, None)
    assert rc == -1 and b'order' in L.lib.gpsig_last_error_string()
    rc = L.lib.gpsig_trunc_diag_f64(desc, buf, 1, 4, 2, None, buf, buf, 8, None)
    assert rc == -3 and b'workspace' in L.lib.gpsig_last_error_string()
    rc = L.lib.gpsig_pde_f64(buf, 2, 1, None, 0, 0, 2, None, 0, 0, 0, 2, 1.0, buf, buf, 1024, None)
    assert rc == -1 and b'short' in L.lib.gpsig_last_error_string()
    with pytest.raises(NotImplementedError):
        L.check(-2)
    with pytest.raises(ValueError):
        L.check(-1)


def test_kernel_constructors_follow_reference():
    from gpsig_b200 import kernels, kernels_pde
    k = kernels.SignatureRBF(1600, 8, 5)
    assert k.len_examples == 200 and k.order == 1 and k.normalization and k.difference
    assert k.variances.shape == (6,) and k.lengthscales.shape == (8,) and k.sigma == 1.
    assert kernels.SignatureLinear(150, 3, 4, order=-1).order == 4       # kernels.py:58
    assert kernels.SignatureLinear(150, 3, 4, order=7).order == 4
    assert kernels.SignatureLinear(150, 3, 4, lengthscales=None).lengthscales is None
    assert kernels.SignatureGauss is kernels.SignatureRBF and kernels.SignatureLaplace is kernels.SignatureMatern12
    with pytest.raises(ValueError):
        kernels.SignatureRBF(1601, 8, 5)
    with pytest.raises(ValueError):
        kernels.SignatureRBF(1600, 8, 5, lengthscales=np.ones(3))
    with pytest.raises(ValueError):
        kernels.SignatureRBF(1600, 8, 5, low_rank=True, sparsity='cube')
    with pytest.raises(NotImplementedError):
        kernels.SignatureRBF(1600, 8, 5, low_rank=True, order=2)
    p = kernels_pde.UntruncSignatureKernel(2000, 4, order=2, implementation='cython')
    assert p.len_examples == 500 and p.order == 2
    with pytest.raises(AssertionError):
        kernels_pde.UntruncSignatureKernel(2000, 4, implementation='tpu')
    assert kernels_pde.SignatureLinear(40, 4, 1, 3).order == 1


def test_ops_refuse_cpu_tensors():
    import torch
    from gpsig_b200 import ops
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.trunc_gram(torch.zeros(2, 3, 2, dtype=torch.float64), None, 2)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.pde(torch.zeros(2, 3, 2, dtype=torch.float64))


def test_tf_shim_compiles_against_the_c_abi():
    """shim/gpsig_tf_ops.cc (REGISTER_OP / OpKernel / REGISTER_KERNEL_BUILDER over the C ABI, the pattern of
    gpsig/covariance_op/untrunc_cov_op_gpu.cc:13-40,61-102,148-149) type-checks against include/gpsig_b200.h.
    TensorFlow is not installed here: its few types come from the declaration-only stubs under tests/tf_stub."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location('build_shim', os.path.join(root, 'shim', 'build_shim.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    r = mod.check()
    assert r.returncode == 0, r.stderr
    src = open(os.path.join(root, 'shim', 'gpsig_tf_ops.cc')).read()
    for op in ('GpsigTruncGram', 'GpsigTruncGramGrad', 'GpsigTensVsSeq', 'GpsigPdeDiag', 'GpsigPdeDiagGrad'):
        assert 'REGISTER_OP("%s")' % op in src and 'Name("%s")' % op in src
