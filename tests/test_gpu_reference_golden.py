"""GPU parity against the REFERENCE's own code: the CUDA path (through the C ABI) vs tests/golden/trunc_ref.npz, i.e. what
gpsig/kernels.py + signature_algs.py + lags.py return when executed where they lie on eager numpy stand-ins for
TensorFlow / GPflow (oracle/refrun; tests/golden/make_golden.py --trunc-ref).  Tolerances: 1e-10 float64, 1e-5 float32."""
import os

import numpy as np
import pytest
import torch

from tests.util import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}
M = 4
BASES = [('linear', 'SignatureLinear', {}), ('rbf', 'SignatureRBF', {}), ('cosine', 'SignatureCosine', {}),
         ('poly', 'SignaturePoly', dict(gamma=0.7, degree=3.0)), ('mix', 'SignatureMix', {}),
         ('matern12', 'SignatureMatern12', {}), ('matern32', 'SignatureMatern32', {}), ('matern52', 'SignatureMatern52', {})]


@pytest.fixture(scope='module')
def g():
    return np.load(os.path.join(GOLD, 'trunc_ref.npz'))


def _dev(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


def _kern(g, cls, extra=None, **kw):
    from gpsig_b200 import kernels
    N, L, d = g['X'].shape
    k = getattr(kernels, cls)(L * d, d, M, lengthscales=g['lengthscales'], **(extra or {}), **kw)
    k.sigma = float(g['sigma'])
    k.variances = g['variances']
    return k


def _np(t):
    return t.cpu().numpy()


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base,cls", [('linear', 'SignatureLinear'), ('rbf', 'SignatureRBF')])
@pytest.mark.parametrize("order", [1, 2, M])
@pytest.mark.parametrize("norm", [False, True])
def test_orders_and_normalisation(g, dtype, base, cls, order, norm):
    k = _kern(g, cls, order=order, normalization=norm)
    N, N2 = g['X'].shape[0], g['Y'].shape[0]
    X, Y = _dev(g['X'].reshape(N, -1), dtype), _dev(g['Y'].reshape(N2, -1), dtype)
    tag = '%s_o%d_n%d' % (base, order, norm)
    assert rel_err(_np(k.K(X, Y, return_levels=True)), g['K_' + tag]) < TOL[dtype]
    assert rel_err(_np(k.K(X, return_levels=True)), g['Ksym_' + tag]) < TOL[dtype]
    assert rel_err(_np(k.Kdiag(X, return_levels=True)), g['Kdiag_' + tag]) < TOL[dtype]
    for inc in (0, 1):
        Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(_dev(g['Z_inc%d' % inc], dtype), X, return_levels=True, increments=bool(inc))
        assert rel_err(_np(Kzz), g['Kzz_%s_i%d' % (tag, inc)]) < TOL[dtype]
        assert rel_err(_np(Kzx), g['Kzx_%s_i%d' % (tag, inc)]) < TOL[dtype]
        assert rel_err(_np(Kxx), g['Kxxd_%s_i%d' % (tag, inc)]) < TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base,cls,extra", BASES)
@pytest.mark.parametrize("norm", [False, True])
def test_every_base_kernel(g, dtype, base, cls, extra, norm):
    k = _kern(g, cls, extra, order=1, normalization=norm)
    N, N2 = g['X'].shape[0], g['Y'].shape[0]
    X, Y, Z = _dev(g['X'].reshape(N, -1), dtype), _dev(g['Y'].reshape(N2, -1), dtype), _dev(g['Z_inc1'], dtype)
    tag = '%s_n%d' % (base, norm)
    # Matern12 on coincident points is only defined to sqrt(eps) (kernels.py:959-961)
    tol = {torch.float64: 1e-6, torch.float32: 1e-2}[dtype] if base == 'matern12' else TOL[dtype]
    assert rel_err(_np(k.K(X, Y)), g['Kall_' + tag]) < tol
    assert rel_err(_np(k.Kdiag(X)), g['Kdiagall_' + tag]) < tol
    assert rel_err(_np(k.K_tens_vs_seq(Z, X, increments=True)), g['Kzxall_' + tag]) < tol
    assert rel_err(_np(k.K_tens(Z, increments=True)), g['Kzzall_' + tag]) < tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_options_no_difference_lags_seq_covs(g, dtype):
    N, N2 = g['X'].shape[0], g['Y'].shape[0]
    X, Y, Y14 = (_dev(g[n].reshape(g[n].shape[0], -1), dtype) for n in ('X', 'Y', 'Y14'))
    tol = TOL[dtype]
    assert rel_err(_np(_kern(g, 'SignatureRBF', order=1, normalization=True, difference=False).K(X, Y, return_levels=True)), g['Knodiff_rbf']) < tol
    assert rel_err(_np(_kern(g, 'SignatureRBF', order=1, normalization=False, difference=False).K(X, Y, return_levels=True)), g['Knodiff_rbf_n0']) < tol
    for norm in (0, 1):
        k = _kern(g, 'SignatureRBF', order=1, normalization=bool(norm), num_lags=2)
        assert rel_err(_np(k.K(X, Y14, return_levels=True)), g['Klags_rbf_n%d' % norm]) < tol
        assert rel_err(_np(k.Kdiag(X, return_levels=True)), g['Kdiaglags_rbf_n%d' % norm]) < tol
        k = _kern(g, 'SignatureRBF', order=1, normalization=bool(norm))
        Kxx, Kxx2, Kx2x2 = k.K_seq_n_seq_covs(X, Y, return_levels=True)
        assert rel_err(_np(Kxx), g['Ksns_xx_n%d' % norm]) < tol
        assert rel_err(_np(Kxx2), g['Ksns_xx2_n%d' % norm]) < tol
        assert rel_err(_np(Kx2x2), g['Ksns_x2x2_n%d' % norm]) < tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("family", ['rbf', 'exp'])
def test_spectral_base_kernel(g, dtype, family):
    from gpsig_b200 import kernels
    N, L, d = g['X'].shape
    k = kernels.SignatureSpectral(L * d, d, M, family=family, Q=3, normalization=False)
    k.sigma, k.variances = float(g['sigma']), g['variances']
    k.alpha, k.omega, k.gamma = (g['spec_%s_%s' % (family, nm)] for nm in ('alpha', 'omega', 'gamma'))
    X, Y = _dev(g['X'].reshape(N, -1), dtype), _dev(g['Y'].reshape(g['Y'].shape[0], -1), dtype)
    assert rel_err(_np(k.K(X, Y, return_levels=True)), g['Kspec_' + family]) < TOL[dtype]


@pytest.mark.parametrize("lw", [0, 1])
def test_svgp_glue_kuu_kuf_kff(g, lw):
    """Kuu_Kuf_Kff / Kuu / Kuf of gpsig/inducing_variables.py:51-137, as the reference's own module returns them."""
    from gpsig_b200.inducing_variables import InducingSequences, InducingTensors, Kuf, Kuu, Kuu_Kuf_Kff
    dt = torch.float64
    k = _kern(g, 'SignatureRBF', order=1, normalization=True)
    N = g['X'].shape[0]
    X = _dev(g['X'].reshape(N, -1), dt)
    feats = [('tens_i%d' % inc, InducingTensors(_dev(g['Z_inc%d' % inc], dt), M, increments=bool(inc), learn_weights=bool(lw)), 'W_tens')
             for inc in (0, 1)]
    feats.append(('seq', InducingSequences(_dev(g['Zseq'], dt), M, learn_weights=bool(lw)), 'W_seq'))
    for name, feat, wkey in feats:
        if lw:
            feat.W = _dev(g[wkey], dt)
        tag = 'glue_%s_w%d' % (name, lw)
        Kzz, Kzx, Kxx = Kuu_Kuf_Kff(feat, k, X, jitter=1e-3)
        assert rel_err(_np(Kzz), g[tag + '_Kzz']) < 1e-10
        assert rel_err(_np(Kzx), g[tag + '_Kzx']) < 1e-10
        assert rel_err(_np(Kxx), g[tag + '_Kxx']) < 1e-10
        assert rel_err(_np(Kuu(feat, k, jitter=1e-3)), g[tag + '_Kuu']) < 1e-10
        assert rel_err(_np(Kuf(feat, k, X)), g[tag + '_Kuf']) < 1e-10


def test_pde_python_layer_and_backward_function():
    """tests/golden/pde_ref_py.npz: the reference's ``UntruncSignatureKernel.Kdiag`` (Cython implementation, with
    lengthscales / sigma / lags) and its own ``_KdiagGrad`` executed on the stand-ins, vs the CUDA kernels."""
    from gpsig_b200 import kernels_pde
    g = np.load(os.path.join(GOLD, 'pde_ref_py.npz'))
    X, ls, gout = g['X'], g['lengthscales'], g['gout']
    N, L, d = X.shape
    Xd = _dev(X.reshape(N, -1), torch.float64)
    for order in (0, 1, 2):
        k = kernels_pde.UntruncSignatureKernel(L * d, d, lengthscales=ls, order=order, implementation='cython')
        k.sigma = float(g['sigma'])
        assert rel_err(_np(k.Kdiag(Xd)), g['Kdiag_class_order%d' % order]) < 1e-10
        # the reference's backward function differentiates w.r.t. the op's input sequences
        k0 = kernels_pde.UntruncSignatureKernel(L * d, d, lengthscales=None, order=order, implementation='cython')
        Xg = Xd.clone().requires_grad_()
        (k0.Kdiag(Xg) * _dev(gout, torch.float64)).sum().backward()
        assert rel_err(_np(Xg.grad).reshape(N, L, d), g['KdiagGrad_fn_order%d' % order]) < 1e-10
    k = kernels_pde.UntruncSignatureKernel(L * d, d, lengthscales=ls, order=1, num_lags=2, implementation='cython')
    assert rel_err(_np(k.Kdiag(Xd)), g['Kdiag_class_lags2_order1']) < 1e-10
