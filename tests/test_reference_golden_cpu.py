"""The CPU oracle against the REFERENCE's own code: tests/golden/trunc_ref.npz holds what gpsig/kernels.py +
signature_algs.py + lags.py return when they are executed where they lie under /root/reference, on eager numpy stand-ins
for TensorFlow / GPflow (oracle/refrun, tests/golden/make_golden.py --trunc-ref).  This is the pin of the truncated path:
the numpy restatement (and the C port, which the GPU tests use at larger sizes) must reproduce the reference to
rounding, for every base kernel, order, normalisation, lags, differencing option and covariance method."""
import os

import numpy as np
import pytest

from oracle import c_oracle, trunc_np
from tests.util import rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
M = 4
BASES = [('linear', {}), ('rbf', {}), ('cosine', {}), ('poly', dict(gamma=0.7, degree=3.0)), ('mix', {}),
         ('matern12', {}), ('matern32', {}), ('matern52', {})]


@pytest.fixture(scope='module')
def g():
    return np.load(os.path.join(GOLD, 'trunc_ref.npz'))


def _oracle(g, base, bp=None, **kw):
    N, L, d = g['X'].shape
    return trunc_np.SignatureKernelOracle(L * d, d, M, base=base, lengthscales=g['lengthscales'], variances=g['variances'],
                                          sigma=float(g['sigma']), **(bp or {}), **kw)


def _flat(a):
    return a.reshape(a.shape[0], -1)


@pytest.mark.parametrize("base", ['linear', 'rbf'])
@pytest.mark.parametrize("order", [1, 2, M])
@pytest.mark.parametrize("norm", [False, True])
def test_orders_and_normalisation(g, base, order, norm):
    o = _oracle(g, base, order=order, normalization=norm)
    X, Y = _flat(g['X']), _flat(g['Y'])
    tag = '%s_o%d_n%d' % (base, order, norm)
    assert rel_err(o.K(X, Y, return_levels=True), g['K_' + tag]) < 1e-13
    assert rel_err(o.K(X, return_levels=True), g['Ksym_' + tag]) < 1e-13
    assert rel_err(o.Kdiag(X, return_levels=True), g['Kdiag_' + tag]) < 1e-13
    for inc in (0, 1):
        Kzz, Kzx, Kxx = o.K_tens_n_seq_covs(g['Z_inc%d' % inc], X, return_levels=True, increments=bool(inc))
        assert rel_err(Kzz, g['Kzz_%s_i%d' % (tag, inc)]) < 1e-13
        assert rel_err(Kzx, g['Kzx_%s_i%d' % (tag, inc)]) < 1e-13
        assert rel_err(Kxx, g['Kxxd_%s_i%d' % (tag, inc)]) < 1e-13


@pytest.mark.parametrize("base,bp", BASES)
@pytest.mark.parametrize("norm", [False, True])
def test_every_base_kernel(g, base, bp, norm):
    o = _oracle(g, base, bp, order=1, normalization=norm)
    X, Y = _flat(g['X']), _flat(g['Y'])
    tag = '%s_n%d' % (base, norm)
    # Matern12 takes sqrt(max(r^2, 1e-40)) of rounding noise on coincident points (kernels.py:959-961): 1e-6 there
    tol = 1e-6 if base == 'matern12' else 1e-12
    assert rel_err(o.K(X, Y), g['Kall_' + tag]) < tol
    assert rel_err(o.Kdiag(X), g['Kdiagall_' + tag]) < tol
    assert rel_err(o.K_tens_vs_seq(g['Z_inc1'], X, increments=True), g['Kzxall_' + tag]) < tol
    assert rel_err(o.K_tens(g['Z_inc1'], increments=True), g['Kzzall_' + tag]) < tol


def test_options_no_difference_lags_seq_covs(g):
    X, Y, Y14 = _flat(g['X']), _flat(g['Y']), _flat(g['Y14'])
    assert rel_err(_oracle(g, 'rbf', order=1, normalization=True, difference=False).K(X, Y, return_levels=True), g['Knodiff_rbf']) < 1e-13
    assert rel_err(_oracle(g, 'rbf', order=1, normalization=False, difference=False).K(X, Y, return_levels=True), g['Knodiff_rbf_n0']) < 1e-13
    for norm in (0, 1):
        o = _oracle(g, 'rbf', order=1, normalization=bool(norm), num_lags=2)
        assert rel_err(o.K(X, Y14, return_levels=True), g['Klags_rbf_n%d' % norm]) < 1e-12
        assert rel_err(o.Kdiag(X, return_levels=True), g['Kdiaglags_rbf_n%d' % norm]) < 1e-12
        o = _oracle(g, 'rbf', order=1, normalization=bool(norm))
        Kxx, Kxx2, Kx2x2 = o.K_seq_n_seq_covs(X, Y, return_levels=True)
        assert rel_err(Kxx, g['Ksns_xx_n%d' % norm]) < 1e-13
        assert rel_err(Kxx2, g['Ksns_xx2_n%d' % norm]) < 1e-13
        assert rel_err(Kx2x2, g['Ksns_x2x2_n%d' % norm]) < 1e-13


@pytest.mark.parametrize("base", ['linear', 'rbf'])
def test_c_port_raw_levels(g, base):
    """The C port (the checker of the large GPU tests and the timed CPU baseline) on the same fixtures: raw first-order
    levels of the scaled sequences = the reference's un-normalised levels divided by sigma * variances."""
    X, Y = g['X'] / g['lengthscales'], g['Y'] / g['lengthscales']
    R = c_oracle.trunc_gram(X, Y, M, base)
    ref = g['K_%s_o1_n0' % base] / (float(g['sigma']) * g['variances'])[:, None, None]
    assert rel_err(R, ref) < 1e-13


@pytest.mark.parametrize("family", ['rbf', 'exp'])
def test_spectral_base_kernel(g, family):
    """SignatureSpectral (kernels.py:1084-1122) with the parameters the reference's constructor drew."""
    N, L, d = g['X'].shape
    o = trunc_np.SignatureKernelOracle(L * d, d, M, base='spectral', lengthscales=None, variances=g['variances'], sigma=float(g['sigma']),
                                       normalization=False, family=family, alpha=g['spec_%s_alpha' % family],
                                       omega=g['spec_%s_omega' % family], gamma=g['spec_%s_gamma' % family])
    assert rel_err(o.K(_flat(g['X']), _flat(g['Y']), return_levels=True), g['Kspec_' + family]) < 1e-13


@pytest.mark.parametrize("lw", [0, 1])
def test_svgp_glue_kuu_kuf_kff(g, lw):
    """gpsig/inducing_variables.py:51-137 executed by the reference itself vs oracle/svgp_np.py (inducing tensors in both
    forms, inducing sequences; with the per-level mixing weights W)."""
    from oracle import svgp_np
    o = _oracle(g, 'rbf', order=1, normalization=True)
    X = _flat(g['X'])
    cases = [('tens_i%d' % inc, g['Z_inc%d' % inc], dict(increments=bool(inc)), g['W_tens']) for inc in (0, 1)]
    cases.append(('seq', _flat(g['Zseq']), dict(sequences=True), g['W_seq']))
    for name, Z, kw, W in cases:
        tag = 'glue_%s_w%d' % (name, lw)
        Kzz, Kzx, Kxx = svgp_np.Kuu_Kuf_Kff(o, Z, X, W=W if lw else None, jitter=1e-3, **kw)
        assert rel_err(Kzz, g[tag + '_Kzz']) < 1e-12
        assert rel_err(Kzx, g[tag + '_Kzx']) < 1e-12
        assert rel_err(Kxx, g[tag + '_Kxx']) < 1e-12


def test_pde_python_layer_and_backward_function():
    """tests/golden/pde_ref_py.npz: the reference's ``UntruncSignatureKernel.Kdiag`` (kernels_pde.py:136-185, Cython
    implementation = the compiled oracle/_ref solver) and its own ``_KdiagGrad`` (kernels_pde.py:465-509) executed on the
    stand-ins.  The backward function's output is BIT-EQUAL to the transcription the CUDA backward is tested against
    (``kdiag_grad_order*`` of pde_ref.npz, oracle/pde_np.kdiag_grad_from_grids)."""
    from oracle import pde_np
    g = np.load(os.path.join(GOLD, 'pde_ref_py.npz'))
    p = np.load(os.path.join(GOLD, 'pde_ref.npz'))
    X, ls, gout = g['X'], g['lengthscales'], g['gout']
    assert np.array_equal(X, p['X'])
    for order in (0, 1, 2):
        assert np.array_equal(float(g['sigma']) * pde_np.pde_kernel(X / ls, None, order), g['Kdiag_class_order%d' % order])
        assert np.array_equal(p['kdiag_grad_order%d' % order] * gout[:, None, None], g['KdiagGrad_fn_order%d' % order])
    o = trunc_np.SignatureKernelOracle(X.shape[1] * X.shape[2], X.shape[2], 2, base='linear', lengthscales=ls, num_lags=2)
    Xl = o._seq(X.reshape(X.shape[0], -1))                 # lags + scaling as kernels.py / kernels_pde.py:136-157 apply them
    assert rel_err(pde_np.pde_kernel(Xl, None, 1), g['Kdiag_class_lags2_order1']) < 1e-13
