"""GPU parity of the higher-order recursion (order > 1; gpsig/signature_algs.py:37-74, 129-160).

This is the variant the reference itself validates (notebooks/signature_kernel.ipynb cells 11-29):
with the linear base kernel, order = num_levels and no normalisation the kernel equals the inner
product of truncated signatures level by level -- an EXTERNAL known answer (computed here with
Chen's identity, oracle/sigs_np.py), so these tests pin the CUDA path independently of the restated
TensorFlow algorithm.
"""
import os

import numpy as np
import pytest
import torch

from tests.util import brownian, inducing_tensors, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}
GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def _dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


def test_linear_full_order_equals_signature_inner_products():
    """notebooks/signature_kernel.ipynb cells 11-13."""
    from gpsig_b200 import kernels
    from oracle import sigs_np
    N, N2, L, L2, d, m = 7, 5, 13, 10, 3, 4
    X, Y = brownian(N, L, d, 0), brownian(N2, L2, d, 1)
    k = kernels.SignatureLinear(L * d, d, m, order=m, normalization=False, lengthscales=None)
    G = k.K(_dev(X.reshape(N, -1)), return_levels=True).cpu().numpy()
    assert rel_err(G, sigs_np.signature_gram_levels(X, None, m)) < 1e-12
    G2 = k.K(_dev(X.reshape(N, -1)), _dev(Y.reshape(N2, -1)), return_levels=True).cpu().numpy()
    assert rel_err(G2, sigs_np.signature_gram_levels(X, Y, m)) < 1e-12


def test_tens_vs_seq_full_order_equals_rank1_tensor_inner_products():
    """notebooks/signature_kernel.ipynb cells 21-23."""
    from gpsig_b200 import kernels
    from oracle import sigs_np
    N, L, d, m, Mz = 5, 11, 3, 3, 6
    X = brownian(N, L, d, 2)
    Z = inducing_tensors(X, m, Mz, False, 3)
    k = kernels.SignatureLinear(L * d, d, m, order=m, normalization=False, lengthscales=None)
    K = k.K_tens_vs_seq(_dev(Z), _dev(X.reshape(N, -1)), return_levels=True).cpu().numpy()
    S = [sigs_np.signature(x, m) for x in X]
    kk = 0
    for lvl in range(1, m + 1):
        for z in range(Mz):
            t = sigs_np.rank1_tensor_levels(Z[kk:kk + lvl, z])
            for n in range(N):
                assert abs(K[lvl, z, n] - np.sum(t * S[n][lvl])) < 1e-11 * max(1.0, abs(K[lvl, z, n]))
        kk += lvl


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base,cls", [('linear', 'SignatureLinear'), ('rbf', 'SignatureRBF')])
@pytest.mark.parametrize("order", [2, 4])
@pytest.mark.parametrize("norm", [False, True])
def test_golden_fixtures_higher_order(dtype, base, cls, order, norm):
    """Committed fixtures of the restated reference algorithm (tests/golden/make_golden.py)."""
    from gpsig_b200 import kernels
    g = np.load(os.path.join(GOLD, 'trunc_oracle.npz'))
    X, Y, ls = g['X'], g['Y'], g['lengthscales']
    N, L, d = X.shape
    m = 4
    k = getattr(kernels, cls)(L * d, d, m, order=order, normalization=norm, lengthscales=ls)
    k.sigma = 1.5
    k.variances = np.linspace(0.5, 1.5, m + 1)
    tag = '%s_o%d_n%d' % (base, order, norm)
    Xd, Yd = _dev(X.reshape(N, -1), dtype), _dev(Y.reshape(Y.shape[0], -1), dtype)
    assert rel_err(k.K(Xd, Yd, return_levels=True).cpu().numpy(), g['K_' + tag]) < TOL[dtype]
    assert rel_err(k.K(Xd, return_levels=True).cpu().numpy(), g['Ksym_' + tag]) < TOL[dtype]
    for inc in (0, 1):
        Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(_dev(g['Z_inc%d' % inc], dtype), Xd, return_levels=True, increments=bool(inc))
        assert rel_err(Kzz.cpu().numpy(), g['Kzz_%s_i%d' % (tag, inc)]) < TOL[dtype]
        assert rel_err(Kzx.cpu().numpy(), g['Kzx_%s_i%d' % (tag, inc)]) < TOL[dtype]
        assert rel_err(Kxx.cpu().numpy(), g['Kxxd_%s_i%d' % (tag, inc)]) < TOL[dtype]


@pytest.mark.parametrize("L,L2,d,m,order", [(30, 41, 2, 5, 3), (65, 33, 4, 3, 2), (20, 300, 3, 4, 4), (9, 9, 8, 8, 8)])
def test_higher_order_vs_oracle_shapes(L, L2, d, m, order):
    from gpsig_b200 import ops
    from oracle.trunc_np import SignatureKernelOracle
    N, N2 = 4, 6
    X, Y = brownian(N, L, d, 4) * 0.7, brownian(N2, L2, d, 5) * 0.7
    o = SignatureKernelOracle(L * d, d, m, base='rbf', order=order, normalization=False, lengthscales=None)
    R = o._K_seq(X, Y)
    G = ops.trunc_gram(_dev(X), _dev(Y), m, base='rbf', order=order).cpu().numpy()
    scale = np.abs(R).max(axis=(1, 2), keepdims=True)
    assert np.max(np.abs(G - R) / scale) < 1e-10
    D = ops.trunc_diag(_dev(X), m, base='rbf', order=order).cpu().numpy()
    assert np.max(np.abs(D - o._K_seq_diag(X)) / np.abs(o._K_seq_diag(X)).max(axis=1, keepdims=True)) < 1e-10


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_golden_fixtures_first_order(dtype):
    """The same fixtures at order 1 (the fused kernels)."""
    from gpsig_b200 import kernels
    g = np.load(os.path.join(GOLD, 'trunc_oracle.npz'))
    X, Y, ls = g['X'], g['Y'], g['lengthscales']
    N, L, d = X.shape
    m = 4
    for base, cls in (('linear', 'SignatureLinear'), ('rbf', 'SignatureRBF')):
        for norm in (False, True):
            k = getattr(kernels, cls)(L * d, d, m, order=1, normalization=norm, lengthscales=ls)
            k.sigma = 1.5
            k.variances = np.linspace(0.5, 1.5, m + 1)
            tag = '%s_o1_n%d' % (base, norm)
            Xd, Yd = _dev(X.reshape(N, -1), dtype), _dev(Y.reshape(Y.shape[0], -1), dtype)
            assert rel_err(k.K(Xd, Yd, return_levels=True).cpu().numpy(), g['K_' + tag]) < TOL[dtype]
            for inc in (0, 1):
                Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(_dev(g['Z_inc%d' % inc], dtype), Xd, return_levels=True, increments=bool(inc))
                assert rel_err(Kzx.cpu().numpy(), g['Kzx_%s_i%d' % (tag, inc)]) < TOL[dtype]
                assert rel_err(Kzz.cpu().numpy(), g['Kzz_%s_i%d' % (tag, inc)]) < TOL[dtype]


@pytest.mark.parametrize("base", ['linear', 'rbf'])
@pytest.mark.parametrize("m,order", [(2, 2), (3, 2), (4, 4), (5, 3)])
@pytest.mark.parametrize("difference", [True, False])
def test_higher_order_reverse_mode(base, m, order, difference):
    """Adjoints of the higher-order recursion (csrc/generic_vjp_kernels.cuh; the reference gets them from TF autodiff of
    signature_algs.py:37-74, 129-160): vector-Jacobian identity against central differences of the oracle's raw levels,
    for Gram pairs, diagonal pairs and tensors vs sequences (with and without increments)."""
    import torch
    from gpsig_b200 import ops
    from oracle.trunc_np import SignatureKernelOracle
    from tests.util import brownian, inducing_tensors
    EPS, TOLV = 1e-5, 2e-6
    N, N2, L, L2, d, Mz = 4, 3, 8, 7, 3, 4
    rng = np.random.default_rng(9)
    X, Y = brownian(N, L, d, 0) + 0.3, brownian(N2, L2, d, 1) - 0.2
    o = SignatureKernelOracle(L * d, d, m, base=base, lengthscales=None, normalization=False, order=order, difference=difference)
    dev = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64, device='cuda')  # noqa: E731

    def check(nm, a, f):
        assert abs(a - f) <= TOLV * max(abs(f), abs(a), 1e-12) + 1e-9, (nm, a, f)

    G = rng.standard_normal((m + 1, N, N2))
    VX, VY = rng.standard_normal(X.shape), rng.standard_normal(Y.shape)
    gX, gY = ops.trunc_vjp(dev(X), dev(Y), dev(G), m, base=base, order=order, difference=difference)
    f = lambda A, B: float(np.sum(G * o._K_seq(A, B)))  # noqa: E731
    check('X', float((gX.cpu().numpy() * VX).sum()), (f(X + EPS * VX, Y) - f(X - EPS * VX, Y)) / (2 * EPS))
    check('X2', float((gY.cpu().numpy() * VY).sum()), (f(X, Y + EPS * VY) - f(X, Y - EPS * VY)) / (2 * EPS))
    Gd = rng.standard_normal((m + 1, N))
    gX, _ = ops.trunc_vjp(dev(X), None, dev(Gd), m, base=base, order=order, difference=difference, diag=True)
    f = lambda A: float(np.sum(Gd * o._K_seq_diag(A)))  # noqa: E731
    check('diag', float((gX.cpu().numpy() * VX).sum()), (f(X + EPS * VX) - f(X - EPS * VX)) / (2 * EPS))
    for increments in (False, True):
        Z = inducing_tensors(X, m, Mz, increments, 2)
        VZ = rng.standard_normal(Z.shape)
        Gz = rng.standard_normal((m + 1, Mz, N))
        gZ, gX = ops.tens_vs_seq_vjp(dev(Z), dev(X), dev(Gz), m, increments=increments, base=base, order=order, difference=difference)
        f = lambda A, B: float(np.sum(Gz * o._K_tens_vs_seq(A, B, increments=increments)))  # noqa: E731
        check('Z', float((gZ.cpu().numpy() * VZ).sum()), (f(Z + EPS * VZ, X) - f(Z - EPS * VZ, X)) / (2 * EPS))
        check('Xz', float((gX.cpu().numpy() * VX).sum()), (f(Z, X + EPS * VX) - f(Z, X - EPS * VX)) / (2 * EPS))


def test_higher_order_kernel_trains():
    """order = num_levels (what train_gpsig_.py:62 / train_gpsig_vosf.py:102 use) through the kernel classes under
    autograd: gradient of a scalar of K_tens_n_seq_covs w.r.t. the lengthscales against central differences of the same
    CUDA forward."""
    import torch
    from gpsig_b200 import kernels
    from tests.util import brownian, inducing_tensors
    N, L, d, m, Mz = 6, 12, 3, 3, 4
    X = brownian(N, L, d, 0)
    Xt = torch.as_tensor(X.reshape(N, -1), dtype=torch.float64, device='cuda')
    Z = torch.as_tensor(inducing_tensors(X, m, Mz, True, 5), dtype=torch.float64, device='cuda')

    def loss(ls):
        k = kernels.SignatureRBF(L * d, d, m, order=m)
        k.lengthscales = ls
        Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Z, Xt, increments=True)
        return (Kzz * Kzz).sum() + Kzx.sum() + (Kxx * Kxx).sum()

    ls = torch.nn.Parameter(torch.tensor([1.1, 0.9, 1.3], dtype=torch.float64, device='cuda'))
    loss(ls).backward()
    g = ls.grad.clone()
    for i in range(d):
        e = torch.zeros(d, dtype=torch.float64, device='cuda'); e[i] = 1e-6
        with torch.no_grad():
            fd = (loss(ls.detach() + e) - loss(ls.detach() - e)) / 2e-6
        assert abs(float(g[i]) - float(fd)) <= 1e-5 * max(abs(float(fd)), 1e-8), (i, float(g[i]), float(fd))
