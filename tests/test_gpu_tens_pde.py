"""GPU parity: inducing-tensor covariances and the PDE wavefront kernel vs the CPU oracle."""
import numpy as np
import pytest
import torch

from tests.util import brownian, inducing_tensors, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}


def _dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("increments", [False, True])
@pytest.mark.parametrize("base,cls", [('linear', 'SignatureLinear'), ('rbf', 'SignatureRBF'), ('matern52', 'SignatureMatern52')])
@pytest.mark.parametrize("normalization", [False, True])
def test_tens_n_seq_covs(dtype, increments, base, cls, normalization):
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, L, d, m, Mz = 37, 21, 3, 4, 13
    X = brownian(N, L, d, 0)
    Z = inducing_tensors(X, m, Mz, increments, 2)
    ls = np.array([1.1, 0.8, 1.4])
    k = getattr(kernels, cls)(L * d, d, m, lengthscales=ls, normalization=normalization)
    k.sigma = 0.6
    o = SignatureKernelOracle(L * d, d, m, base=base, lengthscales=ls, normalization=normalization, sigma=0.6)
    for ret in (True, False):
        got = k.K_tens_n_seq_covs(_dev(Z, dtype), _dev(X.reshape(N, -1), dtype), return_levels=ret, increments=increments)
        ref = o.K_tens_n_seq_covs(Z, X.reshape(N, -1), return_levels=ret, increments=increments)
        for g, r, nm in zip(got, ref, ('Kzz', 'Kzx', 'Kxx')):
            assert tuple(g.shape) == r.shape, nm
            assert rel_err(g.cpu().numpy(), r) < TOL[dtype], (nm, ret, rel_err(g.cpu().numpy(), r))
    got = k.K_tens_n_seq_covs(_dev(Z, dtype), _dev(X.reshape(N, -1), dtype), full_X_cov=True, increments=increments)
    ref = o.K_tens_n_seq_covs(Z, X.reshape(N, -1), full_X_cov=True, increments=increments)
    for g, r in zip(got, ref):
        assert rel_err(g.cpu().numpy(), r) < TOL[dtype]
    g = k.K_tens(_dev(Z, dtype), increments=increments).cpu().numpy()
    assert rel_err(g, o.K_tens(Z, increments=increments)) < TOL[dtype]
    g = k.K_tens_vs_seq(_dev(Z, dtype), _dev(X.reshape(N, -1), dtype), increments=increments).cpu().numpy()
    assert rel_err(g, o.K_tens_vs_seq(Z, X.reshape(N, -1), increments=increments)) < TOL[dtype]


@pytest.mark.parametrize("m,d,L", [(1, 2, 5), (2, 8, 100), (5, 4, 33), (6, 3, 12), (8, 2, 20)])
def test_tens_vs_seq_levels_and_column_blocks(m, d, L):
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, Mz = 70, 11
    X = brownian(N, L, d, 3)
    Z = inducing_tensors(X, m, Mz, True, 4)
    Xd, Zd = _dev(X), _dev(Z)
    G = ops.tens_vs_seq(Zd, Xd, m, increments=True, base='rbf')
    R = c_oracle.tens_vs_seq(Z, X, m, True, 'rbf')
    assert np.max(np.abs(G.cpu().numpy() - R) / np.abs(R).max(axis=(1, 2), keepdims=True)) < 1e-10
    parts = [ops.tens_vs_seq(Zd, Xd, m, increments=True, base='rbf', cols=c) for c in ((0, 33), (33, 70))]
    assert torch.equal(torch.cat(parts, dim=2), G)


def test_seq_n_seq_covs_follows_reference():
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, N2, L, d, m = 9, 14, 15, 2, 3
    X, Y = brownian(N, L, d, 5), brownian(N2, L, d, 6)
    for normalization in (True, False):
        k = kernels.SignatureRBF(L * d, d, m, normalization=normalization)
        o = SignatureKernelOracle(L * d, d, m, base='rbf', normalization=normalization)
        got = k.K_seq_n_seq_covs(_dev(X.reshape(N, -1)), _dev(Y.reshape(N2, -1)), return_levels=True)
        ref = o.K_seq_n_seq_covs(X.reshape(N, -1), Y.reshape(N2, -1), return_levels=True)
        for g, r in zip(got, ref):
            assert rel_err(g.cpu().numpy(), r) < 1e-10
    with pytest.raises(NameError):   # the reference's own behaviour, kernels.py:756
        kernels.SignatureRBF(L * d, d, m).K_seq_n_seq_covs(_dev(X.reshape(N, -1)), _dev(Y.reshape(N2, -1)), full_X2_cov=True)


@pytest.mark.parametrize("L,d,order", [(2, 1, 0), (20, 3, 0), (20, 3, 2), (70, 4, 1), (33, 8, 3), (300, 2, 0), (90, 5, 2), (9, 3, 5)])
def test_pde_diag_vs_oracle(L, d, order):
    from gpsig_b200 import ops
    from oracle import pde_np
    N = 6
    X = brownian(N, L, d, 7)
    ls = np.linspace(0.8, 1.2, d)
    G = ops.pde(_dev(X), None, dyadic_order=order, lengthscales=_dev(ls), sigma=1.3).cpu().numpy()
    R = 1.3 * pde_np.pde_kernel(X / ls, None, order)
    assert np.max(np.abs(G - R) / np.abs(R)) < 1e-10, (np.max(np.abs(G - R) / np.abs(R)))


@pytest.mark.parametrize("N,L,d,order", [(40, 200, 4, 2), (500, 300, 3, 1), (7, 130, 8, 3), (1200, 140, 2, 2)])
def test_pde_diag_lower_triangle_schedule(N, L, d, order, monkeypatch):
    """GPSIG_PDE_SYM=1: the strips of a diagonal pair solve only from their diagonal block down, as sigKer_fast.pyx:33-61
    fills only the lower triangle (opt-in: measured slower).  Same values as the full-grid schedule, with
    few pairs (one CTA per pair) and with more pairs than resident CTAs (persistent CTAs, buffers reused across pairs)."""
    from gpsig_b200 import ops
    X = _dev(brownian(N, L, d, 11))
    full = ops.pde(X, None, dyadic_order=order)
    monkeypatch.setenv('GPSIG_PDE_SYM', '1')
    tri = ops.pde(X, None, dyadic_order=order)
    assert float(((tri - full).abs() / full.abs()).max()) < 1e-11


def test_pde_golden_from_reference_cython():
    """Golden vectors produced by the reference's own compiled sigKer_fast (tests/golden/make_golden.py)."""
    import os
    from gpsig_b200 import kernels_pde
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'pde_ref.npz'))
    for order in (0, 1, 2):
        k = kernels_pde.UntruncSignatureKernel(g['X'].shape[1] * g['X'].shape[2], g['X'].shape[2], order=order,
                                               implementation='cython')
        got = k.Kdiag(_dev(g['X'].reshape(g['X'].shape[0], -1))).cpu().numpy()
        ref = g['kdiag_order%d' % order]
        assert np.max(np.abs(got - ref) / np.abs(ref)) < 1e-10


def test_pde_gram_rectangular_and_rows():
    from gpsig_b200 import kernels_pde, ops
    from oracle import pde_np
    N, N2, L, L2, d = 5, 7, 30, 45, 3
    X, Y = brownian(N, L, d, 8), brownian(N2, L2, d, 9)
    R = pde_np.pde_gram(X, Y, 1)
    G = ops.pde(_dev(X), _dev(Y), dyadic_order=1, gram=True)
    assert np.max(np.abs(G.cpu().numpy() - R) / np.abs(R)) < 1e-10
    parts = [ops.pde(_dev(X), _dev(Y), dyadic_order=1, gram=True, rows=r) for r in ((0, 2), (2, 5))]
    assert torch.equal(torch.cat(parts, dim=0), G)
    k = kernels_pde.SignatureLinear(L * d, d, 1, None)
    Ks = k.K(_dev(X.reshape(N, -1)))
    assert torch.allclose(Ks, Ks.T, rtol=1e-12, atol=0)
    assert torch.allclose(torch.diagonal(Ks), k.Kdiag(_dev(X.reshape(N, -1))), rtol=1e-12, atol=0)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("num_lags", [1, 2])
def test_lags_follow_reference(dtype, num_lags):
    """num_lags > 0: lagged copies by linear interpolation (gpsig/lags.py), scaled by gamma (kernels.py:343-399)."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, L, d, m, Mz = 9, 40, 3, 3, 5
    X = brownian(N, L, d, 11)
    ls = np.array([1.2, 0.7, 0.9])
    k = kernels.SignatureRBF(L * d, d, m, lengthscales=ls, num_lags=num_lags)
    o = SignatureKernelOracle(L * d, d, m, base='rbf', lengthscales=ls, num_lags=num_lags)
    Xd = _dev(X.reshape(N, -1), dtype)
    assert rel_err(k.K(Xd, return_levels=True).cpu().numpy(), o.K(X.reshape(N, -1), return_levels=True)) < TOL[dtype]
    Z = np.random.default_rng(12).standard_normal((m * (m + 1) // 2, Mz, 2, d * (num_lags + 1))) * 0.3
    got = k.K_tens_n_seq_covs(_dev(Z, dtype), Xd, increments=True)
    ref = o.K_tens_n_seq_covs(Z, X.reshape(N, -1), increments=True)
    for g, r in zip(got, ref):
        assert rel_err(g.cpu().numpy(), r) < TOL[dtype]


@pytest.mark.parametrize("num_lags", [1, 2])
def test_pde_lags_follow_reference(num_lags):
    """kernels_pde.py:135-157: lagged copies + lengthscales + gamma before the solve (oracle: trunc_np's restatement of
    gpsig/lags.py applied in the same order, then the restated solver)."""
    from gpsig_b200 import kernels_pde
    from oracle import pde_np, trunc_np
    N, L, d, order = 5, 30, 2, 1
    X = brownian(N, L, d, 13)
    ls = np.array([1.3, 0.8])
    k = kernels_pde.UntruncSignatureKernel(L * d, d, lengthscales=ls, order=order, num_lags=num_lags, implementation='cython')
    got = k.Kdiag(_dev(X.reshape(N, -1))).cpu().numpy()
    Xl = trunc_np.add_lags_to_sequences(X, k.lags) / ls[None, None, None, :] * k.gamma[None, None, :, None]
    ref = pde_np.pde_kernel(Xl.reshape(N, L, -1), None, order)
    assert np.max(np.abs(got - ref) / np.abs(ref)) < 1e-10
    # and under autograd (the reference's backward chained through the interpolation)
    Xt = _dev(X.reshape(N, -1)).requires_grad_(True)
    out = k.Kdiag(Xt)
    out.sum().backward()
    assert np.max(np.abs(out.detach().cpu().numpy() - ref) / np.abs(ref)) < 1e-10
    gs = pde_np.pde_kdiag_grad(Xl.reshape(N, L, -1), order, 0)                     # d / d(lagged, scaled coordinates)
    eps = 1e-6                                                                     # chain through the linear map numerically
    Xp = X.copy(); Xp[2, 7, 1] += eps
    Xlp = trunc_np.add_lags_to_sequences(Xp, k.lags) / ls[None, None, None, :] * k.gamma[None, None, :, None]
    want = float(np.sum(gs * (Xlp.reshape(N, L, -1) - Xl.reshape(N, L, -1)) / eps))
    assert abs(float(Xt.grad.reshape(N, L, d)[2, 7, 1]) - want) < 1e-6 * max(1.0, abs(want))
