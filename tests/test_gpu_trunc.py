"""GPU parity: fused truncated-Gram CUDA kernels (through the C ABI) vs the CPU oracle.

Tolerances are the north star's: 1e-10 relative (to the Gram's scale) in
float64, 1e-5 in float32.
"""
import numpy as np
import pytest
import torch

from tests.util import brownian, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}


def _dev(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


BASES = [('linear', 'SignatureLinear', {}), ('rbf', 'SignatureRBF', {}), ('cosine', 'SignatureCosine', {}),
         ('poly', 'SignaturePoly', dict(gamma=1.0, degree=3.0)), ('mix', 'SignatureMix', {}),
         ('matern12', 'SignatureMatern12', {}), ('matern32', 'SignatureMatern32', {}),
         ('matern52', 'SignatureMatern52', {})]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base,cls,bp", BASES)
@pytest.mark.parametrize("normalization", [False, True])
def test_gram_all_base_kernels(dtype, base, cls, bp, normalization):
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, N2, L, L2, d, m = 11, 9, 23, 17, 3, 4
    X, Y = brownian(N, L, d, 0), brownian(N2, L2, d, 1)
    ls = np.array([0.9, 1.3, 0.7])
    k = getattr(kernels, cls)(L * d, d, m, lengthscales=ls, normalization=normalization, **bp)
    k.sigma = 1.7
    k.variances = np.linspace(0.5, 1.5, m + 1)
    orc_kw = dict(bp)
    if base == 'mix':
        orc_kw = dict(mixing=0.5)
    o = SignatureKernelOracle(L * d, d, m, base=base, lengthscales=ls, normalization=normalization, sigma=1.7,
                              variances=np.linspace(0.5, 1.5, m + 1), **orc_kw)
    tol = TOL[dtype]
    if base == 'matern12':
        # exp(-r) is not differentiable at r = 0: on coincident points (the diagonal blocks of K(X, X))
        # r = sqrt(max(|x|^2 + |y|^2 - 2 x.y, 1e-40)) is sqrt(rounding noise) ~ sqrt(eps), so ANY two
        # evaluation orders (TF's matmul vs FMA chains) differ by ~sqrt(eps) there: the reference itself
        # is only defined to that accuracy (kernels.py:959-961, 1135-1138).
        # The normalised K(X, X2) inherits it through the diagonals k(x_i, x_i).
        tol = {torch.float64: 1e-6, torch.float32: 1e-2}[dtype]
    for ret in (True, False):
        G = k.K(_dev(X.reshape(N, -1), dtype), _dev(Y.reshape(N2, -1), dtype), return_levels=ret).cpu().numpy()
        R = o.K(X.reshape(N, -1), Y.reshape(N2, -1), return_levels=ret)
        assert G.shape == R.shape
        assert rel_err(G, R) < tol, (base, ret, rel_err(G, R))
    Gs = k.K(_dev(X.reshape(N, -1), dtype), return_levels=True).cpu().numpy()
    Rs = o.K(X.reshape(N, -1), return_levels=True)
    assert rel_err(Gs, Rs) < tol, (base, rel_err(Gs, Rs))


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("L,L2,d,m", [(2, 2, 1, 1), (3, 40, 2, 2), (50, 50, 3, 4), (33, 34, 5, 3), (65, 97, 8, 5),
                                      (130, 161, 4, 6), (200, 200, 8, 5), (225, 256, 7, 8), (257, 300, 2, 4),
                                      (100, 520, 9, 5), (40, 30, 16, 4)])
def test_gram_shapes_raw_levels(dtype, L, L2, d, m):
    """Every columns-per-lane shape, the column-chunked path (L2 > 256), all level counts, d padding."""
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, N2 = 5, 10
    X, Y = brownian(N, L, d, 2), brownian(N2, L2, d, 3)
    for base in ('linear', 'rbf'):
        G = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), m, base=base).cpu().numpy()
        R = c_oracle.trunc_gram(X, Y, m, base)
        scale = np.abs(R).max(axis=(1, 2), keepdims=True)
        err = np.max(np.abs(G - R) / scale)
        assert err < TOL[dtype], (base, L, L2, d, m, err)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("L,L2", [(200, 200), (200, 217), (200, 218), (200, 224), (400, 200), (400, 224), (376, 31)])
def test_systolic_schedules(dtype, L, L2):
    """The launch-time variants of the systolic kernel (trunc_skew_kernel.cuh): lane 0 as a dummy lane (L2 <= 31 x
    columns per lane) or as a working lane, the 16384-entry exp2 table (fits next to the x ring up to L = 375 at d = 8)
    or the 2048-entry one; also with points far apart (the exponent clamp of the table exp2)."""
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, N2, d, m = 6, 9, 8, 5
    for scale in ((1.0, 10.0) if dtype == torch.float64 else (1.0,)):      # float32 cannot hold a 1e-5 exponent at that range
        X, Y = scale * brownian(N, L, d, 12), scale * brownian(N2, L2, d, 13)
        for base in ('rbf', 'linear'):
            G = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), m, base=base).cpu().numpy()
            R = c_oracle.trunc_gram(X, Y, m, base)
            ref = np.abs(R).max(axis=(1, 2), keepdims=True)
            err = np.max(np.abs(G - R) / ref)
            # scaled paths: |x|^2 ~ 1e3..1e4, so the exponent x.y - |x|^2/2 - |y|^2/2 carries ~1e-12 of rounding in ANY
            # float64 evaluation order (the oracle's included): 1e-9 there
            assert np.isfinite(G).all() and err < (TOL[dtype] if scale == 1.0 else 1e-9), (base, scale, L, L2, err)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("L,L2,d", [(1000, 200, 8), (2500, 64, 8), (1500, 256, 16)])
def test_rows_longer_than_the_systolic_ring(dtype, L, L2, d):
    """Row sequences too long for the three-slot shared-memory ring of the systolic kernel (about 800 points at d = 8
    in float64) are taken by the general kernel, which streams them: same result, no 'sequence too long' error."""
    from gpsig_b200 import ops
    from oracle import c_oracle
    X, Y = brownian(3, L, d, 21), brownian(4, L2, d, 22)
    for base in ('rbf', 'linear'):
        G = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), 4, base=base).cpu().numpy()
        R = c_oracle.trunc_gram(X, Y, 4, base)
        ref = np.abs(R).max(axis=(1, 2), keepdims=True)
        # float32 sums 2500 x 64 increments per level: 5e-5 at that length (rounding of the float32 recursion, not the kernel choice)
        tol = TOL[dtype] if (dtype == torch.float64 or L < 2000) else 5e-5
        assert np.max(np.abs(G - R) / ref) < tol, (base, L, L2, d)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_diag_matches_gram_diagonal_and_oracle(dtype):
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, L, d, m = 19, 45, 3, 5
    X = brownian(N, L, d, 4)
    Xd = _dev(X, dtype)
    for base in ('linear', 'rbf', 'matern32'):
        D = ops.trunc_diag(Xd, m, base=base)
        G = ops.trunc_gram(Xd, None, m, base=base)
        # the diagonal comes from trunc_gram_kernel, the Gram possibly from the systolic kernel (different summation order)
        assert torch.allclose(D, torch.diagonal(G, dim1=1, dim2=2), rtol=10 * TOL[dtype], atol=0)
        R = c_oracle.trunc_diag(X, m, base)
        assert np.max(np.abs(D.cpu().numpy() - R) / np.abs(R).max(axis=1, keepdims=True)) < TOL[dtype]


def test_no_difference_and_row_blocks():
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, N2, L, L2, d, m = 13, 7, 20, 31, 3, 3
    X, Y = brownian(N, L, d, 5) * 0.3, brownian(N2, L2, d, 6) * 0.3
    Xd, Yd = _dev(X, torch.float64), _dev(Y, torch.float64)
    R = c_oracle.trunc_gram(X, Y, m, 'rbf', difference=False)
    G = ops.trunc_gram(Xd, Yd, m, base='rbf', difference=False).cpu().numpy()
    assert np.max(np.abs(G - R) / np.abs(R).max(axis=(1, 2), keepdims=True)) < 1e-10
    # row-block sharding: blocks stitched together == the full matrix, bit for bit
    full = ops.trunc_gram(Xd, Yd, m, base='rbf')
    parts = [ops.trunc_gram(Xd, Yd, m, base='rbf', rows=(a, b)) for a, b in ((0, 4), (4, 4), (4, 13))]
    assert torch.equal(torch.cat(parts, dim=1), full)


def test_symmetric_gram_is_symmetric_and_unit_diagonal():
    from gpsig_b200 import kernels
    N, L, d, m = 40, 30, 4, 4
    X = _dev(brownian(N, L, d, 7).reshape(N, -1), torch.float64)
    k = kernels.SignatureRBF(L * d, d, m)
    K = k.K(X, return_levels=True)
    assert torch.allclose(K, K.transpose(1, 2), rtol=0, atol=1e-12)
    assert torch.allclose(torch.diagonal(K, dim1=1, dim2=2), torch.ones(m + 1, N, dtype=K.dtype, device='cuda'), atol=1e-12)
    assert torch.equal(k.Kdiag(X), torch.full((N,), float(m + 1), dtype=torch.float64, device='cuda'))


def test_known_answer_level1_telescopes():
    """Level 1 of the differenced kernel telescopes to 4 corner values of the base kernel."""
    from gpsig_b200 import ops
    X, Y = brownian(6, 200, 8, 8), brownian(4, 200, 8, 9)
    G = ops.trunc_gram(_dev(X, torch.float64), _dev(Y, torch.float64), 2, base='linear').cpu().numpy()
    ref = np.einsum('id,jd->ij', X[:, -1] - X[:, 0], Y[:, -1] - Y[:, 0])
    assert np.allclose(G[0], 1.0) and np.max(np.abs(G[1] - ref)) < 1e-11


def test_errors():
    from gpsig_b200 import kernels, ops
    with pytest.raises(ValueError):
        kernels.SignatureLinear(10, 3, 2)
    with pytest.raises(NotImplementedError):
        kernels.SignatureLinear(12, 3, 2, order=2, low_rank=True)
    with pytest.raises(ValueError):
        kernels.SignatureLinear(12, 3, 2, num_lags=-1)
    with pytest.raises(ValueError):
        kernels.SignatureLinear(12, 3, 2, variances=np.ones(5))
    X = torch.zeros(3, 1, 2, dtype=torch.float64, device='cuda')
    with pytest.raises(ValueError):
        ops.trunc_gram(X, None, 2)           # L=1 cannot be differenced
    with pytest.raises(RuntimeError):
        ops.trunc_gram(X.cpu(), None, 2)     # no CPU path
    with pytest.raises(NotImplementedError):
        ops.trunc_gram(torch.zeros(3, 4, 2, dtype=torch.float64, device='cuda'), None, 9)
    with pytest.raises(ValueError):
        ops.trunc_gram(torch.zeros(3, 4, 2, dtype=torch.float64, device='cuda'), None, 3, order=4)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("family", ['rbf', 'exp'])
def test_spectral_base_kernel(dtype, family):
    """SignatureSpectral (kernels.py:1073-1122): K, Kdiag, inducing-tensor covariances vs the oracle."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    from tests.util import inducing_tensors
    N, N2, L, L2, d, m, Mz, Q = 7, 5, 17, 12, 3, 3, 4, 4
    X, Y = brownian(N, L, d, 31), brownian(N2, L2, d, 32)
    rng = np.random.default_rng(33)
    alpha, omega, gamma = np.exp(rng.standard_normal(Q)), np.exp(0.3 * rng.standard_normal((Q, d))), np.exp(rng.standard_normal((Q, d)))
    tol = 1e-10 if dtype == torch.float64 else 2e-5
    for normalization in (False, True):
        k = kernels.SignatureSpectral(L * d, d, m, family=family, Q=Q, normalization=normalization)
        assert k.lengthscales is None
        k.alpha, k.omega, k.gamma = alpha, omega, gamma
        o = SignatureKernelOracle(L * d, d, m, base='spectral', lengthscales=None, normalization=normalization,
                                  family=family, alpha=alpha, omega=omega, gamma=gamma)
        Xd, Yd = _dev(X.reshape(N, -1), dtype), _dev(Y.reshape(N2, -1), dtype)
        k2 = kernels.SignatureSpectral(L2 * d, d, m, family=family, Q=Q, normalization=normalization)
        assert rel_err(k.K(Xd, return_levels=True).cpu().numpy(), o.K(X.reshape(N, -1), return_levels=True)) < tol
        assert rel_err(k.Kdiag(Xd, return_levels=True).cpu().numpy(), o.Kdiag(X.reshape(N, -1), return_levels=True)) < tol
        for inc in (False, True):
            Z = inducing_tensors(X, m, Mz, inc, 34)
            got = k.K_tens_n_seq_covs(_dev(Z, dtype), Xd, increments=inc, return_levels=True)
            ref = o.K_tens_n_seq_covs(Z, X.reshape(N, -1), increments=inc, return_levels=True)
            for g, r in zip(got, ref):
                assert rel_err(g.cpu().numpy(), r) < tol
        del k2
    # rectangular K(X, X2) needs equal L in the GPflow layout only per argument: raw op level
    from gpsig_b200 import ops
    blk = torch.as_tensor(np.concatenate([alpha, omega.ravel(), gamma.ravel()]), dtype=torch.float64, device='cuda')
    got = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), m, base='spectral', base_params=(family, Q, blk))
    o = SignatureKernelOracle(L * d, d, m, base='spectral', lengthscales=None, normalization=False,
                              family=family, alpha=alpha, omega=omega, gamma=gamma)
    assert rel_err(got.cpu().numpy(), o._K_seq(X, Y)) < tol
    with pytest.raises(NameError):
        kernels.SignatureSpectral(L * d, d, m, family='mixed').K(_dev(X.reshape(N, -1), dtype))
    with pytest.raises(ValueError):
        kernels.SignatureSpectral(L * d, d, m, family='bogus')


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("N,L", [(64, 70), (131, 80), (300, 66)])
def test_symmetric_gram_triangular_mode(dtype, N, L):
    """K(X, X) on the systolic kernel computes the CTAs touching i <= j and mirrors the rest: exactly symmetric,
    equal (to tolerance) to K(X, X2 = X) and to the oracle on a sub-block, for levels and for the sum."""
    from gpsig_b200 import kernels, ops
    from oracle.trunc_np import SignatureKernelOracle
    d, m = 3, 4
    Xn = brownian(N, L, d, 70)
    X = _dev(Xn.reshape(N, -1), dtype)
    k = kernels.SignatureRBF(L * d, d, m)
    for ret in (True, False):
        Ks, Kf = k.K(X, return_levels=ret), k.K(X, X.clone(), return_levels=ret)
        assert torch.equal(Ks, Ks.transpose(-1, -2))
        # the jitter on the diagonal exists only in the symmetric call (kernels.py:432)
        off = 1.0 - torch.eye(N, dtype=dtype, device='cuda')
        scale = float(Kf.abs().max())
        assert float(((Ks - Kf) * off).abs().max()) / scale < 10 * TOL[dtype]
    raw = ops.trunc_gram(X.reshape(N, L, d), None, m, base='rbf')
    assert torch.equal(raw, raw.transpose(1, 2))
    o = SignatureKernelOracle(L * d, d, m, base='rbf', normalization=False)
    sub = np.r_[0:5, N - 6:N]
    ref = o._K_seq(Xn[sub], Xn[sub])
    got = raw[:, sub][:, :, sub].cpu().numpy()
    assert rel_err(got, ref) < TOL[dtype]
    # row-sharded symmetric calls still work and agree to tolerance with the mirrored full matrix
    parts = torch.cat([ops.trunc_gram(X.reshape(N, L, d), None, m, base='rbf', rows=r) for r in ((0, N // 3), (N // 3, N))], dim=1)
    assert float((parts - raw).abs().max()) / float(raw.abs().max()) < TOL[dtype]
