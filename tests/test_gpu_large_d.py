"""GPU parity at state dimensions d > 16 (the reference's benchmark shapes: num_lags=1 with a time channel give
d_eff = 26, 28, 46, 126, ...; benchmarks/datasets.json, run_gpsig_benchmarks.py:32, train_gpsig.py:28-29).

K / Kdiag run the lifted path (lifted_kernels.cuh: GEMM-shaped lift + recursion on the lifted tiles; float32 lift on
tcgen05 tensor cores), Kzx / Kzz the general kernels.  Oracle: oracle/trunc_np.py (numpy restatement of
gpsig/signature_algs.py + kernels.py glue) and the C port.  Tolerances: 1e-10 float64, 1e-5 float32.
"""
import numpy as np
import pytest
import torch

from tests.util import brownian, inducing_tensors, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}


def _dev(a, dtype):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("d,L,L2,m", [(28, 40, 33, 4), (46, 70, 100, 5), (126, 50, 50, 4), (17, 130, 200, 3), (64, 9, 257, 2),
                                      (20, 300, 30, 6), (24, 210, 205, 3)])
def test_gram_raw_levels_large_d(dtype, d, L, L2, m):
    from gpsig_b200 import ops
    from oracle import c_oracle
    N, N2 = 7, 5
    X, Y = brownian(N, L, d, 2), brownian(N2, L2, d, 3)
    for base in ('linear', 'rbf'):
        G = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), m, base=base).cpu().numpy()
        R = c_oracle.trunc_gram(X, Y, m, base)
        scale = np.abs(R).max(axis=(1, 2), keepdims=True)
        err = np.max(np.abs(G - R) / scale)
        assert err < TOL[dtype], (base, d, L, L2, m, err)
    # rows = a range of a larger Gram (sharding), and the no-difference mode
    G = ops.trunc_gram(_dev(X, dtype), _dev(Y, dtype), m, base='rbf', rows=(2, 6)).cpu().numpy()
    R = c_oracle.trunc_gram(X, Y, m, 'rbf')[:, 2:6]
    assert np.max(np.abs(G - R) / np.abs(R).max(axis=(1, 2), keepdims=True)) < TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base,cls,bp", [('linear', 'SignatureLinear', {}), ('rbf', 'SignatureRBF', {}),
                                         ('matern32', 'SignatureMatern32', {}), ('poly', 'SignaturePoly', dict(gamma=1.0, degree=3.0))])
@pytest.mark.parametrize("d", [28, 126])
def test_kernel_api_large_d(dtype, base, cls, bp, d):
    """K (normalised, level-summed and per level), K(X, X), Kdiag through the reference-shaped classes."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, N2, L, L2, m = 9, 6, 31, 24, 4
    X, Y = brownian(N, L, d, 0) * 0.5, brownian(N2, L2, d, 1) * 0.5
    # lengthscales ~ sqrt(d): what the reference's initialisation gives (utils.py:90-100 suggest_initial_lengthscales: sqrt(mean squared
    # difference per feature x d)), and what keeps |x_p - y_q|^2 O(1) so that the base kernel is well conditioned in float32
    ls = np.linspace(0.8, 1.4, d) * np.sqrt(d / 8.0)
    var = np.linspace(0.5, 1.5, m + 1)
    k = getattr(kernels, cls)(L * d, d, m, lengthscales=ls, **bp)
    k.sigma = 1.3
    k.variances = var
    o = SignatureKernelOracle(L * d, d, m, base=base, lengthscales=ls, sigma=1.3, variances=var, **bp)
    tol = TOL[dtype]
    for ret in (True, False):
        G = k.K(_dev(X.reshape(N, -1), dtype), _dev(Y.reshape(N2, -1), dtype), return_levels=ret).cpu().numpy()
        R = o.K(X.reshape(N, -1), Y.reshape(N2, -1), return_levels=ret)
        assert rel_err(G, R) < tol, (base, d, ret, rel_err(G, R))
    Gs = k.K(_dev(X.reshape(N, -1), dtype)).cpu().numpy()
    assert rel_err(Gs, o.K(X.reshape(N, -1))) < tol
    assert np.array_equal(Gs, Gs.T)
    Gd = k.Kdiag(_dev(X.reshape(N, -1), dtype), return_levels=True).cpu().numpy()
    assert rel_err(Gd, o.Kdiag(X.reshape(N, -1), return_levels=True)) < tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("d", [28, 46])
@pytest.mark.parametrize("increments", [False, True])
def test_inducing_tensor_covariances_large_d(dtype, d, increments):
    """Kzz, Kzx, Kxx_diag of K_tens_n_seq_covs (the SVGP shape, kernels.py:623-704) at d > 16."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, L, m, Mz = 12, 25, 4, 7
    X = brownian(N, L, d, 0) * 0.5
    Z = inducing_tensors(X, m, Mz, increments, 5)
    k = kernels.SignatureRBF(L * d, d, m)
    o = SignatureKernelOracle(L * d, d, m, base='rbf')
    Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(_dev(Z, dtype), _dev(X.reshape(N, -1), dtype), increments=increments)
    Rzz, Rzx, Rxx = o.K_tens_n_seq_covs(Z, X.reshape(N, -1), increments=increments)
    tol = TOL[dtype] * (10 if dtype == torch.float32 else 1)     # differences of nearby exp values (kernels.py:331) in float32
    assert rel_err(Kzz.cpu().numpy(), Rzz) < tol
    assert rel_err(Kzx.cpu().numpy(), Rzx) < tol
    assert rel_err(Kxx.cpu().numpy(), Rxx) < TOL[dtype]


@pytest.mark.parametrize("cap_mb", [None, 3])
def test_large_d_many_pairs_blocks(cap_mb, monkeypatch):
    """Streams of 8 pairs with a ragged last stream; with the scratch capped at 3 MB the Gram takes many blocks."""
    from gpsig_b200 import ops
    from oracle import c_oracle
    if cap_mb:
        monkeypatch.setenv('GPSIG_LIFT_CAP_MB', str(cap_mb))
    N, N2, L, d, m = 150, 260, 20, 24, 3
    X, Y = brownian(N, L, d, 7), brownian(N2, L, d, 8)
    G = ops.trunc_gram(_dev(X, torch.float64), _dev(Y, torch.float64), m, base='rbf').cpu().numpy()
    R = c_oracle.trunc_gram(X, Y, m, 'rbf')
    assert np.max(np.abs(G - R) / np.abs(R).max(axis=(1, 2), keepdims=True)) < 1e-10


@pytest.mark.parametrize("d,L,L2", [(28, 40, 33), (126, 100, 100), (46, 150, 200)])
def test_tensor_core_lift_matches_cuda_core_lift(d, L, L2, monkeypatch):
    """float32: the tcgen05 (3 x TF32) lift and the register-tiled FFMA lift agree within the float32 tolerance, incl. L > 128
    (two accumulator row tiles per pair) and every columns-per-lane shape."""
    from gpsig_b200 import ops
    N, N2, m = 6, 5, 4
    X, Y = _dev(brownian(N, L, d, 2), torch.float32), _dev(brownian(N2, L2, d, 3), torch.float32)
    ls = _dev(np.full(d, np.sqrt(d / 8.0)), torch.float32)
    for base in ('linear', 'rbf', 'matern52'):
        A = ops.trunc_gram(X, Y, m, base=base, lengthscales=ls).cpu().numpy()
        monkeypatch.setenv('GPSIG_NO_MMA', '1')
        B = ops.trunc_gram(X, Y, m, base=base, lengthscales=ls).cpu().numpy()
        monkeypatch.delenv('GPSIG_NO_MMA')
        assert np.max(np.abs(A - B) / np.abs(B).max(axis=(1, 2), keepdims=True)) < 1e-5, base      # both are float32 evaluations: each is within 1e-5 of the oracle


@pytest.mark.parametrize("base", ['linear', 'rbf', 'matern32'])
@pytest.mark.parametrize("d", [20, 70])
def test_reverse_mode_large_d(base, d):
    """Adjoints at d > 16 (the feature dimension is taken in chunks): vector-Jacobian identity against central
    differences of the oracle's raw levels, as tests/test_gpu_vjp.py does at d <= 16.  Gram pairs, diagonal pairs,
    tensors vs sequences (with increments) and tensors vs tensors -- what K_tens_n_seq_covs differentiates."""
    from gpsig_b200 import ops
    from oracle.trunc_np import SignatureKernelOracle
    EPS, TOLV = 1e-5, 2e-6
    N, N2, L, L2, m, Mz = 4, 3, 8, 7, 3, 4
    rng = np.random.default_rng(5)
    sc = np.sqrt(8.0 / d)
    X, Y = (brownian(N, L, d, 0) + 0.3) * sc, (brownian(N2, L2, d, 1) - 0.2) * sc
    o = SignatureKernelOracle(L * d, d, m, base=base, lengthscales=None, normalization=False)
    dev = lambda a: _dev(a, torch.float64)  # noqa: E731

    def check(nm, a, f):
        assert abs(a - f) <= TOLV * max(abs(f), abs(a), 1e-12) + 1e-9, (nm, a, f)

    G = rng.standard_normal((m + 1, N, N2))
    VX, VY = rng.standard_normal(X.shape), rng.standard_normal(Y.shape)
    gX, gY = ops.trunc_vjp(dev(X), dev(Y), dev(G), m, base=base)
    f = lambda A, B: float(np.sum(G * o._K_seq(A, B)))  # noqa: E731
    check('X', float((gX.cpu().numpy() * VX).sum()), (f(X + EPS * VX, Y) - f(X - EPS * VX, Y)) / (2 * EPS))
    check('X2', float((gY.cpu().numpy() * VY).sum()), (f(X, Y + EPS * VY) - f(X, Y - EPS * VY)) / (2 * EPS))
    Gd = rng.standard_normal((m + 1, N))
    gX, _ = ops.trunc_vjp(dev(X), None, dev(Gd), m, base=base, diag=True)
    f = lambda A: float(np.sum(Gd * o._K_seq_diag(A)))  # noqa: E731
    check('diag', float((gX.cpu().numpy() * VX).sum()), (f(X + EPS * VX) - f(X - EPS * VX)) / (2 * EPS))
    for increments in (False, True):
        Z = inducing_tensors(X, m, Mz, increments, 2)
        VZ = rng.standard_normal(Z.shape)
        Gz = rng.standard_normal((m + 1, Mz, N))
        gZ, gX = ops.tens_vs_seq_vjp(dev(Z), dev(X), dev(Gz), m, increments=increments, base=base)
        f = lambda A, B: float(np.sum(Gz * o._K_tens_vs_seq(A, B, increments=increments)))  # noqa: E731
        check('Z', float((gZ.cpu().numpy() * VZ).sum()), (f(Z + EPS * VZ, X) - f(Z - EPS * VZ, X)) / (2 * EPS))
        check('Xz', float((gX.cpu().numpy() * VX).sum()), (f(Z, X + EPS * VX) - f(Z, X - EPS * VX)) / (2 * EPS))
        G2 = rng.standard_normal((m + 1, Mz, Mz))
        gZ = ops.tens_tens_vjp(dev(Z), dev(G2), m, increments=increments, base=base)
        f = lambda A: float(np.sum(G2 * o._K_tens(A, increments=increments)))  # noqa: E731
        check('Zz', float((gZ.cpu().numpy() * VZ).sum()), (f(Z + EPS * VZ) - f(Z - EPS * VZ)) / (2 * EPS))


def test_svgp_covariances_differentiable_at_large_d():
    """K_tens_n_seq_covs under autograd at d = 28 (ArabicDigits with num_lags=1 and time): gradients reach the inducing
    tensors and the lengthscales (finite and non-zero)."""
    from gpsig_b200 import kernels
    N, L, d, m, Mz = 10, 20, 28, 4, 6
    X = brownian(N, L, d, 0) * 0.5
    Z = torch.nn.Parameter(_dev(inducing_tensors(X, m, Mz, True, 5), torch.float64))
    k = kernels.SignatureRBF(L * d, d, m)
    k.lengthscales = torch.nn.Parameter(torch.full((d,), 2.0, dtype=torch.float64, device='cuda'))
    Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Z, _dev(X.reshape(N, -1), torch.float64), increments=True)
    (Kzz.sum() + (Kzx ** 2).sum() + Kxx.sum()).backward()
    for t in (Z.grad, k.lengthscales.grad):
        assert t is not None and bool(torch.isfinite(t).all()) and float(t.abs().max()) > 0


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base", ['rbf', 'linear', 'matern32'])
@pytest.mark.parametrize("increments", [False, True])
def test_kzx_lifted_blocks_and_column_ranges(dtype, base, increments, monkeypatch):
    """Kzx at d > 16 on the lifted form (tens_lifted_kernels.cuh): several blocks of sequences (scratch capped so that a
    block holds 32 of them, the last one ragged), per-level output, a column sub-range."""
    from gpsig_b200 import ops
    from oracle.trunc_np import SignatureKernelOracle
    monkeypatch.setenv('GPSIG_TENS_LIFT_CAP_KB', '400')
    N, L, d, m, Mz = 75, 9, 20, 3, 7
    X = brownian(N, L, d, 31) / np.sqrt(d / 8.0)
    Z = inducing_tensors(X, m, Mz, increments, 32)
    o = SignatureKernelOracle(L * d, d, m, base=base, normalization=False, lengthscales=None)
    R = o._K_tens_vs_seq(Z, X, increments=increments)                      # [m+1, Mz, N]
    scale = np.abs(R).max(axis=(1, 2), keepdims=True)
    tol = 1e-10 if dtype == torch.float64 else 1e-5
    G = ops.tens_vs_seq(_dev(Z, dtype), _dev(X, dtype), m, increments=increments, base=base).cpu().numpy()
    assert np.max(np.abs(G - R) / scale) < tol
    Gc = ops.tens_vs_seq(_dev(Z, dtype), _dev(X, dtype), m, increments=increments, base=base, cols=(10, 61)).cpu().numpy()
    assert np.max(np.abs(Gc - R[:, :, 10:61]) / scale) < tol
    Gs = ops.tens_vs_seq(_dev(Z, dtype), _dev(X, dtype), m, increments=increments, base=base, return_levels=False).cpu().numpy()
    assert np.max(np.abs(Gs - R.sum(axis=0))) / np.abs(R.sum(axis=0)).max() < tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base", ['rbf', 'linear'])
@pytest.mark.parametrize("increments,difference", [(True, True), (False, True), (True, False)])
def test_kzx_lifted_adjoint_matches_thread_per_pair_adjoint(dtype, base, increments, difference, monkeypatch):
    """Reverse mode of Kzx at d > 16: the lifted form (library: GEMM + reverse time recursion -> dL/dkappa; two plain GEMMs
    for dL/dZ, dL/dX) against the thread-per-pair adjoint kernel (GPSIG_TENS_VJP_NO_LIFT=1), element by element, with
    several blocks of sequences."""
    from gpsig_b200 import ops
    N, L, d, m, Mz = 70, 9, 20, 3, 5
    rng = np.random.default_rng(41)
    X = brownian(N, L, d, 42) / np.sqrt(d / 8.0)
    Z = inducing_tensors(X, m, Mz, increments, 43)
    G = rng.standard_normal((m + 1, Mz, N))
    args = (_dev(Z, dtype), _dev(X, dtype), _dev(G, dtype), m)
    kw = dict(increments=increments, base=base, difference=difference)
    monkeypatch.setattr(ops, '_LIFT_GD_BYTES', (m * (m + 1) // 2) * (2 if increments else 1) * Mz * L * 8 * 32)      # 32 sequences per block
    gZ, gX = ops.tens_vs_seq_vjp(*args, **kw)
    monkeypatch.setenv('GPSIG_TENS_VJP_NO_LIFT', '1')
    rZ, rX = ops.tens_vs_seq_vjp(*args, **kw)
    tol = 1e-9 if dtype == torch.float64 else 2e-4
    assert rel_err(gZ.cpu().numpy(), rZ.cpu().numpy()) < tol
    assert rel_err(gX.cpu().numpy(), rX.cpu().numpy()) < tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("base", ['rbf', 'linear', 'matern32'])
@pytest.mark.parametrize("d,Mz,increments", [(3, 37, True), (8, 50, True), (8, 16, False), (13, 33, False), (28, 45, True),
                                             (70, 21, True), (126, 19, False), (200, 18, True)])
def test_kzz_tiles_match_the_first_kernels(dtype, base, d, Mz, increments, monkeypatch):
    """Kzz and its adjoint on 16 x 16 tiles of tensor pairs (tens_tens_tile_kernels.cuh: any d, ragged tiles; the adjoint up to
    d = 128, above that the first kernel) against the first kernels (GPSIG_TENS_TENS_OLD=1): levels, summed with level scales
    and lengthscales, and dL/dZ."""
    from gpsig_b200 import ops
    m = 3
    rng = np.random.default_rng(d * 100 + Mz)
    X = brownian(6, 12, d, 1) / np.sqrt(d / 8.0)
    # points (not increments) shrunk: with far-apart points dL/dZ is the rounding noise of the cancelling diagonal terms
    Z = _dev(inducing_tensors(X, m, Mz, increments, 2) * (1.0 if increments else 0.3), dtype)
    len_ = _dev(rng.uniform(0.7, 1.4, d), dtype)
    lsc = _dev(rng.uniform(0.5, 1.5, m + 1), dtype)
    G = _dev(rng.standard_normal((m + 1, Mz, Mz)), dtype)

    def run():
        lev = ops.tens_tens(Z, m, increments=increments, base=base)
        tot = ops.tens_tens(Z, m, increments=increments, base=base, lengthscales=len_, level_scale=lsc, return_levels=False)
        gz = ops.tens_tens_vjp(Z, G, m, increments=increments, base=base)
        return [t.cpu().numpy() for t in (lev, tot, gz)]
    new = run()
    monkeypatch.setenv('GPSIG_TENS_TENS_OLD', '1')
    old = run()
    # adjoint: sums of cancelling terms in another order; float32: differences of nearby exp values (kernels.py:278)
    tols = (1e-11, 1e-11, 1e-8) if dtype == torch.float64 else (1e-4, 1e-4, 2e-3)
    for a, b, tol in zip(new, old, tols):
        assert np.isfinite(a).all()
        assert rel_err(a, b) < tol


@pytest.mark.parametrize("m,d", [(1, 5), (5, 8), (6, 20)])
def test_kzz_tiles_other_level_counts(m, d, monkeypatch):
    """The tiled Kzz kernels at m = 1 (no other component), 5 and 6 levels against the first kernels and the oracle."""
    from gpsig_b200 import ops
    from oracle import trunc_np
    Mz = 23
    X = brownian(5, 10, d, 3) / np.sqrt(d / 8.0)
    Zn = inducing_tensors(X, m, Mz, True, 4)
    Z = _dev(Zn, torch.float64)
    G = _dev(np.random.default_rng(5).standard_normal((m + 1, Mz, Mz)), torch.float64)
    lev = ops.tens_tens(Z, m, increments=True, base='rbf').cpu().numpy()
    gz = ops.tens_tens_vjp(Z, G, m, increments=True, base='rbf').cpu().numpy()
    monkeypatch.setenv('GPSIG_TENS_TENS_OLD', '1')
    assert rel_err(lev, ops.tens_tens(Z, m, increments=True, base='rbf').cpu().numpy()) < 1e-11
    assert rel_err(gz, ops.tens_tens_vjp(Z, G, m, increments=True, base='rbf').cpu().numpy()) < 1e-9
    o = trunc_np.SignatureKernelOracle(10 * d, d, m, base='rbf')
    assert rel_err(lev, o.K_tens(Zn, return_levels=True, increments=True)) < 1e-11
