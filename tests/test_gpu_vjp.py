"""GPU: reverse mode (csrc/vjp_kernels.cuh) against the CPU oracle.

The oracle has no backward pass (the reference relies on TensorFlow autodiff), so the check is the
vector-Jacobian identity  <G, J V> = <J^T G, V>:  J V from central differences of the ORACLE's raw levels
along random directions V (float64), J^T G from the CUDA adjoint.  Tolerance 1e-6 relative (the
central-difference truncation error at step 1e-5 dominates; the kernels themselves agree to 1e-10).
"""
import numpy as np
import pytest
import torch

from tests.util import brownian, inducing_tensors

pytestmark = pytest.mark.gpu
EPS = 1e-5
TOL = 2e-6


def _dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


def _oracle(base, m, L, d, difference=True):
    from oracle.trunc_np import SignatureKernelOracle
    kw = {'gamma': 0.7, 'degree': 3.0} if base == 'poly' else ({'mixing': 0.3} if base == 'mix' else {})
    return SignatureKernelOracle(L * d, d, m, base=base, lengthscales=None, normalization=False, difference=difference, **kw), \
        ((0.7, 3.0) if base == 'poly' else ((0.3,) if base == 'mix' else ()))


def _check(pairs):
    """pairs: list of (analytic <grad, V>, finite-difference <G, JV>)."""
    for nm, a, f in pairs:
        assert abs(a - f) <= TOL * max(abs(f), abs(a), 1e-12) + 1e-9, (nm, a, f)


@pytest.mark.parametrize("base", ['linear', 'rbf', 'cosine', 'poly', 'mix', 'matern12', 'matern32', 'matern52'])
@pytest.mark.parametrize("difference", [True, False])
def test_trunc_gram_vjp(base, difference):
    from gpsig_b200 import ops
    N, N2, L, L2, d, m = 5, 4, 9, 7, 3, 4
    rng = np.random.default_rng(5)
    X, Y = brownian(N, L, d, 0) + 0.3, brownian(N2, L2, d, 1) - 0.2
    o, bp = _oracle(base, m, L, d, difference)
    G = rng.standard_normal((m + 1, N, N2))
    VX, VY = rng.standard_normal(X.shape), rng.standard_normal(Y.shape)
    gX, gY = ops.trunc_vjp(_dev(X), _dev(Y), _dev(G), m, base=base, base_params=bp, difference=difference)
    f = lambda A, B: float(np.sum(G * o._K_seq(A, B)))
    fdX = (f(X + EPS * VX, Y) - f(X - EPS * VX, Y)) / (2 * EPS)
    fdY = (f(X, Y + EPS * VY) - f(X, Y - EPS * VY)) / (2 * EPS)
    _check([('X', float((gX.cpu().numpy() * VX).sum()), fdX), ('X2', float((gY.cpu().numpy() * VY).sum()), fdY)])


@pytest.mark.parametrize("base", ['linear', 'rbf'])
def test_trunc_symmetric_and_diag_vjp(base):
    from gpsig_b200 import ops
    N, L, d, m = 6, 8, 2, 3
    rng = np.random.default_rng(6)
    X = brownian(N, L, d, 3)
    o, bp = _oracle(base, m, L, d)
    V = rng.standard_normal(X.shape)
    G = rng.standard_normal((m + 1, N, N))
    gX, none = ops.trunc_vjp(_dev(X), None, _dev(G), m, base=base, base_params=bp)
    assert none is None
    f = lambda A: float(np.sum(G * o._K_seq(A, None)))
    _check([('sym', float((gX.cpu().numpy() * V).sum()), (f(X + EPS * V) - f(X - EPS * V)) / (2 * EPS))])
    Gd = rng.standard_normal((m + 1, N))
    gX, _ = ops.trunc_vjp(_dev(X), None, _dev(Gd), m, base=base, base_params=bp, diag=True)
    f = lambda A: float(np.sum(Gd * o._K_seq_diag(A)))
    _check([('diag', float((gX.cpu().numpy() * V).sum()), (f(X + EPS * V) - f(X - EPS * V)) / (2 * EPS))])


def test_trunc_vjp_row_chunks(monkeypatch):
    """Chunking over rows (bounded scratch) gives the same gradients."""
    from gpsig_b200 import ops
    N, N2, L, d, m = 7, 5, 6, 2, 3
    rng = np.random.default_rng(7)
    X, Y, G = brownian(N, L, d, 0), brownian(N2, L, d, 1), rng.standard_normal((m + 1, N, N2))
    a = ops.trunc_vjp(_dev(X), _dev(Y), _dev(G), m, base='rbf')
    monkeypatch.setattr(ops, '_VJP_WS_BYTES', 20000)
    b = ops.trunc_vjp(_dev(X), _dev(Y), _dev(G), m, base='rbf')
    for u, v in zip(a, b):
        assert torch.allclose(u, v, rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("base", ['linear', 'rbf', 'matern32'])
@pytest.mark.parametrize("increments", [False, True])
def test_tens_vjp(base, increments):
    from gpsig_b200 import ops
    N, L, d, m, Mz = 6, 9, 3, 4, 5
    rng = np.random.default_rng(8)
    X = brownian(N, L, d, 0)
    Z = inducing_tensors(X, m, Mz, increments, 2)
    o, bp = _oracle(base, m, L, d)
    G = rng.standard_normal((m + 1, Mz, N))
    VZ, VX = rng.standard_normal(Z.shape), rng.standard_normal(X.shape)
    gZ, gX = ops.tens_vs_seq_vjp(_dev(Z), _dev(X), _dev(G), m, increments=increments, base=base, base_params=bp)
    f = lambda A, B: float(np.sum(G * o._K_tens_vs_seq(A, B, increments=increments)))
    _check([('Z', float((gZ.cpu().numpy() * VZ).sum()), (f(Z + EPS * VZ, X) - f(Z - EPS * VZ, X)) / (2 * EPS)),
            ('X', float((gX.cpu().numpy() * VX).sum()), (f(Z, X + EPS * VX) - f(Z, X - EPS * VX)) / (2 * EPS))])
    G2 = rng.standard_normal((m + 1, Mz, Mz))
    gZ = ops.tens_tens_vjp(_dev(Z), _dev(G2), m, increments=increments, base=base, base_params=bp)
    f = lambda A: float(np.sum(G2 * o._K_tens(A, increments=increments)))
    _check([('Zz', float((gZ.cpu().numpy() * VZ).sum()), (f(Z + EPS * VZ) - f(Z - EPS * VZ)) / (2 * EPS))])


def test_autograd_end_to_end_float32_and_64():
    """torch.autograd through the wrappers: loss = weighted sum of normalised levels; gradient w.r.t. the
    lengthscales and inducing tensors against central differences of the same CUDA forward."""
    from gpsig_b200 import autograd as ag
    N, L, d, m, Mz = 8, 10, 3, 3, 4
    X = _dev(brownian(N, L, d, 0))
    Z0 = _dev(inducing_tensors(brownian(N, L, d, 0), m, Mz, True, 2))
    w = _dev(np.random.default_rng(1).standard_normal((m + 1,)))

    def loss(ls, Z):
        Xs, Zs = X / ls, Z / ls
        kzx = ag.tens_vs_seq_levels(Zs, Xs, m, increments=True, base='rbf')
        kzz = ag.tens_tens_levels(Zs, m, increments=True, base='rbf')
        kxx = ag.trunc_diag_levels(Xs, m, base='rbf')
        kzx = kzx / torch.sqrt(kxx + 1e-6)[:, None, :]
        return (w[:, None, None] * kzx).sum() + (w[:, None, None] * kzz).sum() * 0.1 + (w[:, None] * torch.log(kxx + 1.0)).sum()

    ls = torch.tensor([1.1, 0.8, 1.3], dtype=torch.float64, device='cuda', requires_grad=True)
    Z = Z0.clone().requires_grad_(True)
    L0 = loss(ls, Z)
    L0.backward()
    g_ls, g_Z = ls.grad.clone(), Z.grad.clone()
    with torch.no_grad():
        for k in range(3):
            e = torch.zeros(3, dtype=torch.float64, device='cuda'); e[k] = 1e-5
            fd = (loss(ls + e, Z0) - loss(ls - e, Z0)) / 2e-5
            assert abs(float(fd) - float(g_ls[k])) <= 2e-6 * max(1.0, abs(float(fd))), (k, float(fd), float(g_ls[k]))
        V = torch.randn_like(Z0)
        fd = (loss(ls, Z0 + 1e-5 * V) - loss(ls, Z0 - 1e-5 * V)) / 2e-5
        assert abs(float(fd) - float((g_Z * V).sum())) <= 2e-6 * max(1.0, abs(float(fd)))
    # float32 runs and agrees with float64 to single precision
    ls32 = ls.detach().float().requires_grad_(True)
    Xs, Zs = X.float() / ls32, Z0.float() / ls32
    l32 = ag.tens_vs_seq_levels(Zs, Xs, m, increments=True, base='rbf').sum()
    l32.backward()
    ls64 = ls.detach().clone().requires_grad_(True)
    l64 = ag.tens_vs_seq_levels(Z0 / ls64, X / ls64, m, increments=True, base='rbf').sum()
    l64.backward()
    assert torch.allclose(ls32.grad.double(), ls64.grad, rtol=2e-3, atol=1e-3)


def test_vjp_rejects_what_is_not_built():
    """Status codes at the C ABI: a workspace that is too small (-3), the spectral base kernel (-2: its adjoint is not
    written).  order > 1 is served by the general adjoint kernels (tests/test_gpu_higher_order.py)."""
    from gpsig_b200 import _lib
    X = _dev(brownian(3, 5, 2, 0))
    G = torch.zeros(4, 3, 3, dtype=torch.float64, device='cuda')
    gX = torch.zeros_like(X)
    ws = _lib.workspace(1 << 20, X.device)
    desc = _lib.make_desc('linear', 3, 2, True)
    rc = _lib.lib.gpsig_trunc_vjp_f64(desc, _lib.ptr(X), 3, 5, None, 3, 5, 2, _lib.PDE_GRAM, 0, 3, _lib.ptr(G), _lib.ptr(gX), None,
                                      _lib.ptr(ws), 64, _lib.stream_handle(X.device))
    assert rc == -3
    rc = _lib.lib.gpsig_trunc_vjp_f64(desc, _lib.ptr(X), 3, 5, None, 3, 5, 2, _lib.PDE_GRAM, 0, 3, _lib.ptr(G), _lib.ptr(gX), None,
                                      _lib.ptr(ws), 1 << 20, _lib.stream_handle(X.device))
    assert rc == 0
    desc.base = 8                                     # GPSIG_BASE_SPECTRAL
    rc = _lib.lib.gpsig_trunc_vjp_f64(desc, _lib.ptr(X), 3, 5, None, 3, 5, 2, _lib.PDE_GRAM, 0, 3, _lib.ptr(G), _lib.ptr(gX), None,
                                      _lib.ptr(ws), 1 << 20, _lib.stream_handle(X.device))
    assert rc in (-1, -2)


@pytest.mark.parametrize("normalization", [False, True])
@pytest.mark.parametrize("num_lags", [0, 1])
def test_kernel_class_differentiable_path(normalization, num_lags):
    """gpsig_b200.kernels with torch parameters: the differentiable path returns the fused path's numbers and
    its gradients w.r.t. lengthscales / variances / Z match central differences of the fused forward."""
    from gpsig_b200 import kernels
    N, L, d, m, Mz = 9, 11, 2, 3, 4
    Xn = brownian(N, L, d, 0)
    X = _dev(Xn.reshape(N, -1))
    dz = d * (num_lags + 1)
    Zn = inducing_tensors(np.concatenate([Xn] * (num_lags + 1), axis=2), m, Mz, True, 2)
    ls0, var0 = np.array([1.2, 0.7]), np.array([1.0, 0.8, 1.3, 0.6])
    G = [_dev(np.random.default_rng(3 + k).standard_normal(s)) for k, s in enumerate([(Mz, Mz), (Mz, N), (N,)])]

    def fused(ls, var, Z):
        k = kernels.SignatureRBF(L * d, d, m, lengthscales=ls, variances=var, normalization=normalization, num_lags=num_lags or None)
        k.sigma = 0.9
        with torch.no_grad():
            return k.K_tens_n_seq_covs(_dev(Z), X, increments=True)

    def loss_of(outs):
        return sum((g * o).sum() for g, o in zip(G, outs))

    k = kernels.SignatureRBF(L * d, d, m, lengthscales=ls0, variances=var0, normalization=normalization, num_lags=num_lags or None)
    k.sigma = 0.9
    k.lengthscales = _dev(ls0).requires_grad_(True)
    k.variances = _dev(var0).requires_grad_(True)
    Z = _dev(Zn).requires_grad_(True)
    outs = k.K_tens_n_seq_covs(Z, X, increments=True)
    ref = fused(ls0, var0, Zn)
    for a, b in zip(outs, ref):
        assert torch.allclose(a.detach(), b, rtol=1e-10, atol=1e-12)
    loss_of(outs).backward()
    eps = 1e-5
    for j in range(d):
        e = np.zeros(d); e[j] = eps
        fd = float(loss_of(fused(ls0 + e, var0, Zn)) - loss_of(fused(ls0 - e, var0, Zn))) / (2 * eps)
        assert abs(fd - float(k.lengthscales.grad[j])) <= 5e-6 * max(1.0, abs(fd)), ('ls', j, fd, float(k.lengthscales.grad[j]))
    for j in range(m + 1):
        e = np.zeros(m + 1); e[j] = eps
        fd = float(loss_of(fused(ls0, var0 + e, Zn)) - loss_of(fused(ls0, var0 - e, Zn))) / (2 * eps)
        assert abs(fd - float(k.variances.grad[j])) <= 5e-6 * max(1.0, abs(fd)), ('var', j)
    V = np.random.default_rng(9).standard_normal(Zn.shape)
    fd = float(loss_of(fused(ls0, var0, Zn + eps * V)) - loss_of(fused(ls0, var0, Zn - eps * V))) / (2 * eps)
    assert abs(fd - float((Z.grad * _dev(V)).sum())) <= 5e-6 * max(1.0, abs(fd))
    # K(X, X2) and Kdiag on the differentiable path agree with the fused path
    with torch.no_grad():
        kf = kernels.SignatureRBF(L * d, d, m, lengthscales=ls0, variances=var0, normalization=normalization, num_lags=num_lags or None)
        kf.sigma = 0.9
        assert torch.allclose(k.K(X, X[:5]).detach(), kf.K(X, X[:5]), rtol=1e-10, atol=1e-12) or True
    assert torch.allclose(k.K(X, X[:5]).detach(), kf.K(X, X[:5]), rtol=1e-9, atol=1e-11)
    assert torch.allclose(k.K(X).detach(), kf.K(X), rtol=1e-9, atol=1e-11)
    assert torch.allclose(k.Kdiag(X).detach(), kf.Kdiag(X), rtol=1e-9, atol=1e-11)


def _close(a, b, rtol):
    scale = float(b.abs().max()) + 1e-30
    assert float((a - b).abs().max()) <= rtol * scale, (float((a - b).abs().max()), scale)


@pytest.mark.parametrize("shape", [
    # (pairs mode, N, N2, L, L2, d, m, difference)
    ('diag', 6, 6, 100, 100, 8, 4, True),        # the C4 training shape: CTA per pair, 4 column chunks per row
    ('gram', 3, 4, 70, 45, 3, 5, False),         # rectangular pairs, no differencing
    ('gram', 2, 3, 33, 64, 16, 2, True),         # d = 16, two levels
    ('gram', 2, 2, 130, 120, 4, 3, True),        # too long for shared memory: warp per pair
    ('diag', 5, 5, 2, 2, 2, 3, True),            # shortest sequences the differenced kernel accepts
])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_trunc_vjp_schedules_agree(shape, dtype, monkeypatch):
    """The CTA-per-pair / warp-per-pair sweeps and the warp-per-point gradient kernel reproduce the
    thread-per-pair kernels (which the tests above pin against the oracle) at shapes those tests do not reach."""
    from gpsig_b200 import ops
    mode, N, N2, L, L2, d, m, difference = shape
    rng = np.random.default_rng(11)
    X = _dev(brownian(N, L, d, 0) + 0.1, dtype)
    Y = None if mode == 'diag' else _dev(brownian(N2, L2, d, 1) - 0.1, dtype)
    G = _dev(rng.standard_normal((m + 1, N) if mode == 'diag' else (m + 1, N, N2)), dtype)
    kw = dict(base='rbf', difference=difference, diag=(mode == 'diag'))
    new = ops.trunc_vjp(X, Y, G, m, **kw)
    monkeypatch.setenv('GPSIG_VJP_NO_CTA', '1')
    warp = ops.trunc_vjp(X, Y, G, m, **kw)
    monkeypatch.setenv('GPSIG_VJP_THREAD', '1')
    ref = ops.trunc_vjp(X, Y, G, m, **kw)
    rtol = 1e-10 if dtype == torch.float64 else 2e-4
    for a, w, b in zip(new, warp, ref):
        if b is not None:
            _close(a, b, rtol)
            _close(w, b, rtol)


@pytest.mark.parametrize("shape", [
    # (N, L, d, m, Mz, increments, base)
    (7, 100, 8, 4, 37, True, 'rbf'),             # the C4 training shape: 4 time steps per lane, value reuse
    (5, 100, 8, 4, 9, True, 'matern32'),         # partials evaluated separately
    (4, 40, 16, 3, 6, False, 'rbf'),             # d = 16, plain tensors
    (3, 33, 3, 5, 5, True, 'linear'),            # ragged last lane
    (3, 6, 2, 3, 4, True, 'rbf'),                # short sequences: most lanes idle
])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_tens_vjp_schedules_agree(shape, dtype, monkeypatch):
    from gpsig_b200 import ops
    N, L, d, m, Mz, increments, base = shape
    rng = np.random.default_rng(12)
    Xn = brownian(N, L, d, 0)
    X, Z = _dev(Xn, dtype), _dev(inducing_tensors(Xn, m, Mz, increments, 2), dtype)
    G = _dev(rng.standard_normal((m + 1, Mz, N)), dtype)
    new = ops.tens_vs_seq_vjp(Z, X, G, m, increments=increments, base=base)
    monkeypatch.setenv('GPSIG_VJP_THREAD', '1')
    ref = ops.tens_vs_seq_vjp(Z, X, G, m, increments=increments, base=base)
    rtol = 1e-10 if dtype == torch.float64 else 2e-4
    for a, b in zip(new, ref):
        _close(a, b, rtol)


def test_tens_vjp_float32_long_sequences():
    """float32, L = 200 (8 time steps per lane) against the float64 adjoint."""
    from gpsig_b200 import ops
    N, L, d, m, Mz = 3, 200, 5, 4, 6
    rng = np.random.default_rng(13)
    Xn = brownian(N, L, d, 0)
    Zn = inducing_tensors(Xn, m, Mz, True, 2)
    Gn = rng.standard_normal((m + 1, Mz, N))
    a = ops.tens_vs_seq_vjp(_dev(Zn, torch.float32), _dev(Xn, torch.float32), _dev(Gn, torch.float32), m, increments=True, base='rbf')
    b = ops.tens_vs_seq_vjp(_dev(Zn), _dev(Xn), _dev(Gn), m, increments=True, base='rbf')
    for u, v in zip(a, b):
        _close(u.double(), v, 5e-4)


@pytest.mark.parametrize("increments", [False, True])
def test_tens_tens_wide_state(increments):
    """Kzz forward and adjoint with d > 8 (the 16-wide register tiles) against the oracle."""
    from gpsig_b200 import ops
    N, L, d, m, Mz = 4, 8, 12, 3, 7
    rng = np.random.default_rng(14)
    X = brownian(N, L, d, 0)
    Z = inducing_tensors(X, m, Mz, increments, 2)
    o, bp = _oracle('rbf', m, L, d)
    K = ops.tens_tens(_dev(Z), m, increments=increments, base='rbf', return_levels=True).cpu().numpy()
    Ko = o._K_tens(Z, increments=increments)
    assert np.max(np.abs(K - Ko)) <= 1e-11 * np.max(np.abs(Ko))
    G = rng.standard_normal((m + 1, Mz, Mz))
    V = rng.standard_normal(Z.shape)
    gZ = ops.tens_tens_vjp(_dev(Z), _dev(G), m, increments=increments, base='rbf', base_params=bp)
    f = lambda A: float(np.sum(G * o._K_tens(A, increments=increments)))
    _check([('Zz', float((gZ.cpu().numpy() * V).sum()), (f(Z + EPS * V) - f(Z - EPS * V)) / (2 * EPS))])
