"""GPU parity of the SVGP glue around the hot path: ``Kuu`` / ``Kuf`` / ``Kuu_Kuf_Kff``
(gpsig/inducing_variables.py:51-137) and the conditional of ``SVGP._build_predict`` (gpsig/models.py:109-131)."""
import numpy as np
import pytest
import torch

from tests.util import brownian, inducing_tensors, rel_err

pytestmark = pytest.mark.gpu


def _dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


@pytest.mark.parametrize("increments", [False, True])
@pytest.mark.parametrize("learn_weights", [False, True])
def test_inducing_tensors_covariances(increments, learn_weights):
    from gpsig_b200 import kernels
    from gpsig_b200.inducing_variables import InducingTensors, Kuf, Kuu, Kuu_Kuf_Kff
    from oracle import svgp_np
    from oracle.trunc_np import SignatureKernelOracle
    N, L, d, m, Mz = 23, 19, 3, 3, 6
    X = brownian(N, L, d, 50)
    Z = inducing_tensors(X, m, Mz, increments, 51)
    ls = np.array([0.9, 1.1, 1.3])
    k = kernels.SignatureRBF(L * d, d, m, lengthscales=ls)
    o = SignatureKernelOracle(L * d, d, m, base='rbf', lengthscales=ls)
    feat = InducingTensors(_dev(Z), m, increments=increments, learn_weights=learn_weights)
    W = None
    if learn_weights:
        W = np.eye(Mz)[None] + 0.1 * np.random.default_rng(52).standard_normal((m, Mz, Mz))
        feat.W = _dev(W)
    Xd = _dev(X.reshape(N, -1))
    rzz, rzx, rxx = svgp_np.Kuu_Kuf_Kff(o, Z, X.reshape(N, -1), increments=increments, W=W, jitter=1e-6)
    Kzz, Kzx, Kxx = Kuu_Kuf_Kff(feat, k, Xd, jitter=1e-6)
    for g, r in ((Kzz, rzz), (Kzx, rzx), (Kxx, rxx)):
        assert rel_err(g.cpu().numpy(), r) < 1e-10
    assert rel_err(Kuu(feat, k, jitter=1e-6).cpu().numpy(), rzz) < 1e-10
    assert rel_err(Kuf(feat, k, Xd).cpu().numpy(), rzx) < 1e-10
    with pytest.raises(NameError):      # the reference's own defect, inducing_variables.py:60
        Kuu_Kuf_Kff(feat, k, Xd, jitter=1e-6, full_f_cov=True)


@pytest.mark.parametrize("learn_weights", [False, True])
def test_inducing_sequences_covariances(learn_weights):
    from gpsig_b200 import kernels
    from gpsig_b200.inducing_variables import InducingSequences, Kuf, Kuu, Kuu_Kuf_Kff
    from oracle import svgp_np
    from oracle.trunc_np import SignatureKernelOracle
    N, L, Lz, d, m, Mz = 11, 16, 16, 2, 3, 5
    X, Zs = brownian(N, L, d, 53), brownian(Mz, Lz, d, 54)
    k = kernels.SignatureLinear(L * d, d, m)
    o = SignatureKernelOracle(L * d, d, m, base='linear')
    feat = InducingSequences(_dev(Zs), m, learn_weights=learn_weights)
    W = None
    if learn_weights:
        W = np.eye(Mz)[None] + 0.1 * np.random.default_rng(55).standard_normal((m, Mz, Mz))
        feat.W = _dev(W)
    Xd = _dev(X.reshape(N, -1))
    rzz, rzx, rxx = svgp_np.Kuu_Kuf_Kff(o, Zs.reshape(Mz, -1), X.reshape(N, -1), W=W, jitter=1e-6, sequences=True)
    Kzz, Kzx, Kxx = Kuu_Kuf_Kff(feat, k, Xd, jitter=1e-6)
    for g, r in ((Kzz, rzz), (Kzx, rzx), (Kxx, rxx)):
        assert rel_err(g.cpu().numpy(), r) < 1e-10
    # Kuu / Kuf go through ``kern.K`` (:101-121), which normalises once -- unlike K_seq_n_seq_covs, whose Kzx and
    # Kzz the reference divides twice (kernels.py:746,783; kept as is)
    Zf = Zs.reshape(Mz, -1)
    if learn_weights:
        ezz = svgp_np.mix_levels(W, Kzz=o.K(Zf, return_levels=True)) + 1e-6 * np.eye(Mz)
        ezx = svgp_np.mix_levels(W, Kzx=o.K(Zf, X.reshape(N, -1), return_levels=True))
    else:
        ezz, ezx = o.K(Zf) + 1e-6 * np.eye(Mz), o.K(Zf, X.reshape(N, -1))
    assert rel_err(Kuu(feat, k, jitter=1e-6).cpu().numpy(), ezz) < 1e-10
    assert rel_err(Kuf(feat, k, Xd).cpu().numpy(), ezx) < 1e-10


@pytest.mark.parametrize("whiten", [True, False])
@pytest.mark.parametrize("q_diag", [False, True])
def test_build_predict_matches_restated_conditional(whiten, q_diag):
    """``SVGP._build_predict`` (models.py:109-131): covariances from the CUDA kernels + the conditional."""
    from gpsig_b200 import conditionals, kernels, settings
    from gpsig_b200.inducing_variables import InducingTensors
    from oracle import svgp_np
    from oracle.trunc_np import SignatureKernelOracle
    N, L, d, m, Mz, R = 31, 25, 4, 4, 9, 3
    X = brownian(N, L, d, 56)
    Z = inducing_tensors(X, m, Mz, True, 57)
    k = kernels.SignatureRBF(L * d, d, m)
    o = SignatureKernelOracle(L * d, d, m, base='rbf')
    rng = np.random.default_rng(58)
    q_mu = rng.standard_normal((Mz, R))
    q_sqrt = np.abs(rng.standard_normal((Mz, R))) if q_diag else np.tril(rng.standard_normal((R, Mz, Mz)))
    feat = InducingTensors(_dev(Z), m, increments=True)
    mean, var, Kzz = conditionals.build_predict(feat, k, _dev(X.reshape(N, -1)), _dev(q_mu), None, whiten=whiten, return_Kzz=True)
    rzz, rzx, rxx = svgp_np.Kuu_Kuf_Kff(o, Z, X.reshape(N, -1), increments=True, jitter=settings.jitter)
    rm, rv = svgp_np.base_conditional(rzx, rzz, rxx, q_mu, white=whiten)
    assert rel_err(mean.cpu().numpy(), rm) < 1e-8 and rel_err(var.cpu().numpy(), rv) < 1e-8
    mean, var = conditionals.base_conditional(_dev(rzx), _dev(rzz), _dev(rxx), _dev(q_mu), q_sqrt=_dev(q_sqrt), white=whiten)
    rm, rv = svgp_np.base_conditional(rzx, rzz, rxx, q_mu, q_sqrt=q_sqrt, white=whiten)
    assert rel_err(mean.cpu().numpy(), rm) < 1e-9 and rel_err(var.cpu().numpy(), rv) < 1e-9
    if not q_diag:
        mean, var = conditionals.base_conditional(_dev(rzx), _dev(rzz), _dev(o.K(X.reshape(N, -1))), _dev(q_mu), q_sqrt=_dev(q_sqrt),
                                                  white=whiten, full_cov=True)
        rm, rv = svgp_np.base_conditional(rzx, rzz, o.K(X.reshape(N, -1)), q_mu, q_sqrt=q_sqrt, white=whiten, full_cov=True)
        assert rel_err(mean.cpu().numpy(), rm) < 1e-9 and rel_err(var.cpu().numpy(), rv) < 1e-9


def test_elbo_style_gradients_flow_to_inducing_tensors():
    """A training-style scalar of the predictive moments back-propagates to Z, W, lengthscales and q_mu
    through the CUDA adjoints."""
    from gpsig_b200 import conditionals, kernels
    from gpsig_b200.inducing_variables import InducingTensors
    N, L, d, m, Mz, R = 12, 14, 2, 3, 4, 1
    X = brownian(N, L, d, 59)
    Z = _dev(inducing_tensors(X, m, Mz, True, 60)).requires_grad_(True)
    k = kernels.SignatureRBF(L * d, d, m)
    k.lengthscales = _dev(np.ones(d)).requires_grad_(True)
    feat = InducingTensors(Z, m, increments=True, learn_weights=True)
    feat.W.requires_grad_(True)
    q_mu = _dev(np.random.default_rng(61).standard_normal((Mz, R))).requires_grad_(True)
    mean, var = conditionals.build_predict(feat, k, _dev(X.reshape(N, -1)), q_mu, None, whiten=True)
    loss = (mean ** 2).sum() + var.sum()
    loss.backward()
    for t in (Z, feat.W, k.lengthscales, q_mu):
        assert t.grad is not None and torch.isfinite(t.grad).all() and float(t.grad.abs().max()) > 0


def test_inducing_sequences_gradients_reach_Z_and_kernel_parameters():
    """K_seq_n_seq_covs (kernels.py:706-794) on the differentiable path: same values as the fused forward, and
    gradients w.r.t. the inducing sequences, lengthscales, variances and sigma (checked by central differences of
    the non-differentiable CUDA forward)."""
    from gpsig_b200 import kernels
    from gpsig_b200.inducing_variables import InducingSequences, Kuu_Kuf_Kff
    N, L, d, m, Mz = 7, 12, 2, 3, 4
    X, Zs = brownian(N, L, d, 70), brownian(Mz, L, d, 71)
    Xd = _dev(X.reshape(N, -1))
    for norm in (True, False):
        k = kernels.SignatureRBF(L * d, d, m, normalization=norm, lengthscales=np.array([0.9, 1.2]))
        with torch.no_grad():
            ref = k.K_seq_n_seq_covs(_dev(Zs.reshape(Mz, -1)), Xd)
        Zt = _dev(Zs.reshape(Mz, -1)).requires_grad_(True)
        k.lengthscales = _dev(np.array([0.9, 1.2])).requires_grad_(True)
        k.variances = _dev(np.ones(m + 1)).requires_grad_(True)
        k.sigma = _dev(np.array(1.0)).requires_grad_(True)
        got = k.K_seq_n_seq_covs(Zt, Xd)
        for g, r in zip(got, ref):
            assert g.requires_grad and rel_err(g.detach().cpu().numpy(), r.cpu().numpy()) < 1e-10
        rng = np.random.default_rng(72)
        w = [_dev(rng.standard_normal(tuple(g.shape))) for g in got]
        loss = sum((g * ww).sum() for g, ww in zip(got, w))
        loss.backward()
        for t in (Zt, k.lengthscales, k.variances, k.sigma):
            assert t.grad is not None and torch.isfinite(t.grad).all() and float(t.grad.abs().max()) > 0

        def f(Zv, ls):
            kk = kernels.SignatureRBF(L * d, d, m, normalization=norm, lengthscales=ls)
            with torch.no_grad():
                o = kk.K_seq_n_seq_covs(_dev(Zv), Xd)
            return float(sum((g * ww).sum() for g, ww in zip(o, w)))
        # directional derivative w.r.t. Z and w.r.t. the lengthscales
        dZ = rng.standard_normal((Mz, L * d))
        h = 1e-5
        fd = (f(Zs.reshape(Mz, -1) + h * dZ, np.array([0.9, 1.2])) - f(Zs.reshape(Mz, -1) - h * dZ, np.array([0.9, 1.2]))) / (2 * h)
        an = float((Zt.grad.cpu().numpy() * dZ).sum())
        assert abs(fd - an) < 2e-6 * max(1.0, abs(fd)), (norm, fd, an)
        dl = np.array([0.3, -0.7])
        fd = (f(Zs.reshape(Mz, -1), np.array([0.9, 1.2]) + h * dl) - f(Zs.reshape(Mz, -1), np.array([0.9, 1.2]) - h * dl)) / (2 * h)
        an = float((k.lengthscales.grad.cpu().numpy() * dl).sum())
        assert abs(fd - an) < 2e-6 * max(1.0, abs(fd)), (norm, fd, an)
    # through the SVGP dispatch
    feat = InducingSequences(_dev(Zs).requires_grad_(True), m)
    k = kernels.SignatureRBF(L * d, d, m)
    Kzz, Kzx, Kxx = Kuu_Kuf_Kff(feat, k, Xd, jitter=1e-6)
    (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()
    assert feat.Z.grad is not None and float(feat.Z.grad.abs().max()) > 0


def test_tensor_variances_at_prediction_time():
    """Kdiag / K_norms with trained (tensor) variances and sigma under no_grad (np.sum on a tensor)."""
    from gpsig_b200 import kernels
    N, L, d, m = 5, 10, 2, 3
    X = _dev(brownian(N, L, d, 73).reshape(N, -1))
    k = kernels.SignatureRBF(L * d, d, m)
    k.variances = _dev(np.linspace(0.5, 1.5, m + 1)).requires_grad_(True)
    k.sigma = _dev(np.array(1.3)).requires_grad_(True)
    with torch.no_grad():
        kd = k.Kdiag(X)
        const, _ = k.K_norms(X)
    assert rel_err(kd.cpu().numpy(), np.full(N, 1.3 * np.linspace(0.5, 1.5, m + 1).sum())) < 1e-12
    assert rel_err(const.cpu().numpy(), np.full(N, 1.3 * np.linspace(0.5, 1.5, m + 1).sum())) < 1e-12
    const, lv = k.K_norms(X)                      # differentiable path: reaches sigma / variances
    const.sum().backward()
    assert k.variances.grad is not None and float(k.variances.grad.abs().max()) > 0


def test_poly_and_mix_scalars_are_refused_when_trainable():
    from gpsig_b200 import kernels
    N, L, d, m = 4, 8, 2, 2
    X = _dev(brownian(N, L, d, 74).reshape(N, -1))
    k = kernels.SignaturePoly(L * d, d, m, gamma=1.0, degree=3)
    ref = k.K(X)
    k.gamma = torch.tensor(1.0, device='cuda')                 # a plain tensor is fine
    assert rel_err(k.K(X).cpu().numpy(), ref.cpu().numpy()) < 1e-14
    k.gamma = torch.tensor(1.0, device='cuda', requires_grad=True)
    with pytest.raises(NotImplementedError):
        k.K(X)
    k = kernels.SignatureMix(L * d, d, m)
    k.mixing = torch.tensor(0.5, device='cuda', requires_grad=True)
    with pytest.raises(NotImplementedError):
        k.K(X)
    # spectral kernel with lags: the lag weights are not clobbered by the [Q, d] gamma block (kernels.py:82 vs :1094)
    ks = kernels.SignatureSpectral(L * d, d, m, Q=2, num_lags=1)
    assert np.asarray(ks._lag_gamma()).shape == (2,) and np.asarray(ks.gamma).shape == (2, d)


class _DLPackOnly:
    """A device array that is NOT a torch tensor: exposes only the DLPack protocol (what tf.experimental.dlpack /
    CuPy / JAX arrays offer)."""

    def __init__(self, t):
        self._t = t
        self.shape = tuple(t.shape)

    def __dlpack__(self, *a, **kw):
        return self._t.__dlpack__(*a, **kw)

    def __dlpack_device__(self):
        return self._t.__dlpack_device__()


def test_dlpack_producers_are_ingested_zero_copy():
    """north star: 'DLPack zero-copy' -- any __dlpack__ producer is accepted by the kernels' methods and read in place."""
    from gpsig_b200 import kernels, kernels_pde
    from gpsig_b200.kernels import as_device_tensor
    N, L, d, m, Mz = 6, 12, 3, 3, 4
    Xn = brownian(N, L, d, 75)
    Xt = _dev(Xn.reshape(N, -1))
    Xp = _DLPackOnly(Xt)
    assert not torch.is_tensor(Xp)
    assert as_device_tensor(Xp).data_ptr() == Xt.data_ptr()            # zero copy: same device memory
    k = kernels.SignatureRBF(L * d, d, m)
    assert torch.equal(k.K(Xp), k.K(Xt))
    assert torch.equal(k.K(Xp, _DLPackOnly(Xt[:3])), k.K(Xt, Xt[:3]))
    assert torch.equal(k.Kdiag(Xp), k.Kdiag(Xt))
    Z = _dev(inducing_tensors(Xn, m, Mz, True, 76))
    for a, b in zip(k.K_tens_n_seq_covs(_DLPackOnly(Z), Xp, increments=True), k.K_tens_n_seq_covs(Z, Xt, increments=True)):
        assert torch.equal(a, b)
    kp = kernels_pde.UntruncSignatureKernel(L * d, d, order=1)
    assert torch.equal(kp.Kdiag(Xp), kp.Kdiag(Xt))
    with pytest.raises(TypeError):
        k.K(Xn.reshape(N, -1))                                          # numpy goes through compute_*


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_covariance_step_replays_as_a_cuda_graph(dtype):
    """Forward + backward of K_tens_n_seq_covs (Kzz, Kzx, Kxx_diag on three streams, hand-written adjoints) captures into
    one CUDA graph: no entry point synchronises or allocates outside torch's pool.  The replay on NEW inputs (copied into
    the captured buffers) reproduces the eager step."""
    import warnings
    from gpsig_b200 import kernels
    warnings.filterwarnings('ignore', message="The AccumulateGrad node's stream")
    N, L, d, m, Mz = 10, 20, 3, 3, 6
    k = kernels.SignatureRBF(L * d, d, m, lengthscales=np.full(d, 1.3))
    X = _dev(brownian(N, L, d, 70).reshape(N, -1), dtype).requires_grad_(True)
    Z = _dev(inducing_tensors(brownian(N, L, d, 70), m, Mz, True, 71), dtype).requires_grad_(True)

    def step():
        Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Z, X, full_X_cov=False, increments=True)
        ((Kzz ** 2).sum() + (Kzx ** 2).sum() + Kxx.sum()).backward()

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    X.grad = None
    Z.grad = None
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    cX, cZ = X.grad, Z.grad
    with torch.no_grad():                                     # new values in the captured input buffers
        X.copy_(_dev(brownian(N, L, d, 72).reshape(N, -1), dtype))
        Z.copy_(_dev(inducing_tensors(brownian(N, L, d, 72), m, Mz, True, 73), dtype))
    cX.zero_()
    cZ.zero_()
    graph.replay()
    torch.cuda.synchronize()
    gX, gZ = cX.clone(), cZ.clone()
    X.grad = None
    Z.grad = None
    step()
    tol = 1e-10 if dtype == torch.float64 else 1e-4
    assert float(gX.abs().max()) > 0 and float(gZ.abs().max()) > 0
    assert rel_err(gX.cpu().numpy(), X.grad.cpu().numpy()) < tol
    assert rel_err(gZ.cpu().numpy(), Z.grad.cpu().numpy()) < tol


def test_numpy_hyperparameters_are_copied_to_the_device_once_per_value():
    from gpsig_b200 import kernels
    d = 3
    k = kernels.SignatureRBF(10 * d, d, 2, lengthscales=np.full(d, 1.3))
    like = _dev(np.zeros(1))
    a, b = k._param(k.lengthscales, like), k._param(k.lengthscales, like)
    assert a.data_ptr() == b.data_ptr()
    k.lengthscales = np.full(d, 0.7)                         # a new value is a new copy, the old one is untouched
    c = k._param(k.lengthscales, like)
    assert c.data_ptr() != a.data_ptr() and float(c[0]) == 0.7 and float(a[0]) == 1.3
    assert k._param(k.lengthscales, like.float()).dtype == torch.float32


def test_graphed_covariances_match_the_eager_method():
    """``graphed_tens_n_seq_covs``: forward and backward as CUDA graphs (make_graphed_callables), fed new values."""
    import warnings
    from gpsig_b200 import kernels
    warnings.filterwarnings('ignore', message="The AccumulateGrad node's stream")
    N, L, d, m, Mz = 9, 16, 3, 3, 5
    k = kernels.SignatureRBF(L * d, d, m, lengthscales=np.full(d, 1.2))
    X = _dev(brownian(N, L, d, 80).reshape(N, -1)).requires_grad_(True)
    Z = _dev(inducing_tensors(brownian(N, L, d, 80), m, Mz, True, 81)).requires_grad_(True)
    f = k.graphed_tens_n_seq_covs(Z, X, increments=True)
    X2 = _dev(brownian(N, L, d, 82).reshape(N, -1)).requires_grad_(True)
    Z2 = _dev(inducing_tensors(brownian(N, L, d, 82), m, Mz, True, 83)).requires_grad_(True)
    outs = f(Z2, X2)
    sum((o ** 2).sum() for o in outs).backward()
    gZ, gX = Z2.grad.clone(), X2.grad.clone()
    Z2.grad = None
    X2.grad = None
    ref = k.K_tens_n_seq_covs(Z2, X2, increments=True)
    sum((o ** 2).sum() for o in ref).backward()
    for a, b in zip(outs, ref):
        assert rel_err(a.detach().cpu().numpy(), b.detach().cpu().numpy()) < 1e-12
    assert rel_err(gZ.cpu().numpy(), Z2.grad.cpu().numpy()) < 1e-10
    assert rel_err(gX.cpu().numpy(), X2.grad.cpu().numpy()) < 1e-10
