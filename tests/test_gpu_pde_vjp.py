"""GPU parity: the reference's hand-written backward of the PDE ``Kdiag`` (``_KdiagGrad``,
gpsig/kernels_pde.py:465-509; ``_untrunc_cov_grad.py:25-77``) through ``gpsig_pde_kdiag_vjp_f64``."""
import os

import numpy as np
import pytest
import torch

from tests.util import brownian, rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def _dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float64, device='cuda')


@pytest.mark.parametrize("rev_solver", [0, 1])
@pytest.mark.parametrize("L,d,order", [(2, 1, 0), (5, 2, 0), (20, 3, 0), (20, 3, 2), (33, 4, 1), (34, 8, 0), (70, 5, 1),
                                       (9, 16, 3), (130, 2, 0)])
def test_pde_kdiag_vjp_vs_oracle(L, d, order, rev_solver):
    from gpsig_b200 import ops
    from oracle import pde_np
    N = 5
    X = brownian(N, L, d, 21)
    g = np.linspace(0.5, 1.5, N)
    got = ops.pde_kdiag_vjp(_dev(X), _dev(g), dyadic_order=order, rev_solver=rev_solver).cpu().numpy()
    ref = g[:, None, None] * pde_np.pde_kdiag_grad(X, order, rev_solver)
    assert rel_err(got, ref) < 1e-10


def test_pde_kdiag_vjp_golden_from_reference_cython():
    """Golden: the reference's backward formula on the grids of its own compiled solver (make_golden.py)."""
    from gpsig_b200 import ops
    g = np.load(os.path.join(GOLD, 'pde_ref.npz'))
    X = g['X']
    for order in (0, 1, 2):
        got = ops.pde_kdiag_vjp(_dev(X), torch.ones(X.shape[0], dtype=torch.float64, device='cuda'),
                                dyadic_order=order, rev_solver=0).cpu().numpy()
        assert rel_err(got, g['kdiag_grad_order%d' % order]) < 1e-10


def test_pde_kdiag_vjp_rows_and_chunks():
    """Row ranges and a workspace that only holds 2 sequences at a time give the same numbers."""
    from gpsig_b200 import ops
    N, L, d, order = 7, 25, 3, 1
    X = _dev(brownian(N, L, d, 22))
    g = _dev(np.linspace(-1., 1., N))
    full = ops.pde_kdiag_vjp(X, g, dyadic_order=order)
    part = ops.pde_kdiag_vjp(X, g[2:5].contiguous(), dyadic_order=order, rows=(2, 5))
    assert torch.allclose(part[2:5], full[2:5], rtol=1e-12, atol=1e-14)
    assert float(part[:2].abs().max()) == 0. and float(part[5:].abs().max()) == 0.
    old = ops._PDE_VJP_WS_BYTES
    try:
        ops._PDE_VJP_WS_BYTES = 2 * ops.lib.gpsig_pde_kdiag_vjp_workspace_bytes(1, L, d, order)
        chunked = ops.pde_kdiag_vjp(X, g, dyadic_order=order)
    finally:
        ops._PDE_VJP_WS_BYTES = old
    assert torch.allclose(chunked, full, rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("implementation,rev_solver", [('cython', 0), ('gpu_op', 1)])
def test_pde_kdiag_autograd_lengthscales_sigma(implementation, rev_solver):
    """``Kdiag`` under autograd: gradients w.r.t. X, the lengthscales and sigma chain the reference's
    backward through the scaling (kernels_pde.py:150-151) and ``sigma *`` (:185)."""
    from gpsig_b200 import kernels_pde
    from oracle import pde_np
    N, L, d, order = 4, 18, 3, 1
    Xn = brownian(N, L, d, 23)
    ls = np.array([0.8, 1.1, 1.3])
    k = kernels_pde.UntruncSignatureKernel(L * d, d, order=order, implementation=implementation)
    k.lengthscales = _dev(ls).requires_grad_(True)
    k.sigma = torch.tensor(1.7, dtype=torch.float64, device='cuda', requires_grad=True)
    X = _dev(Xn.reshape(N, -1)).requires_grad_(True)
    w = np.linspace(1., 2., N)
    out = k.Kdiag(X)
    (out * _dev(w)).sum().backward()
    kd = pde_np.pde_kernel(Xn / ls, None, order)
    assert rel_err(out.detach().cpu().numpy(), 1.7 * kd) < 1e-10
    gs = 1.7 * w[:, None, None] * pde_np.pde_kdiag_grad(Xn / ls, order, rev_solver)      # d / d(scaled X)
    assert rel_err(X.grad.cpu().numpy().reshape(N, L, d), gs / ls) < 1e-10
    assert rel_err(k.lengthscales.grad.cpu().numpy(), -(gs * Xn / ls ** 2).sum(axis=(0, 1))) < 1e-10
    assert rel_err(k.sigma.grad.cpu().numpy(), (w * kd).sum()) < 1e-10
    # without autograd the fused forward is used and agrees
    with torch.no_grad():
        assert torch.allclose(k.Kdiag(X), out.detach(), rtol=1e-12, atol=0)
