"""Differentiable raw levels: ``torch.autograd.Function`` wrappers around the forward kernels and their
hand-written CUDA adjoints (``csrc/vjp_kernels.cuh``, C ABI ``gpsig_*_vjp_*``).

The reference trains through TensorFlow autodiff of ``gpsig/signature_algs.py`` (``training.py`` optimises
the kernel hyper-parameters and the inducing tensors through ``Kuu_Kuf_Kff``, ``inducing_variables.py:51-137``).
Here the CUDA ops return RAW levels ``K_l`` and their gradients w.r.t. the coordinates they are called with;
everything around them -- division by the lengthscales (``kernels.py:343-399``), per-level normalisation with
jitter and the ``sigma * variances`` weighting (``kernels.py:431-477, 509-541, 543-704``) -- is elementwise
torch code on small tensors, differentiated by torch.  Any ``order`` (order > 1: csrc/generic_vjp_kernels.cuh).
"""
import torch

from . import ops


class _TruncLevels(torch.autograd.Function):
    """Raw levels [m+1, N, N2] of K(X, X2) (X2 may be None) or [m+1, N] of the diagonal."""

    @staticmethod
    def forward(ctx, X, X2, m, base, base_params, difference, diag, order=1):
        ctx.save_for_backward(X, X2)
        ctx.cfg = (m, base, tuple(base_params), difference, diag, order)
        Xd = X.detach().contiguous()
        if diag:
            return ops.trunc_diag(Xd, m, base=base, base_params=base_params, difference=difference, order=order)
        X2d = None if X2 is None else X2.detach().contiguous()
        return ops.trunc_gram(Xd, X2d, m, base=base, base_params=base_params, difference=difference, return_levels=True, order=order)

    @staticmethod
    def backward(ctx, g):
        X, X2 = ctx.saved_tensors
        m, base, bp, difference, diag, order = ctx.cfg
        gX, gX2 = ops.trunc_vjp(X.detach().contiguous(), None if X2 is None else X2.detach().contiguous(), g.contiguous(), m,
                                base=base, base_params=bp, difference=difference, diag=diag, order=order)
        return gX, gX2, None, None, None, None, None, None


class _TensSeqLevels(torch.autograd.Function):
    """Raw levels [m+1, Mz, N] of inducing tensors vs sequences."""

    @staticmethod
    def forward(ctx, Z, X, m, increments, base, base_params, difference, order=1):
        ctx.save_for_backward(Z, X)
        ctx.cfg = (m, increments, base, tuple(base_params), difference, order)
        return ops.tens_vs_seq(Z.detach().contiguous(), X.detach().contiguous(), m, increments=increments, base=base,
                               base_params=base_params, difference=difference, return_levels=True, order=order)

    @staticmethod
    def backward(ctx, g):
        Z, X = ctx.saved_tensors
        m, increments, base, bp, difference, order = ctx.cfg
        gZ, gX = ops.tens_vs_seq_vjp(Z.detach().contiguous(), X.detach().contiguous(), g.contiguous(), m, increments=increments,
                                     base=base, base_params=bp, difference=difference, order=order)
        return gZ, gX, None, None, None, None, None, None


class _TensTensLevels(torch.autograd.Function):
    """Raw levels [m+1, Mz, Mz] of inducing tensors vs themselves."""

    @staticmethod
    def forward(ctx, Z, m, increments, base, base_params):
        ctx.save_for_backward(Z)
        ctx.cfg = (m, increments, base, tuple(base_params))
        return ops.tens_tens(Z.detach().contiguous(), m, increments=increments, base=base, base_params=base_params,
                             return_levels=True)

    @staticmethod
    def backward(ctx, g):
        (Z,) = ctx.saved_tensors
        m, increments, base, bp = ctx.cfg
        return ops.tens_tens_vjp(Z.detach().contiguous(), g.contiguous(), m, increments=increments, base=base, base_params=bp), \
            None, None, None, None


class _PdeKdiag(torch.autograd.Function):
    """k(x_i, x_i) of the untruncated kernel with the REFERENCE's gradient (``_KdiagGrad``,
    gpsig/kernels_pde.py:465-509 / ``_untrunc_cov_grad.py:25-77``): a continuous-adjoint formula, not the
    derivative of the stencil -- kept as is so that training follows the reference's trajectory."""

    @staticmethod
    def forward(ctx, X, dyadic_order, rev_solver):
        ctx.save_for_backward(X)
        ctx.cfg = (dyadic_order, rev_solver)
        return ops.pde(X.detach().contiguous(), None, dyadic_order=dyadic_order, gram=False)

    @staticmethod
    def backward(ctx, g):
        (X,) = ctx.saved_tensors
        dyadic_order, rev_solver = ctx.cfg
        return ops.pde_kdiag_vjp(X.detach().contiguous(), g.contiguous(), dyadic_order=dyadic_order,
                                 rev_solver=rev_solver), None, None


def pde_kdiag(X, dyadic_order=0, rev_solver=0):
    return _PdeKdiag.apply(X, int(dyadic_order), int(rev_solver))


def trunc_levels(X, X2, m, base='linear', base_params=(), difference=True, order=1):
    return _TruncLevels.apply(X, X2, int(m), base, tuple(base_params), bool(difference), False, int(order))


def trunc_diag_levels(X, m, base='linear', base_params=(), difference=True, order=1):
    return _TruncLevels.apply(X, None, int(m), base, tuple(base_params), bool(difference), True, int(order))


def tens_vs_seq_levels(Z, X, m, increments=False, base='linear', base_params=(), difference=True, order=1):
    return _TensSeqLevels.apply(Z, X, int(m), bool(increments), base, tuple(base_params), bool(difference), int(order))


def tens_tens_levels(Z, m, increments=False, base='linear', base_params=()):
    return _TensTensLevels.apply(Z, int(m), bool(increments), base, tuple(base_params))
