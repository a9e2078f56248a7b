"""Functional layer: one Python function per C-ABI entry point.

Inputs are contiguous CUDA tensors ``[N, L, d]`` (float32 or float64); the
functions allocate the output and the workspace with torch, pass raw device
pointers plus the current CUDA stream through ctypes and return the output
tensor.  Nothing here computes: if the CUDA library is missing the import of
``_lib`` has already failed.
"""
import os

import torch

from . import _lib
from ._lib import check, lib, make_desc, make_epilogue, ptr, require_cuda, sfx, stream_handle, workspace


def _same(t, ref, name):
    if t is not None:
        require_cuda(t, name)
        if t.dtype != ref.dtype:
            raise TypeError("gpsig_b200: %s has dtype %s, expected %s" % (name, t.dtype, ref.dtype))
        if t.device != ref.device:
            raise RuntimeError("gpsig_b200: %s is on %s, expected %s" % (name, t.device, ref.device))


def trunc_gram(X, X2, num_levels, base='linear', base_params=(), order=1, difference=True, lengthscales=None,
               norm_x=None, norm_x2=None, level_scale=None, sym_jitter=0.0, return_levels=True, rows=None,
               upper_only=False, out=None):
    """Rows ``rows=(begin, end)`` (default all) of the truncated signature Gram.

    Replaces ``SignatureKernel._K_seq`` + epilogue (gpsig/kernels.py:209-238, 431-477).
    Returns ``[m+1, R, N2]`` (``return_levels``) or ``[R, N2]``.  ``upper_only`` (K(X, X), row-sharded callers):
    entries below the diagonal may be left unwritten; complete the assembled matrix with ``mirror_lower``.
    ``out``: write into this contiguous tensor of the result's shape (a row slice of a larger matrix: sharding).
    """
    require_cuda(X, 'X')
    for t, n in ((X2, 'X2'), (lengthscales, 'lengthscales'), (norm_x, 'norm_x'), (norm_x2, 'norm_x2'),
                 (level_scale, 'level_scale')):
        _same(t, X, n)
    N, L, d = X.shape
    if X2 is not None and X2.shape[2] != d:
        raise ValueError("X and X2 must have the same state dimension")
    N2, L2 = (N, L) if X2 is None else X2.shape[:2]
    rb, re = (0, N) if rows is None else rows
    m = int(num_levels)
    shape = (m + 1, re - rb, N2) if return_levels else (re - rb, N2)
    if out is None:
        out = torch.empty(shape, dtype=X.dtype, device=X.device)
    else:
        _same(out, X, 'out')
        if tuple(out.shape) != shape or not out.is_contiguous():
            raise ValueError("gpsig_b200: out must be a contiguous tensor of shape %s, got %s" % (shape, tuple(out.shape)))
    desc = make_desc(base, m, order, difference, base_params)
    epi = make_epilogue(norm_x, norm_x2, level_scale, sym_jitter, return_levels, upper_only and X2 is None)
    esz = X.element_size()
    wsb = lib.gpsig_trunc_gram_workspace_bytes(N, L, 0 if X2 is None else N2, L2, d, esz)
    ws = workspace(wsb, X.device)
    fn = getattr(lib, 'gpsig_trunc_gram_' + sfx(X.dtype))
    with torch.cuda.device(X.device):
        check(fn(desc, ptr(X), N, L, ptr(X2), N2, L2, d, ptr(lengthscales), rb, re, epi, ptr(out), ptr(ws), wsb,
                 stream_handle(X.device)))
    return out


def mirror_lower(K):
    """In place ``K[..., i, j] = K[..., j, i]`` for i > j on a contiguous ``[N, N]`` or ``[P, N, N]`` CUDA tensor."""
    require_cuda(K, 'K')
    if K.dim() not in (2, 3) or K.shape[-1] != K.shape[-2]:
        raise ValueError("mirror_lower needs [N, N] or [P, N, N]")
    fn = getattr(lib, 'gpsig_mirror_lower_' + sfx(K.dtype))
    with torch.cuda.device(K.device):
        check(fn(ptr(K), K.shape[-1], 1 if K.dim() == 2 else K.shape[0], stream_handle(K.device)))
    return K


def trunc_diag(X, num_levels, base='linear', base_params=(), order=1, difference=True, lengthscales=None):
    """Raw per-level diagonals ``[m+1, N]`` (``_K_seq_diag``, gpsig/kernels.py:190-207)."""
    require_cuda(X, 'X')
    _same(lengthscales, X, 'lengthscales')
    N, L, d = X.shape
    m = int(num_levels)
    out = torch.empty((m + 1, N), dtype=X.dtype, device=X.device)
    desc = make_desc(base, m, order, difference, base_params)
    wsb = lib.gpsig_trunc_diag_workspace_bytes(N, L, d, X.element_size())
    ws = workspace(wsb, X.device)
    fn = getattr(lib, 'gpsig_trunc_diag_' + sfx(X.dtype))
    with torch.cuda.device(X.device):
        check(fn(desc, ptr(X), N, L, d, ptr(lengthscales), ptr(out), ptr(ws), wsb, stream_handle(X.device)))
    return out


def tens_vs_seq(Z, X, num_levels, increments=False, base='linear', base_params=(), order=1, difference=True,
                lengthscales=None, norm_x=None, level_scale=None, return_levels=True, cols=None):
    """Inducing tensors vs sequences (``_K_tens_vs_seq``, gpsig/kernels.py:314-341 + epilogue :604-620).

    Z ``[T, Mz, d]`` or ``[T, Mz, 2, d]``; returns ``[m+1, Mz, C]`` or ``[Mz, C]`` for columns ``cols``.
    """
    require_cuda(X, 'X')
    for t, n in ((Z, 'Z'), (lengthscales, 'lengthscales'), (norm_x, 'norm_x'), (level_scale, 'level_scale')):
        _same(t, X, n)
    N, L, d = X.shape
    m = int(num_levels)
    T = m * (m + 1) // 2
    if Z.shape[0] != T or Z.shape[-1] != d or Z.dim() != (4 if increments else 3) or (increments and Z.shape[2] != 2):
        raise ValueError("Z must be [%d, Mz, %sd=%d], got %s" % (T, '2, ' if increments else '', d, tuple(Z.shape)))
    Mz = Z.shape[1]
    cb, ce = (0, N) if cols is None else cols
    shape = (m + 1, Mz, ce - cb) if return_levels else (Mz, ce - cb)
    out = torch.empty(shape, dtype=X.dtype, device=X.device)
    desc = make_desc(base, m, order, difference, base_params)
    epi = make_epilogue(None, norm_x, level_scale, 0.0, return_levels)
    wsb = lib.gpsig_tens_vs_seq_workspace_bytes(Mz, int(increments), N, L, d, m, X.element_size())
    ws = workspace(wsb, X.device)
    fn = getattr(lib, 'gpsig_tens_vs_seq_' + sfx(X.dtype))
    with torch.cuda.device(X.device):
        check(fn(desc, ptr(Z), Mz, int(increments), ptr(X), N, L, d, ptr(lengthscales), cb, ce, epi, ptr(out),
                 ptr(ws), wsb, stream_handle(X.device)))
    return out


def tens_tens(Z, num_levels, increments=False, base='linear', base_params=(), lengthscales=None, level_scale=None,
              return_levels=True):
    """Inducing tensors vs themselves (``_K_tens``, gpsig/kernels.py:264-284)."""
    require_cuda(Z, 'Z')
    _same(lengthscales, Z, 'lengthscales')
    _same(level_scale, Z, 'level_scale')
    m = int(num_levels)
    T = m * (m + 1) // 2
    if Z.shape[0] != T or Z.dim() != (4 if increments else 3) or (increments and Z.shape[2] != 2):
        raise ValueError("Z must be [%d, Mz, %sd], got %s" % (T, '2, ' if increments else '', tuple(Z.shape)))
    Mz, d = Z.shape[1], Z.shape[-1]
    shape = (m + 1, Mz, Mz) if return_levels else (Mz, Mz)
    out = torch.empty(shape, dtype=Z.dtype, device=Z.device)
    desc = make_desc(base, m, 1, True, base_params)
    epi = make_epilogue(None, None, level_scale, 0.0, return_levels)
    fn = getattr(lib, 'gpsig_tens_tens_' + sfx(Z.dtype))
    with torch.cuda.device(Z.device):
        check(fn(desc, ptr(Z), Mz, int(increments), d, ptr(lengthscales), epi, ptr(out), stream_handle(Z.device)))
    return out


def pde(X, Y=None, dyadic_order=0, gram=False, lengthscales=None, sigma=1.0, rows=None):
    """Untruncated (Goursat PDE) kernel, float64 (``sig_kern_diag``, gpsig/sigKer_fast.pyx:15-62).

    ``gram=False``: ``[R]`` values ``sigma * k(x_a, y_a)`` (``Y=None`` -> the reference's Kdiag);
    ``gram=True``: ``[R, N2]``.
    """
    require_cuda(X, 'X')
    if X.dtype != torch.float64:
        raise TypeError("the PDE kernel is float64 only (as the reference: sigKer_fast.pyx:15)")
    _same(Y, X, 'Y')
    _same(lengthscales, X, 'lengthscales')
    N, L, d = X.shape
    N2, L2 = (N, L) if Y is None else Y.shape[:2]
    rb, re = (0, N) if rows is None else rows
    out = torch.empty((re - rb, N2) if gram else (re - rb,), dtype=X.dtype, device=X.device)
    mode = _lib.PDE_GRAM if gram else _lib.PDE_DIAG
    wsb = lib.gpsig_pde_workspace_bytes(N, L, 0 if Y is None else N2, L2, d, int(dyadic_order), mode)
    ws = workspace(wsb, X.device)
    with torch.cuda.device(X.device):
        check(lib.gpsig_pde_f64(ptr(X), N, L, ptr(Y), N2, L2, d, ptr(lengthscales), int(dyadic_order), mode, rb, re,
                                float(sigma), ptr(out), ptr(ws), wsb, stream_handle(X.device)))
    return out


# ---- reverse mode (include/gpsig_b200.h: gpsig_*_vjp_*) -------------------------------------------
_VJP_WS_BYTES = 4 << 30      # scratch budget of one trunc_vjp launch; larger problems are chunked over rows


def trunc_vjp(X, X2, grad_levels, num_levels, base='linear', base_params=(), difference=True, diag=False, order=1):
    """dL/dX, dL/dX2 of the raw levels of ``trunc_gram`` (``diag=False``, grad_levels [m+1, N, N2]) or
    ``trunc_diag`` (``diag=True``, grad_levels [m+1, N]).  X, X2 are the coordinates the forward was called
    with (already scaled).  Returns (gradX, gradX2) (gradX2 is None when X2 is None)."""
    require_cuda(X, 'X')
    _same(X2, X, 'X2')
    _same(grad_levels, X, 'grad_levels')
    N, L, d = X.shape
    N2, L2 = (N, L) if X2 is None else X2.shape[:2]
    m = int(num_levels)
    gradX = torch.zeros_like(X)
    gradX2 = None if X2 is None else torch.zeros_like(X2)
    order = max(1, min(int(order), m))
    desc = make_desc(base, m, order, difference, base_params)
    esz = X.element_size()
    fn = getattr(lib, 'gpsig_trunc_vjp_' + sfx(X.dtype))
    per_row = 1 if diag else N2
    rows_per_call = max(1, int(_VJP_WS_BYTES // max(1, lib.gpsig_trunc_vjp_workspace_bytes_order(per_row, L, L2, m, order, int(difference), esz))))
    with torch.cuda.device(X.device):
        for rb in range(0, N, rows_per_call):
            re = min(N, rb + rows_per_call)
            g = grad_levels[:, rb:re].contiguous()
            wsb = lib.gpsig_trunc_vjp_workspace_bytes_order((re - rb) * per_row, L, L2, m, order, int(difference), esz)
            ws = workspace(wsb, X.device)
            check(fn(desc, ptr(X), N, L, ptr(X2), N2, L2, d, _lib.PDE_DIAG if diag else _lib.PDE_GRAM, rb, re, ptr(g),
                     ptr(gradX), ptr(gradX2), ptr(ws), wsb, stream_handle(X.device)))
    return gradX, gradX2


def tens_vs_seq_vjp(Z, X, grad_levels, num_levels, increments=False, base='linear', base_params=(), difference=True, order=1):
    """dL/dZ, dL/dX of the raw levels [m+1, Mz, N] of ``tens_vs_seq``."""
    require_cuda(X, 'X')
    _same(Z, X, 'Z')
    _same(grad_levels, X, 'grad_levels')
    N, L, d = X.shape
    m = int(num_levels)
    Mz = Z.shape[1]
    gradZ, gradX = torch.zeros_like(Z), torch.zeros_like(X)
    desc = make_desc(base, m, max(1, min(int(order), m)), difference, base_params)
    esz = X.element_size()
    if (d > 16 or (d > 8 and esz == 8)) and int(order) == 1 and base in ('rbf', 'linear') and not os.environ.get('GPSIG_TENS_VJP_NO_LIFT'):
        return _tens_vs_seq_vjp_lifted(Z, X, grad_levels, m, increments, base, desc)
    fn = getattr(lib, 'gpsig_tens_vs_seq_vjp_' + sfx(X.dtype))
    cols_per_call = max(1, int(_VJP_WS_BYTES // max(1, lib.gpsig_tens_vs_seq_vjp_workspace_bytes(Mz, 1, L, m, esz))))
    with torch.cuda.device(X.device):
        for cb in range(0, N, cols_per_call):
            ce = min(N, cb + cols_per_call)
            g = grad_levels[:, :, cb:ce].contiguous()
            wsb = lib.gpsig_tens_vs_seq_vjp_workspace_bytes(Mz, ce - cb, L, m, esz)
            ws = workspace(wsb, X.device)
            check(fn(desc, ptr(Z), Mz, int(increments), ptr(X), N, L, d, cb, ce, ptr(g), ptr(gradZ), ptr(gradX),
                     ptr(ws), wsb, stream_handle(X.device)))
    return gradZ, gradX


_LIFT_GD_BYTES = 1 << 30     # lifted adjoint values of one block of sequences


def _tens_vs_seq_vjp_lifted(Z, X, grad_levels, m, increments, base, desc):
    """Reverse mode of ``tens_vs_seq`` at state dimensions above 16 (float64: above 8), lifted form: the library returns
    ``Gd[r, e] = dL/dkappa(z point r, x point e)`` (times kappa for RBF) for a block of sequences
    (``gpsig_tens_vs_seq_lifted_gd_*``: GEMM + reverse time recursion, csrc/tens_lifted_kernels.cuh); the two remaining
    contractions are plain GEMMs, ``dL/dZ = Gd X - rowsum(Gd) Z`` and ``dL/dX = Gd^T Z - colsum(Gd) X`` (linear: without
    the sums), done with ``torch.matmul``."""
    N, L, d = X.shape
    T, Mz = Z.shape[0], Z.shape[1]
    zs = 2 if increments else 1
    R = T * zs * Mz
    esz = X.element_size()
    gradZ, gradX = torch.zeros_like(Z), torch.zeros_like(X)
    Zr = (Z.permute(0, 2, 1, 3) if increments else Z).reshape(R, d).contiguous()          # row r = (c zs + s) Mz + z
    gZr = torch.zeros_like(Zr)
    fn = getattr(lib, 'gpsig_tens_vs_seq_lifted_gd_' + sfx(X.dtype))
    cols_per_call = max(1, int(_LIFT_GD_BYTES // max(1, R * L * esz)))
    if cols_per_call >= 32:
        cols_per_call = cols_per_call // 32 * 32
    wsb = lib.gpsig_tens_vs_seq_lifted_gd_workspace_bytes(Mz, int(increments), N, L, d, m, esz)
    rbf = base == 'rbf'
    with torch.cuda.device(X.device):
        ws = workspace(wsb, X.device)
        if rbf:
            Zr1 = torch.cat([Zr, torch.ones((R, 1), dtype=X.dtype, device=X.device)], dim=1)
            ones_e = torch.ones((L * min(N, cols_per_call), 1), dtype=X.dtype, device=X.device)
        for cb in range(0, N, cols_per_call):
            ce = min(N, cb + cols_per_call)
            C = ce - cb
            g = grad_levels[:, :, cb:ce].contiguous()
            Gd = torch.empty((R, L * C), dtype=X.dtype, device=X.device)
            check(fn(desc, ptr(Z), Mz, int(increments), ptr(X), N, L, d, cb, ce, ptr(g), ptr(Gd), ptr(ws), wsb,
                     stream_handle(X.device)))
            Xe = X[cb:ce].permute(1, 0, 2).reshape(L * C, d)                              # column e = l C + n
            if rbf:                                     # the row / column sums ride along as one more GEMM column of ones
                P = Gd @ torch.cat([Xe, ones_e[:L * C]], dim=1)
                Q = Gd.t() @ Zr1
                dZ = P[:, :d] - P[:, d:] * Zr
                dX = Q[:, :d] - Q[:, d:] * Xe
            else:
                dZ = Gd @ Xe
                dX = Gd.t() @ Zr
            gZr += dZ
            gradX[cb:ce] = dX.reshape(L, C, d).permute(1, 0, 2)
            del Gd
    gradZ.copy_(gZr.reshape(T, zs, Mz, d).permute(0, 2, 1, 3) if increments else gZr.reshape(T, Mz, d))
    return gradZ, gradX


def tens_tens_vjp(Z, grad_levels, num_levels, increments=False, base='linear', base_params=()):
    """dL/dZ of the raw levels [m+1, Mz, Mz] of ``tens_tens``."""
    require_cuda(Z, 'Z')
    _same(grad_levels, Z, 'grad_levels')
    m = int(num_levels)
    Mz, d = Z.shape[1], Z.shape[-1]
    gradZ = torch.empty_like(Z)
    desc = make_desc(base, m, 1, True, base_params)
    fn = getattr(lib, 'gpsig_tens_tens_vjp_' + sfx(Z.dtype))
    with torch.cuda.device(Z.device):
        check(fn(desc, ptr(Z), Mz, int(increments), d, ptr(grad_levels.contiguous()), ptr(gradZ), stream_handle(Z.device)))
    return gradZ


_PDE_VJP_WS_BYTES = 96 << 30   # cap on the grid scratch of one pde_kdiag_vjp call (8 P^2 bytes per sequence in flight)


def pde_kdiag_vjp(X, grad_out, dyadic_order=0, rev_solver=0, rows=None):
    """The reference's gradient of the PDE ``Kdiag`` w.r.t. the (scaled) sequences
    (``_KdiagGrad``, gpsig/kernels_pde.py:465-509; ``rev_solver=1``: ``_untrunc_cov_grad.py:25-77``).
    X ``[N, L, d]`` float64, grad_out ``[R]``; returns gradX ``[N, L, d]`` (zero outside ``rows``)."""
    require_cuda(X, 'X')
    if X.dtype != torch.float64:
        raise TypeError("the PDE kernel is float64 only (as the reference: sigKer_fast.pyx:15)")
    _same(grad_out, X, 'grad_out')
    N, L, d = X.shape
    rb, re = (0, N) if rows is None else rows
    if grad_out.numel() != re - rb:
        raise ValueError("grad_out must have one entry per row")
    full = rb == 0 and re == N
    gradX = torch.empty_like(X) if full else torch.zeros_like(X)
    one = lib.gpsig_pde_kdiag_vjp_workspace_bytes(1, L, d, int(dyadic_order))
    # at most half of the device memory (a function of the device, not of what is free now: the same call must ask
    # the caching allocator for the same block every time)
    budget = min(_PDE_VJP_WS_BYTES, torch.cuda.get_device_properties(X.device).total_memory // 2)
    nseq = max(1, min(re - rb, budget // max(1, one)))
    wsb = one * nseq
    ws = workspace(wsb, X.device)
    with torch.cuda.device(X.device):
        check(lib.gpsig_pde_kdiag_vjp_f64(ptr(X), N, L, d, int(dyadic_order), int(rev_solver), rb, re, ptr(grad_out),
                                          ptr(gradX), ptr(ws), wsb, stream_handle(X.device)))
    return gradX
