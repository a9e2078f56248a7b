"""Functional layer: one Python function per C-ABI entry point.

Inputs are contiguous CUDA tensors ``[N, L, d]`` (float32 or float64); the
functions allocate the output and the workspace with torch, pass raw device
pointers plus the current CUDA stream through ctypes and return the output
tensor.  Nothing here computes: if the CUDA library is missing the import of
``_lib`` has already failed.
"""
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
               norm_x=None, norm_x2=None, level_scale=None, sym_jitter=0.0, return_levels=True, rows=None):
    """Rows ``rows=(begin, end)`` (default all) of the truncated signature Gram.

    Replaces ``SignatureKernel._K_seq`` + epilogue (gpsig/kernels.py:209-238, 431-477).
    Returns ``[m+1, R, N2]`` (``return_levels``) or ``[R, N2]``.
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
    out = torch.empty(shape, dtype=X.dtype, device=X.device)
    desc = make_desc(base, m, order, difference, base_params)
    epi = make_epilogue(norm_x, norm_x2, level_scale, sym_jitter, return_levels)
    esz = X.element_size()
    wsb = lib.gpsig_trunc_gram_workspace_bytes(N, L, 0 if X2 is None else N2, L2, d, esz)
    ws = workspace(wsb, X.device)
    fn = getattr(lib, 'gpsig_trunc_gram_' + sfx(X.dtype))
    with torch.cuda.device(X.device):
        check(fn(desc, ptr(X), N, L, ptr(X2), N2, L2, d, ptr(lengthscales), rb, re, epi, ptr(out), ptr(ws), wsb,
                 stream_handle(X.device)))
    return out


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
