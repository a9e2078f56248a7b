"""numpy restatement of the reference's Goursat-PDE solver.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows ``gpsig/sigKer_fast.pyx:15-62`` (``sig_kern_diag``): the explicit
second-order stencil (``pyx:48``) on the dyadically refined grid with
``inc = <dx_{i>>n}, dy_{j>>n}> / 4^n`` (``pyx:41-43``) and boundary
``K[.,0] = K[0,.] = 1`` (``pyx:29-31``).  The reference only ever solves the
diagonal pairs (x_i, x_i), lower triangle; here the same stencil is written
for the full rectangular grid of an arbitrary pair (x, y).  On y == x the two
agree bit for bit (the stencil is symmetric in its two neighbours), which
``tests/test_oracle_pde.py`` checks against the compiled reference
(``oracle/_ref/sigKer_fast``) -- parity PINNED.
"""
import numpy as np


def _increments(x, y, n):
    """inc[a, I, J] = sum_k dx[a,I,k]*dy[a,J,k]/4^n on the coarse grid (pyx:41-43)."""
    dx = x[:, 1:, :] - x[:, :-1, :]
    dy = y[:, 1:, :] - y[:, :-1, :]
    factor = float(2 ** (2 * n))
    inc = np.zeros((x.shape[0], dx.shape[1], dy.shape[1]), dtype=np.float64)
    for k in range(x.shape[2]):                      # same accumulation order as the pyx loop
        inc = inc + (dx[:, :, None, k] * dy[:, None, :, k]) / factor
    return inc


def pde_grid(x, y=None, n=0, solver=1):
    """Full solution grids for paired sequences.

    x: [A, L, d], y: [A, L2, d] (y=None -> y=x).  Returns K [A, P+1, Q+1] with
    P = 2^n (L-1), Q = 2^n (L2-1).  ``solver`` 1 = explicit 2nd order
    (pyx:48), 0 = 1st order (pyx:46).
    """
    x = np.asarray(x, dtype=np.float64)
    y = x if y is None else np.asarray(y, dtype=np.float64)
    A = x.shape[0]
    inc = _increments(x, y, n)
    P, Q = (2 ** n) * (x.shape[1] - 1), (2 ** n) * (y.shape[1] - 1)
    K = np.ones((A, P + 1, Q + 1), dtype=np.float64)
    # anti-diagonal sweep, vectorised over the cells of one diagonal
    for s in range(P + Q - 1):
        i = np.arange(max(0, s - Q + 1), min(P - 1, s) + 1)
        j = s - i
        e = inc[:, i >> n, j >> n]
        k00, k01, k10 = K[:, i, j], K[:, i, j + 1], K[:, i + 1, j]
        if solver == 1:
            K[:, i + 1, j + 1] = (k01 + k10) * (1. + 0.5 * e + (1. / 12) * e ** 2) - k00 * (1. - (1. / 12) * e ** 2)
        else:
            K[:, i + 1, j + 1] = (k01 + k10) + k00 * (e - 1.)
    return K


def pde_kernel(x, y=None, n=0, solver=1):
    """k(x_a, y_a) = K[a, -1, -1] (what ``kernels_pde.py:185`` keeps)."""
    return pde_grid(x, y, n, solver)[:, -1, -1]


def pde_grid_rev(x, n=0):
    """The reference's second output ``K_rev`` (pyx:50,61): first-order stencil
    on the time-reversed path, diagonal pairs only."""
    xr = np.asarray(x, dtype=np.float64)[:, ::-1, :]
    return pde_grid(xr, None, n, solver=0)


def pde_gram(X, Y=None, n=0, solver=1, block=64):
    """[N, N2] untruncated Gram by pairing every x_i with every y_j."""
    X = np.asarray(X, dtype=np.float64)
    Y = X if Y is None else np.asarray(Y, dtype=np.float64)
    out = np.empty((X.shape[0], Y.shape[0]))
    for i in range(X.shape[0]):
        for c in range(0, Y.shape[0], block):
            yb = Y[c:c + block]
            xb = np.broadcast_to(X[i][None], (yb.shape[0],) + X[i].shape)
            out[i, c:c + block] = pde_kernel(xb, yb, n, solver)
    return out


def kdiag_grad_from_grids(x, K, K_rev, n=0):
    """Transcription of the reference's hand-written gradient of ``Kdiag`` w.r.t. the (scaled) sequences,
    given the two solution grids: ``_KdiagGrad`` (gpsig/kernels_pde.py:465-509, Cython path; the identical
    tensor algebra is in ``covariance_op/_untrunc_cov_grad.py:52-77`` for the TF op).

    x [A, L, d]; K, K_rev [A, P+1, P+1] FULL (already symmetrised, ``kernels_pde.py:481-486``) grids.
    Returns ``grad_points`` [A, L, d] for an upstream gradient of 1 on every ``K[a, -1, -1]``.
    """
    x = np.asarray(x, dtype=np.float64)
    A, M, D = x.shape
    c = 2 ** n
    inc_X = np.repeat((x[:, 1:, :] - x[:, :-1, :]) / float(c), c, axis=1)        # :489-490
    K_rev_rev = K_rev[:, ::-1, ::-1]                                             # :493-494
    KK = K[:, :-1, :-1] * K_rev_rev[:, 1:, 1:]                                   # :496
    K_grad = (1. / c) * np.einsum('apq,aqd->apd', KK, inc_X)                     # :498-502
    K_grad = K_grad.reshape(A, M - 1, c, D).sum(axis=2)                          # :504
    z = np.zeros((A, 1, D))
    return -2. * np.concatenate([K_grad, z], axis=1) + 2. * np.concatenate([z, K_grad], axis=1)   # :508


def symmetrise_lower(K):
    """``K += K^T - diag`` on a lower-triangular grid (kernels_pde.py:481-486)."""
    K = np.array(K, dtype=np.float64)
    d = np.einsum('aii->ai', K)
    return K + K.transpose(0, 2, 1) - d[:, :, None] * np.eye(K.shape[1])[None]


def pde_kdiag_grad(x, n=0, rev_solver=0):
    """d Kdiag / d x as the reference computes it: forward grid from the 2nd-order stencil, reversed-path grid
    from the 1st-order stencil (``rev_solver=0``: Cython path, pyx:50,61) or from the 2nd-order one
    (``rev_solver=1``: the TF op solves the reversed increments with the same ``untrunc_cov`` kernel,
    _untrunc_cov_grad.py:44-49)."""
    x = np.asarray(x, dtype=np.float64)
    K = pde_grid(x, None, n, solver=1)
    K_rev = pde_grid(x[:, ::-1, :], None, n, solver=rev_solver)
    return kdiag_grad_from_grids(x, K, K_rev, n)
