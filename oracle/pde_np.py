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
