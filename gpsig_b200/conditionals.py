"""The Gaussian conditional downstream of the hot path.

``gpsig.models.SVGP._build_predict`` (models.py:109-131) feeds ``Kuu_Kuf_Kff`` into GPflow's
``base_conditional``.  GPflow is a third-party dependency that is not vendored in the reference
(``requirements.txt:8``, gpflow==1.5.1); this module restates ``gpflow/conditionals.py:base_conditional`` of that
release in torch on the device (Cholesky and triangular solves are cuSOLVER / cuBLAS library calls):

    Lm   = chol(Kmm);  A = Lm^-1 Kmn
    fvar = Knn - A^T A   (full_cov)   |   Knn - colsum(A^2)
    A    = Lm^-T A                      unless ``white``
    fmean = A^T f
    fvar += (L_q^T A)^T (L_q^T A)       with q_sqrt [R, M, M] lower-triangular or [M, R] diagonal
"""
import torch

from . import settings
from .inducing_variables import Kuu_Kuf_Kff


def base_conditional(Kmn, Kmm, Knn, f, *, full_cov=False, q_sqrt=None, white=False):
    """Kmn [M, N], Kmm [M, M], Knn [N] (or [N, N] with ``full_cov``), f [M, R], q_sqrt None | [M, R] | [R, M, M].
    Returns fmean [N, R] and fvar [N, R] (or [R, N, N])."""
    num_func = f.shape[1]
    Lm = torch.linalg.cholesky(Kmm)
    A = torch.linalg.solve_triangular(Lm, Kmn, upper=False)
    if full_cov:
        fvar = (Knn - A.transpose(0, 1) @ A)[None].repeat(num_func, 1, 1)           # R x N x N
    else:
        fvar = (Knn - (A * A).sum(dim=0))[None].repeat(num_func, 1)                 # R x N
    if not white:
        A = torch.linalg.solve_triangular(Lm.transpose(0, 1), A, upper=True)
    fmean = A.transpose(0, 1) @ f
    if q_sqrt is not None:
        if q_sqrt.dim() == 2:
            LTA = A * q_sqrt.transpose(0, 1)[:, :, None]                            # R x M x N
        elif q_sqrt.dim() == 3:
            L = torch.tril(q_sqrt)
            LTA = L.transpose(-1, -2) @ A[None].expand(num_func, -1, -1)
        else:
            raise ValueError("Bad dimension for q_sqrt: %s" % str(q_sqrt.dim()))
        if full_cov:
            fvar = fvar + LTA.transpose(-1, -2) @ LTA
        else:
            fvar = fvar + (LTA * LTA).sum(dim=1)
    if not full_cov:
        fvar = fvar.transpose(0, 1)                                                  # N x R
    return fmean, fvar


def build_predict(feat, kern, X_new, q_mu, q_sqrt, *, whiten=True, full_cov=False, return_Kzz=False):
    """``SVGP._build_predict`` for tensor / sequence inducing variables with a zero mean function
    (models.py:109-131): covariances from the CUDA kernels, then the conditional."""
    Kzz, Kzx, Kxx = Kuu_Kuf_Kff(feat, kern, X_new, jitter=settings.jitter, full_f_cov=full_cov)
    f_mean, f_var = base_conditional(Kzx, Kzz, Kxx, q_mu, full_cov=full_cov,
                                     q_sqrt=None if q_sqrt is None else torch.tril(q_sqrt), white=whiten)
    return (f_mean, f_var, Kzz) if return_Kzz else (f_mean, f_var)
