"""numpy restatement of the SVGP glue around the hot path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

* ``Kuu_Kuf_Kff`` with ``learn_weights``: gpsig/inducing_variables.py:51-63, 123-137.
* ``base_conditional``: the algorithm lives in the reference's un-vendored dependency **gpflow==1.5.1**
  (``requirements.txt:8``; call site gpsig/models.py:116); restated from that release's
  ``gpflow/conditionals.py:base_conditional`` (Cholesky of Kmm, two triangular solves, ``fvar = Knn - A^T A +
  (L_q^T A)^T (L_q^T A)``).  That one function is unpinned in the golden-vector sense (GPflow is not under
  /root/reference and cannot be installed here); it is checked against the closed-form Gaussian conditional instead
  (tests/test_oracle_cpu.py).  ``Kuu_Kuf_Kff`` above IS pinned: tests/golden/trunc_ref.npz holds what the reference's own
  gpsig/inducing_variables.py returns (oracle/refrun), tests/test_reference_golden_cpu.py.
"""
import numpy as np
import scipy.linalg as sla


def mix_levels(W, Kzz=None, Kzx=None):
    out = []
    if Kzz is not None:
        out.append(Kzz[0] + np.sum(np.matmul(np.matmul(W, Kzz[1:]), np.swapaxes(W, -1, -2)), axis=0))
    if Kzx is not None:
        out.append(Kzx[0] + np.sum(np.matmul(W, Kzx[1:]), axis=0))
    return out[0] if len(out) == 1 else tuple(out)


def Kuu_Kuf_Kff(kern, Z, X, increments=False, W=None, jitter=0.0, sequences=False):
    """``kern``: oracle.trunc_np.SignatureKernelOracle.  ``sequences``: Z are inducing sequences [Mz, Lz*d]."""
    rl = W is not None
    if sequences:
        Kzz, Kzx, Kxx = kern.K_seq_n_seq_covs(Z, X, return_levels=rl)
    else:
        Kzz, Kzx, Kxx = kern.K_tens_n_seq_covs(Z, X, return_levels=rl, increments=increments)
    if rl:
        Kzz, Kzx = mix_levels(W, Kzz, Kzx)
        Kxx = np.sum(Kxx, axis=0)
    return Kzz + jitter * np.eye(Kzz.shape[0]), Kzx, Kxx + jitter


def base_conditional(Kmn, Kmm, Knn, f, full_cov=False, q_sqrt=None, white=False):
    R = f.shape[1]
    Lm = np.linalg.cholesky(Kmm)
    A = sla.solve_triangular(Lm, Kmn, lower=True)
    if full_cov:
        fvar = np.tile((Knn - A.T @ A)[None], (R, 1, 1))
    else:
        fvar = np.tile((Knn - np.sum(np.square(A), axis=0))[None], (R, 1))
    if not white:
        A = sla.solve_triangular(Lm.T, A, lower=False)
    fmean = A.T @ f
    if q_sqrt is not None:
        if q_sqrt.ndim == 2:
            LTA = A * q_sqrt.T[:, :, None]
        else:
            LTA = np.matmul(np.swapaxes(np.tril(q_sqrt), -1, -2), np.tile(A[None], (R, 1, 1)))
        if full_cov:
            fvar = fvar + np.matmul(np.swapaxes(LTA, -1, -2), LTA)
        else:
            fvar = fvar + np.sum(np.square(LTA), axis=1)
    if not full_cov:
        fvar = fvar.T
    return fmean, fvar
