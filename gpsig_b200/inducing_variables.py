"""Inducing variables of the sparse variational GP -- drop-in for ``gpsig.inducing_variables``
(the callers of the hot path).

The reference registers ``Kuu`` / ``Kuf`` / ``Kuu_Kuf_Kff`` through GPflow's multiple dispatch on
``(InducingTensors | InducingSequences, SignatureKernel)`` (inducing_variables.py:51-137).  Here they are
plain functions that switch on the feature class; every covariance comes from the CUDA kernels through
``gpsig_b200.kernels`` (differentiable when ``Z`` / ``W`` / kernel parameters require grad), and the
per-level mixing with ``W`` (``learn_weights``) is torch ``matmul`` on ``[m, Mz, .]`` tensors.
"""
import numpy as np
import torch

from . import settings
from .kernels import SignatureKernel


class SignatureInducing:
    """inducing_variables.py:14-26.  ``Z`` (and ``W``) are torch tensors on the device; set
    ``requires_grad_()`` on them to train them."""

    def __init__(self, Z, num_levels, learn_weights=False):
        self.Z = Z if torch.is_tensor(Z) else torch.as_tensor(np.asarray(Z), dtype=settings.float_type, device=settings.device())
        self.learn_weights = learn_weights
        if learn_weights:                                   # :25-26  identity mixing per level
            n = len(self)
            self.W = torch.eye(n, dtype=self.Z.dtype, device=self.Z.device)[None].repeat(num_levels, 1, 1)


class InducingTensors(SignatureInducing):
    """inducing_variables.py:28-49: ``Z [m(m+1)/2, Mz, d]`` or ``[m(m+1)/2, Mz, 2, d]`` with ``increments``."""

    def __init__(self, Z, num_levels, increments=False, **kwargs):
        len_tensors = int(num_levels * (num_levels + 1) / 2)
        assert Z.shape[0] == len_tensors
        if increments:
            assert Z.ndim == 4
            assert Z.shape[2] == 2
        super().__init__(Z, num_levels, **kwargs)
        self.len_tensors = len_tensors
        self.increments = increments

    def __len__(self):
        return self.Z.shape[1]


class InducingSequences(SignatureInducing):
    """inducing_variables.py:88-98: ``Z [num_inducing, len_inducing, num_features]``."""

    def __init__(self, Z, num_levels, **kwargs):
        super().__init__(Z, num_levels, **kwargs)
        self.len_inducing = self.Z.shape[1]

    def __len__(self):
        return self.Z.shape[0]

    def _flat(self):
        return self.Z.reshape(self.Z.shape[0], -1)


def _mix_zz(W, Kzz):
    """Kzz[0] + sum_l W_l Kzz[l] W_l^T   (inducing_variables.py:56,82,105,128)."""
    return Kzz[0] + torch.matmul(torch.matmul(W, Kzz[1:]), W.transpose(-1, -2)).sum(dim=0)


def _mix_zx(W, Kzx):
    """Kzx[0] + sum_l W_l Kzx[l]   (inducing_variables.py:57,72,119,129)."""
    return Kzx[0] + torch.matmul(W, Kzx[1:]).sum(dim=0)


def _eye_like(K):
    return torch.eye(K.shape[0], dtype=K.dtype, device=K.device)


def _check(feat, kern):
    if not isinstance(kern, SignatureKernel):
        raise NotImplementedError("Kuu/Kuf are defined for SignatureKernel only (inducing_variables.py:51,65,77)")
    if not isinstance(feat, (InducingTensors, InducingSequences)):
        raise NotImplementedError("unknown inducing-variable class %r" % type(feat))


def Kuu(feat, kern, *, jitter=0.0, full_f_cov=False):
    """inducing_variables.py:77-86 (tensors), :101-110 (sequences)."""
    _check(feat, kern)
    if isinstance(feat, InducingTensors):
        if feat.learn_weights:
            Kzz = _mix_zz(feat.W, kern.K_tens(feat.Z, return_levels=True, increments=feat.increments))
        else:
            Kzz = kern.K_tens(feat.Z, increments=feat.increments)
    else:
        if feat.learn_weights:
            Kzz = _mix_zz(feat.W, kern.K(feat._flat(), return_levels=True, presliced=True))
        else:
            Kzz = kern.K(feat._flat(), presliced=True)
    return Kzz + jitter * _eye_like(Kzz)


def Kuf(feat, kern, X_new):
    """inducing_variables.py:65-75 (tensors), :112-121 (sequences)."""
    _check(feat, kern)
    if isinstance(feat, InducingTensors):
        if feat.learn_weights:
            return _mix_zx(feat.W, kern.K_tens_vs_seq(feat.Z, X_new, return_levels=True, increments=feat.increments))
        return kern.K_tens_vs_seq(feat.Z, X_new, increments=feat.increments)
    if feat.learn_weights:
        return _mix_zx(feat.W, kern.K(feat._flat(), X_new, presliced_X=True, return_levels=True))
    return kern.K(feat._flat(), X_new, presliced_X=True)


def Kuu_Kuf_Kff(feat, kern, X_new, *, jitter=0.0, full_f_cov=False):
    """inducing_variables.py:51-63 (tensors), :123-137 (sequences): the three covariances of one SVGP step."""
    _check(feat, kern)
    if isinstance(feat, InducingTensors):
        covs = lambda **kw: kern.K_tens_n_seq_covs(feat.Z, X_new, full_X_cov=full_f_cov, increments=feat.increments, **kw)  # noqa: E731
    else:
        covs = lambda **kw: kern.K_seq_n_seq_covs(feat._flat(), X_new, full_X2_cov=full_f_cov, **kw)  # noqa: E731
    if feat.learn_weights:
        Kzz, Kzx, Kxx = covs(return_levels=True)
        Kzz, Kzx, Kxx = _mix_zz(feat.W, Kzz), _mix_zx(feat.W, Kzx), Kxx.sum(dim=0)
    else:
        Kzz, Kzx, Kxx = covs()
    Kzz = Kzz + jitter * _eye_like(Kzz)
    if full_f_cov:
        # the reference adds ``jitter * tf.eye(tf.shape(X)[0])`` with an undefined name ``X`` (:60, :134)
        raise NameError("name 'X' is not defined")
    return Kzz, Kzx, Kxx + jitter
