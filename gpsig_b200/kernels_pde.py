"""Untruncated (PDE) signature kernel -- drop-in for ``gpsig.kernels_pde``.

``UntruncSignatureKernel.Kdiag`` (gpsig/kernels_pde.py:160-185) solves the
Goursat PDE of each sequence against itself and keeps ``sigma * K[-1, -1]``.
The reference does this either on the host through ``tf.py_func`` + Cython
(``implementation='cython'``, gpsig/sigKer_fast.pyx) or with its TF custom op
(``implementation='gpu_op'``, limited to 2^order (L-1) + 1 < 1024).  Here both
spellings run the same sm_100a wavefront kernel (``csrc/pde_kernel.cuh``);
the length limit does not apply.

As in the reference, ``Kdiag`` never evaluates ``_base_kern`` (kernels_pde.py:177
is commented out): the PDE path is linear-base-kernel only, so ``SignatureRBF``
and ``SignatureLinear`` give the same ``Kdiag`` (SURVEY.md note N7).
``K(X, X2)`` (extension; the reference declares ``compute_K`` but never defines
``K``) returns the full untruncated Gram.
"""
import numpy as np
import torch

from . import autograd, ops, settings
from . import lags as _lags
from .kernels import as_device_tensor


class UntruncSignatureKernel:
    """gpsig/kernels_pde.py:28-101.  Keywords ``order=``, ``lengthscales=`` (SURVEY.md note N6)."""

    def __init__(self, input_dim, num_features, lengthscales=1, order=0, num_lags=None, implementation='gpu_op',
                 num_levels=None, active_dims=None, name=None):
        self.input_dim = int(input_dim)
        self.name = name
        self.active_dims = slice(input_dim) if active_dims is None else active_dims
        self.num_features = num_features
        self.len_examples = self._validate_number_of_features(input_dim, num_features)
        assert implementation in ['cython', 'gpu_op'], "implementation should be 'cython' or 'gpu_op'"
        self.implementation = implementation
        self.order = int(order)
        if self.order < 0:
            raise ValueError('order must be a nonnegative integer')
        self.sigma = 1.
        self.num_levels = num_levels
        if num_lags is None:
            self.num_lags = 0
        else:
            if not isinstance(num_lags, int) or num_lags < 0:
                raise ValueError('The variable num_lags most be a nonnegative integer or None.')
            self.num_lags = int(num_lags)
            if num_lags > 0:                                            # kernels_pde.py:64-69
                self.lags = 0.1 * np.asarray(range(1, num_lags + 1), dtype=np.float64)
                gamma = 1. / np.asarray(range(1, self.num_lags + 2), dtype=np.float64)
                self.gamma = gamma / np.sum(gamma)
        if lengthscales is not None:
            self.lengthscales = self._validate_signature_param("lengthscales", lengthscales, self.num_features)
        else:
            self.lengthscales = None

    def _validate_number_of_features(self, input_dim, num_features):
        if input_dim % num_features == 0:
            return int(input_dim / num_features)
        raise ValueError("The arguments num_features and input_dim are not consistent.")

    def _validate_signature_param(self, name, value, length):
        value = value * np.ones(length, dtype=np.float64)
        correct_shape = () if length == 1 else (length,)
        if np.asarray(value).squeeze().shape != correct_shape:
            raise ValueError("shape of parameter {} is not what is expected ({})".format(name, length))
        return value

    def _slice(self, X):
        if isinstance(self.active_dims, slice):
            return X[:, self.active_dims]
        return X[:, torch.as_tensor(np.asarray(self.active_dims), device=X.device)]

    def _seq(self, X):
        return X.reshape(X.shape[0], -1, self.num_features).contiguous()

    def _t(self, v, like):
        return v.to(dtype=like.dtype, device=like.device) if torch.is_tensor(v) else \
            torch.as_tensor(np.asarray(v, dtype=np.float64), dtype=like.dtype, device=like.device)

    def _apply_scaling_and_lags_to_sequences(self, Xs):
        """kernels_pde.py:135-157: lagged copies by linear interpolation, division by the lengthscales, ``gamma``
        weights; [N, L, d] -> [N, L, d (num_lags + 1)].  Torch preprocessing on the device (differentiable)."""
        N, L, _ = Xs.shape
        X = _lags.add_lags_to_sequences(Xs, self._t(self.lags, Xs)) if self.num_lags > 0 else Xs[:, :, None, :]
        if self.lengthscales is not None:
            X = X / self._t(self.lengthscales, Xs)[None, None, None, :]
        if self.num_lags > 0:
            X = X * self._t(self.gamma, Xs)[None, None, :, None]
        return X.reshape(N, L, -1).contiguous()

    def _ls(self, like):
        if self.lengthscales is None:
            return None
        return torch.as_tensor(self.lengthscales, dtype=like.dtype, device=like.device)

    def Kdiag(self, X, presliced=False, name=None, rows=None):
        """sigma * k(x_i, x_i) for every sequence (kernels_pde.py:160-185)."""
        X = as_device_tensor(X, "X")
        if not presliced:
            X = self._slice(X)
        Xs = self._seq(X)
        params = (self.lengthscales, self.sigma) + ((self.lags, self.gamma) if self.num_lags > 0 else ())
        if torch.is_grad_enabled() and (Xs.requires_grad or any(torch.is_tensor(p) and p.requires_grad for p in params)):
            return self._Kdiag_differentiable(Xs, rows)
        if self.num_lags > 0:
            return ops.pde(self._apply_scaling_and_lags_to_sequences(Xs), None, dyadic_order=self.order, gram=False,
                           sigma=float(self.sigma), rows=rows)
        return ops.pde(Xs, None, dyadic_order=self.order, gram=False, lengthscales=self._ls(Xs),
                       sigma=float(self.sigma), rows=rows)

    def _Kdiag_differentiable(self, Xs, rows=None):
        """Training path: the scaling by the lengthscales (kernels_pde.py:150-151) and ``sigma *`` (:185) are
        torch expressions; the solve carries the reference's own backward (``autograd._PdeKdiag``):
        1st-order reversed solve for ``implementation='cython'`` (sigKer_fast.pyx:50,61), 2nd-order for
        ``'gpu_op'`` (_untrunc_cov_grad.py:44-49)."""
        if rows is not None:
            Xs = Xs[rows[0]:rows[1]]
        Xs = self._apply_scaling_and_lags_to_sequences(Xs)
        rev_solver = 0 if self.implementation == 'cython' else 1
        return self.sigma * autograd.pde_kdiag(Xs, self.order, rev_solver)

    def K(self, X, X2=None, presliced=False, rows=None):
        """Full untruncated Gram sigma * k(x_i, y_j) (extension, see module docstring)."""
        X = as_device_tensor(X, "X")
        X2 = None if X2 is None else as_device_tensor(X2, "X2")
        if not presliced:
            X = self._slice(X)
            X2 = None if X2 is None else self._slice(X2)
        Xs = self._seq(X)
        X2s = None if X2 is None else self._seq(X2)
        if self.num_lags > 0:
            Xs = self._apply_scaling_and_lags_to_sequences(Xs)
            X2s = None if X2s is None else self._apply_scaling_and_lags_to_sequences(X2s)
            return ops.pde(Xs, X2s, dyadic_order=self.order, gram=True, sigma=float(self.sigma), rows=rows)
        return ops.pde(Xs, X2s, dyadic_order=self.order, gram=True, lengthscales=self._ls(Xs),
                       sigma=float(self.sigma), rows=rows)

    @staticmethod
    def _dev(A):
        return torch.as_tensor(np.ascontiguousarray(A), dtype=torch.float64, device=settings.device())

    def compute_K(self, X, Y):
        return self.K(self._dev(X), self._dev(Y)).cpu().numpy()

    def compute_K_symm(self, X):
        """The reference's ``compute_K_symm`` returns ``Kdiag`` (kernels_pde.py:110-112)."""
        return self.Kdiag(self._dev(X)).cpu().numpy()


class SignatureRBF(UntruncSignatureKernel):
    """kernels_pde.py:404-417 (positional ``order, num_levels`` as in the reference's subclass)."""

    def __init__(self, input_dim, num_features, order, num_levels, **kwargs):
        UntruncSignatureKernel.__init__(self, input_dim, num_features, order=order, num_levels=num_levels, **kwargs)


class SignatureLinear(UntruncSignatureKernel):
    """kernels_pde.py:420-437."""

    def __init__(self, input_dim, num_features, order, num_levels, **kwargs):
        UntruncSignatureKernel.__init__(self, input_dim, num_features, order=order, num_levels=num_levels, **kwargs)
