"""Truncated signature kernels -- drop-in for ``gpsig.kernels`` on B200.

Same class names, constructor keywords, method names, argument meaning, result
shapes and error behaviour as the reference (``gpsig/kernels.py``); the
numbers come from the hand-written sm_100a kernels behind the C ABI
(``include/gpsig_b200.h``) instead of a TensorFlow graph.  Inputs are torch
CUDA tensors in GPflow's 2-D layout ``[N, L*d]`` (``compute_*`` methods take
and return numpy, like the reference's autoflow wrappers).

Differences that are deliberate:
  * parameters are plain values (the constrained values GPflow's
    ``params_as_tensors`` would hand to the methods), no TF variables;
  * ``low_rank=True`` raises NotImplementedError (out of the hot-path scope,
    SURVEY.md 2 C8); ``num_lags > 0`` is applied as torch preprocessing
    (``gpsig_b200/lags.py``) before the CUDA kernels;
  * ``order > 1`` (higher-order recursion, gpsig/signature_algs.py:37-74, 129-160)
    runs the general kernels of ``csrc/generic_kernels.cuh`` (num_levels <= 8).
"""
import numpy as np
import torch

from . import lags as _lags
from . import ops, settings


_SIDE_STREAMS = {}


def _side_streams(device):
    """Two non-default streams per device for the independent covariance blocks of the SVGP step
    (``GPSIG_NO_STREAMS=1`` keeps everything on the caller's stream)."""
    import os
    if os.environ.get('GPSIG_NO_STREAMS') or device.type != 'cuda':
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = (torch.cuda.Stream(device), torch.cuda.Stream(device))
    return _SIDE_STREAMS[key]


def _scalar_param(value, name):
    """Base-kernel scalars (SignaturePoly.gamma, SignatureMix.mixing: trainable Parameters in the reference,
    kernels.py:1016, 1056).  They enter the CUDA kernels as constants; a tensor that requires grad would be
    detached silently, so it is refused instead."""
    if torch.is_tensor(value):
        if value.requires_grad and torch.is_grad_enabled():
            raise NotImplementedError('gpsig_b200: %s is not trainable (no adjoint w.r.t. the base-kernel scalars); '
                                      'pass a plain number or a tensor without requires_grad' % name)
        return float(value.detach().cpu())
    return float(value)


def as_device_tensor(A, what="input"):
    """Zero-copy ingestion of a device array: a torch tensor is used as it is, any other object that implements
    the DLPack protocol (``__dlpack__`` / ``__dlpack_device__``: TensorFlow / CuPy / JAX / Numba device arrays) is
    wrapped with ``torch.from_dlpack`` -- the CUDA kernels then read the producer's own memory through the raw
    pointer (north star: "TF custom op with DLPack zero-copy"; the TF side is ``tf.experimental.dlpack``)."""
    if torch.is_tensor(A):
        return A
    if not isinstance(A, np.ndarray) and hasattr(A, '__dlpack__') and hasattr(A, '__dlpack_device__'):
        return torch.from_dlpack(A)
    raise TypeError("gpsig_b200 kernels take torch CUDA tensors or DLPack-capable device arrays for %s "
                    "(use compute_* for numpy)" % what)


class SignatureKernel:
    """gpsig/kernels.py:16-134 (constructor and validators)."""

    _base = None          # name of the static kernel, set by the subclasses
    _base_params = ()

    def __init__(self, input_dim, num_features, num_levels, active_dims=None, variances=1, lengthscales=1, order=1,
                 normalization=True, difference=True, num_lags=None, low_rank=False, num_components=50,
                 rank_bound=None, sparsity='sqrt', name=None):
        self.input_dim = int(input_dim)
        self.name = name
        if active_dims is None:                                     # gpflow Kernel.__init__
            self.active_dims = slice(input_dim)
        elif isinstance(active_dims, slice):
            self.active_dims = active_dims
        else:
            self.active_dims = np.array(active_dims, dtype=np.int64)
        self.num_features = num_features
        self.num_levels = num_levels
        self.len_examples = self._validate_number_of_features(input_dim, num_features)
        self.order = num_levels if (order <= 0 or order >= num_levels) else order      # kernels.py:58

        if self.order != 1 and low_rank:                            # kernels.py:60-61
            raise NotImplementedError('Higher-order algorithms not compatible with low-rank mode (yet).')

        self.normalization = normalization
        self.difference = difference
        self.variances = self._validate_signature_param("variances", variances, num_levels + 1)
        self.sigma = 1.
        self.low_rank, self.num_components, self.rank_bound, self.sparsity = \
            self._validate_low_rank_params(low_rank, num_components, rank_bound, sparsity)
        if self.low_rank:
            raise NotImplementedError('gpsig_b200: the low-rank (Nystrom) mode is outside the B200 hot path.')

        if num_lags is None:
            self.num_lags = 0
        else:
            if not isinstance(num_lags, int) or num_lags < 0:       # kernels.py:75-76
                raise ValueError('The variable num_lags most be a nonnegative integer or None.')
            self.num_lags = int(num_lags)
            if num_lags > 0:                                        # kernels.py:78-83
                self.lags = 0.1 * np.asarray(range(1, num_lags + 1), dtype=np.float64)
                gamma = 1. / np.asarray(range(1, self.num_lags + 2), dtype=np.float64)
                self.gamma = gamma / np.sum(gamma)

        if lengthscales is not None:
            self.lengthscales = self._validate_signature_param("lengthscales", lengthscales, self.num_features)
        else:
            self.lengthscales = None

    # ---- validators (kernels.py:95-134) -------------------------------------
    def _validate_number_of_features(self, input_dim, num_features):
        if input_dim % num_features == 0:
            return int(input_dim / num_features)
        raise ValueError("The arguments num_features and input_dim are not consistent.")

    def _validate_low_rank_params(self, low_rank, num_components, rank_bound, sparsity):
        if low_rank is not None and low_rank == True:  # noqa: E712  (as the reference)
            if not type(low_rank) == bool:
                raise ValueError("Unknown low-rank argument: %s. It should be True of False." % low_rank)
            if sparsity not in ['log', 'sqrt', 'lin']:
                raise ValueError("Unknown sparsity argument %s. Possible values are 'sqrt', 'log', 'lin'" % sparsity)
            if rank_bound is not None and rank_bound <= 0:
                raise ValueError("The rank-bound in the low-rank algorithm must be either None or a positiv integer.")
            if num_components is None or num_components <= 0:
                raise ValueError("The number of components in the kernel approximation must be a positive integer.")
            if rank_bound is None:
                rank_bound = num_components
        else:
            low_rank = False
        return low_rank, num_components, rank_bound, sparsity

    def _validate_signature_param(self, name, value, length):
        value = value * np.ones(length, dtype=np.float64)
        correct_shape = () if length == 1 else (length,)
        if np.asarray(value).squeeze().shape != correct_shape:
            raise ValueError("shape of parameter {} is not what is expected ({})".format(name, length))
        return value

    # ---- helpers ---------------------------------------------------------------
    def _slice(self, X, X2):
        """gpflow 1.5 ``Kernel._slice``: keep the active input columns."""
        def sl(A):
            if A is None:
                return None
            if isinstance(self.active_dims, slice):
                return A[:, self.active_dims]
            return A[:, torch.as_tensor(self.active_dims, device=A.device)]
        return sl(X), sl(X2)

    def _seq(self, X):
        """[N, L*d] -> contiguous [N, L, d]  (kernels.py:418-420).  With lags: the scaled, lagged
        sequences [N, L, d*(num_lags+1)] of ``_apply_scaling_and_lags_to_sequences`` (kernels.py:343-365)."""
        X = as_device_tensor(X, "sequences")
        X = X.reshape(X.shape[0], -1, self.num_features)
        if self.num_lags > 0:
            X = _lags.add_lags_to_sequences(X, self._param(self.lags, X))           # [N, L, num_lags+1, d]
            if self.lengthscales is not None:
                X = X / self._param(self.lengthscales, X)[None, None, None, :]
            X = X * self._param(self._lag_gamma(), X)[None, None, :, None]
            X = X.reshape(X.shape[0], X.shape[1], -1)
        return X.contiguous()

    def _tens(self, Z, increments):
        """Inducing tensors as the CUDA kernels take them.  Without lags the lengthscales are applied
        inside the library; with lags the scaling of kernels.py:367-399 is done here."""
        Z = as_device_tensor(Z, "inducing tensors").contiguous()
        if self.num_lags > 0 and self.lengthscales is not None:
            shp = Z.shape
            Zr = Z.reshape(shp[:-1] + (self.num_lags + 1, self.num_features))
            Zr = Zr / self._param(self.lengthscales, Z) * self._param(self._lag_gamma(), Z)[:, None]
            Z = Zr.reshape(shp).contiguous()
        return Z

    def _lag_gamma(self):
        """Lag weights of kernels.py:82-83 (``self.gamma``; SignatureSpectral re-uses that name for its own block)."""
        return self.gamma

    def _param(self, value, like):
        if value is None:
            return None
        if torch.is_tensor(value):                      # a trainable parameter used on the non-differentiable path
            return value.detach().to(dtype=like.dtype, device=like.device).contiguous()
        # numpy / scalar hyper-parameters: one host -> device copy per VALUE, not per call (a pageable cudaMemcpy costs
        # 0.3 ms of host time: three of them were half of the SVGP covariance step, which is launch bound at minibatch 50)
        arr = np.ascontiguousarray(np.asarray(value, dtype=np.float64))
        cache = self.__dict__.setdefault('_param_cache', {})
        key = (arr.tobytes(), arr.shape, like.dtype, like.device)
        hit = cache.get(key)
        if hit is None:
            if len(cache) >= 64:
                cache.clear()
            hit = cache[key] = torch.as_tensor(arr, dtype=like.dtype, device=like.device).contiguous()
        return hit

    def _level_scale(self, like):
        """sigma * variances  (kernels.py:472)."""
        if torch.is_tensor(self.variances) or torch.is_tensor(self.sigma):
            return (float(self.sigma) * self._param(self.variances, like).reshape(-1)).contiguous()
        return self._param(float(self.sigma) * np.asarray(self.variances, dtype=np.float64).reshape(-1), like)

    def _kw(self, like):
        # with lags the scaling has already been applied by _seq / _tens
        return dict(base=self._base, base_params=self._get_base_params(like), order=self.order,
                    difference=self.difference,
                    lengthscales=None if self.num_lags > 0 else self._param(self.lengthscales, like))

    def _get_base_params(self, like=None):
        return self._base_params

    def _K_seq_diag(self, X):
        """Raw per-level diagonals [m+1, N] of [N, L, d] sequences (kernels.py:190-207)."""
        return ops.trunc_diag(X, self.num_levels, **self._kw(X))

    def _K_seq_diag_rows(self, X, rows):
        """Diagonals for a row block only (the other rows' entries are never read by the Gram epilogue): a rank of
        a row-sharded Gram does not repeat the whole X-side diagonal."""
        if rows is None or (rows[0] == 0 and rows[1] == X.shape[0]):
            return self._K_seq_diag(X)
        full = torch.zeros((self.num_levels + 1, X.shape[0]), dtype=X.dtype, device=X.device)
        if rows[1] > rows[0]:
            full[:, rows[0]:rows[1]] = self._K_seq_diag(X[rows[0]:rows[1]])
        return full

    # ---- differentiable path (training) ------------------------------------------
    # The reference trains lengthscales, variances, sigma and the inducing tensors through TensorFlow
    # autodiff of these methods (training.py; inducing_variables.py:51-137).  Here: when any input or
    # parameter is a torch tensor that requires grad, the methods below route through the CUDA ops'
    # hand-written adjoints (gpsig_b200/autograd.py -> csrc/vjp_kernels.cuh); scaling, normalisation and
    # the sigma * variances weighting are the same formulas written as differentiable torch expressions.
    # Any order (order > 1: the general adjoint kernels, csrc/generic_vjp_kernels.cuh); not the spectral base kernel.
    def _wants_grad(self, *tensors):
        if not torch.is_grad_enabled():
            return False
        return any(torch.is_tensor(t) and t.requires_grad
                   for t in tensors + (self.variances, self.lengthscales, self.sigma))

    def _tparam(self, value, like):
        if torch.is_tensor(value):
            return value.to(dtype=like.dtype, device=like.device)
        return self._param(value, like)

    def _d_check(self):
        pass

    def _d_seq(self, X):
        X = X.reshape(X.shape[0], -1, self.num_features)
        ls = None if self.lengthscales is None else self._tparam(self.lengthscales, X)
        if self.num_lags > 0:
            X = _lags.add_lags_to_sequences(X, self._param(self.lags, X))
            if ls is not None:
                X = X / ls[None, None, None, :]
            X = X * self._tparam(self._lag_gamma(), X)[None, None, :, None]
            return X.reshape(X.shape[0], X.shape[1], -1)
        return X if ls is None else X / ls

    def _d_tens(self, Z):
        if self.lengthscales is None:
            return Z
        ls = self._tparam(self.lengthscales, Z)
        if self.num_lags > 0:
            shp = Z.shape
            Zr = Z.reshape(shp[:-1] + (self.num_lags + 1, self.num_features))
            return (Zr / ls * self._tparam(self._lag_gamma(), Z)[:, None]).reshape(shp)
        return Z / ls

    def _d_scale(self, like):
        sigma = self.sigma if torch.is_tensor(self.sigma) else float(self.sigma)
        return sigma * self._tparam(self.variances, like).reshape(-1)

    def _d_bk(self, order=True):
        kw = dict(base=self._base, base_params=self._get_base_params())
        if order:
            kw['order'] = self.order
        return kw

    def _d_gram(self, Xs, X2s, return_levels):
        """Differentiable K(X, X2) from scaled sequences (kernels.py:431-477)."""
        from . import autograd as ag
        sc = self._d_scale(Xs)
        lv = ag.trunc_levels(Xs, X2s, self.num_levels, difference=self.difference, **self._d_bk())
        if self.normalization:
            jitter = settings.jitter
            n1 = ag.trunc_diag_levels(Xs, self.num_levels, difference=self.difference, **self._d_bk()) + jitter
            if X2s is None:
                lv = lv + jitter * torch.eye(Xs.shape[0], dtype=Xs.dtype, device=Xs.device)[None]
                n2 = n1
            else:
                n2 = ag.trunc_diag_levels(X2s, self.num_levels, difference=self.difference, **self._d_bk()) + jitter
            lv = lv / (torch.sqrt(n1)[:, :, None] * torch.sqrt(n2)[:, None, :])
        lv = lv * sc[:, None, None]
        return lv if return_levels else lv.sum(dim=0)

    def _d_tens_n_seq(self, Z, Xs, increments, full_X_cov, return_levels, want=('zz', 'zx', 'xx')):
        from . import autograd as ag
        self._d_check()
        Zs = self._d_tens(Z)
        sc = self._d_scale(Xs if Xs is not None else Zs)
        m, jitter = self.num_levels, settings.jitter
        out = {}
        want_diag = Xs is not None and ('zx' in want or 'xx' in want) and (self.normalization or not full_X_cov)
        # Kzz, the diagonal and Kzx are independent and, at minibatch sizes, each too small to fill the GPU: issue the
        # first two on side streams.  Autograd replays every node on the stream of its forward, so the three adjoints
        # overlap as well.
        side = _side_streams(Zs.device) if ('zz' in want and 'zx' in want and want_diag) else None
        cur = torch.cuda.current_stream(Zs.device) if side else None
        diag = None
        if side:
            for st in side:
                st.wait_stream(cur)
            with torch.cuda.stream(side[0]):
                Zs.record_stream(side[0]); sc.record_stream(side[0])
                out['zz'] = ag.tens_tens_levels(Zs, m, increments=increments, **self._d_bk(order=False)) * sc[:, None, None]
            with torch.cuda.stream(side[1]):
                Xs.record_stream(side[1])
                diag = ag.trunc_diag_levels(Xs, m, difference=self.difference, **self._d_bk())
        else:
            if 'zz' in want:
                out['zz'] = ag.tens_tens_levels(Zs, m, increments=increments, **self._d_bk(order=False)) * sc[:, None, None]
            if want_diag:
                diag = ag.trunc_diag_levels(Xs, m, difference=self.difference, **self._d_bk())
        if 'zx' in want:
            kzx = ag.tens_vs_seq_levels(Zs, Xs, m, increments=increments, difference=self.difference, **self._d_bk())
            if side:
                for st in side:
                    cur.wait_stream(st)
                out['zz'].record_stream(cur); diag.record_stream(cur)
            if self.normalization:
                kzx = kzx / torch.sqrt(diag + jitter)[:, None, :]
            out['zx'] = kzx * sc[:, None, None]
        if 'xx' in want:
            if full_X_cov:
                out['xx'] = self._d_gram(Xs, None, True)
            elif self.normalization:
                out['xx'] = sc[:, None].repeat(1, Xs.shape[0])
            else:
                out['xx'] = diag * sc[:, None]
        if not return_levels:
            out = {k: v.sum(dim=0) for k, v in out.items()}
        return out

    def _d_seq_n_seq(self, X, X2, full_X2_cov, return_levels):
        """Differentiable K_seq_n_seq_covs (kernels.py:706-794): gradients reach the inducing sequences, the inputs,
        lengthscales, variances and sigma.  Same order of operations as the reference, including the division of
        the cross term by sqrt(diag Kxx) twice (kernels.py:746 and :783)."""
        from . import autograd as ag
        self._d_check()
        Xs, X2s = self._d_seq(X), self._d_seq(X2)
        sc = self._d_scale(Xs)
        m, jitter, bk = self.num_levels, settings.jitter, self._d_bk()
        Kxx = ag.trunc_levels(Xs, None, m, difference=self.difference, **bk)
        Kxx2 = ag.trunc_levels(Xs, X2s, m, difference=self.difference, **bk)
        if self.normalization:
            Kxx = Kxx + jitter * torch.eye(Xs.shape[0], dtype=Xs.dtype, device=Xs.device)[None]       # :742
            dsq = torch.sqrt(torch.diagonal(Kxx, dim1=1, dim2=2))                                       # :744
            Kxx = Kxx / (dsq[:, :, None] * dsq[:, None, :])
            Kxx2 = Kxx2 / dsq[:, :, None]                                                                # :746
        if full_X2_cov:                                            # only reachable without normalisation (:756)
            Kx2x2 = ag.trunc_levels(X2s, None, m, difference=self.difference, **bk) * sc[:, None, None]
        else:
            d2 = ag.trunc_diag_levels(X2s, m, difference=self.difference, **bk)
            if self.normalization:
                Kxx2 = Kxx2 / (dsq[:, :, None] * torch.sqrt(d2 + jitter)[:, None, :])                    # :783
                Kx2x2 = sc[:, None].repeat(1, X2s.shape[0])
            else:
                Kx2x2 = d2 * sc[:, None]
        Kxx, Kxx2 = Kxx * sc[:, None, None], Kxx2 * sc[:, None, None]
        if return_levels:
            return Kxx, Kxx2, Kx2x2
        return Kxx.sum(dim=0), Kxx2.sum(dim=0), Kx2x2.sum(dim=0)

    # ---- public API ------------------------------------------------------------
    supports_out = True          # K(..., out=) writes a row block in place (gpsig_b200.sharding)

    def K(self, X, X2=None, presliced=False, return_levels=False, presliced_X=False, presliced_X2=False, rows=None,
          upper_only=False, out=None):
        """Signature kernel between sequences (kernels.py:401-477).

        ``rows=(begin, end)`` (extension) restricts the computation to a row
        block of the Gram -- the unit of multi-GPU sharding.  ``upper_only`` (with ``X2=None``): the caller only
        needs the entries on and above the diagonal of its rows (``sharding.sharded_gram`` mirrors the rest).
        ``out`` (extension): preallocated result, e.g. a row slice of the full matrix.
        """
        X = as_device_tensor(X, "X")
        X2 = None if X2 is None else as_device_tensor(X2, "X2")
        if presliced:
            presliced_X = presliced_X2 = True
        if not presliced_X and not presliced_X2:
            X, X2 = self._slice(X, X2)
        elif not presliced_X:
            X, _ = self._slice(X, None)
        elif not presliced_X2 and X2 is not None:
            X2, _ = self._slice(X2, None)
        if self._wants_grad(X, X2):
            self._d_check()
            if rows is not None or out is not None:
                raise NotImplementedError('gpsig_b200: rows= / out= are not supported on the differentiable path')
            return self._d_gram(self._d_seq(X), None if X2 is None else self._d_seq(X2), return_levels)
        Xs = self._seq(X)
        jitter = settings.jitter
        kw = self._kw(Xs)
        if X2 is None:
            norm = None
            if self.normalization:                                   # kernels.py:431-434
                norm = self._K_seq_diag(Xs) + jitter
            return ops.trunc_gram(Xs, None, self.num_levels, norm_x=norm, norm_x2=norm,
                                  level_scale=self._level_scale(Xs), sym_jitter=jitter if self.normalization else 0.0,
                                  return_levels=return_levels, rows=rows, upper_only=upper_only, out=out, **kw)
        X2s = self._seq(X2)
        n1 = n2 = None
        if self.normalization:                                       # kernels.py:456-470
            n1 = self._K_seq_diag_rows(Xs, rows) + jitter
            n2 = self._K_seq_diag(X2s) + jitter
        return ops.trunc_gram(Xs, X2s, self.num_levels, norm_x=n1, norm_x2=n2, level_scale=self._level_scale(Xs),
                              return_levels=return_levels, rows=rows, out=out, **kw)

    def K_norms(self, X, presliced=False):
        """kernels.py:480-506."""
        X = as_device_tensor(X, "X")
        if not presliced:
            X, _ = self._slice(X, None)
        if self._wants_grad(X):
            from . import autograd as ag
            self._d_check()
            Xs = self._d_seq(X)
            const = self._d_scale(Xs).sum().expand(Xs.shape[0])
            return const, ag.trunc_diag_levels(Xs, self.num_levels, difference=self.difference, **self._d_bk())
        Xs = self._seq(X)
        const = self._level_scale(Xs).sum().expand(Xs.shape[0]).contiguous()
        return const, self._K_seq_diag(Xs)

    def Kdiag(self, X, presliced=False, return_levels=False):
        """kernels.py:509-541."""
        X = as_device_tensor(X, "X")
        N = X.shape[0]
        if self._wants_grad(X):
            self._d_check()
            if self.normalization:
                K = self._d_scale(X)[:, None].repeat(1, N)
            else:
                from . import autograd as ag
                if not presliced:
                    X, _ = self._slice(X, None)
                Xs = self._d_seq(X)
                K = ag.trunc_diag_levels(Xs, self.num_levels, difference=self.difference, **self._d_bk()) * self._d_scale(Xs)[:, None]
            return K if return_levels else K.sum(dim=0)
        if self.normalization:
            sc = self._level_scale(X)
            if return_levels:
                return sc[:, None].repeat(1, N)
            return sc.sum().expand(N).contiguous()
        if not presliced:
            X, _ = self._slice(X, None)
        Xs = self._seq(X)
        K = self._K_seq_diag(Xs) * self._level_scale(Xs)[:, None]
        return K if return_levels else K.sum(dim=0)

    def K_tens(self, Z, return_levels=False, increments=False):
        """Square covariance of inducing tensors (kernels.py:543-567)."""
        Z = as_device_tensor(Z, "Z")
        if self._wants_grad(Z):
            return self._d_tens_n_seq(Z, None, increments, False, return_levels, want=('zz',))['zz']
        Z = self._tens(Z, increments)   # Kzz does not depend on `order` (signature_algs.py:76-99)
        return ops.tens_tens(Z, self.num_levels, increments=increments, base=self._base,
                             base_params=self._get_base_params(Z), lengthscales=self._kw(Z)['lengthscales'],
                             level_scale=self._level_scale(Z), return_levels=return_levels)

    def K_tens_vs_seq(self, Z, X, return_levels=False, increments=False, presliced=False, cols=None):
        """Inducing tensors vs sequences (kernels.py:570-620)."""
        Z, X = as_device_tensor(Z, "Z"), as_device_tensor(X, "X")
        if not presliced:
            X, _ = self._slice(X, None)
        if self._wants_grad(Z, X):
            if cols is not None:
                raise NotImplementedError('gpsig_b200: cols= is not supported on the differentiable path')
            return self._d_tens_n_seq(Z, self._d_seq(X), increments, False, return_levels, want=('zx',))['zx']
        Xs = self._seq(X)
        norm = self._K_seq_diag(Xs) + settings.jitter if self.normalization else None      # kernels.py:604-613
        return ops.tens_vs_seq(self._tens(Z, increments), Xs, self.num_levels, increments=increments, norm_x=norm,
                               level_scale=self._level_scale(Xs), return_levels=return_levels, cols=cols,
                               **self._kw(Xs))

    def graphed_tens_n_seq_covs(self, Z, X, full_X_cov=False, increments=False):
        """``K_tens_n_seq_covs`` with its forward and its backward each captured into a CUDA graph
        (``torch.cuda.make_graphed_callables``) for a training loop with a FIXED minibatch shape: the step is launch
        bound at the reference's minibatch of 50 (about 45 kernels, 1.9 ms eager against 1.3 ms replayed).
        ``Z`` [T, Mz, (2,) d] and ``X`` [N, L * d] are sample tensors on the device that require grad (their values may
        change between calls, their shapes may not); returns ``f(Z, X) -> (Kzz, Kzx, Kxx)``, differentiable.  Tensor
        hyper-parameters that require grad are not traced by the capture (pass plain numbers / numpy arrays, or use the
        eager method)."""
        if any(torch.is_tensor(t) and t.requires_grad for t in (self.variances, self.lengthscales, self.sigma)):
            raise NotImplementedError('gpsig_b200: graphed covariances take constant hyper-parameters; '
                                      'use K_tens_n_seq_covs for trainable lengthscales / variances / sigma')
        if not (Z.requires_grad and X.requires_grad):
            raise ValueError('graphed_tens_n_seq_covs: the sample Z and X must require grad')

        def covs(Zt, Xt):
            return self.K_tens_n_seq_covs(Zt, Xt, full_X_cov=full_X_cov, increments=increments)
        return torch.cuda.make_graphed_callables(covs, (Z, X))

    def K_tens_n_seq_covs(self, Z, X, full_X_cov=False, return_levels=False, increments=False, presliced=False):
        """Kzz, Kzx, Kxx in one call (kernels.py:623-704)."""
        Z, X = as_device_tensor(Z, "Z"), as_device_tensor(X, "X")
        if not presliced:
            X, _ = self._slice(X, None)
        if self._wants_grad(Z, X):
            o = self._d_tens_n_seq(Z, self._d_seq(X), increments, full_X_cov, return_levels)
            return o['zz'], o['zx'], o['xx']
        Xs = self._seq(X)
        Z = self._tens(Z, increments)
        N = Xs.shape[0]
        jitter = settings.jitter
        sc = self._level_scale(Xs)
        kw = self._kw(Xs)
        Kzz = ops.tens_tens(Z, self.num_levels, increments=increments, base=self._base,
                            base_params=self._get_base_params(Z), lengthscales=kw['lengthscales'], level_scale=sc,
                            return_levels=return_levels)
        diag = self._K_seq_diag(Xs)
        norm = diag + jitter if self.normalization else None
        Kzx = ops.tens_vs_seq(Z, Xs, self.num_levels, increments=increments, norm_x=norm, level_scale=sc,
                              return_levels=return_levels, **kw)
        if full_X_cov:                                               # kernels.py:659-680
            Kxx = ops.trunc_gram(Xs, None, self.num_levels, norm_x=norm, norm_x2=norm, level_scale=sc,
                                 sym_jitter=jitter if self.normalization else 0.0, return_levels=return_levels, **kw)
        else:                                                        # kernels.py:682-704
            if self.normalization:
                Kxx = sc[:, None].repeat(1, N)
            else:
                Kxx = diag * sc[:, None]
            if not return_levels:
                Kxx = Kxx.sum(dim=0)
        return Kzz, Kzx, Kxx

    def K_seq_n_seq_covs(self, X, X2, full_X2_cov=False, return_levels=False, presliced=False):
        """Kxx, Kxx2, Kx2x2 for inducing sequences (kernels.py:706-794).

        Follows the reference to the letter, including that (a) with
        ``normalization`` and ``full_X2_cov=True`` it raises NameError
        (kernels.py:756-761 use undefined names) and (b) with ``normalization``
        and ``full_X2_cov=False`` the cross term is divided by sqrt(diag Kxx)
        twice (kernels.py:746 and :783).
        """
        X, X2 = as_device_tensor(X, "X"), as_device_tensor(X2, "X2")
        if not presliced:
            X2, _ = self._slice(X2, None)
        if self.normalization and full_X2_cov:
            raise NameError("name 'K_x2x2_lvls' is not defined")   # kernels.py:756
        if self._wants_grad(X, X2):
            return self._d_seq_n_seq(X, X2, full_X2_cov, return_levels)
        Xs, X2s = self._seq(X), self._seq(X2)
        N2 = X2s.shape[0]
        jitter = settings.jitter
        sc = self._level_scale(Xs)
        kw = self._kw(Xs)
        if self.normalization:
            nx = self._K_seq_diag(Xs) + jitter
            Kxx = ops.trunc_gram(Xs, None, self.num_levels, norm_x=nx, norm_x2=nx, level_scale=sc, sym_jitter=jitter,
                                 return_levels=return_levels, **kw)
            d2 = self._K_seq_diag(X2s) + jitter
            # kernels.py:746 then :783 -> divided by sqrt(nx) twice and sqrt(d2) once
            Kxx2 = ops.trunc_gram(Xs, X2s, self.num_levels, norm_x=nx, norm_x2=d2, level_scale=None,
                                  return_levels=True, **kw)
            Kxx2 = Kxx2 / torch.sqrt(nx)[:, :, None] * sc[:, None, None]
            if not return_levels:
                Kxx2 = Kxx2.sum(dim=0)
            Kx2x2 = sc[:, None].repeat(1, N2)
            if not return_levels:
                Kx2x2 = Kx2x2.sum(dim=0)
            return Kxx, Kxx2, Kx2x2
        Kxx = ops.trunc_gram(Xs, None, self.num_levels, level_scale=sc, return_levels=return_levels, **kw)
        Kxx2 = ops.trunc_gram(Xs, X2s, self.num_levels, level_scale=sc, return_levels=return_levels, **kw)
        if full_X2_cov:
            Kx2x2 = ops.trunc_gram(X2s, None, self.num_levels, level_scale=sc, return_levels=return_levels, **kw)
        else:
            Kx2x2 = self._K_seq_diag(X2s) * sc[:, None]
            if not return_levels:
                Kx2x2 = Kx2x2.sum(dim=0)
        return Kxx, Kxx2, Kx2x2

    # ---- numpy-in / numpy-out wrappers (the reference's autoflow methods, kernels.py:142-187)
    @staticmethod
    def _dev(A):
        return torch.as_tensor(np.ascontiguousarray(A), dtype=settings.float_type, device=settings.device())

    def compute_K(self, X, Y):
        return self.K(self._dev(X), self._dev(Y)).cpu().numpy()

    def compute_K_symm(self, X):
        return self.K(self._dev(X)).cpu().numpy()

    def compute_K_level_diags(self, X):
        return self.Kdiag(self._dev(X), return_levels=True).cpu().numpy()

    def compute_K_levels(self, X, X2):
        return self.K(self._dev(X), self._dev(X2), return_levels=True).cpu().numpy()

    def compute_Kdiag(self, X):
        return self.Kdiag(self._dev(X)).cpu().numpy()

    def compute_K_tens(self, Z):
        return self.K_tens(self._dev(Z), return_levels=False).cpu().numpy()

    def compute_K_tens_vs_seq(self, Z, X):
        return self.K_tens_vs_seq(self._dev(Z), self._dev(X), return_levels=False).cpu().numpy()

    def compute_K_incr_tens(self, Z):
        return self.K_tens(self._dev(Z), increments=True, return_levels=False).cpu().numpy()

    def compute_K_incr_tens_vs_seq(self, Z, X):
        return self.K_tens_vs_seq(self._dev(Z), self._dev(X), increments=True, return_levels=False).cpu().numpy()


class SignatureLinear(SignatureKernel):
    """Identity state-space embedding (kernels.py:966-986)."""
    _base = 'linear'


class SignatureCosine(SignatureKernel):
    """Cosine similarity (kernels.py:988-1008)."""
    _base = 'cosine'


class SignaturePoly(SignatureKernel):
    """Polynomial kernel (kernels.py:1011-1028)."""
    _base = 'poly'

    def __init__(self, input_dim, num_features, num_levels, gamma=1, degree=3, **kwargs):
        SignatureKernel.__init__(self, input_dim, num_features, num_levels, **kwargs)
        self._lag_weights = getattr(self, 'gamma', None)     # kernels.py:82-83 vs :1016 use the same attribute name
        self.gamma = float(gamma)
        self.degree = float(degree)

    def _lag_gamma(self):
        return self._lag_weights

    def _get_base_params(self, like=None):
        return (_scalar_param(self.gamma, 'SignaturePoly.gamma'), float(self.degree))


class SignatureRBF(SignatureKernel):
    """Gauss / RBF / squared-exponential (kernels.py:1030-1044)."""
    _base = 'rbf'


SignatureGauss = SignatureRBF


class SignatureMix(SignatureKernel):
    """Convex combination of RBF and identity (kernels.py:1050-1072)."""
    _base = 'mix'

    def __init__(self, input_dim, num_features, num_levels, **kwargs):
        SignatureKernel.__init__(self, input_dim, num_features, num_levels, **kwargs)
        self.mixing = 0.5

    def _get_base_params(self, like=None):
        return (_scalar_param(self.mixing, 'SignatureMix.mixing'),)


class SignatureSpectral(SignatureKernel):
    """Spectral-mixture state-space embedding (kernels.py:1073-1122):
    ``kappa(x, y) = sum_q alpha_q * k_q(gamma_q * (x - y)) * cos(2 pi omega_q . (x - y))`` with ``k_q`` Gaussian
    (``family='gauss'|'gaussian'|'rbf'``) or exponential (``'exp'|'exponential'``).  As in the reference the
    constructor passes ``lengthscales=None`` and draws ``alpha [Q]``, ``omega, gamma [Q, num_features]`` as
    ``exp(randn)`` (:1092-1094).  Forward only, on the general (un-specialised) kernels.

    Two reference defects are NOT reproduced: ``family='mixed'`` evaluates an undefined name ``Q`` (:1112) --
    here it raises the same ``NameError`` when the kernel is evaluated; and ``_spectral`` broadcasts a batched
    3-D input ``[N, L, d]`` as if it were 2-D (:1103, then ``tf.tile`` with 4 multiples on a 5-D tensor fails), so
    in the reference ``Kdiag`` / ``normalization=True`` cannot be evaluated for this class -- here the batched
    call means what it means for every other base kernel (each sequence against itself)."""
    _base = 'spectral'

    def __init__(self, input_dim, num_features, num_levels, family='gauss', Q=5, **kwargs):
        SignatureKernel.__init__(self, input_dim, num_features, num_levels, lengthscales=None, **kwargs)
        if family in ('exp', 'exponential'):
            self.family = 'exp'
        elif family in ('gauss', 'gaussian', 'rbf'):
            self.family = 'rbf'
        elif family in ('mixed', 'mix'):
            self.family = 'mixed'
        else:
            raise ValueError("Unrecognized spectral family name.")
        self.Q = Q
        self._lag_weights = getattr(self, 'gamma', None)     # kernels.py:82-83 vs :1094 use the same attribute name
        self.alpha = np.exp(np.random.randn(Q))
        self.omega = np.exp(np.random.randn(Q, self.num_features))
        self.gamma = np.exp(np.random.randn(Q, self.num_features))

    def _get_base_params(self, like=None):
        if self.family == 'mixed':
            raise NameError("name 'Q' is not defined")           # kernels.py:1112
        dev = settings.device() if like is None else like.device
        d = self.num_features * (self.num_lags + 1)
        omega, gamma = np.asarray(self.omega, dtype=np.float64), np.asarray(self.gamma, dtype=np.float64)
        if omega.shape != (self.Q, d) or gamma.shape != (self.Q, d) or np.asarray(self.alpha).shape != (self.Q,):
            raise ValueError("alpha / omega / gamma must have shapes [Q], [Q, d], [Q, d]")
        block = np.concatenate([np.asarray(self.alpha, dtype=np.float64).ravel(), omega.ravel(), gamma.ravel()])
        return (self.family, self.Q, torch.as_tensor(block, dtype=torch.float64, device=dev).contiguous())

    def _lag_gamma(self):
        return self._lag_weights

    def _d_check(self):
        raise NotImplementedError('gpsig_b200: gradients are not implemented for the spectral base kernel')


class SignatureMatern12(SignatureKernel):
    """kernels.py:1124-1138."""
    _base = 'matern12'


SignatureLaplace = SignatureMatern12
SignatureExponential = SignatureMatern12


class SignatureMatern32(SignatureKernel):
    """kernels.py:1144-1157."""
    _base = 'matern32'


class SignatureMatern52(SignatureKernel):
    """kernels.py:1161-1173."""
    _base = 'matern52'
