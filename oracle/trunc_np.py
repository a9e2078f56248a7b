"""numpy restatement of the reference's truncated signature-kernel path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Every function follows one function of the reference (cited file:line relative
to ``/root/reference``) and deliberately keeps the reference's *shape* of the
computation: the lifted base-kernel tensor ``M`` is materialised (in row
blocks, so it fits in RAM) and the recursions are whole-tensor exclusive
cumulative sums, exactly as the TensorFlow graph does.  Parity status: PINNED --
``tests/golden/trunc_ref.npz`` holds what the reference's own ``kernels.py`` /
``signature_algs.py`` / ``lags.py`` return when executed on eager numpy stand-ins
for TensorFlow / GPflow (``oracle/refrun``); this restatement reproduces them to
1e-13 (``tests/test_reference_golden_cpu.py``), next to the brute-force definition
and the signature known answers of ``tests/test_oracle_cpu.py``.
"""
import numpy as np

JITTER = 1e-6  # gpflow 1.5.1 default ``settings.jitter`` (numerics.jitter_level)


# ----------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------
def _cumsum_excl(A, axis):
    """``tf.cumsum(A, exclusive=True, axis=axis)``: out[0]=0, out[k]=sum_{<k}."""
    C = np.cumsum(A, axis=axis)
    out = np.zeros_like(A)
    src = [slice(None)] * A.ndim
    dst = [slice(None)] * A.ndim
    src[axis] = slice(0, -1)
    dst[axis] = slice(1, None)
    out[tuple(dst)] = C[tuple(src)]
    return out


# ----------------------------------------------------------------------------
# base kernels -- gpsig/kernels.py:946-1173
# ----------------------------------------------------------------------------
def square_dist(X, X2=None):
    """gpsig/kernels.py:946-957 (batched over leading axes)."""
    Xs = np.sum(np.square(X), axis=-1)
    if X2 is None:
        dist = -2 * np.matmul(X, np.swapaxes(X, -1, -2))
        dist += Xs[..., :, None] + Xs[..., None, :]
        return dist
    X2s = np.sum(np.square(X2), axis=-1)
    dist = -2 * np.matmul(X, np.swapaxes(X2, -1, -2))
    dist += Xs[..., :, None] + X2s[..., None, :]
    return dist


def euclid_dist(X, X2=None):
    """gpsig/kernels.py:959-961."""
    return np.sqrt(np.maximum(square_dist(X, X2), 1e-40))


def base_lin(X, X2=None, **kw):
    """gpsig/kernels.py:979-986."""
    Y = X if X2 is None else X2
    return np.matmul(X, np.swapaxes(Y, -1, -2))


def base_cos(X, X2=None, **kw):
    """gpsig/kernels.py:1000-1008."""
    Xn = np.sqrt(np.sum(np.square(X), axis=-1))
    Y = X if X2 is None else X2
    Yn = Xn if X2 is None else np.sqrt(np.sum(np.square(X2), axis=-1))
    return np.matmul(X, np.swapaxes(Y, -1, -2)) / (Xn[..., :, None] * Yn[..., None, :])


def base_poly(X, X2=None, gamma=1.0, degree=3.0, **kw):
    """gpsig/kernels.py:1024-1028."""
    Y = X if X2 is None else X2
    return (np.matmul(X, np.swapaxes(Y, -1, -2)) + gamma) ** degree


def base_rbf(X, X2=None, **kw):
    """gpsig/kernels.py:1042-1044."""
    return np.exp(-square_dist(X, X2) / 2)


def base_mix(X, X2=None, mixing=0.5, **kw):
    """gpsig/kernels.py:1061-1072."""
    Xs = np.sum(np.square(X), axis=-1)
    Y = X if X2 is None else X2
    Ys = Xs if X2 is None else np.sum(np.square(X2), axis=-1)
    inner = np.matmul(X, np.swapaxes(Y, -1, -2))
    ds = Xs[..., :, None] + Ys[..., None, :] - 2 * inner
    return mixing * np.exp(-ds / 2) + (1. - mixing) * inner


def base_matern12(X, X2=None, **kw):
    """gpsig/kernels.py:1135-1138."""
    return np.exp(-euclid_dist(X, X2))


def base_matern32(X, X2=None, **kw):
    """gpsig/kernels.py:1154-1157."""
    r = euclid_dist(X, X2)
    return (1. + np.sqrt(3.) * r) * np.exp(-np.sqrt(3.) * r)


def base_matern52(X, X2=None, **kw):
    """gpsig/kernels.py:1171-1173."""
    r = euclid_dist(X, X2)
    return (1.0 + np.sqrt(5.) * r + 5. / 3. * np.square(r)) * np.exp(-np.sqrt(5.) * r)


def base_spectral(X, X2=None, family='rbf', alpha=None, omega=None, gamma=None, **kw):
    """gpsig/kernels.py:1101-1122 (``_spectral``), families 'rbf' and 'exp'; alpha [Q], omega/gamma [Q, d].
    2-D inputs ``[n, d] x [n2, d]`` exactly as the reference broadcasts them (:1103-1105); a batched 3-D input
    ``[N, L, d]`` is evaluated per batch entry (the reference's expression does not support it, see
    gpsig_b200.kernels.SignatureSpectral)."""
    Y = X if X2 is None else X2
    diff = X[..., :, None, :] - Y[..., None, :, :]                                   # [..., n, n2, d]
    out = np.zeros(diff.shape[:-1], dtype=X.dtype)
    for q in range(len(alpha)):
        quad = np.sum(np.square(diff * gamma[q]), axis=-1)
        kern = np.exp(-np.sqrt(quad) / 2) if family == 'exp' else np.exp(-quad / 2)   # :1108, :1110
        out = out + kern * np.cos(2. * np.pi * np.sum(diff * omega[q], axis=-1)) * alpha[q]   # :1117, :1122
    return out


BASE_KERNELS = {
    'spectral': base_spectral,
    'linear': base_lin, 'cosine': base_cos, 'poly': base_poly, 'rbf': base_rbf,
    'mix': base_mix, 'matern12': base_matern12, 'matern32': base_matern32,
    'matern52': base_matern52,
}


# ----------------------------------------------------------------------------
# lags -- gpsig/lags.py
# ----------------------------------------------------------------------------
def lin_interp(time, X, time_query, jitter=JITTER):
    """gpsig/lags.py:7-38."""
    pairwise = time[:, None, None] - time_query[None, :, :]
    left_idx = np.argmax(np.where(pairwise > jitter, -np.inf, pairwise), axis=0)
    right_idx = left_idx + 1
    X_left, X_right = np.take(X, left_idx, axis=-2), np.take(X, np.minimum(right_idx, X.shape[-2] - 1), axis=-2)
    t_left, t_right = time[left_idx], time[np.minimum(right_idx, time.shape[0] - 1)]
    return X_left + (time_query[None, ..., None] - t_left[None, ..., None]) * (X_right - X_left) / (t_right[None, ..., None] - t_left[None, ..., None])


def add_lags_to_sequences(X, lags):
    """gpsig/lags.py:41-63.  [N, L, d] -> [N, L, num_lags+1, d]."""
    L = X.shape[1]
    time = np.arange(L, dtype=np.float64) / (L - 1)
    time_lags = np.maximum(time[:, None] - lags[None, :], 0.)
    return np.concatenate((X[:, :, None, :], lin_interp(time, X, time_lags)), axis=2)


# ----------------------------------------------------------------------------
# recursions -- gpsig/signature_algs.py
# ----------------------------------------------------------------------------
def _difference_2d(M):
    """signature_algs.py:26 / :56 (works for the 4-D and the 3-D layout)."""
    return M[:, 1:, ..., 1:] + M[:, :-1, ..., :-1] - M[:, :-1, ..., 1:] - M[:, 1:, ..., :-1]


def signature_kern_first_order(M, num_levels, difference=True):
    """gpsig/signature_algs.py:8-35.

    M: [N, L, N2, L2] -> [num_levels+1, N, N2]; or [N, L, L] -> [num_levels+1, N].
    """
    if M.ndim == 4:
        K = [np.ones((M.shape[0], M.shape[2]), dtype=M.dtype)]
    else:
        K = [np.ones((M.shape[0],), dtype=M.dtype)]
    if difference:
        M = _difference_2d(M)
    K.append(np.sum(M, axis=(1, -1)))
    R = M
    for i in range(2, num_levels + 1):
        R = M * _cumsum_excl(_cumsum_excl(R, axis=1), axis=-1)
        K.append(np.sum(R, axis=(1, -1)))
    return np.stack(K, axis=0)


def signature_kern_higher_order(M, num_levels, order=2, difference=True):
    """gpsig/signature_algs.py:37-74."""
    if M.ndim == 4:
        K = [np.ones((M.shape[0], M.shape[2]), dtype=M.dtype)]
    else:
        K = [np.ones((M.shape[0],), dtype=M.dtype)]
    if difference:
        M = _difference_2d(M)
    K.append(np.sum(M, axis=(1, -1)))
    R = [[M]]
    for i in range(2, num_levels + 1):
        d = min(i, order)
        dprev = len(R)
        R_next = [[None] * d for _ in range(d)]
        tot = sum(R[a][b] for a in range(dprev) for b in range(dprev))
        R_next[0][0] = M * _cumsum_excl(_cumsum_excl(tot, axis=1), axis=-1)
        for j in range(2, d + 1):
            col = sum(R[a][j - 2] for a in range(dprev))
            row = sum(R[j - 2][b] for b in range(dprev))
            R_next[0][j - 1] = 1. / j * M * _cumsum_excl(col, axis=1)
            R_next[j - 1][0] = 1. / j * M * _cumsum_excl(row, axis=-1)
            for k in range(2, d + 1):
                R_next[j - 1][k - 1] = 1. / (j * k) * M * R[j - 2][k - 2]
        K.append(np.sum(sum(R_next[a][b] for a in range(d) for b in range(d)), axis=(1, -1)))
        R = R_next
    return np.stack(K, axis=0)


def tensor_kern(M, num_levels):
    """gpsig/signature_algs.py:76-99.  M: [T, Mz, Mz2] -> [num_levels+1, Mz, Mz2]."""
    K = [np.ones(M.shape[1:], dtype=M.dtype)]
    k = 0
    for i in range(1, num_levels + 1):
        R = M[k]
        k += 1
        for j in range(1, i):
            R = M[k] * R
            k += 1
        K.append(R)
    return np.stack(K, axis=0)


def signature_kern_tens_vs_seq_first_order(M, num_levels, difference=True):
    """gpsig/signature_algs.py:101-127.  M: [T, Mz, N, L] -> [num_levels+1, Mz, N]."""
    if difference:
        M = M[..., 1:] - M[..., :-1]
    K = [np.ones(M.shape[1:3], dtype=M.dtype)]
    k = 0
    for i in range(1, num_levels + 1):
        R = M[k]
        k += 1
        for j in range(1, i):
            R = M[k] * _cumsum_excl(R, axis=2)
            k += 1
        K.append(np.sum(R, axis=2))
    return np.stack(K, axis=0)


def signature_kern_tens_vs_seq_higher_order(M, num_levels, order=2, difference=True):
    """gpsig/signature_algs.py:129-160."""
    if difference:
        M = M[..., 1:] - M[..., :-1]
    K = [np.ones(M.shape[1:3], dtype=M.dtype)]
    k = 0
    for i in range(1, num_levels + 1):
        R = [M[k]]
        k += 1
        for j in range(1, i):
            d = min(j + 1, order)
            R_next = [None] * d
            R_next[0] = M[k] * _cumsum_excl(sum(R), axis=2)
            for l in range(1, d):
                R_next[l] = 1. / (l + 1) * M[k] * R[l - 1]
            R = R_next
            k += 1
        K.append(np.sum(sum(R), axis=2))
    return np.stack(K, axis=0)


# ----------------------------------------------------------------------------
# kernel wrappers -- gpsig/kernels.py:190-794
# ----------------------------------------------------------------------------
class SignatureKernelOracle:
    """Restatement of ``gpsig.kernels.SignatureKernel`` (kernels.py:16-794) for
    ``low_rank=False``, ``num_lags=0``.  Parameters are the *constrained*
    values (the reference stores them through a softplus transform,
    kernels.py:66-67,87; ``params_as_tensors`` hands the constrained values to
    every method)."""

    def __init__(self, input_dim, num_features, num_levels, base='linear', variances=1, lengthscales=1,
                 order=1, normalization=True, difference=True, sigma=1.0, jitter=JITTER,
                 block=16, dtype=np.float64, num_lags=0, **base_params):
        if input_dim % num_features != 0:  # kernels.py:95-103
            raise ValueError("The arguments num_features and input_dim are not consistent.")
        self.num_features = num_features
        self.num_levels = num_levels
        self.len_examples = input_dim // num_features
        self.order = num_levels if (order <= 0 or order >= num_levels) else order  # kernels.py:58
        self.normalization = normalization
        self.difference = difference
        self.dtype = dtype
        self.variances = (variances * np.ones(num_levels + 1)).astype(dtype)
        self.sigma = dtype(sigma)
        self.lengthscales = None if lengthscales is None else (lengthscales * np.ones(num_features)).astype(dtype)
        self.jitter = jitter
        self.num_lags = int(num_lags or 0)
        if self.num_lags > 0:                                        # kernels.py:78-83
            self.lags = 0.1 * np.arange(1, self.num_lags + 1, dtype=np.float64)
            gamma = 1. / np.arange(1, self.num_lags + 2, dtype=np.float64)
            self.gamma = gamma / np.sum(gamma)
        self.block = block
        self.base = base
        self.base_params = base_params

    # -- helpers ------------------------------------------------------------
    def _base_kern(self, X, X2=None):
        return BASE_KERNELS[self.base](X, X2, **self.base_params)

    def _seq(self, X):
        X = np.asarray(X, dtype=self.dtype)
        X = X.reshape(X.shape[0], -1, self.num_features)           # kernels.py:419
        if self.num_lags > 0:                                        # kernels.py:343-365
            X = add_lags_to_sequences(X, self.lags)
            if self.lengthscales is not None:
                X = X / self.lengthscales[None, None, None, :]
            X = X * self.gamma[None, None, :, None]
            return X.reshape(X.shape[0], X.shape[1], -1)
        if self.lengthscales is not None:                            # kernels.py:358-359
            X = X / self.lengthscales[None, None, :]
        return X

    def _tens(self, Z, increments):
        Z = np.asarray(Z, dtype=self.dtype)                          # kernels.py:367-399
        if self.lengthscales is not None:
            if self.num_lags > 0:
                shp = Z.shape
                Zr = Z.reshape(shp[:-1] + (self.num_lags + 1, self.num_features))
                return (Zr / self.lengthscales * self.gamma[:, None]).reshape(shp)
            Z = Z / self.lengthscales
        return Z

    def _recursion(self, M):
        if self.order == 1:
            return signature_kern_first_order(M, self.num_levels, difference=self.difference)
        return signature_kern_higher_order(M, self.num_levels, order=self.order, difference=self.difference)

    def _K_seq_diag(self, X):
        """kernels.py:190-207 (batched 3-D base kernel [N, L, L])."""
        out = []
        for a in range(0, X.shape[0], self.block * 8):
            Xa = X[a:a + self.block * 8]
            out.append(self._recursion(self._base_kern(Xa)))
        return np.concatenate(out, axis=1)

    def _K_seq(self, X, X2=None):
        """kernels.py:209-238, evaluated in [block x block] tiles of sequences."""
        Y = X if X2 is None else X2
        N, L, d = X.shape
        N2, L2, _ = Y.shape
        out = np.empty((self.num_levels + 1, N, N2), dtype=self.dtype)
        b = self.block
        for a in range(0, N, b):
            Xa = X[a:a + b].reshape(-1, d)
            for c in range(0, N2, b):
                Yc = Y[c:c + b].reshape(-1, d)
                M = self._base_kern(Xa, Yc).reshape(-1, L, Yc.shape[0] // L2, L2)
                out[:, a:a + b, c:c + b] = self._recursion(M)
        return out

    def _K_tens(self, Z, increments=False):
        """kernels.py:264-284."""
        T, Mz, d = Z.shape[0], Z.shape[1], Z.shape[-1]
        if increments:
            Zr = Z.reshape(T, 2 * Mz, d)
            M = self._base_kern(Zr).reshape(T, Mz, 2, Mz, 2)
            M = M[:, :, 1, :, 1] + M[:, :, 0, :, 0] - M[:, :, 1, :, 0] - M[:, :, 0, :, 1]
        else:
            M = self._base_kern(Z)
        return tensor_kern(M, self.num_levels)

    def _K_tens_vs_seq(self, Z, X, increments=False):
        """kernels.py:314-341, evaluated in blocks of sequences."""
        T, Mz, d = Z.shape[0], Z.shape[1], Z.shape[-1]
        N, L, _ = X.shape
        out = np.empty((self.num_levels + 1, Mz, N), dtype=self.dtype)
        b = self.block * 8
        for a in range(0, N, b):
            Xa = X[a:a + b]
            na = Xa.shape[0]
            Xf = Xa.reshape(na * L, d)
            if increments:
                M = self._base_kern(Z.reshape(2 * Mz * T, d), Xf).reshape(T, Mz, 2, na, L)
                M = M[:, :, 1] - M[:, :, 0]
            else:
                M = self._base_kern(Z.reshape(Mz * T, d), Xf).reshape(T, Mz, na, L)
            if self.order == 1:
                out[:, :, a:a + b] = signature_kern_tens_vs_seq_first_order(M, self.num_levels, difference=self.difference)
            else:
                out[:, :, a:a + b] = signature_kern_tens_vs_seq_higher_order(M, self.num_levels, order=self.order, difference=self.difference)
        return out

    def _scale(self):
        return (self.sigma * self.variances)

    @staticmethod
    def _ret(K_lvls, return_levels):
        return K_lvls if return_levels else np.sum(K_lvls, axis=0)

    # -- public API (same names as the reference) ---------------------------
    def K(self, X, X2=None, return_levels=False):
        """kernels.py:401-477."""
        Xs = self._seq(X)
        if X2 is None:
            K_lvls = self._K_seq(Xs)
            if self.normalization:                                   # kernels.py:431-434
                K_lvls = K_lvls + self.jitter * np.eye(Xs.shape[0], dtype=self.dtype)[None]
                dsq = np.sqrt(np.diagonal(K_lvls, axis1=1, axis2=2))
                K_lvls = K_lvls / (dsq[:, :, None] * dsq[:, None, :])
        else:
            X2s = self._seq(X2)
            K_lvls = self._K_seq(Xs, X2s)
            if self.normalization:                                   # kernels.py:456-470
                d1 = np.sqrt(self._K_seq_diag(Xs) + self.jitter)
                d2 = np.sqrt(self._K_seq_diag(X2s) + self.jitter)
                K_lvls = K_lvls / (d1[:, :, None] * d2[:, None, :])
        K_lvls = K_lvls * self._scale()[:, None, None]               # kernels.py:472
        return self._ret(K_lvls, return_levels)

    def K_norms(self, X):
        """kernels.py:480-506."""
        Xs = self._seq(X)
        return np.full((Xs.shape[0],), self.sigma * np.sum(self.variances)), self._K_seq_diag(Xs)

    def Kdiag(self, X, return_levels=False):
        """kernels.py:509-541."""
        N = np.asarray(X).shape[0]
        if self.normalization:
            if return_levels:
                return np.tile(self._scale()[:, None], [1, N])
            return np.full((N,), self.sigma * np.sum(self.variances))
        K = self._K_seq_diag(self._seq(X)) * self._scale()[:, None]
        return self._ret(K, return_levels)

    def K_tens(self, Z, return_levels=False, increments=False):
        """kernels.py:543-567."""
        K = self._K_tens(self._tens(Z, increments), increments=increments) * self._scale()[:, None, None]
        return self._ret(K, return_levels)

    def K_tens_vs_seq(self, Z, X, return_levels=False, increments=False):
        """kernels.py:570-620."""
        Xs = self._seq(X)
        K = self._K_tens_vs_seq(self._tens(Z, increments), Xs, increments=increments)
        if self.normalization:                                       # kernels.py:604-613
            K = K / np.sqrt(self._K_seq_diag(Xs) + self.jitter)[:, None, :]
        K = K * self._scale()[:, None, None]
        return self._ret(K, return_levels)

    def K_tens_n_seq_covs(self, Z, X, full_X_cov=False, return_levels=False, increments=False):
        """kernels.py:623-704."""
        Xs = self._seq(X)
        Zs = self._tens(Z, increments)
        N = Xs.shape[0]
        Kzz = self._K_tens(Zs, increments=increments)
        Kzx = self._K_tens_vs_seq(Zs, Xs, increments=increments)
        sc = self._scale()
        if full_X_cov:
            Kxx = self._K_seq(Xs)
            if self.normalization:                                   # kernels.py:665-671
                Kxx = Kxx + self.jitter * np.eye(N, dtype=self.dtype)[None]
                dsq = np.sqrt(np.diagonal(Kxx, axis1=1, axis2=2))
                Kxx = Kxx / (dsq[:, :, None] * dsq[:, None, :])
                Kzx = Kzx / dsq[:, None, :]
            Kxx = Kxx * sc[:, None, None]
        else:
            Kxx = self._K_seq_diag(Xs)
            if self.normalization:                                   # kernels.py:688-694
                Kzx = Kzx / np.sqrt(Kxx + self.jitter)[:, None, :]
                Kxx = np.tile(sc[:, None], [1, N])
            else:
                Kxx = Kxx * sc[:, None]
        Kzz = Kzz * sc[:, None, None]
        Kzx = Kzx * sc[:, None, None]
        return self._ret(Kzz, return_levels), self._ret(Kzx, return_levels), self._ret(Kxx, return_levels)

    def K_seq_n_seq_covs(self, X, X2, full_X2_cov=False, return_levels=False):
        """kernels.py:706-794 (the live ``full_X2_cov=False`` branch; the
        ``True`` branch raises NameError in the reference, kernels.py:756-761)."""
        if full_X2_cov and self.normalization:
            raise NameError("reference kernels.py:756 references undefined K_x2x2_lvls")
        Xs, X2s = self._seq(X), self._seq(X2)
        N2 = X2s.shape[0]
        sc = self._scale()
        Kxx = self._K_seq(Xs)
        Kxx2 = self._K_seq(Xs, X2s)
        if self.normalization:                                       # kernels.py:740-746
            Kxx = Kxx + self.jitter * np.eye(Xs.shape[0], dtype=self.dtype)[None]
            dsq = np.sqrt(np.diagonal(Kxx, axis1=1, axis2=2))
            Kxx = Kxx / (dsq[:, :, None] * dsq[:, None, :])
            Kxx2 = Kxx2 / dsq[:, :, None]
        if full_X2_cov:
            Kx2x2 = self._K_seq(X2s) * sc[:, None, None]
        else:
            Kx2x2 = self._K_seq_diag(X2s)
            if self.normalization:                                   # kernels.py:778-784
                d2 = np.sqrt(Kx2x2 + self.jitter)
                Kxx2 = Kxx2 / (dsq[:, :, None] * d2[:, None, :])
                Kx2x2 = np.tile(sc[:, None], [1, N2])
            else:
                Kx2x2 = Kx2x2 * sc[:, None]
        Kxx = Kxx * sc[:, None, None]
        Kxx2 = Kxx2 * sc[:, None, None]
        return self._ret(Kxx, return_levels), self._ret(Kxx2, return_levels), self._ret(Kx2x2, return_levels)
