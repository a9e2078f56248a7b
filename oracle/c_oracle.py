"""ctypes front end of ``oracle/libgpsig_oracle.so`` (the plain-C port).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, 'libgpsig_oracle.so')
BASE_IDS = {'linear': 0, 'rbf': 1, 'cosine': 2, 'poly': 3, 'mix': 4,
            'matern12': 5, 'matern32': 6, 'matern52': 7}
_lib = None


def build(force=False):
    src = os.path.join(HERE, 'csrc', 'gpsig_oracle.c')
    if force or not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
        subprocess.check_call(['make', '-C', HERE, '-B', 'libgpsig_oracle.so'], stdout=subprocess.DEVNULL)
    return SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            build()
        _lib = ctypes.CDLL(SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _prm(params):
    return np.ascontiguousarray(np.asarray(list(params) + [0.0] * 4, dtype=np.float64)[:4])


def trunc_gram(X, X2, m, base='linear', params=(), difference=True):
    """X [N, L, d] (already scaled) -> [m+1, N, N2] unnormalised levels."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    X2c = None if X2 is None else np.ascontiguousarray(X2, dtype=np.float64)
    N, L, d = X.shape
    N2, L2 = (N, L) if X2c is None else X2c.shape[:2]
    out = np.empty((m + 1, N, N2))
    prm = _prm(params)
    lib().oracle_trunc_gram_f64(_p(X), _p(X2c), N, N2, L, L2, d, m, BASE_IDS[base], _p(prm), int(difference), _p(out))
    return out


def trunc_diag(X, m, base='linear', params=(), difference=True):
    X = np.ascontiguousarray(X, dtype=np.float64)
    N, L, d = X.shape
    out = np.empty((m + 1, N))
    prm = _prm(params)
    lib().oracle_trunc_diag_f64(_p(X), N, L, d, m, BASE_IDS[base], _p(prm), int(difference), _p(out))
    return out


def tens_vs_seq(Z, X, m, increments, base='linear', params=(), difference=True):
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    X = np.ascontiguousarray(X, dtype=np.float64)
    N, L, d = X.shape
    Mz = Z.shape[1]
    out = np.empty((m + 1, Mz, N))
    prm = _prm(params)
    lib().oracle_tens_vs_seq_f64(_p(Z), Mz, int(increments), _p(X), N, L, d, m, BASE_IDS[base], _p(prm),
                                 int(difference), _p(out))
    return out


def pde_pairs(X, Y, ix, iy, n):
    X = np.ascontiguousarray(X, dtype=np.float64)
    Y = X if Y is None else np.ascontiguousarray(Y, dtype=np.float64)
    ix = np.ascontiguousarray(ix, dtype=np.int32)
    iy = np.ascontiguousarray(iy, dtype=np.int32)
    out = np.empty(ix.shape[0])
    lib().oracle_pde_pairs_f64(_p(X), _p(Y), _p(ix), _p(iy), ix.shape[0], X.shape[1], Y.shape[1], X.shape[2], n, _p(out))
    return out
