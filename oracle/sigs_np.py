"""Brute-force truncated path signatures (known-answer oracle).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference validates its truncated kernel against ``esig.tosig.stream2sig``
(``notebooks/signature_kernel.ipynb`` cells 4-13, 15-23, 25-29).  ``esig`` is
not installed here; the signature of a piecewise-linear path is restated from
its definition with Chen's identity: S(x) = exp(dx_1) (x) ... (x) exp(dx_{L-1}),
where exp(v) = sum_k v^{(x)k}/k! is the tensor exponential.
"""
import numpy as np


def tensor_exp(v, m):
    """Levels 0..m of exp(v) for v in R^d: [1, v, v(x)v/2!, ...]."""
    out = [np.ones(())]
    cur = np.ones(())
    for k in range(1, m + 1):
        cur = np.multiply.outer(cur, v) / k
        out.append(cur)
    return out


def tensor_mul(A, B, m):
    """Truncated tensor-algebra product of two level lists."""
    out = []
    for k in range(m + 1):
        acc = 0
        for a in range(k + 1):
            acc = acc + np.multiply.outer(A[a], B[k - a])
        out.append(acc)
    return out


def signature(x, m):
    """Levels 0..m of the signature of the piecewise-linear path x[L, d]."""
    d = x.shape[1]
    S = [np.ones(())] + [np.zeros((d,) * k) for k in range(1, m + 1)]
    for p in range(x.shape[0] - 1):
        S = tensor_mul(S, tensor_exp(x[p + 1] - x[p], m), m)
    return S


def signature_gram_levels(X, X2, m):
    """[m+1, N, N2] level-wise inner products <S_k(x_i), S_k(y_j)>."""
    SX = [signature(x, m) for x in X]
    SY = SX if X2 is None else [signature(y, m) for y in X2]
    out = np.empty((m + 1, len(SX), len(SY)))
    for i, a in enumerate(SX):
        for j, b in enumerate(SY):
            for k in range(m + 1):
                out[k, i, j] = np.sum(a[k] * b[k])
    return out


def rank1_tensor_levels(Zk):
    """Inducing tensor of level i given by i vectors Zk[i, d]: z_1 (x) ... (x) z_i."""
    cur = np.ones(())
    for v in Zk:
        cur = np.multiply.outer(cur, v)
    return cur
