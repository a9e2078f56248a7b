"""Shared helpers for the test-suite (synthetic inputs of SURVEY.md 8d)."""
import numpy as np


def brownian(N, L, d, seed):
    """X = cumsum(randn(N, L, d) / sqrt(L), axis=1) with default_rng(seed)."""
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


def inducing_tensors(X, m, Mz, increments, seed):
    """Z as gpsig/utils.py:8-53 builds it: observations of sampled sequences
    at sorted random times (+ their successors for increments) + 0.4*randn."""
    rng = np.random.default_rng(seed)
    N, L, d = X.shape
    T = m * (m + 1) // 2
    Z = np.empty((T, Mz, 2, d) if increments else (T, Mz, d))
    seqs = X[rng.integers(0, N, size=Mz)]
    k = 0
    for lvl in range(1, m + 1):
        for z in range(Mz):
            idx = np.sort(rng.choice(L - 1, size=lvl, replace=False))
            for j, t in enumerate(idx):
                if increments:
                    Z[k + j, z, 0] = seqs[z, t]
                    Z[k + j, z, 1] = seqs[z, t + 1]
                else:
                    Z[k + j, z] = seqs[z, t]
        k += lvl
    return Z + 0.4 * rng.standard_normal(Z.shape)


def rel_err(a, b):
    """max |a-b| / max |b|  -- the parity metric of SURVEY.md 8d / H3."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
