"""Lagged copies of the sequences as extra state dimensions (host-side torch preprocessing).

Mirror of ``gpsig/lags.py:7-63`` (``lin_interp``, ``add_lags_to_sequences``): the
lagged versions are computed by linear interpolation in time on the device,
before the sequences reach the CUDA kernels (which then see ``d * (num_lags+1)``
state dimensions).  Elementwise / gather work on ``N*L*d`` values: not a hot path.
"""
import torch

from . import settings


def lin_interp(time, X, time_query):
    """gpsig/lags.py:7-38.  time [L], X [N, L, d], time_query [L, num_lags] -> [N, L, num_lags, d]."""
    pairwise = time[:, None, None] - time_query[None, :, :]
    masked = torch.where(pairwise > settings.jitter, torch.full_like(pairwise, -float('inf')), pairwise)
    left_idx = torch.argmax(masked, dim=0)                       # [L, num_lags]: last time point <= query
    right_idx = torch.clamp(left_idx + 1, max=time.shape[0] - 1)
    X_left, X_right = X[:, left_idx, :], X[:, right_idx, :]      # [N, L, num_lags, d]
    t_left, t_right = time[left_idx], time[right_idx]
    w = (time_query - t_left) / (t_right - t_left)
    return X_left + w[None, :, :, None] * (X_right - X_left)


def add_lags_to_sequences(X, lags):
    """gpsig/lags.py:41-63.  X [N, L, d], lags [num_lags] -> [N, L, num_lags + 1, d]."""
    L = X.shape[1]
    time = torch.arange(L, dtype=X.dtype, device=X.device) / (L - 1)
    time_lags = torch.clamp(time[:, None] - lags[None, :], min=0.)
    return torch.cat((X[:, :, None, :], lin_interp(time, X, time_lags)), dim=2)
