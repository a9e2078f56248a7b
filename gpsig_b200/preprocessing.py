"""Sequence preprocessing on the device -- drop-in for ``gpsig.preprocessing`` (the step upstream of the kernels,
SURVEY.md 8(f4)).  The reference works sequence by sequence in numpy (``np.apply_along_axis`` / Python loops,
preprocessing.py:5-156); here ragged lists are tabulated with one gather and the augmentations are batched torch
expressions on ``[N, L, d]`` tensors.  Same function names, arguments and output shapes.
"""
import numpy as np
import torch

from . import settings


def _as_tensor(a, device=None, dtype=None):
    if torch.is_tensor(a):
        return a
    return torch.as_tensor(np.asarray(a), dtype=dtype or settings.float_type, device=device or settings.device())


def pad_sequence(max_length, pre, seq):
    """preprocessing.py:5-18: repeat the first (``pre``) or the last element up to ``max_length``."""
    seq = _as_tensor(seq)
    n = max_length - seq.shape[0]
    if bool(pre):
        return torch.cat((seq[:1].expand(n, -1), seq), dim=0)
    return torch.cat((seq, seq[-1:].expand(n, -1)), dim=0)


def tabulate_list_of_sequences(sequences_list, orient_ax=0, pad_with=None, pre=False, device=None, dtype=None):
    """preprocessing.py:21-62: list of ``[len_i, d]`` (or ``[d, len_i]`` with ``orient_ax=1``) sequences ->
    ``[N, max_len, d]``, padded by repeating the last / first element (which leaves the signature unchanged) or with
    the constant ``pad_with``.  One concatenation + one gather on the device."""
    seqs = [_as_tensor(s, device, dtype) for s in sequences_list]
    if not all(s.dim() == 2 for s in seqs):
        raise ValueError('Make sure ndim == 2 for all sequences in the list!')
    if orient_ax == 1:
        seqs = [s.transpose(0, 1) for s in seqs]
    feats = [s.shape[1] for s in seqs]
    if not all(f == feats[0] for f in feats):
        raise ValueError('Different path dimensions found. Please preprocess sequences beforehand so that all paths '
                         'contain the same number of features.')
    lens = torch.as_tensor([s.shape[0] for s in seqs], device=seqs[0].device)
    max_length = int(lens.max())
    flat = torch.cat(seqs, dim=0)                                            # [sum len_i, d]
    start = torch.cumsum(lens, 0) - lens
    pos = torch.arange(max_length, device=flat.device)[None, :]              # position in the padded sequence
    src = pos - (max_length - lens)[:, None] if pre else pos                 # position in the original sequence
    inside = (src >= 0) & (src < lens[:, None])
    table = flat[start[:, None] + src.clamp(min=0).minimum(lens[:, None] - 1)]
    if pad_with is not None:
        table = torch.where(inside[:, :, None], table, torch.full_like(table, float(pad_with)))
    return table


def _table3(sequences_array, num_features):
    sequences_array = _as_tensor(sequences_array)
    if sequences_array.dim() == 3:
        if num_features is None:
            num_features = sequences_array.shape[2]
        else:
            assert num_features == sequences_array.shape[2]
    else:
        num_features = num_features or 1
    return sequences_array.reshape(sequences_array.shape[0], -1, num_features)


def add_time_to_table(sequences_array, num_features=None):
    """preprocessing.py:64-104: prepend a time coordinate that runs 0..1 over the NON-repeated part of each sequence
    and stays at 1 over the trailing repetitions of the last element; returns ``[N, L * (d + 1)]``."""
    X = _table3(sequences_array, num_features)
    N, L, _ = X.shape
    same = (X == X[:, -1:, :]).all(dim=2)                                    # equals the last element
    # trailing run of `same` (it always contains the last element itself)
    run = torch.cumprod(same.flip(1).to(torch.int64), dim=1).sum(dim=1)
    unique_length = (L - (run - 1)).to(X.dtype)                              # preprocessing.py:74-78
    pos = torch.arange(L, device=X.device, dtype=X.dtype)[None, :]
    time = torch.minimum(pos, unique_length[:, None] - 1) / (unique_length[:, None] - 1)
    return torch.cat((time[:, :, None], X), dim=2).reshape(N, -1)


def add_natural_parametrization_to_table(sequences_array, num_features=None):
    """preprocessing.py:106-131: prepend the arc length; returns ``[N, L, d + 1]`` (3-D, as the reference does)."""
    X = _table3(sequences_array, num_features)
    steps = torch.linalg.norm(X[:, 1:] - X[:, :-1], dim=2)
    nat = torch.cumsum(torch.cat((torch.zeros_like(steps[:, :1]), steps), dim=1), dim=1)
    return torch.cat((nat[:, :, None], X), dim=2)


def add_time_to_list(sequences_list):
    """preprocessing.py:133-144: time coordinate (1..len)/len for every sequence of a ragged list."""
    out = []
    for x in sequences_list:
        x = _as_tensor(x)
        t = torch.arange(1, x.shape[0] + 1, device=x.device, dtype=x.dtype) / x.shape[0]
        out.append(torch.cat((t[:, None], x), dim=1))
    return out


def add_natural_parametrization_to_list(sequences_list):
    """preprocessing.py:146-156."""
    out = []
    for x in sequences_list:
        x = _as_tensor(x)
        steps = torch.linalg.norm(x[1:] - x[:-1], dim=1)
        nat = torch.cumsum(torch.cat((torch.zeros_like(steps[:1]), steps), dim=0), dim=0)
        out.append(torch.cat((nat[:, None], x), dim=1))
    return out
