"""Plain-Python restatement of ``gpsig/preprocessing.py`` (sequence by sequence, as the reference works).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""
import numpy as np


def tabulate(seqs, pad_with=None, pre=False):
    """preprocessing.py:5-62 (orient_ax = 0)."""
    Lmax = max(s.shape[0] for s in seqs)
    out = np.empty((len(seqs), Lmax, seqs[0].shape[1]))
    for a, s in enumerate(seqs):
        n = Lmax - s.shape[0]
        fill = np.full((n, s.shape[1]), float(pad_with)) if pad_with is not None else np.repeat((s[:1] if pre else s[-1:]), n, axis=0)
        out[a] = np.concatenate((fill, s) if pre else (s, fill), axis=0)
    return out


def add_time_to_table(table):
    """preprocessing.py:64-104."""
    N, L, d = table.shape
    out = np.empty((N, L, d + 1))
    for a in range(N):
        s = table[a]
        rep = 0
        while rep + 1 < L and np.array_equal(s[-2 - rep], s[-1]):               # trailing copies of the last element
            rep += 1
        uniq = L - rep
        with np.errstate(invalid='ignore', divide='ignore'):
            t = np.arange(uniq, dtype=np.float64) / (uniq - 1)
        out[a, :, 0] = np.concatenate((t, np.full(rep, t[-1])))
        out[a, :, 1:] = s
    return out.reshape(N, -1)


def add_natural_parametrization_to_table(table):
    """preprocessing.py:106-131."""
    N, L, d = table.shape
    out = np.empty((N, L, d + 1))
    for a in range(N):
        arc = 0.0
        for l in range(L):
            if l > 0:
                arc += float(np.sqrt(np.sum((table[a, l] - table[a, l - 1]) ** 2)))
            out[a, l, 0] = arc
        out[a, :, 1:] = table[a]
    return out
