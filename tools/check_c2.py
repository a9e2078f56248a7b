"""Quick parity check at the c2 shape (L=200, d=8, m=5, RBF): a 48x40 block of pairs against the C oracle."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402
from oracle import c_oracle  # noqa: E402


def brownian(N, L, d, seed):
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


ok = True
for (L, L2, d, m) in ((200, 200, 8, 5), (100, 64, 3, 4), (67, 33, 8, 5), (200, 199, 4, 4)):
    X, Y = brownian(48, L, d, 0), brownian(40, L2, d, 1)
    for base in ('rbf', 'linear'):
        R = c_oracle.trunc_gram(X, Y, m, base)
        scale = np.abs(R).max(axis=(1, 2), keepdims=True)
        for dt, tol in ((torch.float64, 1e-10), (torch.float32, 1e-5)):
            G = ops.trunc_gram(torch.as_tensor(X, dtype=dt, device='cuda'), torch.as_tensor(Y, dtype=dt, device='cuda'), m,
                               base=base).cpu().numpy()
            err = float(np.max(np.abs(G - R) / scale))
            flag = 'ok' if err < tol else 'FAIL'
            ok &= err < tol
            print("L=%d L2=%d d=%d m=%d %-6s %s max rel err %.2e %s" % (L, L2, d, m, base, str(dt)[6:], err, flag))
sys.exit(0 if ok else 1)
