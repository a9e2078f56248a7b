"""Kernel-only timing of the fused Gram (c2 shape, N=N2=n) for both dtypes: pairs/s from the library's event hooks."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402
from gpsig_b200._lib import lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
rng = np.random.default_rng(0)
X = np.cumsum(rng.standard_normal((n, 200, 8)) / np.sqrt(200), axis=1)
Y = np.cumsum(rng.standard_normal((n, 200, 8)) / np.sqrt(200), axis=1)
for dt in (torch.float64, torch.float32):
    Xd, Yd = torch.as_tensor(X, dtype=dt, device='cuda'), torch.as_tensor(Y, dtype=dt, device='cuda')
    ops.trunc_gram(Xd, Yd, 5, base='rbf', return_levels=False)
    lib.gpsig_profile_enable(1)
    ms = []
    for _ in range(3):
        K = ops.trunc_gram(Xd, Yd, 5, base='rbf', return_levels=False)
        ms.append(lib.gpsig_profile_last_kernel_ms())
    lib.gpsig_profile_enable(0)
    print("%s %s: %.2f ms, %.3f M pairs/s  (sum %.6e)" % (os.environ.get('GPSIG_LIB_NAME', 'default'), str(dt)[6:], min(ms), n * n / min(ms) / 1e3, float(K.sum())))
