"""Kernel time of the PDE Kdiag (c3 shape by default): python tools/time_pde.py [N] [L] [d] [order]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402
from gpsig_b200._lib import lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 500
d = int(sys.argv[3]) if len(sys.argv) > 3 else 4
order = int(sys.argv[4]) if len(sys.argv) > 4 else 2
rng = np.random.default_rng(0)
X = torch.as_tensor(np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1), dtype=torch.float64, device='cuda')
sym = os.environ.get('GPSIG_PDE_SYM') is not None
K = ops.pde(X, None, dyadic_order=order)
lib.gpsig_profile_enable(1)
ms = []
for _ in range(5):
    K = ops.pde(X, None, dyadic_order=order)
    ms.append(lib.gpsig_profile_last_kernel_ms())
lib.gpsig_profile_enable(0)
os.environ.pop('GPSIG_PDE_SYM', None)
K0 = ops.pde(X, None, dyadic_order=order)
print("N=%d L=%d d=%d order=%d sym=%s: %.3f ms, %.1f k pairs/s, max rel diff to the full-grid solve %.2e" % (
    N, L, d, order, 'on' if sym else 'off', min(ms), N / min(ms), float(((K - K0).abs() / K0.abs()).max())))
