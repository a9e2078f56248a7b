"""float64 large-d path only (DFMA lift + recursion): python tools/time_large_d_f64.py [N] [L] [d] [m]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 100
d = int(sys.argv[3]) if len(sys.argv) > 3 else 126
m = int(sys.argv[4]) if len(sys.argv) > 4 else 4
rng = np.random.default_rng(0)
X = np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)
Y = np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)
ls = torch.as_tensor(np.full(d, np.sqrt(d / 8.0)), dtype=torch.float64, device='cuda')
Xd, Yd = torch.as_tensor(X, device='cuda'), torch.as_tensor(Y, device='cuda')
f = lambda: ops.trunc_gram(Xd, Yd, m, base='rbf', lengthscales=ls, return_levels=False)  # noqa: E731
K = f()
best = 1e30
for _ in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); K = f(); e1.record(); e1.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("f64 DFMA lift N=%d L=%d d=%d m=%d: %8.2f ms  %7.3f M pairs/s  (sum %.10e)" % (N, L, d, m, best, N * N / best / 1e3, float(K.sum())))
