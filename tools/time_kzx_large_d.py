"""Kzx (inducing tensors vs sequences) at state dimensions above 16: python tools/time_kzx_large_d.py [d] [N] [Mz]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402

d = int(sys.argv[1]) if len(sys.argv) > 1 else 28
N = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
Mz = int(sys.argv[3]) if len(sys.argv) > 3 else 500
L, m = 100, 4
rng = np.random.default_rng(0)
Xn = np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1) / np.sqrt(d / 8.0)
Zn = rng.standard_normal((m * (m + 1) // 2, Mz, 2, d)) * 0.5 / np.sqrt(d / 8.0)
for dt in (torch.float64, torch.float32):
    X = torch.as_tensor(Xn, dtype=dt, device='cuda')
    Z = torch.as_tensor(Zn, dtype=dt, device='cuda')
    f = lambda: ops.tens_vs_seq(Z, X, m, increments=True, base='rbf', return_levels=False)  # noqa: E731
    f()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); K = f(); e1.record(); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("Kzx %s d=%d N=%d Mz=%d L=%d m=%d: %8.2f ms  %7.2f M (z,x) pairs/s" % (str(dt)[6:], d, N, Mz, L, m, best, Mz * N / best / 1e3))
