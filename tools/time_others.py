"""Kernel-only timings of the PDE wavefront (c3) and the inducing-tensor kernels (c4)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402
from gpsig_b200._lib import lib  # noqa: E402


def brownian(N, L, d, seed):
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


def timed(fn, reps=3):
    fn()
    lib.gpsig_profile_enable(1)
    ms = []
    for _ in range(reps):
        out = fn()
        ms.append(lib.gpsig_profile_last_kernel_ms())
    lib.gpsig_profile_enable(0)
    return min(ms), out


what = sys.argv[1:] or ['pde', 'tens']
if 'pde' in what:
    for (N, L, d, order) in ((1000, 500, 4, 2), (1000, 500, 4, 0), (4000, 200, 8, 1)):
        X = torch.as_tensor(brownian(N, L, d, 0), dtype=torch.float64, device='cuda')
        ms, K = timed(lambda: ops.pde(X, None, dyadic_order=order))
        P = (L - 1) << order
        print("pde diag N=%d L=%d d=%d order=%d: %.3f ms, %.1f k pairs/s, %.2f G cells/s (sum %.8e)" %
              (N, L, d, order, ms, N / ms, N * P * P / ms / 1e6, float(K.sum())))
    X = torch.as_tensor(brownian(200, 200, 8, 0), dtype=torch.float64, device='cuda')
    ms, K = timed(lambda: ops.pde(X, X, dyadic_order=0, gram=True))
    print("pde gram 200x200 L=200 d=8 order=0: %.3f ms, %.1f k pairs/s, %.2f G cells/s" % (ms, 40000 / ms, 40000 * 199 * 199 / ms / 1e6))
if 'tens' in what:
    N, L, d, m, Mz = 10000, 100, 8, 4, 500
    Xn = brownian(N, L, d, 0)
    Zn = np.random.default_rng(2).standard_normal((m * (m + 1) // 2, Mz, 2, d)) * 0.5
    for dt in (torch.float64, torch.float32):
        X = torch.as_tensor(Xn, dtype=dt, device='cuda')
        Z = torch.as_tensor(Zn, dtype=dt, device='cuda')
        ms, K = timed(lambda: ops.tens_vs_seq(Z, X, m, increments=True, base='rbf', return_levels=False))
        print("tens_vs_seq c4 %s: %.3f ms, %.2f M (z,x) pairs/s" % (str(dt)[6:], ms, Mz * N / ms / 1e3))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.trunc_diag(X, m, base='rbf')
        e0.record(); D = ops.trunc_diag(X, m, base='rbf'); e1.record(); e1.synchronize()
        print("trunc_diag c4 %s (N=10000, L=100): %.3f ms" % (str(dt)[6:], e0.elapsed_time(e1)))
        e0.record(); Kzz = ops.tens_tens(Z, m, increments=True, base='rbf'); e1.record(); e1.synchronize()
        print("tens_tens c4 %s: %.3f ms" % (str(dt)[6:], e0.elapsed_time(e1)))
