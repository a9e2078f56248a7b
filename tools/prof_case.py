"""Small driver for ncu / compute-sanitizer: one call of one hot-path entry point.

    python tools/prof_case.py trunc f64 1000      # K(X, X2) raw levels, N=N2=1000, L=200, d=8, m=5
    python tools/prof_case.py pde 256 500 2       # PDE Kdiag, N=256, L=500, dyadic order 2
    python tools/prof_case.py pde_vjp 296 500 2   # PDE Kdiag backward (two sweeps)
    python tools/prof_case.py tens f64 2000       # Kzx, N=2000, L=100, d=8, m=4, Mz=500
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402


def brownian(N, L, d, seed):
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


what = sys.argv[1]
if what == 'trunc':
    dt = torch.float64 if sys.argv[2] == 'f64' else torch.float32
    N = int(sys.argv[3])
    X = torch.as_tensor(brownian(N, 200, 8, 0), dtype=dt, device='cuda')
    Y = torch.as_tensor(brownian(N, 200, 8, 1), dtype=dt, device='cuda')
    for _ in range(2):
        K = ops.trunc_gram(X, Y, 5, base='rbf', return_levels=False)
    torch.cuda.synchronize()
    print(float(K.sum()))
elif what == 'pde':
    N, L, order = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    X = torch.as_tensor(brownian(N, L, 4, 0), dtype=torch.float64, device='cuda')
    for _ in range(2):
        K = ops.pde(X, None, dyadic_order=order)
    torch.cuda.synchronize()
    print(float(K.sum()))
elif what == 'pde_vjp':
    N, L, order = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    X = torch.as_tensor(brownian(N, L, 4, 0), dtype=torch.float64, device='cuda')
    g = torch.ones(N, dtype=torch.float64, device='cuda')
    for _ in range(2):
        K = ops.pde_kdiag_vjp(X, g, dyadic_order=order)
    torch.cuda.synchronize()
    print(float(K.sum()))
elif what == 'tens':
    dt = torch.float64 if sys.argv[2] == 'f64' else torch.float32
    N = int(sys.argv[3])
    X = torch.as_tensor(brownian(N, 100, 8, 0), dtype=dt, device='cuda')
    Z = torch.as_tensor(np.random.default_rng(2).standard_normal((10, 500, 2, 8)), dtype=dt, device='cuda')
    for _ in range(2):
        K = ops.tens_vs_seq(Z, X, 4, increments=True, base='rbf', return_levels=False)
    torch.cuda.synchronize()
    print(float(K.sum()))
