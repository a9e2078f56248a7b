"""Times the large-state-dimension path (lifted_kernels.cuh) on the reference's benchmark shapes.
    python tools/time_large_d.py [N] [L] [d] [m]
Prints ms and pairs/s for float32 (tensor-core lift), float32 with GPSIG_NO_MMA=1 (FFMA lift) and float64 (DFMA lift)."""
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
ls = np.full(d, np.sqrt(d / 8.0))


def run(dt, tag):
    Xd, Yd = torch.as_tensor(X, dtype=dt, device='cuda'), torch.as_tensor(Y, dtype=dt, device='cuda')
    lsd = torch.as_tensor(ls, dtype=dt, device='cuda')
    f = lambda: ops.trunc_gram(Xd, Yd, m, base='rbf', lengthscales=lsd, return_levels=False)  # noqa: E731
    f()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        f()
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("%-28s N=%d L=%d d=%d m=%d: %8.2f ms  %7.2f M pairs/s" % (tag, N, L, d, m, best, N * N / best / 1e3))


run(torch.float32, 'f32 tcgen05 lift')
os.environ['GPSIG_NO_MMA'] = '1'
run(torch.float32, 'f32 FFMA lift')
del os.environ['GPSIG_NO_MMA']
run(torch.float64, 'f64 DFMA lift')
