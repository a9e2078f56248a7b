import os, sys
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from gpsig_b200 import ops
from gpsig_b200._lib import lib
n = 1000
rng = np.random.default_rng(0)
X = np.cumsum(rng.standard_normal((n, 200, 8)) / np.sqrt(200), axis=1)
Y = np.cumsum(rng.standard_normal((n, 200, 8)) / np.sqrt(200), axis=1)
for dt in (torch.float64, torch.float32):
    Xd, Yd = torch.as_tensor(X, dtype=dt, device='cuda'), torch.as_tensor(Y, dtype=dt, device='cuda')
    for base in ('rbf', 'linear'):
        for m in (5, 4, 2):
            ops.trunc_gram(Xd, Yd, m, base=base, return_levels=False)
            lib.gpsig_profile_enable(1)
            ms = []
            for _ in range(3):
                ops.trunc_gram(Xd, Yd, m, base=base, return_levels=False)
                ms.append(lib.gpsig_profile_last_kernel_ms())
            lib.gpsig_profile_enable(0)
            print(str(dt)[6:], base, 'm=%d' % m, '%.2f ms %.2f M pairs/s' % (min(ms), n * n / min(ms) / 1e3))
