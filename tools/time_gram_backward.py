"""K(X, X2) raw levels forward + backward at a Gram size (inducing sequences), for ncu launch lists."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import autograd as ag  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dt = torch.float32 if 'f32' in sys.argv else torch.float64
L, d, m = 100, 8, 4
rng = np.random.default_rng(0)
X = torch.as_tensor(np.cumsum(rng.standard_normal((n, L, d)) / 10, axis=1), dtype=dt, device='cuda').requires_grad_(True)
Y = torch.as_tensor(np.cumsum(rng.standard_normal((n, L, d)) / 10, axis=1), dtype=dt, device='cuda').requires_grad_(True)
for _ in range(3):
    X.grad = None; Y.grad = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ag.trunc_levels(X, Y, m, base='rbf').sum().backward()
    e1.record(); e1.synchronize()
    print("K(X, X2) %dx%d fwd+bwd %.3f ms" % (n, n, e0.elapsed_time(e1)))
