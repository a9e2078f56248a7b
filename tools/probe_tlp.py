"""TLP probe for the float64 systolic kernel: shape with 2 columns per lane (L2 = 64: 158 registers, so 12 warps per CTA fit
without spills).  Run with the default library and with a GPSIG_SKEW_NW_F64=12 variant (GPSIG_LIB_NAME)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402
from gpsig_b200._lib import lib  # noqa: E402

n, L, L2 = 2000, 200, 64
rng = np.random.default_rng(0)
X = torch.as_tensor(np.cumsum(rng.standard_normal((n, L, 8)) / np.sqrt(L), axis=1), dtype=torch.float64, device='cuda')
Y = torch.as_tensor(np.cumsum(rng.standard_normal((n, L2, 8)) / np.sqrt(L2), axis=1), dtype=torch.float64, device='cuda')
ops.trunc_gram(X, Y, 5, base='rbf', return_levels=False)
lib.gpsig_profile_enable(1)
ms = []
for _ in range(3):
    ops.trunc_gram(X, Y, 5, base='rbf', return_levels=False)
    ms.append(lib.gpsig_profile_last_kernel_ms())
print(os.environ.get('GPSIG_LIB_NAME', 'default'), "L2=64 f64: %.2f ms" % min(ms))
