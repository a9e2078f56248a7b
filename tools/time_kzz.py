"""Kzz (inducing tensors against themselves) forward and adjoint alone, tiled kernels against the first kernels
(GPSIG_TENS_TENS_OLD=1), CUDA events on torch's stream.  Usage: python tools/time_kzz.py [d ...]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpsig_b200 import ops  # noqa: E402


def ev_time(fn, reps=20):
    fn()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


m, Mz = 4, 500
for d in [int(a) for a in sys.argv[1:]] or [8, 28]:
    for dt in (torch.float64, torch.float32):
        Z = torch.as_tensor(np.random.default_rng(2).standard_normal((m * (m + 1) // 2, Mz, 2, d)) * 0.5 / np.sqrt(d / 8.0), dtype=dt, device='cuda')
        G = torch.randn(m + 1, Mz, Mz, dtype=dt, device='cuda')
        row = []
        for old in (False, True):
            if old:
                os.environ['GPSIG_TENS_TENS_OLD'] = '1'
            else:
                os.environ.pop('GPSIG_TENS_TENS_OLD', None)
            row += [ev_time(lambda: ops.tens_tens(Z, m, increments=True, base='rbf')),
                    ev_time(lambda: ops.tens_tens_vjp(Z, G, m, increments=True, base='rbf'))]
        os.environ.pop('GPSIG_TENS_TENS_OLD', None)
        print("d = %3d %s: tiles forward %.3f ms, adjoint %.3f ms; first kernels forward %.3f ms, adjoint %.3f ms"
              % (d, str(dt)[6:], row[0], row[1], row[2], row[3]), flush=True)
