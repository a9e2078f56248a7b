"""Forward + backward of the SVGP covariances at the C4 training shape (train_gpsig.py:20,67-68: minibatch 50).

Times, with CUDA events on torch's stream, each reverse-mode op alone and the whole
`K_tens_n_seq_covs` forward + backward (Kzz, Kzx, Kxx_diag -> scalar -> gradients of Z and X).
Usage: python tools/time_training_step.py [float64|float32 ...] [minibatch ...]
"""
import os
import sys

import warnings

import numpy as np
import torch

warnings.filterwarnings('ignore', message="The AccumulateGrad node's stream")

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpsig_b200  # noqa: E402
from gpsig_b200 import autograd as ag  # noqa: E402


def brownian(N, L, d, seed):
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


def ev_time(fn, reps=5):
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


L, d, m, Mz = 100, int(os.environ.get('GPSIG_TIME_D', '8')), 4, 500      # GPSIG_TIME_D=28: the reference's ArabicDigits shape
TT = m * (m + 1) // 2
Zn = np.random.default_rng(2).standard_normal((TT, Mz, 2, d)) * 0.5 / np.sqrt(d / 8.0)
DTYPES = [getattr(torch, a) for a in sys.argv[1:] if a.startswith('float')] or [torch.float64, torch.float32]
BATCHES = [int(a) for a in sys.argv[1:] if a.isdigit()] or [50, 1000]
for dt in DTYPES:
    name = str(dt)[6:]
    for nb in BATCHES:
        X = torch.as_tensor(brownian(nb, L, d, 0) / np.sqrt(d / 8.0), dtype=dt, device='cuda').requires_grad_(True)
        Z = torch.as_tensor(Zn, dtype=dt, device='cuda').requires_grad_(True)

        def diag_step():
            X.grad = None
            ag.trunc_diag_levels(X, m, base='rbf').sum().backward()

        def kzx_step():
            X.grad = None; Z.grad = None
            ag.tens_vs_seq_levels(Z, X, m, increments=True, base='rbf').sum().backward()

        def kzz_step():
            Z.grad = None
            ag.tens_tens_levels(Z, m, increments=True, base='rbf').sum().backward()

        print("%s minibatch %d: Kxx_diag fwd+bwd %.3f ms, Kzx fwd+bwd %.3f ms, Kzz fwd+bwd %.3f ms" %
              (name, nb, ev_time(diag_step), ev_time(kzx_step), ev_time(kzz_step)), flush=True)

        k = gpsig_b200.kernels.SignatureRBF(d * L, d, m, lengthscales=np.ones(d))
        X2 = X.detach().reshape(nb, L * d).clone().requires_grad_(True)
        Zt = Z.detach().clone().requires_grad_(True)

        def full_step():
            X2.grad = None; Zt.grad = None
            Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Zt, X2, full_X_cov=False, increments=True)
            (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()

        print("%s minibatch %d: K_tens_n_seq_covs fwd+bwd %.3f ms" % (name, nb, ev_time(full_step)), flush=True)

        # the same step captured once into a CUDA graph and replayed (static shapes: what a training loop with a fixed
        # minibatch does): removes the host-side launch cost of its ~45 kernels
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    full_step()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            X2.grad = None; Zt.grad = None
            with torch.cuda.graph(graph):
                Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Zt, X2, full_X_cov=False, increments=True)
                (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()
            cX, cZ = X2.grad, Zt.grad             # the buffers the captured backward writes
            full_step()                           # eager, into fresh .grad tensors
            eX, eZ = X2.grad.clone(), Zt.grad.clone()
            cX.zero_(); cZ.zero_()
            graph.replay()
            torch.cuda.synchronize()
            err = max(float((cX - eX).abs().max() / eX.abs().max()), float((cZ - eZ).abs().max() / eZ.abs().max()))
            print("%s minibatch %d: K_tens_n_seq_covs fwd+bwd as a CUDA graph %.3f ms (grads against eager: %.1e)" %
                  (name, nb, ev_time(graph.replay, reps=20), err), flush=True)
            del graph
            fgr = k.graphed_tens_n_seq_covs(Zt, X2, increments=True)       # forward / backward graphs behind autograd

            def graphed_step():
                X2.grad = None; Zt.grad = None
                Kzz, Kzx, Kxx = fgr(Zt, X2)
                (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()
            print("%s minibatch %d: kern.graphed_tens_n_seq_covs fwd+bwd %.3f ms" % (name, nb, ev_time(graphed_step, reps=20)), flush=True)
        except Exception as e:       # noqa: BLE001
            print("CUDA graph capture failed: %r" % (e,), flush=True)

# Gram-sized backward (inducing sequences: K_seq_n_seq_covs): K(X, X2) raw levels, forward + backward
for dt in (DTYPES if 'GPSIG_TIME_D' not in os.environ else []):
    for n in (64, 256):
        X = torch.as_tensor(brownian(n, L, d, 0), dtype=dt, device='cuda').requires_grad_(True)
        Y = torch.as_tensor(brownian(n, L, d, 1), dtype=dt, device='cuda').requires_grad_(True)

        def gram_step():
            X.grad = None; Y.grad = None
            ag.trunc_levels(X, Y, m, base='rbf').sum().backward()

        print("%s K(X, X2) %dx%d L=%d fwd+bwd %.3f ms" % (str(dt)[6:], n, n, L, ev_time(gram_step)), flush=True)
