#!/usr/bin/env python
"""Headline benchmark: signature-kernel Gram pairs/sec (L=200, M=5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--dtype f64|f32] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[4] "c5", per-pair shape of configs[1] "c2"):
SignatureRBF, num_levels=5, order=1, normalization=True, synthetic Brownian
paths, L = 200, d = 8.  One step = ONE row slab of the large Gram,
K(X[slab], X2) with 2000 slab rows against N2 = 20 000 columns = 4e7 ordered
pairs, computed through gpsig_b200.sharding.sharded_gram: the slab's rows are
split over the N ranks, every rank computes its rows straight into its slice of
the [2000, 20000] result and the slices are exchanged with ONE in-place NCCL
all-gather INSIDE the timed region, so the finished slab is on every rank
(strong scaling: the total work per step is fixed).  The full c2 Gram
(N = N2 = 2000) is timed as other_configs.c2_f64 / c2_f32.

Prints ONE JSON line (rank 0).  `value` = pairs/s with the inputs resident in
HBM; `e2e` = the same through the numpy-in/numpy-out call (pinned host
buffers, H2D + D2H inside the timed region); `roofline` = the fused Gram
kernel against the measured CUDA-core FMA peak (it is compute bound -- see
DESIGN.md); `cpu_baseline` = the C port of the reference algorithm
(oracle/) on the host cores, bounded sample.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CFG = dict(workload="c5 slab: SignatureRBF m=5 order=1 normalization=True, K(X[slab], X2) = 2000 rows x 20000 columns of the large "
                    "Gram per step (per-pair shape of c2: L=200, d=8), rows sharded over the ranks + in-place NCCL all-gather, Brownian paths",
           N=2000, N2=20000, L=200, d=8, m=5)
METRIC = "signature-kernel Gram pairs/sec (L=200,M=5)"
UNIT = "pairs/s"


def brownian(N, L, d, seed):
    rng = np.random.default_rng(seed)
    return np.cumsum(rng.standard_normal((N, L, d)) / np.sqrt(L), axis=1)


def flops_per_pair(L, L2, d, m, rbf=True):
    """2 d L L2 + c_k L L2 + 4 m (L-1)(L2-1); exp counted as part of c_k = 4."""
    return 2 * d * L * L2 + (4 if rbf else 0) * L * L2 + 4 * m * (L - 1) * (L2 - 1)


def bytes_per_pair(N, N2, L, L2, d, s, levels_out=1):
    """s * [(N L d + N2 L2 d) / (N N2) + outputs]."""
    return s * ((N * L * d + N2 * L2 * d) / (N * N2) + levels_out)


# --------------------------------------------------------------------------------------
# reference arm / CPU baseline: the C port of the reference algorithm (oracle/), all cores
# --------------------------------------------------------------------------------------
def cpu_pairs_per_sec(n_side, steps=1, warmup=0):
    from oracle import c_oracle
    c_oracle.build()
    L, d, m = CFG['L'], CFG['d'], CFG['m']
    X, Y = brownian(n_side, L, d, 0), brownian(n_side, L, d, 1)
    for _ in range(warmup):
        c_oracle.trunc_gram(X[:8], Y[:8], m, 'rbf')
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        # normalised K(X, X2) = Gram levels + the two diagonals (kernels.py:456-470)
        c_oracle.trunc_gram(X, Y, m, 'rbf')
        c_oracle.trunc_diag(X, m, 'rbf')
        c_oracle.trunc_diag(Y, m, 'rbf')
        times.append(time.perf_counter() - t0)
    return n_side * n_side / statistics.mean(times), statistics.mean(times)


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count()
    os.environ['OMP_NUM_THREADS'] = str(cores)     # torchrun pins it to 1; the CPU arm uses every host core
    # calibrate so that one step is a few seconds
    cpu_pairs_per_sec(16)
    v0, t0 = cpu_pairs_per_sec(32)
    side = int(max(16, min(512, (v0 * 4.0) ** 0.5)))
    v, t = cpu_pairs_per_sec(side, steps=args.steps, warmup=args.warmup)
    line = {"metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic Brownian paths (seeded)", "impl": "reference", "config": dict(CFG, sample="%dx%d pairs per step" % (side, side)),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%dx%d sequence pairs of the c2 shape per step, C port of signature_algs.py:8-35 + kernels.py glue, OpenMP over pairs" % (side, side)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------
# ours
# --------------------------------------------------------------------------------------
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits',
                                       '-lms', '100'], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smmax, reasons, power = [], [], set(), []
        for ln in self.f.read().splitlines():
            c = [x.strip() for x in ln.split(',')]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); smmax.append(float(c[2])); power.append(float(c[3]))
            except ValueError:
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), c[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smmax), "reasons": sorted(reasons),
                   "power_w_max": max(power), "samples": len(sm)}
        return out


def measured_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the headline kernel at the bench shape (c2), from the
# committed ncu --set full captures (algorithmic bytes: 83.2 MB f64 / 41.6 MB f32 per launch)
NCU_TRAFFIC_BYTES = {"f64": 77.79e6 + 17.83e6, "f32": 70.71e6 + 7.09e6}
# the slab launch of the headline step (rank 0's 2000 rows x 20000 columns, float64): dram read + write of that one launch
NCU_TRAFFIC_SLAB_F64 = 598322176 + 312179456
NCU_TRAFFIC_SLAB_SOURCE = "profiles/r02zc_slab_traffic.csv (ncu dram__bytes_read.sum + dram__bytes_write.sum of the slab launch at N=1, final round-2 kernel: 0.60 GB read, 0.31 GB written, against 0.60 GB algorithmic; the x rows are re-streamed from L2 by every CTA)"
NCU_TRAFFIC_SOURCE = "profiles/r02zc_c2_traffic.csv (ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, mean of 4, N=N2=2000 = c2, final round-2 kernels; part of the 16-32 MB output still sits in L2 when the kernel ends; float32 reads rows stored with every coordinate twice for the packed contraction: 1.7x its algorithmic input bytes)"


def fma_peak_tflops(torch, lib, esz):
    """CUDA-core FMA peak measured with the library's FMA-chain microbenchmark."""
    from gpsig_b200._lib import check, stream_handle
    dev = torch.device('cuda', torch.cuda.current_device())
    scratch = torch.zeros(64, dtype=torch.float64, device=dev)
    blocks, iters = 148 * 16, 20000
    best = 0.0
    for rep in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        check(lib.gpsig_fma_peak(esz, blocks, iters, ctypes.c_void_p(scratch.data_ptr()), stream_handle(dev)))
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        if rep > 0:
            best = max(best, 2.0 * blocks * 256 * 8 * iters / (ms * 1e-3) / 1e12)
    return best


def other_configs(torch, ops, lib):
    """Secondary kernels of the path on their BASELINE configs (kernel time from the library's events), rank 0 only."""
    out = {}

    def timed(fn, reps=3):
        fn()
        lib.gpsig_profile_enable(1)
        ms = []
        for _ in range(reps):
            fn()
            ms.append(lib.gpsig_profile_last_kernel_ms())
        lib.gpsig_profile_enable(0)
        return sum(ms) / len(ms)

    # c3: PDE, N=1000, L=500, d=4, dyadic order 2 (diagonal pairs, what the reference computes)
    N, L, d, order = 1000, 500, 4, 2
    X = torch.as_tensor(brownian(N, L, d, 0), dtype=torch.float64, device='cuda')
    ms = timed(lambda: ops.pde(X, None, dyadic_order=order))
    P = (L - 1) << order
    cells = float(N) * P * P
    out["c3_pde"] = {"workload": "kernels_pde Kdiag, N=1000, L=500, d=4, dyadic order 2 (P=1996)", "kernel": "pde_block_kernel",
                     "kernel_ms": ms, "pairs_per_s": N / (ms * 1e-3), "cells_per_s": cells / (ms * 1e-3),
                     "fp64_instr_per_cell": 3, "dtype": "f64"}
    # c3 backward: the reference's _KdiagGrad formula (reversed grid through an HBM scratch, 16 P^2 bytes per sequence)
    g = torch.ones(N, dtype=torch.float64, device='cuda')
    ops.pde_kdiag_vjp(X, g, dyadic_order=order)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    ops.pde_kdiag_vjp(X, g, dyadic_order=order)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    hbm_peak = measured_peaks()[0].get("hbm_gbs")
    gbs = 16.0 * P * P * N / (ms * 1e-3) / 1e9
    out["c3_pde_vjp"] = {"workload": "kernels_pde Kdiag backward (_KdiagGrad), N=1000, L=500, d=4, dyadic order 2", "kernel": "pde_vjp_sweep_kernel<0>+<1>",
                         "ms": ms, "seqs_per_s": N / (ms * 1e-3), "cells_per_s": 2 * cells / (ms * 1e-3),
                         "bound": "hbm", "algorithmic_bytes_per_seq": 16.0 * P * P, "achieved_gbs": gbs, "peak_gbs": hbm_peak,
                         "frac": (gbs / hbm_peak) if hbm_peak else None, "dtype": "f64"}
    del g
    torch.cuda.empty_cache()
    out.update(large_d(torch, ops, lib))
    # c1 with the higher-order recursion (order = num_levels = 4: the variant the reference validates against signatures,
    # notebooks/signature_kernel.ipynb cells 11-13): the general kernel (generic_kernels.cuh), one CTA per pair
    Xc = torch.as_tensor(brownian(100, 50, 3, 0), dtype=torch.float64, device='cuda')
    for order in (1, 4):
        ms = timed(lambda: ops.trunc_gram(Xc, None, 4, base='linear', order=order, return_levels=False), reps=5)
        out["c1_order%d_f64" % order] = {"workload": "c1: SignatureLinear m=4 order=%d K(X, X), N=100, L=50, d=3" % order,
                                         "kernel": "trunc_generic_kernel" if order > 1 else "trunc_skew_kernel (tri) + mirror_lower_kernel",
                                         "kernel_ms": ms, "delivered_pairs_per_s": 100 * 100 / (ms * 1e-3)}
    ref_cuda = reference_cuda_pde(torch, ops, timed)
    if ref_cuda:
        out["c3_vs_reference_cuda"] = ref_cuda
    # c5 (symmetric sweep) on a 4000-row block: K(X, X) computes the CTAs touching i <= j and mirrors the rest;
    # delivered = N^2 ordered pairs in the output, computed = what the kernel evaluated
    Ns = 4000
    Xs = brownian(Ns, 200, 8, 0)
    for dt, nm in ((torch.float64, 'f64'), (torch.float32, 'f32')):
        Xd = torch.as_tensor(Xs, dtype=dt, device='cuda')
        ms = timed(lambda: ops.trunc_gram(Xd, None, 5, base='rbf', return_levels=False))
        out["c5_sym_block_" + nm] = {"workload": "SignatureRBF m=5 K(X, X) raw, N=4000 of the 50k rows, L=200, d=8 (upper triangle + mirror)",
                                     "kernel": "trunc_skew_kernel (tri) + mirror_lower_kernel", "kernel_ms": ms,
                                     "delivered_pairs_per_s": Ns * Ns / (ms * 1e-3)}
        del Xd
    # c4: inducing tensors vs sequences, N=10k, L=100, d=8, m=4, Mz=500 (increments)
    N, L, d, m, Mz = 10000, 100, 8, 4, 500
    Xn = brownian(N, L, d, 0)
    Zn = np.random.default_rng(2).standard_normal((m * (m + 1) // 2, Mz, 2, d)) * 0.5
    for dt, nm in ((torch.float64, 'f64'), (torch.float32, 'f32')):
        Xd = torch.as_tensor(Xn, dtype=dt, device='cuda')
        Zd = torch.as_tensor(Zn, dtype=dt, device='cuda')
        ms = timed(lambda: ops.tens_vs_seq(Zd, Xd, m, increments=True, base='rbf', return_levels=False))
        out["c4_kzx_" + nm] = {"workload": "K_tens_vs_seq (increments), Mz=500, N=10000, L=100, d=8, m=4", "kernel": "tens_vs_seq_kernel2",
                               "kernel_ms": ms, "pairs_per_s": Mz * N / (ms * 1e-3), "flops_per_pair": 43574}
        del Xd, Zd
    # the same at d = 28 (ArabicDigits with num_lags=1 and a time channel, the reference's own experiment shape): the
    # lifted Kzx (tens_lift_gemm_kernel + tens_lift_recur_kernel)
    d28 = 28
    rng28 = np.random.default_rng(8)
    X28 = np.cumsum(rng28.standard_normal((N, L, d28)) / np.sqrt(L), axis=1) / np.sqrt(d28 / 8.0)
    Z28 = rng28.standard_normal((m * (m + 1) // 2, Mz, 2, d28)) * 0.5 / np.sqrt(d28 / 8.0)
    for dt, nm in ((torch.float64, 'f64'), (torch.float32, 'f32')):
        Xd = torch.as_tensor(X28, dtype=dt, device='cuda')
        Zd = torch.as_tensor(Z28, dtype=dt, device='cuda')
        ms = timed(lambda: ops.tens_vs_seq(Zd, Xd, m, increments=True, base='rbf', return_levels=False))
        out["c4_kzx_d28_" + nm] = {"workload": "K_tens_vs_seq (increments), Mz=500, N=10000, L=100, d=28, m=4",
                                   "kernel": "tens_lift_gemm_kernel + tens_lift_recur_kernel", "kernel_ms": ms,
                                   "pairs_per_s": Mz * N / (ms * 1e-3)}
        del Xd, Zd
    # c4 training step (train_gpsig.py:20,67-68): Kzz, Kzx, Kxx_diag of a minibatch of 50, forward + backward
    from gpsig_b200 import kernels as _k
    nb = 50
    for dt, nm in ((torch.float64, 'f64'), (torch.float32, 'f32')):
        kern = _k.SignatureRBF(L * d, d, m, lengthscales=np.ones(d))
        Xb = torch.as_tensor(Xn[:nb].reshape(nb, L * d), dtype=dt, device='cuda').requires_grad_(True)
        Zb = torch.as_tensor(Zn, dtype=dt, device='cuda').requires_grad_(True)

        def step():
            Xb.grad = None
            Zb.grad = None
            Kzz, Kzx, Kxx = kern.K_tens_n_seq_covs(Zb, Xb, full_X_cov=False, increments=True)
            (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()

        step()
        best = 1e30
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step()
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        out["c4_svgp_step_" + nm] = {"workload": "K_tens_n_seq_covs forward + backward (grads of Z and X), minibatch 50, Mz=500, L=100, d=8, m=4",
                                     "kernels": "tens_tens_tile(+vjp), trunc_gram diag (+vjp tables/grad), tens_vs_seq_kernel2 + tens_seq_vjp_warp_kernel",
                                     "ms": best}
        # the same step captured into a CUDA graph (fixed minibatch shape) and replayed: no host-side launch cost
        try:
            import warnings
            warnings.filterwarnings('ignore', message="The AccumulateGrad node's stream")
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step()
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            Xb.grad = None
            Zb.grad = None
            with torch.cuda.graph(graph):
                Kzz, Kzx, Kxx = kern.K_tens_n_seq_covs(Zb, Xb, full_X_cov=False, increments=True)
                (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()
            graph.replay()
            best = 1e30
            for _ in range(10):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                graph.replay()
                e1.record()
                e1.synchronize()
                best = min(best, e0.elapsed_time(e1))
            out["c4_svgp_step_" + nm]["ms_cuda_graph"] = best
            del graph
        except Exception as e:       # noqa: BLE001
            out["c4_svgp_step_" + nm]["ms_cuda_graph"] = "capture failed: %r" % (e,)
    return out


def large_d(torch, ops, lib):
    """State dimension d > 16 (the reference's benchmark shapes have d_eff = 26..1928: benchmarks/datasets.json with
    num_lags=1 and a time channel, run_gpsig_benchmarks.py:32): the lifted path, float32 lift on tcgen05 tensor cores
    (3 x TF32), float64 lift on DFMA.  N = N2 = 1000, L = 100, d = 126, m = 4, RBF with lengthscales sqrt(d/8)."""
    out = {}
    N, L, d, m = 1000, 100, 126, 4
    X, Y = brownian(N, L, d, 0), brownian(N, L, d, 1)
    F = flops_per_pair(L, L, d, m)
    LP2, KP = 128, 128
    for dt, nm in ((torch.float32, 'f32'), (torch.float64, 'f64')):
        Xd, Yd = torch.as_tensor(X, dtype=dt, device='cuda'), torch.as_tensor(Y, dtype=dt, device='cuda')
        ls = torch.full((d,), float(np.sqrt(d / 8.0)), dtype=dt, device='cuda')
        fn = lambda: ops.trunc_gram(Xd, Yd, m, base='rbf', lengthscales=ls, return_levels=False)  # noqa: E731
        fn()
        best = 1e30
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        rec = {"workload": "SignatureRBF m=4 K(X, X2) levels summed, N=N2=1000, L=100, d=126 (lengthscales sqrt(d/8))", "dtype": nm,
               "kernels": ("lift_mma_kernel (tcgen05.mma kind::tf32, 3-term split) + lift_recur_kernel" if nm == 'f32'
                           else "lift_tile_kernel (DFMA) + lift_recur_kernel"),
               "ms": best, "pairs_per_s": N * N / (best * 1e-3), "flops_per_pair": F,
               "algorithmic_tflops": N * N * F / (best * 1e-3) / 1e12}
        if nm == 'f32':
            rec["mma_tflops_issued"] = N * N * 3 * 2.0 * 128 * LP2 * KP / (best * 1e-3) / 1e12
            rec["note"] = ("whole call (pack + operand split + lift + recursion); the same shape on the FFMA lift (GPSIG_NO_MMA=1) "
                           "runs at ~5.2 M pairs/s; per-kernel split: profiles/r02f_launches_large_d.csv")
        out["c_large_d_" + nm] = rec
        del Xd, Yd
    torch.cuda.empty_cache()
    return out


def reference_cuda_pde(torch, ops, timed):
    """SURVEY.md's bar for the PDE kernel: the reference's own UntruncCovKernel (untrunc_cov_op_gpu.cu:5-36, launcher :71-81),
    compiled unmodified for sm_100a by oracle/build_ref.py into oracle/_ref/, timed beside pde_block_kernel on a shape it can
    run (P + 1 <= 1024 threads: L=250, dyadic order 2, P=996).  Only its KERNEL is timed (inputs resident: the [A,P,P] increment
    tensor that kernels_pde.py:176-180 materialises with matmul + tf.repeat, and the [A,P+2,P+2] grid of ones) - the
    materialisation itself is left out, which favours the reference."""
    so = os.path.join(ROOT, 'oracle', '_ref', 'libuntrunc_cov_ref.so')
    if not os.path.exists(so):
        return None
    ref = ctypes.CDLL(so)
    ref.gpsig_ref_untrunc_cov.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    ref.gpsig_ref_untrunc_cov.restype = ctypes.c_int
    A, L, d, order = 1000, 250, 4, 2
    P = (L - 1) << order
    X = torch.as_tensor(brownian(A, L, d, 0), dtype=torch.float64, device='cuda')
    dX = X[:, 1:] - X[:, :-1]
    E = torch.matmul(dX, dX.transpose(1, 2))                                   # [A, L-1, L-1]
    E = (E.repeat_interleave(1 << order, 1) / (1 << order)).repeat_interleave(1 << order, 2) / (1 << order)
    # the kernel also sweeps grid row/column P+1 and reads increments one row past the [P,P] block of the last sequence
    # (its results there are discarded by kernels_pde.py:183): pad so that the read stays inside the allocation
    incr = torch.zeros(A * P * P + 2 * P + 2, dtype=torch.float64, device='cuda')
    incr[:A * P * P] = E.reshape(-1)
    del E
    sol = torch.ones((A, P + 2, P + 2), dtype=torch.float64, device='cuda')
    best, got = 1e30, None
    for rep in range(3):
        sol.fill_(1.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # the reference launches on the legacy default stream (untrunc_cov_op_gpu.cu:80); so are these events
        with torch.cuda.stream(torch.cuda.default_stream()):
            e0.record()
            rc = ref.gpsig_ref_untrunc_cov(ctypes.c_void_p(incr.data_ptr()), A, P, ctypes.c_void_p(sol.data_ptr()))
            e1.record()
        e1.synchronize()
        if rc != 0:
            return {"error": "reference kernel launch failed with cudaError %d" % rc}
        best = min(best, e0.elapsed_time(e1))
        got = sol[:, P, P].clone()
    del incr, sol
    torch.cuda.empty_cache()
    ms = timed(lambda: ops.pde(X, None, dyadic_order=order))
    mine = ops.pde(X, None, dyadic_order=order)
    # the reference kernel rounds every increment to float32 (untrunc_cov_op_gpu.cu:27): agreement is ~1e-7, not 1e-10
    rel = float(((mine - got).abs() / got.abs()).max().item())
    return {"workload": "kernels_pde Kdiag, N=1000, L=250, d=4, dyadic order 2 (P=996; the largest the reference kernel runs is P+1 <= 1024)",
            "reference_kernel": "UntruncCovKernel, unmodified, nvcc -arch=sm_100a (oracle/_ref/libuntrunc_cov_ref.so)",
            "reference_kernel_ms": best, "reference_pairs_per_s": A / (best * 1e-3),
            "kernel": "pde_block_kernel", "kernel_ms": ms, "pairs_per_s": A / (ms * 1e-3), "speedup": best / ms,
            "max_rel_diff": rel, "note": "reference increments pass through a float (cu:27); its input tensors (8 GB incr + 8 GB grid) are resident and their construction is not timed"}


def c2_full(torch, kernels, lib, dt, nm, fpeak, steps=3):
    """The full c2 Gram (N = N2 = 2000, normalised K(X, X2)) on one GPU: value + roofline of trunc_skew_kernel."""
    N = 2000
    L, d, m = CFG['L'], CFG['d'], CFG['m']
    esz = 8 if nm == 'f64' else 4
    X = torch.as_tensor(brownian(N, L, d, 0).reshape(N, L * d), dtype=dt, device='cuda')
    X2 = torch.as_tensor(brownian(N, L, d, 1).reshape(N, L * d), dtype=dt, device='cuda')
    kern = kernels.SignatureRBF(L * d, d, m)
    kern.K(X, X2)
    lib.gpsig_profile_enable(1)
    ms, kms = [], []
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        kern.K(X, X2)
        e1.record()
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
        kms.append(lib.gpsig_profile_last_kernel_ms())
    lib.gpsig_profile_enable(0)
    t, kt = statistics.mean(ms), statistics.mean(kms)
    F = flops_per_pair(L, L, d, m)
    ach = N * N * F / (kt * 1e-3) / 1e12
    return {"workload": "c2: SignatureRBF m=5 normalised K(X, X2), N=N2=2000, L=200, d=8 (one GPU)", "dtype": nm, "ms_per_step": t,
            "pairs_per_s": N * N / (t * 1e-3),
            "roofline": {"bound": "cuda-core-" + nm, "kernel": "trunc_skew_kernel", "kernel_ms": kt, "achieved": ach, "peak": fpeak,
                         "unit": "TFLOP/s", "frac": ach / fpeak if fpeak else None, "flops_per_pair": F,
                         "traffic": NCU_TRAFFIC_BYTES.get(nm),
                         "algorithmic_bytes_per_launch": N * N * bytes_per_pair(N, N, L, L, d, esz)}}


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; there is no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    from gpsig_b200 import kernels, ops, settings, sharding
    from gpsig_b200._lib import lib
    dtype = torch.float64 if args.dtype == 'f64' else torch.float32
    esz = 8 if args.dtype == 'f64' else 4
    settings.float_type = dtype
    N, N2, L, d, m = CFG['N'], CFG['N2'], CFG['L'], CFG['d'], CFG['m']
    # the slab (rows) and the columns of the large Gram: the same on every rank (seeded), replicated as SURVEY 8(e) says
    X_host = torch.from_numpy(brownian(N, L, d, 0).reshape(N, L * d)).to(dtype).pin_memory()
    X2_host = torch.from_numpy(brownian(N2, L, d, 1).reshape(N2, L * d)).to(dtype).pin_memory()
    out_host = torch.empty((N, N2), dtype=dtype).pin_memory()
    X, X2 = X_host.to(dev), X2_host.to(dev)
    full = torch.empty((N, N2), dtype=dtype, device=dev)          # the gathered slab, on every rank
    kern = kernels.SignatureRBF(L * d, d, m)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    rb, re = sharding.row_range(N, world, rank)

    def step_resident():
        return sharding.sharded_gram(kern, X, X2, gather=True, out=full)

    def step_rows_only():                                         # the same without the collective (for `exposed`)
        return kern.K(X, X2, rows=(rb, re), out=full[rb:re])

    def step_e2e():
        x = X_host.to(dev, non_blocking=True)
        x2 = X2_host.to(dev, non_blocking=True)
        K = sharding.sharded_gram(kern, x, x2, gather=True, out=full)
        out_host.copy_(K, non_blocking=True)
        return K

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, profile_kernel=False):
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kms = []
        n0 = lib.gpsig_launch_count()
        lib.gpsig_profile_enable(1 if profile_kernel else 0)
        for a, b in ev:
            flush.fill_(1)                      # L2 flush between timed iterations (outside the events)
            a.record()
            fn()
            b.record()
            if profile_kernel:
                kms.append(lib.gpsig_profile_last_kernel_ms())
        barrier()
        lib.gpsig_profile_enable(0)
        launches = lib.gpsig_launch_count() - n0
        tot_ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([tot_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, kms

    sampler = ClockSampler(local) if rank == 0 else None
    tot_ms, launches, kms = timed(step_resident, args.steps, args.warmup, profile_kernel=True)
    clocks = sampler.stop() if sampler else None
    e2e_steps = args.steps if world > 1 else min(args.steps, 5)
    e2e_ms, _, _ = timed(step_e2e, e2e_steps, max(1, args.warmup // 2))
    rows_ms = None
    if world > 1:
        rsteps = min(args.steps, 5)
        rows_ms, _, _ = timed(step_rows_only, rsteps, 1)
        rows_ms /= rsteps
    pairs_per_step = N * N2
    value = pairs_per_step * args.steps / (tot_ms * 1e-3)
    e2e_value = pairs_per_step * e2e_steps / (e2e_ms * 1e-3)
    # parity of the gathered slab with a single-rank compute: the last rank's first rows, recomputed by rank 0 alone
    step_resident()
    gather_check = None
    if world > 1:
        b_last, e_last = sharding.row_range(N, world, world - 1)
        nchk = min(16, e_last - b_last)
        if rank == 0:
            alone = kern.K(X, X2, rows=(b_last, b_last + nchk))
            gather_check = {"rows": [b_last, b_last + nchk], "bit_equal_to_single_rank": bool(torch.equal(alone, full[b_last:b_last + nchk]))}
    torch.cuda.synchronize()

    if rank == 0:
        peaks, how = measured_peaks()
        fpeak = fma_peak_tflops(torch, lib, esz)
        kern_ms = statistics.mean(kms) if kms else None
        F = flops_per_pair(L, L, d, m)
        pairs_launch = (re - rb) * N2
        achieved = pairs_launch * F / (kern_ms * 1e-3) / 1e12 if kern_ms else None
        Bp = bytes_per_pair(re - rb, N2, L, L, d, esz)
        nominal = 37.2 if esz == 8 else 74.4
        step_ms = tot_ms / args.steps
        roof = {"bound": "cuda-core-" + args.dtype, "achieved": achieved, "peak": fpeak, "unit": "TFLOP/s",
                "frac": (achieved / fpeak) if (achieved and fpeak) else None,
                "traffic": (NCU_TRAFFIC_SLAB_F64 if (args.dtype == 'f64' and world == 1) else None),
                "traffic_source": NCU_TRAFFIC_SLAB_SOURCE if (args.dtype == 'f64' and world == 1) else None,
                "algorithmic_bytes_per_launch": pairs_launch * Bp,
                "peak_source": "FMA-chain microbenchmark in this run (gpsig_fma_peak); nominal %.1f TFLOP/s at 1965 MHz" % nominal,
                "frac_of_nominal": achieved / nominal if achieved else None,
                "kernel": "trunc_skew_kernel", "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / step_ms if kern_ms else None,
                "flops_per_pair": F, "pairs_per_launch": pairs_launch, "launch": "rank 0's rows [%d, %d) x %d columns" % (rb, re, N2),
                "hbm": {"algorithmic_bytes_per_pair": Bp, "achieved_gbs": pairs_launch * Bp / (kern_ms * 1e-3) / 1e9 if kern_ms else None,
                        "peak_gbs": peaks.get("hbm_gbs"), "peak_source": how}}
        coll = None
        if world > 1:
            gathered = (N - (re - rb)) * N2 * esz
            coll = {"op": "ncclAllGather, in place (torch.distributed.all_gather_into_tensor of the rank's row slice into the full slab)",
                    "calls_per_step": 1, "received_bytes_per_rank_per_step": gathered, "slab_bytes": N * N2 * esz,
                    "rows_only_ms_per_step": rows_ms, "exposed_ms_per_step": step_ms - rows_ms,
                    "exposed_frac_of_step": (step_ms - rows_ms) / step_ms, "gather_check": gather_check}
        # CPU baseline on a bounded sample (N=1 only: under torchrun the host threads are pinned to 1 per rank)
        cpu = None
        if world == 1:
            os.environ.setdefault('OMP_NUM_THREADS', str(os.cpu_count()))
            cpu_pairs_per_sec(16)
            v0, _ = cpu_pairs_per_sec(32)
            side = int(max(16, min(512, (v0 * 12.0) ** 0.5)))
            cv, ct = cpu_pairs_per_sec(side)
            cpu = {"value": cv, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                   "sample": "%dx%d pairs of the same shape in %.1f s (C port of the reference algorithm, OpenMP)" % (side, side, ct)}
        others = None
        if world == 1 and not args.no_others:
            del X, X2, full
            torch.cuda.empty_cache()
            try:                                  # the headline line above is already measured: never lose it to an extra config
                others = other_configs(torch, ops, lib)
                fp64 = fpeak if esz == 8 else fma_peak_tflops(torch, lib, 8)
                fp32 = fpeak if esz == 4 else fma_peak_tflops(torch, lib, 4)
                others["c2_f64"] = c2_full(torch, kernels, lib, torch.float64, 'f64', fp64)
                others["c2_f32"] = c2_full(torch, kernels, lib, torch.float32, 'f32', fp32)
                post_other_configs(others, fp64)
            except Exception as e:       # noqa: BLE001
                others = dict(others or {}, error="other_configs stopped: %r" % (e,))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": args.dtype, "data": "synthetic Brownian paths (seeded, SURVEY.md 8d)",
                "config": dict(CFG, parallelism="slab rows split over %d rank(s), %d rows each; X2 replicated" % (world, re - rb),
                               l2="flushed between timed iterations (256 MB write)", pairs_per_step=pairs_per_step),
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                        "h2d_bytes_per_step": int(X_host.numel() + X2_host.numel()) * esz,
                        "d2h_bytes_per_step": int(out_host.numel()) * esz,
                        "note": "per rank: pinned host X slab + X2 -> device, sharded_gram, gathered slab -> pinned host"},
                "gpu_launches": int(launches),
                "roofline": roof,
                "collective": coll,
                "other_configs": others,
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def post_other_configs(others, fp64):
    """Fractions that need the FP64 FMA peak of this run."""
    if others and fp64:
        for key in ("c3_pde",):
            if key in others:
                c3 = others[key]
                c3["frac_of_fp64_bound"] = c3["cells_per_s"] * 3 / (fp64 * 1e12 / 2)   # 3 FP64 instructions per cell vs measured DFMA rate


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--dtype', choices=['f64', 'f32'], default='f64')
    ap.add_argument('--impl', choices=['ours', 'reference'], default='ours')
    ap.add_argument('--no-others', action='store_true', help='skip other_configs (developer runs)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        world = int(os.environ.get('WORLD_SIZE', '1'))
        if args.gpus > 1 and world == 1:
            # convenience: relaunch under torchrun
            cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(args.gpus),
                   '--master-addr', '127.0.0.1', '--master-port', '29517', os.path.abspath(__file__)] + sys.argv[1:]
            raise SystemExit(subprocess.call(cmd))
        run_ours(args)


if __name__ == '__main__':
    main()
