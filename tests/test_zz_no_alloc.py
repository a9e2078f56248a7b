"""gpsig_trunc_diag_static_f64 -- the entry point that allocates nothing (scratch in a static __device__ array,
csrc/static_scratch.cu; no counterpart in the reference).

Called through the C ABI (ctypes) and compared with the existing entry point gpsig_trunc_diag_f64 on two calls
in a row (the second one smaller, so it finds the first call's contents in the shared array).  Collected last
on purpose (file name): a failure here must not hide the other tests under ``pytest -x``.
"""
import ctypes
import re
import os

import numpy as np
import pytest
import torch

from tests.util import brownian, rel_err

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _max_bytes():
    hdr = open(os.path.join(ROOT, 'include', 'gpsig_b200.h')).read()
    m = re.search(r'#define\s+GPSIG_STATIC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)', hdr)
    return int(m.group(1)) << int(m.group(2))


def _static_diag(X, m, base, out, ls=None):
    """Raw call of the new entry point: inputs and output only, no workspace argument."""
    from gpsig_b200._lib import lib, make_desc, ptr, stream_handle
    N, L, d = X.shape
    desc = make_desc(base, m, 1, True, ())
    with torch.cuda.device(X.device):
        return lib.gpsig_trunc_diag_static_f64(desc, ptr(X), N, L, d, ptr(ls), ptr(out), stream_handle(X.device))


def test_static_scratch_diag_matches_workspace_entry_point_twice_in_a_row():
    from gpsig_b200 import ops
    from gpsig_b200._lib import lib
    dev = torch.device('cuda', 0)
    m = 4
    # first call: a larger problem; second call: a smaller one, other base kernel, with lengthscales
    cases = [(brownian(300, 60, 5, 90), 'rbf', None), (brownian(37, 23, 3, 91), 'linear', np.array([0.8, 1.1, 1.4]))]
    prepared = []
    for Xn, base, ls in cases:
        X = torch.as_tensor(Xn, dtype=torch.float64, device=dev).contiguous()
        lst = None if ls is None else torch.as_tensor(ls, dtype=torch.float64, device=dev)
        out = torch.full((m + 1, X.shape[0]), float('nan'), dtype=torch.float64, device=dev)
        ref = ops.trunc_diag(X, m, base=base, lengthscales=lst)            # existing entry point gpsig_trunc_diag_f64
        prepared.append((X, base, lst, out, ref))
    torch.cuda.synchronize()
    free = []
    for X, base, lst, out, ref in prepared:                                # two calls in a row, nothing allocated between them
        rc = _static_diag(X, m, base, out, lst)
        assert rc == 0, lib.gpsig_last_error_string()
        torch.cuda.synchronize()
        free.append(torch.cuda.mem_get_info(dev)[0])
    # no cudaMalloc / cudaFree inside a call: the device's free memory does not move between the two calls (the very
    # first call of a process may make the CUDA runtime load the module that holds the 16 MiB array -- lazy loading --
    # which is why the comparison starts after it)
    assert free[1] == free[0]
    for X, base, lst, out, ref in prepared:
        assert torch.isfinite(out).all()
        assert rel_err(out.cpu().numpy(), ref.cpu().numpy()) < 1e-10      # the f64 tolerance of the other GPU tests
        assert torch.equal(out, ref)                                       # same kernels, same inputs: bit for bit
    # and against the CPU oracle, like every other parity test
    from oracle import c_oracle
    Xn, base, ls = cases[1]
    assert rel_err(prepared[1][3].cpu().numpy(), c_oracle.trunc_diag(Xn / ls, m, base)) < 1e-10


def test_static_scratch_refuses_shapes_beyond_the_array():
    from gpsig_b200._lib import lib
    dev = torch.device('cuda', 0)
    L, d, m = 100, 8, 3
    per_seq = lib.gpsig_trunc_diag_workspace_bytes(1, L, d, 8)
    N = _max_bytes() // (per_seq - 128 if per_seq > 256 else per_seq) + 64   # certainly too many sequences
    assert lib.gpsig_trunc_diag_workspace_bytes(N, L, d, 8) > _max_bytes()
    X = torch.zeros((N, L, d), dtype=torch.float64, device=dev)
    out = torch.zeros((m + 1, N), dtype=torch.float64, device=dev)
    rc = _static_diag(X, m, 'rbf', out)
    assert rc == -3                                                        # GPSIG_ERR_WORKSPACE
    msg = lib.gpsig_last_error_string().decode()
    assert 'static' in msg and str(_max_bytes()) in msg
    # the largest shape that fits still runs
    Nfit = N
    while lib.gpsig_trunc_diag_workspace_bytes(Nfit, L, d, 8) > _max_bytes():
        Nfit -= 1
    Xf = torch.as_tensor(brownian(Nfit, L, d, 92), dtype=torch.float64, device=dev)
    outf = torch.empty((m + 1, Nfit), dtype=torch.float64, device=dev)
    assert _static_diag(Xf, m, 'rbf', outf) == 0
    from gpsig_b200 import ops
    assert torch.equal(outf, ops.trunc_diag(Xf, m, base='rbf'))
