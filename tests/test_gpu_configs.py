"""GPU parity at the FULL sizes of the five BASELINE.json configs: the CUDA path runs the whole
config, the oracle checks seeded sub-blocks it can afford, size-independent properties cover the rest."""
import numpy as np
import pytest
import torch

from tests.util import brownian, rel_err

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-10, torch.float32: 1e-5}


def _dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype, device='cuda')


def test_c1_signature_linear_m4_full():
    """C1: SignatureLinear m=4, N=100, L=50, d=3, float64, K(X, X) -- every entry against the oracle."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    N, L, d, m = 100, 50, 3, 4
    X = brownian(N, L, d, 0)
    for normalization in (True, False):
        k = kernels.SignatureLinear(L * d, d, m, normalization=normalization)
        o = SignatureKernelOracle(L * d, d, m, base='linear', normalization=normalization, block=25)
        got = k.K(_dev(X.reshape(N, -1)), return_levels=True)
        assert torch.equal(got, got.transpose(1, 2))
        assert rel_err(got.cpu().numpy(), o.K(X.reshape(N, -1), return_levels=True)) < 1e-10


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_c2_rbf_m5_2000x2000(dtype):
    """C2: SignatureRBF m=5, N=N2=2000, L=200, d=8: the full Gram; a seeded 24x24 sub-block against the C port of
    the reference algorithm; K(X, X2) = K(X2, X)^T; row blocks stitch bit for bit."""
    from gpsig_b200 import kernels, ops
    from oracle import c_oracle
    N, L, d, m = 2000, 200, 8, 5
    Xn, Yn = brownian(N, L, d, 0), brownian(N, L, d, 1)
    X, Y = _dev(Xn, dtype), _dev(Yn, dtype)
    raw = ops.trunc_gram(X, Y, m, base='rbf')                                  # [m+1, N, N]
    rng = np.random.default_rng(5)
    ri, ci = np.sort(rng.choice(N, 24, replace=False)), np.sort(rng.choice(N, 24, replace=False))
    ref = c_oracle.trunc_gram(Xn[ri], Yn[ci], m, 'rbf')
    got = raw[:, torch.as_tensor(ri, device='cuda')][:, :, torch.as_tensor(ci, device='cuda')].cpu().numpy()
    assert np.max(np.abs(got - ref) / np.abs(ref).max(axis=(1, 2), keepdims=True)) < TOL[dtype]
    rawT = ops.trunc_gram(Y, X, m, base='rbf')
    assert float((raw - rawT.transpose(1, 2)).abs().max() / raw.abs().max()) < TOL[dtype]
    assert torch.isfinite(raw).all()
    del rawT
    k = kernels.SignatureRBF(L * d, d, m)
    K = k.K(X.reshape(N, -1), Y.reshape(N, -1))
    parts = torch.cat([k.K(X.reshape(N, -1), Y.reshape(N, -1), rows=r) for r in ((0, 700), (700, 2000))], dim=0)
    assert torch.equal(parts, K)
    # normalised kernel: |K_l| <= 1 per level (Cauchy-Schwarz), so |K| <= m + 1
    assert float(K.abs().max()) <= (m + 1) * (1 + 1e-4)
    # the NORMALISED output (what bench.py times), same seeded sub-block, against the oracle's K (kernels.py:456-477):
    # the oracle's diagonals are those of the sub-block's own sequences, as in the full call
    from oracle.trunc_np import SignatureKernelOracle
    o = SignatureKernelOracle(L * d, d, m, base='rbf', block=8)
    refn = o.K(Xn[ri].reshape(24, -1), Yn[ci].reshape(24, -1), return_levels=True)
    Kl = k.K(X.reshape(N, -1), Y.reshape(N, -1), return_levels=True)
    gotn = Kl[:, torch.as_tensor(ri, device='cuda')][:, :, torch.as_tensor(ci, device='cuda')].cpu().numpy()
    assert rel_err(gotn, refn) < TOL[dtype]
    assert rel_err(K[torch.as_tensor(ri, device='cuda')][:, torch.as_tensor(ci, device='cuda')].cpu().numpy(), refn.sum(0)) < TOL[dtype]
    del Kl
    # row block written in place into a slice of a larger matrix (the sharded path's out=)
    full = torch.full((N, N), float('nan'), dtype=dtype, device='cuda')
    k.K(X.reshape(N, -1), Y.reshape(N, -1), rows=(700, 2000), out=full[700:2000])
    assert torch.equal(full[700:2000], K[700:2000]) and bool(torch.isnan(full[:700]).all())


def test_c3_pde_n1000_l500_order2():
    """C3: kernels_pde Kdiag, N=1000, L=500, d=4, dyadic order 2 (P = 1996): all 1000 diagonal pairs; three of them
    and two off-diagonal pairs against the restated solver; diag(K(X, X)) rows == Kdiag."""
    from gpsig_b200 import kernels_pde, ops
    from oracle import pde_np
    N, L, d, order = 1000, 500, 4, 2
    Xn = brownian(N, L, d, 0)
    X = _dev(Xn)
    k = kernels_pde.UntruncSignatureKernel(L * d, d, order=order, implementation='cython')
    kd = k.Kdiag(X.reshape(N, -1))
    assert kd.shape == (N,) and torch.isfinite(kd).all() and float(kd.min()) >= 1.0
    pick = [0, 517, 999]
    ref = pde_np.pde_kernel(Xn[pick], None, order)
    assert np.max(np.abs(kd[pick].cpu().numpy() - ref) / np.abs(ref)) < 1e-10
    off = ops.pde(X, _dev(Xn[::-1].copy()), dyadic_order=order, gram=False)      # 1000 off-diagonal pairs (i, N-1-i)
    ref = pde_np.pde_kernel(Xn[[3, 640]], Xn[[N - 4, N - 641]], order)
    assert np.max(np.abs(off[[3, 640]].cpu().numpy() - ref) / np.abs(ref)) < 1e-10
    rows = ops.pde(X, None, dyadic_order=order, gram=True, rows=(10, 12))          # 2 x 1000 Gram rows
    assert torch.allclose(torch.stack([rows[0, 10], rows[1, 11]]), kd[10:12], rtol=1e-12, atol=0)
    # the backward at full length on a few sequences (the reference's formula, oracle restatement)
    g = ops.pde_kdiag_vjp(X[:2].contiguous(), torch.ones(2, dtype=torch.float64, device='cuda'), dyadic_order=order)
    assert rel_err(g.cpu().numpy(), pde_np.pde_kdiag_grad(Xn[:2], order, 0)) < 1e-10


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_c4_inducing_tensor_covariances_n10k(dtype):
    """C4: SignatureRBF m=4, N=10 000, L=100, d=8, 500 inducing tensors (increments): Kzz in full, Kzx on 16 seeded
    columns, Kxx diagonal on the same sequences, against the oracle."""
    from gpsig_b200 import kernels
    from oracle.trunc_np import SignatureKernelOracle
    from tests.util import inducing_tensors
    N, L, d, m, Mz = 10000, 100, 8, 4, 500
    Xn = brownian(N, L, d, 0)
    Zn = inducing_tensors(Xn, m, Mz, True, 2)
    k = kernels.SignatureRBF(L * d, d, m)
    o = SignatureKernelOracle(L * d, d, m, base='rbf')
    Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(_dev(Zn, dtype), _dev(Xn.reshape(N, -1), dtype), increments=True)
    assert Kzz.shape == (Mz, Mz) and Kzx.shape == (Mz, N) and Kxx.shape == (N,)
    cols = np.sort(np.random.default_rng(6).choice(N, 16, replace=False))
    rzz, rzx, rxx = o.K_tens_n_seq_covs(Zn, Xn[cols].reshape(16, -1), increments=True)
    ct = torch.as_tensor(cols, device='cuda')
    assert rel_err(Kzz.cpu().numpy(), rzz) < TOL[dtype]
    assert rel_err(Kzx[:, ct].cpu().numpy(), rzx) < TOL[dtype]
    assert rel_err(Kxx[ct].cpu().numpy(), rxx) < TOL[dtype]
    # column shards (the multi-GPU unit of Kzx) stitch bit for bit
    parts = torch.cat([k.K_tens_vs_seq(_dev(Zn, dtype), _dev(Xn.reshape(N, -1), dtype), increments=True, cols=c)
                       for c in ((0, 3000), (3000, N))], dim=1)
    assert torch.equal(parts, k.K_tens_vs_seq(_dev(Zn, dtype), _dev(Xn.reshape(N, -1), dtype), increments=True))


def test_c5_symmetric_sweep_block():
    """C5 (SignatureRBF m=5, L=200, d=8, K(X, X) row-sharded) at N = 4000 of its 50 000 rows: the symmetric
    (upper triangle + mirror) result, the two row shards of a 2-rank split and the oracle on a sub-block agree."""
    from gpsig_b200 import kernels, ops
    from oracle import c_oracle
    N, L, d, m = 4000, 200, 8, 5
    Xn = brownian(N, L, d, 0)
    X = _dev(Xn.reshape(N, -1), torch.float32)
    k = kernels.SignatureRBF(L * d, d, m, normalization=False)
    K = k.K(X)
    assert torch.equal(K, K.T)
    shards = torch.cat([k.K(X, rows=r) for r in ((0, 2000), (2000, 4000))], dim=0)
    assert float((shards - K).abs().max() / K.abs().max()) < 1e-5
    up = torch.cat([k.K(X, rows=r, upper_only=True) for r in ((0, 1000), (1000, 2000), (2000, 3000), (3000, 4000))], dim=0)
    assert torch.equal(ops.mirror_lower(up), K)
    sub = np.r_[0:8, 1996:2004, 3992:4000]
    ref = c_oracle.trunc_gram(Xn[sub], Xn[sub], m, 'rbf').sum(axis=0)
    assert rel_err(K[torch.as_tensor(sub, device='cuda')][:, torch.as_tensor(sub, device='cuda')].cpu().numpy(), ref) < 1e-5
