"""Ragged inputs: ``gpsig_b200.preprocessing`` (device mirror of gpsig/preprocessing.py) against a plain-Python
restatement, and the property the reference relies on (preprocessing.py:23-25): padding a path by repeating its
last element does not change the signature kernel."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ragged(seed, n=9, d=3, lmin=4, lmax=23):
    rng = np.random.default_rng(seed)
    return [np.cumsum(rng.standard_normal((int(rng.integers(lmin, lmax + 1)), d)), axis=0) / 3 for _ in range(n)]


@pytest.mark.parametrize("pre", [False, True])
@pytest.mark.parametrize("pad_with", [None, 0.5])
def test_tabulate_and_augment(pre, pad_with):
    from gpsig_b200 import preprocessing as pp
    from oracle import preprocessing_np as ref
    seqs = _ragged(80)
    tab = pp.tabulate_list_of_sequences(seqs, pad_with=pad_with, pre=pre)
    want = ref.tabulate(seqs, pad_with=pad_with, pre=pre)
    assert tab.is_cuda and np.array_equal(tab.cpu().numpy(), want)
    assert np.array_equal(pp.tabulate_list_of_sequences([s.T for s in seqs], orient_ax=1, pad_with=pad_with, pre=pre).cpu().numpy(), want)
    if not pre:
        assert np.allclose(pp.add_time_to_table(tab).cpu().numpy(), ref.add_time_to_table(want), rtol=0, atol=1e-15, equal_nan=True)
    assert np.allclose(pp.add_natural_parametrization_to_table(tab).cpu().numpy(), ref.add_natural_parametrization_to_table(want),
                       rtol=1e-14, atol=1e-14)
    for got, s in zip(pp.add_time_to_list(seqs), seqs):
        assert np.allclose(got.cpu().numpy()[:, 0], np.arange(1, len(s) + 1) / len(s)) and np.array_equal(got.cpu().numpy()[:, 1:], s)
    for got, s in zip(pp.add_natural_parametrization_to_list(seqs), seqs):
        arc = np.concatenate(([0.], np.cumsum(np.linalg.norm(np.diff(s, axis=0), axis=1))))
        assert np.allclose(got.cpu().numpy()[:, 0], arc, rtol=1e-14, atol=1e-14)
    with pytest.raises(ValueError):
        pp.tabulate_list_of_sequences([seqs[0], seqs[1][:, :2]])
    with pytest.raises(ValueError):
        pp.tabulate_list_of_sequences([seqs[0], seqs[1][0]])


def test_padding_by_repetition_leaves_the_kernels_unchanged():
    """Each ragged sequence against each other, one pair at a time at their own lengths (oracle), equals the
    tabulated batch (CUDA): truncated levels and the PDE kernel."""
    from gpsig_b200 import kernels, kernels_pde, preprocessing as pp
    from oracle import c_oracle, pde_np
    seqs = _ragged(81, n=7, d=2, lmin=5, lmax=40)
    m = 4
    tab = pp.tabulate_list_of_sequences(seqs)
    N, L, d = tab.shape
    k = kernels.SignatureRBF(L * d, d, m, normalization=False)
    K = k.K(tab.reshape(N, -1), return_levels=True).cpu().numpy()
    for i in range(N):
        for j in range(N):
            ref = c_oracle.trunc_gram(seqs[i][None], seqs[j][None], m, 'rbf')[:, 0, 0]
            assert np.max(np.abs(K[:, i, j] - ref)) < 1e-10 * max(1.0, np.max(np.abs(ref)))
    kp = kernels_pde.UntruncSignatureKernel(L * d, d, order=1, implementation='cython')
    kd = kp.Kdiag(tab.reshape(N, -1)).cpu().numpy()
    for i in range(N):
        ref = pde_np.pde_kernel(seqs[i][None], None, 1)[0]
        assert abs(kd[i] - ref) < 1e-10 * abs(ref)
