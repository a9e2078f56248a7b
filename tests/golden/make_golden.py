"""Generates the committed golden fixtures.  Run in the BUILD container only
(it needs /root/reference for the PDE vectors):

    python tests/golden/make_golden.py

* ``pde_ref.npz``    outputs of the REFERENCE's own compiled Cython solver
                     (gpsig/sigKer_fast.pyx via oracle/build_ref.py) -- pins the PDE path; the
                     ``kdiag_grad_*`` entries are the reference's backward formula (a transcription of
                     the TF algebra of ``_KdiagGrad``) evaluated on the reference's own two grids.
* ``trunc_oracle.npz`` outputs of the numpy restatement of the TensorFlow path
                     (oracle/trunc_np.py) on seeded inputs -- pins later refactors of the
                     oracle and the CUDA path to today's numbers (the reference's TF-1.15
                     path itself cannot run here: "parity unpinned" in the strict sense).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import build_ref, pde_np, trunc_np  # noqa: E402
from tests.util import brownian, inducing_tensors  # noqa: E402


def make_pde():
    build_ref.build()
    ref = build_ref.load()
    if ref is None:
        raise SystemExit("reference Cython module could not be built (no /root/reference?)")
    X = brownian(8, 24, 3, 100)
    out = {'X': X}
    for order in (0, 1, 2):
        K, K_rev = ref.sig_kern_diag(X, order, 1)
        K, K_rev = np.array(K), np.array(K_rev)
        out['kdiag_order%d' % order] = K[:, -1, -1]
        out['krev_corner_order%d' % order] = K_rev[:, -1, -1]
        # the reference's hand-written backward (_KdiagGrad, kernels_pde.py:465-509) applied to ITS grids
        out['kdiag_grad_order%d' % order] = pde_np.kdiag_grad_from_grids(
            X, pde_np.symmetrise_lower(K), pde_np.symmetrise_lower(K_rev), order)
        if order < 2:
            out['grid_order%d' % order] = K[:2]
            out['grid_rev_order%d' % order] = K_rev[:2]
    np.savez_compressed(os.path.join(HERE, 'pde_ref.npz'), **out)


def make_trunc():
    N, N2, L, L2, d, m, Mz = 6, 5, 14, 11, 3, 4, 4
    X, Y = brownian(N, L, d, 101), brownian(N2, L2, d, 102)
    ls = np.array([0.9, 1.2, 0.8])
    out = {'X': X, 'Y': Y, 'lengthscales': ls}
    for inc in (False, True):
        out['Z_inc%d' % inc] = inducing_tensors(X, m, Mz, inc, 103)
    for base in ('linear', 'rbf'):
        for order in (1, 2, m):
            for norm in (False, True):
                o = trunc_np.SignatureKernelOracle(L * d, d, m, base=base, order=order, normalization=norm,
                                                   lengthscales=ls, sigma=1.5, variances=np.linspace(0.5, 1.5, m + 1))
                tag = '%s_o%d_n%d' % (base, order, norm)
                out['K_' + tag] = o.K(X.reshape(N, -1), Y.reshape(N2, -1), return_levels=True)
                out['Ksym_' + tag] = o.K(X.reshape(N, -1), return_levels=True)
                for inc in (False, True):
                    Kzz, Kzx, Kxx = o.K_tens_n_seq_covs(out['Z_inc%d' % inc], X.reshape(N, -1), return_levels=True, increments=inc)
                    out['Kzz_%s_i%d' % (tag, inc)] = Kzz
                    out['Kzx_%s_i%d' % (tag, inc)] = Kzx
                    out['Kxxd_%s_i%d' % (tag, inc)] = Kxx
    np.savez_compressed(os.path.join(HERE, 'trunc_oracle.npz'), **out)


if __name__ == '__main__':
    make_pde()
    make_trunc()
    print('wrote', sorted(f for f in os.listdir(HERE) if f.endswith('.npz')))
