"""Generates the committed golden fixtures.  Run in the BUILD container only
(it needs /root/reference for the PDE vectors):

    python tests/golden/make_golden.py
    python tests/golden/make_golden.py --trunc-ref

* ``pde_ref_py.npz`` the reference's PDE Python layer on the same stand-ins: ``UntruncSignatureKernel.Kdiag`` (lengthscales,
                     sigma, lags) and its own backward function ``_KdiagGrad``.
* ``trunc_ref.npz``  outputs of the REFERENCE's own Python code for the truncated path (gpsig/kernels.py,
                     signature_algs.py, lags.py executed where they lie, on the eager numpy stand-ins for
                     TensorFlow / GPflow of oracle/refrun) -- pins the truncated path to the reference itself.

* ``pde_ref.npz``    outputs of the REFERENCE's own compiled Cython solver
                     (gpsig/sigKer_fast.pyx via oracle/build_ref.py) -- pins the PDE path; the
                     ``kdiag_grad_*`` entries are the reference's backward formula (a transcription of
                     the TF algebra of ``_KdiagGrad``) evaluated on the reference's own two grids.
* ``trunc_oracle.npz`` outputs of the numpy restatement of the TensorFlow path
                     (oracle/trunc_np.py) on seeded inputs -- pins later refactors of the
                     oracle and the CUDA path to today's numbers (kept next to trunc_ref.npz,
                     which holds the reference's own outputs for the same inputs).
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


def _val(t):
    while hasattr(t, 'a'):
        t = t.a
    return np.asarray(t, dtype=np.float64)


# (oracle base name, reference class, reference constructor extras, oracle base parameters)
REF_BASES = [('linear', 'SignatureLinear', {}, {}), ('rbf', 'SignatureRBF', {}, {}), ('cosine', 'SignatureCosine', {}, {}),
             ('poly', 'SignaturePoly', dict(gamma=0.7, degree=3), dict(gamma=0.7, degree=3.0)), ('mix', 'SignatureMix', {}, {}),
             ('matern12', 'SignatureMatern12', {}, {}), ('matern32', 'SignatureMatern32', {}, {}),
             ('matern52', 'SignatureMatern52', {}, {})]


def make_trunc_ref():
    """Outputs of the REFERENCE's own gpsig/kernels.py + signature_algs.py + lags.py, executed where they lie under
    /root/reference on the eager numpy stand-ins for TensorFlow / GPflow (oracle/refrun): pins the truncated path --
    oracle and CUDA kernels -- to the reference's code itself.  Same inputs and keys as trunc_oracle.npz, plus every
    base kernel, lags, difference = False and the sequence-vs-sequence covariances."""
    from oracle import refrun
    ref = refrun.load_reference()
    N, N2, L, L2, d, m, Mz = 6, 5, 14, 11, 3, 4, 4
    X, Y = brownian(N, L, d, 101), brownian(N2, L2, d, 102)
    Xf, Yf = X.reshape(N, -1), Y.reshape(N2, -1)
    Y14 = brownian(N2, L, d, 104)                      # same length as X (the lag interpolation needs a common time grid)
    ls = np.array([0.9, 1.2, 0.8])
    var = np.linspace(0.5, 1.5, m + 1)
    out = {'X': X, 'Y': Y, 'Y14': Y14, 'lengthscales': ls, 'variances': var, 'sigma': np.float64(1.5)}
    for inc in (False, True):
        out['Z_inc%d' % inc] = inducing_tensors(X, m, Mz, inc, 103)

    def kern(cls, extra, **kw):
        k = getattr(ref, cls)(L * d, d, m, lengthscales=ls, variances=var, **extra, **kw)
        k.sigma = k.sigma * 0. + 1.5
        return k

    for base, cls in (('linear', 'SignatureLinear'), ('rbf', 'SignatureRBF')):
        for order in (1, 2, m):
            for norm in (False, True):
                k = kern(cls, {}, order=order, normalization=norm)
                tag = '%s_o%d_n%d' % (base, order, norm)
                out['K_' + tag] = _val(k.K(Xf, Yf, return_levels=True))
                out['Ksym_' + tag] = _val(k.K(Xf, return_levels=True))
                out['Kdiag_' + tag] = _val(k.Kdiag(Xf, return_levels=True))
                for inc in (False, True):
                    Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(out['Z_inc%d' % inc], Xf, return_levels=True, increments=inc)
                    out['Kzz_%s_i%d' % (tag, inc)] = _val(Kzz)
                    out['Kzx_%s_i%d' % (tag, inc)] = _val(Kzx)
                    out['Kxxd_%s_i%d' % (tag, inc)] = _val(Kxx)
    # every base kernel, first order (the BASELINE configs), summed levels and the per-level diagonals
    for base, cls, extra, _ in REF_BASES:
        for norm in (False, True):
            k = kern(cls, extra, order=1, normalization=norm)
            tag = '%s_n%d' % (base, norm)
            out['Kall_' + tag] = _val(k.K(Xf, Yf))
            out['Kdiagall_' + tag] = _val(k.Kdiag(Xf))
            out['Kzxall_' + tag] = _val(k.K_tens_vs_seq(out['Z_inc1'], Xf, increments=True))
            out['Kzzall_' + tag] = _val(k.K_tens(out['Z_inc1'], increments=True))
    # options: no differencing, lags, the sequence-vs-sequence covariance triple
    k = kern('SignatureRBF', {}, order=1, normalization=True, difference=False)
    out['Knodiff_rbf'] = _val(k.K(Xf, Yf, return_levels=True))
    k = kern('SignatureRBF', {}, order=1, normalization=False, difference=False)
    out['Knodiff_rbf_n0'] = _val(k.K(Xf, Yf, return_levels=True))
    for norm in (False, True):
        k = kern('SignatureRBF', {}, order=1, normalization=norm, num_lags=2)
        out['Klags_rbf_n%d' % norm] = _val(k.K(Xf, Y14.reshape(N2, -1), return_levels=True))
        out['Kdiaglags_rbf_n%d' % norm] = _val(k.Kdiag(Xf, return_levels=True))
        k = kern('SignatureRBF', {}, order=1, normalization=norm)
        Kxx, Kxx2, Kx2x2 = k.K_seq_n_seq_covs(Xf, Yf, return_levels=True)
        out['Ksns_xx_n%d' % norm], out['Ksns_xx2_n%d' % norm], out['Ksns_x2x2_n%d' % norm] = _val(Kxx), _val(Kxx2), _val(Kx2x2)
    # spectral base kernel: the constructor draws alpha / omega / gamma itself (kernels.py:1092-1094), read back here.
    # Only the un-normalised K(X, X2) can be evaluated in the reference (its batched 3-D call fails, see gpsig_b200/kernels.py)
    for fam in ('rbf', 'exp'):
        np.random.seed(7)
        k = getattr(ref, 'SignatureSpectral')(L * d, d, m, family=fam, Q=3, variances=var, normalization=False)
        k.sigma = k.sigma * 0. + 1.5
        for nm in ('alpha', 'omega', 'gamma'):
            out['spec_%s_%s' % (fam, nm)] = _val(getattr(k, nm))
        out['Kspec_' + fam] = _val(k.K(Xf, Yf, return_levels=True))
    # the SVGP glue around the path (gpsig/inducing_variables.py:51-137): Kuu_Kuf_Kff / Kuu / Kuf for inducing tensors
    # (both forms, with and without the per-level mixing weights) and inducing sequences
    iv = refrun.load_reference_inducing()
    import tensorflow as tf_stub
    rng = np.random.default_rng(106)
    Zs = brownian(4, 9, d, 105)
    out['Zseq'] = Zs
    out['W_tens'] = np.eye(Mz)[None] + 0.3 * rng.standard_normal((m, Mz, Mz))
    out['W_seq'] = np.eye(4)[None] + 0.3 * rng.standard_normal((m, 4, 4))
    k = kern('SignatureRBF', {}, order=1, normalization=True)
    for lw in (0, 1):
        feats = [('tens_i%d' % inc, iv.InducingTensors(out['Z_inc%d' % inc], m, increments=bool(inc), learn_weights=bool(lw)), 'W_tens')
                 for inc in (0, 1)] + [('seq', iv.InducingSequences(Zs, m, learn_weights=bool(lw)), 'W_seq')]
        for name, feat, wkey in feats:
            if lw:
                feat.W = tf_stub.Tensor(out[wkey])
            Kzz, Kzx, Kxx = iv.Kuu_Kuf_Kff(feat, k, Xf, jitter=1e-3)
            tag = 'glue_%s_w%d' % (name, lw)
            out[tag + '_Kzz'], out[tag + '_Kzx'], out[tag + '_Kxx'] = _val(Kzz), _val(Kzx), _val(Kxx)
            out[tag + '_Kuu'] = _val(iv.Kuu(feat, k, jitter=1e-3))
            out[tag + '_Kuf'] = _val(iv.Kuf(feat, k, Xf))
    np.savez_compressed(os.path.join(HERE, 'trunc_ref.npz'), **out)


def make_pde_ref_py():
    """The reference's PDE PYTHON layer executed on the stand-ins (oracle/refrun.load_reference_pde; its Cython solver is the
    compiled oracle/_ref module): ``UntruncSignatureKernel.Kdiag`` with lengthscales / sigma / lags (kernels_pde.py:136-185)
    and its hand-written backward ``_KdiagGrad`` (kernels_pde.py:465-509), called with the forward op's inputs and outputs
    as TensorFlow would.  Same X as pde_ref.npz."""
    from oracle import refrun
    kp = refrun.load_reference_pde()
    import tensorflow as tf_stub
    X = np.load(os.path.join(HERE, 'pde_ref.npz'))['X']
    N, L, d = X.shape
    ls = np.array([0.9, 1.2, 0.8])
    gout = np.random.default_rng(107).standard_normal(N)
    out = {'X': X, 'lengthscales': ls, 'sigma': np.float64(1.7), 'gout': gout}

    class Op:
        pass

    for order in (0, 1, 2):
        k = kp.UntruncSignatureKernel(L * d, d, lengthscales=ls, order=order, implementation='cython')
        k.sigma = k.sigma * 0. + 1.7
        out['Kdiag_class_order%d' % order] = _val(k.Kdiag(X.reshape(N, -1)))
        K, K_rev = (np.array(a) for a in kp.sig_kern_diag(X, order))
        op = Op()
        op.inputs = [tf_stub.Tensor(X), tf_stub.Tensor(np.int64(order))]
        op.outputs = [tf_stub.Tensor(K.copy()), tf_stub.Tensor(K_rev.copy())]
        G = np.zeros_like(K)
        G[:, -1, -1] = gout
        grad, _ = kp._KdiagGrad(op, tf_stub.Tensor(G), None)
        out['KdiagGrad_fn_order%d' % order] = _val(grad)
    k = kp.UntruncSignatureKernel(L * d, d, lengthscales=ls, order=1, num_lags=2, implementation='cython')
    out['Kdiag_class_lags2_order1'] = _val(k.Kdiag(X.reshape(N, -1)))
    np.savez_compressed(os.path.join(HERE, 'pde_ref_py.npz'), **out)


if __name__ == '__main__':
    if '--trunc-ref' in sys.argv:       # in a process of its own: the stand-in `tensorflow` must be the first one imported
        make_trunc_ref()
        make_pde_ref_py()
    else:
        make_pde()
        make_trunc()
    print('wrote', sorted(f for f in os.listdir(HERE) if f.endswith('.npz')))
