"""CPU tests of the oracle itself (no GPU): known-answer properties of the
reference's notebooks, the committed golden vectors, and numpy <-> C port agreement."""
import os

import numpy as np
import pytest

from oracle import c_oracle, pde_np, sigs_np, trunc_np
from tests.util import brownian, inducing_tensors, rel_err

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def test_higher_order_linear_equals_signature_inner_products():
    """notebooks/signature_kernel.ipynb cells 11-13: SignatureLinear(order=num_levels,
    normalization=False) == <S_k(x), S_k(y)> level by level."""
    N, L, d, m = 6, 12, 3, 4
    X, Y = brownian(N, L, d, 0), brownian(5, 9, d, 1)
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base='linear', order=m, normalization=False, lengthscales=None)
    assert rel_err(o._K_seq(X), sigs_np.signature_gram_levels(X, None, m)) < 1e-13
    assert rel_err(o._K_seq(X, Y), sigs_np.signature_gram_levels(X, Y, m)) < 1e-13


def test_tens_vs_seq_equals_rank1_tensor_inner_products():
    """notebooks/signature_kernel.ipynb cells 15-23 (higher order, linear, no increments)."""
    N, L, d, m, Mz = 4, 10, 3, 3, 5
    X = brownian(N, L, d, 2)
    Z = inducing_tensors(X, m, Mz, False, 3)
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base='linear', order=m, normalization=False, lengthscales=None)
    K = o._K_tens_vs_seq(Z, X)
    S = [sigs_np.signature(x, m) for x in X]
    k = 0
    for lvl in range(1, m + 1):
        for z in range(Mz):
            t = sigs_np.rank1_tensor_levels(Z[k:k + lvl, z])
            for n in range(N):
                assert abs(K[lvl, z, n] - np.sum(t * S[n][lvl])) < 1e-12 * max(1, abs(K[lvl, z, n]))
        k += lvl


def test_tens_tens_equals_rank1_tensor_inner_products():
    """notebooks/signature_kernel.ipynb cells 25-29."""
    d, m, Mz = 3, 4, 6
    Z = np.random.default_rng(4).standard_normal((m * (m + 1) // 2, Mz, d))
    o = trunc_np.SignatureKernelOracle(30, d, m, base='linear', normalization=False, lengthscales=None)
    K = o._K_tens(Z)
    k = 0
    for lvl in range(1, m + 1):
        T = [sigs_np.rank1_tensor_levels(Z[k:k + lvl, z]) for z in range(Mz)]
        ref = np.array([[np.sum(a * b) for b in T] for a in T])
        assert rel_err(K[lvl], ref) < 1e-13
        k += lvl


def test_first_order_equals_higher_order_for_two_levels_only():
    """order=1 is a different (cheaper) kernel: it agrees with the signature only up to level 1."""
    X = brownian(5, 15, 2, 5)
    o1 = trunc_np.SignatureKernelOracle(30, 2, 3, base='linear', order=1, normalization=False, lengthscales=None)
    o3 = trunc_np.SignatureKernelOracle(30, 2, 3, base='linear', order=3, normalization=False, lengthscales=None)
    A, B = o1._K_seq(X), o3._K_seq(X)
    assert rel_err(A[:2], B[:2]) < 1e-13 and rel_err(A[2:], B[2:]) > 1e-3


@pytest.mark.parametrize("base", sorted(c_oracle.BASE_IDS))
def test_c_port_matches_numpy_restatement(base):
    N, N2, L, L2, d, m = 7, 6, 13, 10, 3, 4
    X, Y = brownian(N, L, d, 6), brownian(N2, L2, d, 7)
    prm = {'poly': dict(gamma=1.0, degree=3.0), 'mix': dict(mixing=0.5)}.get(base, {})
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base=base, normalization=False, lengthscales=None, **prm)
    tol = 1e-12
    assert rel_err(c_oracle.trunc_gram(X, Y, m, base, list(prm.values())), o._K_seq(X, Y)) < tol
    # Matern kernels take sqrt(max(r2, 1e-40)) of the CANCELLING expansion |x|^2+|y|^2-2x.y
    # (kernels.py:959-961): at coincident points r = sqrt(rounding noise) ~ 1e-8, so the
    # reference's own diagonal is only defined to ~1e-7 (it depends on the BLAS summation order).
    dtol = 1e-6 if base.startswith('matern') else tol
    assert rel_err(c_oracle.trunc_diag(X, m, base, list(prm.values())), o._K_seq_diag(X)) < dtol
    for inc in (False, True):
        Z = inducing_tensors(X, m, 5, inc, 8)
        assert rel_err(c_oracle.tens_vs_seq(Z, X, m, inc, base, list(prm.values())), o._K_tens_vs_seq(Z, X, inc)) < tol


def test_normalised_symmetric_gram_has_unit_diagonal():
    X = brownian(8, 12, 3, 9).reshape(8, -1)
    o = trunc_np.SignatureKernelOracle(36, 3, 4, base='rbf')
    K = o.K(X, return_levels=True)
    assert np.allclose(np.diagonal(K, axis1=1, axis2=2), 1.0, atol=1e-12)
    assert np.allclose(o.Kdiag(X), 5.0)


def test_trunc_golden_vectors():
    g = np.load(os.path.join(GOLD, 'trunc_oracle.npz'))
    X, Y, ls = g['X'], g['Y'], g['lengthscales']
    N, L, d = X.shape
    m = 4
    for base in ('linear', 'rbf'):
        for order in (1, 2, m):
            for norm in (False, True):
                o = trunc_np.SignatureKernelOracle(L * d, d, m, base=base, order=order, normalization=norm,
                                                   lengthscales=ls, sigma=1.5, variances=np.linspace(0.5, 1.5, m + 1))
                tag = '%s_o%d_n%d' % (base, order, norm)
                assert rel_err(o.K(X.reshape(N, -1), Y.reshape(Y.shape[0], -1), return_levels=True), g['K_' + tag]) < 1e-13
                assert rel_err(o.K(X.reshape(N, -1), return_levels=True), g['Ksym_' + tag]) < 1e-13


def test_pde_numpy_and_c_match_reference_cython_golden():
    """tests/golden/pde_ref.npz was produced by the reference's compiled sigKer_fast."""
    g = np.load(os.path.join(GOLD, 'pde_ref.npz'))
    X = g['X']
    ix = np.arange(X.shape[0], dtype=np.int32)
    for order in (0, 1, 2):
        ref = g['kdiag_order%d' % order]
        assert np.array_equal(pde_np.pde_kernel(X, None, order), ref)          # bit for bit
        assert np.max(np.abs(c_oracle.pde_pairs(X, None, ix, ix, order) - ref) / np.abs(ref)) < 1e-14
        assert np.array_equal(pde_np.pde_grid_rev(X, order)[:, -1, -1], g['krev_corner_order%d' % order])
    for order in (0, 1):
        G = pde_np.pde_grid(X[:2], None, order)
        tril = np.tril(np.ones(G.shape[1:], bool))
        assert np.array_equal(G[:, tril], g['grid_order%d' % order][:, tril])   # the reference fills the lower triangle only


def test_pde_live_reference_module_if_built():
    from oracle import build_ref
    ref = build_ref.load()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    X = brownian(5, 17, 4, 10)
    for order in (0, 2):
        K, _ = ref.sig_kern_diag(X, order, 1)
        assert np.array_equal(np.array(K)[:, -1, -1], pde_np.pde_kernel(X, None, order))


def test_pde_rectangular_reduces_to_symmetric():
    X = brownian(3, 9, 2, 11)
    assert np.array_equal(pde_np.pde_kernel(X, X.copy(), 1), pde_np.pde_kernel(X, None, 1))
    G = pde_np.pde_gram(X, None, 0)
    assert np.allclose(G, G.T, rtol=1e-13)


def test_pde_gradient_restatement_matches_reference_grids():
    """The restated backward (oracle/pde_np.pde_kdiag_grad: grids from the numpy solver) against the same
    formula evaluated on the grids of the reference's compiled Cython solver (golden)."""
    g = np.load(os.path.join(GOLD, 'pde_ref.npz'))
    for order in (0, 1, 2):
        got = pde_np.pde_kdiag_grad(g['X'], order, rev_solver=0)
        assert np.array_equal(got, g['kdiag_grad_order%d' % order])


def test_base_conditional_restatement_is_the_gaussian_conditional():
    """oracle/svgp_np.base_conditional (GPflow 1.5.1 restated) against the closed form: for u ~ N(m, S),
    f | u Gaussian => mean Knm Kmm^-1 m, cov Knn - Knm Kmm^-1 (Kmm - S) Kmm^-1 Kmn."""
    from oracle import svgp_np
    rng = np.random.default_rng(40)
    M, N, R = 6, 9, 2
    B = rng.standard_normal((M + N, M + N + 3))
    Kall = B @ B.T
    Kmm, Kmn, Knn = Kall[:M, :M], Kall[:M, M:], Kall[M:, M:]
    f = rng.standard_normal((M, R))
    q = np.tril(rng.standard_normal((R, M, M)))
    mean, var = svgp_np.base_conditional(Kmn, Kmm, Knn, f, full_cov=True, q_sqrt=q, white=False)
    Ki = np.linalg.inv(Kmm)
    assert np.allclose(mean, Kmn.T @ Ki @ f, rtol=1e-9, atol=1e-9)
    for r in range(R):
        S = q[r] @ q[r].T
        assert np.allclose(var[r], Knn - Kmn.T @ Ki @ (Kmm - S) @ Ki @ Kmn, rtol=1e-8, atol=1e-8)
    md, vd = svgp_np.base_conditional(Kmn, Kmm, np.diag(Knn), f, full_cov=False, q_sqrt=q, white=False)
    assert np.allclose(md, mean) and np.allclose(vd, np.stack([np.diag(v) for v in var], axis=1))
    # whitened: u = Lm v, v ~ N(f, q q^T)
    Lm = np.linalg.cholesky(Kmm)
    mw, vw = svgp_np.base_conditional(Kmn, Kmm, np.diag(Knn), f, full_cov=False, q_sqrt=q, white=True)
    mu, vu = svgp_np.base_conditional(Kmn, Kmm, np.diag(Knn), Lm @ f, full_cov=False, q_sqrt=np.matmul(Lm[None], q), white=False)
    assert np.allclose(mw, mu) and np.allclose(vw, vu)


# ---------------------------------------------------------------------------------------------------
# Independent definitional pin of the FIRST-ORDER recursions (the reference's TF graph cannot run here).
# Unrolling R_i = D * cumsum_excl_p cumsum_excl_q R_{i-1}, K_i = sum R_i  (signature_algs.py:26-33) gives
#     K_i(x, y) = sum over p_1 < ... < p_i, q_1 < ... < q_i of prod_k D[p_k, q_k],
# and R = M[k'] * cumsum_excl_l R, K_i = sum_l R  (signature_algs.py:114-125) gives
#     K_i(z, x) = sum over l_1 < ... < l_i of prod_j Md[k_i + j - 1][l_j].
# Both are enumerated with itertools.combinations from base-kernel values computed point by point
# (no code shared with oracle/trunc_np.py or the C port).
# ---------------------------------------------------------------------------------------------------
import itertools


def _kappa(base, x, y):
    if base == 'linear':
        return float(np.dot(x, y))
    return float(np.exp(-0.5 * np.sum((x - y) ** 2)))


def _brute_first_order(x, y, m, base, difference=True):
    L, L2 = len(x), len(y)
    M = np.array([[_kappa(base, x[p], y[q]) for q in range(L2)] for p in range(L)])
    D = M[1:, 1:] - M[:-1, 1:] - M[1:, :-1] + M[:-1, :-1] if difference else M
    P, Q = D.shape
    out = [1.0]
    for i in range(1, m + 1):
        tot = 0.0
        for ps in itertools.combinations(range(P), i):
            for qs in itertools.combinations(range(Q), i):
                term = 1.0
                for p, q in zip(ps, qs):
                    term *= D[p, q]
                tot += term
        out.append(tot)
    return np.array(out)


def _brute_tens_vs_seq(zrows, x, m, base, increments, difference=True):
    """zrows: [T, d] (or [T, 2, d] with increments) components of ONE inducing tensor, x: [L, d]."""
    L = len(x)
    out = [1.0]
    k = 0
    for i in range(1, m + 1):
        comps = []
        for j in range(i):
            if increments:
                v = np.array([_kappa(base, zrows[k + j, 1], x[l]) - _kappa(base, zrows[k + j, 0], x[l]) for l in range(L)])
            else:
                v = np.array([_kappa(base, zrows[k + j], x[l]) for l in range(L)])
            comps.append(v[1:] - v[:-1] if difference else v)
        k += i
        n = len(comps[0])
        tot = 0.0
        for ls in itertools.combinations(range(n), i):
            term = 1.0
            for j, l in enumerate(ls):
                term *= comps[j][l]
            tot += term
        out.append(tot)
    return np.array(out)


@pytest.mark.parametrize("base", ["linear", "rbf"])
@pytest.mark.parametrize("difference", [True, False])
def test_first_order_gram_equals_sum_over_increasing_index_tuples(base, difference):
    """Pins signature_kern_first_order (recursion, exclusive-cumsum convention, 2-D differencing) of the numpy
    restatement AND of the C port against brute-force enumeration, L <= 7, m <= 4."""
    N, N2, L, L2, d, m = 3, 2, 7, 6, 2, 4
    X, Y = brownian(N, L, d, 21) * 1.5, brownian(N2, L2, d, 22) * 1.5
    ref = np.array([[_brute_first_order(X[i], Y[j], m, base, difference) for j in range(N2)] for i in range(N)]).transpose(2, 0, 1)
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base=base, normalization=False, lengthscales=None, difference=difference)
    assert rel_err(o._K_seq(X, Y), ref) < 1e-12
    assert rel_err(c_oracle.trunc_gram(X, Y, m, base, difference=difference), ref) < 1e-12
    # the diagonal form M[N, L, L] (signature_algs.py:13-18) and the symmetric call
    refd = np.array([_brute_first_order(X[i], X[i], m, base, difference) for i in range(N)]).T
    assert rel_err(o._K_seq_diag(X), refd) < 1e-12
    assert rel_err(c_oracle.trunc_diag(X, m, base, difference=difference), refd) < 1e-12
    assert rel_err(np.diagonal(o._K_seq(X), axis1=1, axis2=2), refd) < 1e-12


@pytest.mark.parametrize("base", ["linear", "rbf"])
@pytest.mark.parametrize("increments", [False, True])
def test_first_order_tens_vs_seq_equals_sum_over_increasing_times(base, increments):
    N, L, d, m, Mz = 3, 7, 2, 4, 3
    X = brownian(N, L, d, 23) * 1.5
    Z = inducing_tensors(X, m, Mz, increments, 24)
    ref = np.array([[_brute_tens_vs_seq(Z[:, z], X[n], m, base, increments) for n in range(N)] for z in range(Mz)]).transpose(2, 0, 1)
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base=base, normalization=False, lengthscales=None)
    assert rel_err(o._K_tens_vs_seq(Z, X, increments), ref) < 1e-12
    assert rel_err(c_oracle.tens_vs_seq(Z, X, m, increments, base), ref) < 1e-12


def test_normalised_wrapper_equals_definition_on_brute_force_levels():
    """kernels.py:456-477, the epilogue, from brute-force levels: non-symmetric call = no jitter on the Gram,
    jitter (gpflow default 1e-6) added to both diagonals before the square root (kernels.py:466-470)."""
    N, N2, L, d, m = 3, 2, 6, 2, 3
    X, Y = brownian(N, L, d, 25) * 1.5, brownian(N2, L, d, 26) * 1.5
    var, sigma = np.linspace(0.5, 1.5, m + 1), 1.3
    raw = np.array([[_brute_first_order(X[i], Y[j], m, 'rbf') for j in range(N2)] for i in range(N)]).transpose(2, 0, 1)
    dx = np.array([_brute_first_order(X[i], X[i], m, 'rbf') for i in range(N)]).T
    dy = np.array([_brute_first_order(Y[j], Y[j], m, 'rbf') for j in range(N2)]).T
    ref = sigma * var[:, None, None] * raw / (np.sqrt(dx + 1e-6)[:, :, None] * np.sqrt(dy + 1e-6)[:, None, :])
    o = trunc_np.SignatureKernelOracle(L * d, d, m, base='rbf', lengthscales=None, variances=var, sigma=sigma)
    assert rel_err(o.K(X.reshape(N, -1), Y.reshape(N2, -1), return_levels=True), ref) < 1e-12
    assert rel_err(o.K(X.reshape(N, -1), Y.reshape(N2, -1)), ref.sum(0)) < 1e-12
    # symmetric call (kernels.py:431-434): jitter on the Gram's diagonal, normalised by its own diagonal
    raws = np.array([[_brute_first_order(X[i], X[j], m, 'rbf') for j in range(N)] for i in range(N)]).transpose(2, 0, 1)
    raws = raws + 1e-6 * np.eye(N)[None]
    ds = np.sqrt(np.diagonal(raws, axis1=1, axis2=2))
    refs = sigma * var[:, None, None] * raws / (ds[:, :, None] * ds[:, None, :])
    assert rel_err(o.K(X.reshape(N, -1), return_levels=True), refs) < 1e-12
