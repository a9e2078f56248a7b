/* Plain-C restatement of the reference's signature-kernel algorithms (CPU).
 *
 * TEST INFRASTRUCTURE ONLY: the parity checker and the timed CPU baseline
 * ("port") of bench.py.  Never linked into, imported by, or called from the
 * product library (gpsig_b200/csrc).
 *
 * Each routine does, per pair of sequences, exactly what the reference's
 * TensorFlow graph does on the whole lifted tensor:
 *   - base kernel on points            gpsig/kernels.py:946-957, 979-986, 1042-1044 ...
 *   - 2-D differencing                 gpsig/signature_algs.py:26
 *   - K_1 = sum(M)                     gpsig/signature_algs.py:28
 *   - R <- M * cumsum_excl(cumsum_excl(R, rows), cols); K_i = sum(R)   :31-33
 *   - tensors-vs-sequence recursion    gpsig/signature_algs.py:101-127
 *   - Goursat PDE stencil              gpsig/sigKer_fast.pyx:41-48
 * Pairs are independent, so the outer loops are OpenMP-parallel (the
 * reference's own parallelism is TF intra-op threads / prange over paths).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

enum { BASE_LINEAR = 0, BASE_RBF = 1, BASE_COSINE = 2, BASE_POLY = 3, BASE_MIX = 4,
       BASE_MATERN12 = 5, BASE_MATERN32 = 6, BASE_MATERN52 = 7 };

/* kappa(x, y) from (x.y, |x|^2, |y|^2) -- gpsig/kernels.py:946-1173 */
static double base_eval(int base, const double* prm, double dot, double xs, double ys) {
  switch (base) {
    case BASE_LINEAR: return dot;
    case BASE_RBF: { double dist = -2.0 * dot; dist += xs + ys; return exp(-dist / 2); }
    case BASE_COSINE: return dot / (sqrt(xs) * sqrt(ys));
    case BASE_POLY: return pow(dot + prm[0], prm[1]);
    case BASE_MIX: { double ds = xs + ys - 2 * dot; return prm[0] * exp(-ds / 2) + (1. - prm[0]) * dot; }
    default: {
      double dist = -2.0 * dot; dist += xs + ys;
      double r = sqrt(fmax(dist, 1e-40));
      if (base == BASE_MATERN12) return exp(-r);
      if (base == BASE_MATERN32) return (1. + sqrt(3.) * r) * exp(-sqrt(3.) * r);
      return (1.0 + sqrt(5.) * r + 5. / 3. * (r * r)) * exp(-sqrt(5.) * r);
    }
  }
}

static double dotp(const double* a, const double* b, int d) {
  double s = 0; for (int k = 0; k < d; ++k) s += a[k] * b[k]; return s;
}

/* first-order recursion on one lifted tile M[L x L2] -> K[0..m] (signature_algs.py:8-35).
 * work: 3*L*L2 doubles of scratch */
static void first_order_tile(const double* M, int L, int L2, int m, int difference,
                             double* work, double* K) {
  int R_ = difference ? L - 1 : L, C_ = difference ? L2 - 1 : L2;
  double* D = work; double* R = work + (size_t)L * L2; double* S = R + (size_t)L * L2;
  for (int p = 0; p < R_; ++p)
    for (int q = 0; q < C_; ++q)
      D[p * C_ + q] = difference
        ? M[(p + 1) * L2 + q + 1] + M[p * L2 + q] - M[p * L2 + q + 1] - M[(p + 1) * L2 + q]
        : M[p * L2 + q];
  K[0] = 1.0;
  double s = 0; for (int e = 0; e < R_ * C_; ++e) s += D[e];
  K[1] = s;
  memcpy(R, D, sizeof(double) * R_ * C_);
  for (int i = 2; i <= m; ++i) {
    /* S = cumsum_excl over rows (axis=1) */
    for (int q = 0; q < C_; ++q) { double a = 0; for (int p = 0; p < R_; ++p) { S[p * C_ + q] = a; a += R[p * C_ + q]; } }
    /* R = M * cumsum_excl over cols (axis=-1) of S */
    s = 0;
    for (int p = 0; p < R_; ++p) { double a = 0; for (int q = 0; q < C_; ++q) { double v = S[p * C_ + q]; R[p * C_ + q] = D[p * C_ + q] * a; a += v; s += R[p * C_ + q]; } }
    K[i] = s;
  }
}

/* out[(m+1), N, N2] unnormalised level matrices (kernels.py:209-238).  X2==NULL -> X. */
int oracle_trunc_gram_f64(const double* X, const double* X2, int N, int N2, int L, int L2, int d,
                          int m, int base, const double* prm, int difference, double* out) {
  if (!X2) { X2 = X; N2 = N; L2 = L; }
  #pragma omp parallel
  {
    double* M = (double*)malloc(sizeof(double) * (size_t)L * L2 * 4);
    double* xs = (double*)malloc(sizeof(double) * (L + L2));
    double* ys = xs + L;
    double K[64];
    #pragma omp for schedule(dynamic, 4) collapse(2)
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N2; ++j) {
        const double* x = X + (size_t)i * L * d; const double* y = X2 + (size_t)j * L2 * d;
        for (int p = 0; p < L; ++p) xs[p] = dotp(x + p * d, x + p * d, d);
        for (int q = 0; q < L2; ++q) ys[q] = dotp(y + q * d, y + q * d, d);
        for (int p = 0; p < L; ++p)
          for (int q = 0; q < L2; ++q)
            M[p * L2 + q] = base_eval(base, prm, dotp(x + p * d, y + q * d, d), xs[p], ys[q]);
        first_order_tile(M, L, L2, m, difference, M + (size_t)L * L2, K);
        for (int l = 0; l <= m; ++l) out[((size_t)l * N + i) * N2 + j] = K[l];
      }
    free(M); free(xs);
  }
  return 0;
}

/* out[(m+1), N] diagonal levels k_l(x_i, x_i) (kernels.py:190-207) */
int oracle_trunc_diag_f64(const double* X, int N, int L, int d, int m, int base, const double* prm,
                          int difference, double* out) {
  #pragma omp parallel
  {
    double* M = (double*)malloc(sizeof(double) * (size_t)L * L * 4);
    double* xs = (double*)malloc(sizeof(double) * L);
    double K[64];
    #pragma omp for schedule(dynamic, 4)
    for (int i = 0; i < N; ++i) {
      const double* x = X + (size_t)i * L * d;
      for (int p = 0; p < L; ++p) xs[p] = dotp(x + p * d, x + p * d, d);
      for (int p = 0; p < L; ++p)
        for (int q = 0; q < L; ++q)
          M[p * L + q] = base_eval(base, prm, dotp(x + p * d, x + q * d, d), xs[p], xs[q]);
      first_order_tile(M, L, L, m, difference, M + (size_t)L * L, K);
      for (int l = 0; l <= m; ++l) out[(size_t)l * N + i] = K[l];
    }
    free(M); free(xs);
  }
  return 0;
}

/* out[(m+1), Mz, N] (kernels.py:314-341 + signature_algs.py:101-127).
 * Z: [T, Mz, (2,) d] with T = m(m+1)/2; increments -> kappa(z1,.) - kappa(z0,.) */
int oracle_tens_vs_seq_f64(const double* Z, int Mz, int increments, const double* X, int N, int L, int d,
                           int m, int base, const double* prm, int difference, double* out) {
  int T = m * (m + 1) / 2, zs = increments ? 2 : 1;
  int Ld = difference ? L - 1 : L;
  #pragma omp parallel
  {
    double* M = (double*)malloc(sizeof(double) * (size_t)T * L * 2);
    double* R = M + (size_t)T * L;
    double* xs = (double*)malloc(sizeof(double) * L);
    #pragma omp for schedule(dynamic, 8) collapse(2)
    for (int n = 0; n < N; ++n)
      for (int z = 0; z < Mz; ++z) {
        const double* x = X + (size_t)n * L * d;
        for (int p = 0; p < L; ++p) xs[p] = dotp(x + p * d, x + p * d, d);
        for (int t = 0; t < T; ++t) {
          const double* z0 = Z + (((size_t)t * Mz + z) * zs) * d;
          for (int p = 0; p < L; ++p) {
            double v = base_eval(base, prm, dotp(z0, x + p * d, d), dotp(z0, z0, d), xs[p]);
            if (increments) {
              const double* z1 = z0 + d;
              v = base_eval(base, prm, dotp(z1, x + p * d, d), dotp(z1, z1, d), xs[p]) - v;
            }
            M[t * L + p] = v;
          }
          if (difference) for (int p = 0; p < Ld; ++p) M[t * L + p] = M[t * L + p + 1] - M[t * L + p];
        }
        out[((size_t)0 * Mz + z) * N + n] = 1.0;
        int k = 0;
        for (int i = 1; i <= m; ++i) {
          for (int p = 0; p < Ld; ++p) R[p] = M[k * L + p];
          ++k;
          for (int j = 1; j < i; ++j) {
            double a = 0;
            for (int p = 0; p < Ld; ++p) { double v = R[p]; R[p] = M[k * L + p] * a; a += v; }
            ++k;
          }
          double s = 0; for (int p = 0; p < Ld; ++p) s += R[p];
          out[((size_t)i * Mz + z) * N + n] = s;
        }
      }
    free(M); free(xs);
  }
  return 0;
}

/* Goursat PDE, explicit 2nd-order stencil (sigKer_fast.pyx:41-48) for `npairs`
 * pairs (x[ix[a]], y[iy[a]]); out[a] = K[P, Q].  Two rolling rows of state. */
int oracle_pde_pairs_f64(const double* X, const double* Y, const int* ix, const int* iy, int npairs,
                         int L, int L2, int d, int n, double* out) {
  int P = (1 << n) * (L - 1), Q = (1 << n) * (L2 - 1);
  double factor = (double)(1 << (2 * n));
  #pragma omp parallel
  {
    double* row0 = (double*)malloc(sizeof(double) * (Q + 1) * 2);
    double* row1 = row0 + Q + 1;
    double* incr = (double*)malloc(sizeof(double) * (L2 - 1));
    #pragma omp for schedule(dynamic, 1)
    for (int a = 0; a < npairs; ++a) {
      const double* x = X + (size_t)ix[a] * L * d; const double* y = Y + (size_t)iy[a] * L2 * d;
      for (int q = 0; q <= Q; ++q) row0[q] = 1.0;
      double* prev = row0; double* cur = row1;
      for (int i = 0; i < P; ++i) {
        int ii = i >> n;
        if ((i & ((1 << n) - 1)) == 0)
          for (int jj = 0; jj < L2 - 1; ++jj) {
            double e = 0;
            for (int k = 0; k < d; ++k)
              e = e + (x[(ii + 1) * d + k] - x[ii * d + k]) * (y[(jj + 1) * d + k] - y[jj * d + k]) / factor;
            incr[jj] = e;
          }
        cur[0] = 1.0;
        for (int j = 0; j < Q; ++j) {
          double e = incr[j >> n];
          cur[j + 1] = (prev[j + 1] + cur[j]) * (1. + 0.5 * e + (1. / 12) * (e * e)) - prev[j] * (1. - (1. / 12) * (e * e));
        }
        double* t = prev; prev = cur; cur = t;
      }
      out[a] = prev[Q];
    }
    free(row0); free(incr);
  }
  return 0;
}
