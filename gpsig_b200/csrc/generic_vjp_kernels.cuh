// Reverse mode of the HIGHER-ORDER recursion (order > 1) for sm_100a -- the general, unspecialised path.
//
// The reference differentiates gpsig/signature_algs.py:37-74 (signature_kern_higher_order) and :129-160
// (signature_kern_tens_vs_seq_higher_order) with TensorFlow autodiff; train_gpsig_.py:62 and
// train_gpsig_vosf.py:102 train with order = num_levels.  Here, as for order 1 (vjp_kernels.cuh), the adjoint of the
// RAW levels is hand written and ends in W = dL/dD (sequences) resp. dL/dM (tensors vs sequences); the chain through
// the base kernel into the coordinates is shared with the first-order path (trunc_vjp_grad_* kernels).
//
// Level l+1 of the forward is, per tile element, R'[a][b] = D * F[a][b] with the FACTORS (signature_algs.py:61-72)
//   F[0][0] = EE(sum R)                 exclusive 2-D prefix of the previous level's total      (:64)
//   F[0][j] = EP(sum_a R[a][j-1])/(j+1) exclusive prefix over rows, same column                 (:66)
//   F[j][0] = EQ(sum_b R[j-1][b])/(j+1) exclusive prefix over columns, same row                 (:67)
//   F[j][k] = R[j-1][k-1]/((j+1)(k+1))                                                           (:69)
// A forward sweep (same schedule as trunc_generic_kernel: one CTA per pair, one thread per column, rows in order)
// stores D and every F in a scratch; a backward sweep over the rows in REVERSE order runs the levels top down: the
// adjoints of EE / EP are running sums over the rows already visited (+ a block-wide reverse scan over the columns
// for EE), the adjoint of EQ is a reverse scan within the row, the adjoint of F[j][k] is local.
#pragma once
#include "generic_kernels.cuh"
#include "vjp_kernels.cuh"

namespace gpsig {

// exclusive SUFFIX sum over the threads of the block: sum of v over threads with a larger index
template <typename T>
__device__ __forceinline__ T block_excl_scan_rev(T v, T* warp_tot /* [32] shared */) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  T incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const T t = __shfl_down_sync(0xffffffffu, incl, o);
    if (lane + o < 32) incl += t;
  }
  __syncthreads();
  if (lane == 0) warp_tot[w] = incl;
  __syncthreads();
  T base = T(0);
  for (int k = w + 1; k < nw; ++k) base += warp_tot[k];
  return base + incl - v;
}

// tiles of factors stored per pair: sum over levels 2..m of min(level, order)^2
__host__ __device__ inline int generic_vjp_tiles(int m, int order) {
  int n = 0;
  for (int l = 2; l <= m; ++l) { const int dn = l < order ? l : order; n += dn * dn; }
  return n;
}

struct GenericVjpParams {
  TruncVjpParams V;        // X, Y, sizes, g, W (cell-major [npairs][P*Q]), pair mapping
  int order;
  void* Dt;                // [npairs][P*Q]
  void* Ft;                // [npairs][tiles][P*Q]
};

template <typename T>
__global__ void __launch_bounds__(1024) trunc_generic_vjp_kernel(const GenericVjpParams G) {
  __shared__ T warp_tot[32];
  const TruncVjpParams& P = G.V;
  const int diff = P.difference ? 1 : 0;
  const int PR = P.L - diff, QC = P.L2 - diff;
  const int d = P.d, m = P.m, order = G.order;
  const size_t pair = blockIdx.x;
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * P.L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * P.L2 * d;
  const int q = threadIdx.x;
  const bool colok = q < QC;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const size_t cells = (size_t)PR * QC;
  const int NTL = generic_vjp_tiles(m, order);
  T* Dt = static_cast<T*>(G.Dt) + pair * cells;
  T* Ft = static_cast<T*>(G.Ft) + pair * (size_t)NTL * cells;
  T* W = static_cast<T*>(P.W) + pair * cells;
  auto kap = [&](int p, int qq) { return kappa_raw_call<T>(x + (size_t)p * d, y + (size_t)qq * d, d, P.base, p0, p1); };

  // ---------------- forward: store D and the factors ----------------
  {
    T Ctot[kGenMax], Ccol[kGenMax][kGenMax];
    for (int l = 0; l < kGenMax; ++l) { Ctot[l] = T(0); for (int b = 0; b < kGenMax; ++b) Ccol[l][b] = T(0); }
    T Mlo = T(0), Mhi = T(0);
    if (diff && colok) { Mlo = kap(0, q); Mhi = kap(0, q + 1); }
    for (int p = 0; p < PR; ++p) {
      T Dv = T(0);
      if (colok) {
        if (diff) { const T a = kap(p + 1, q), b = kap(p + 1, q + 1); Dv = b + Mlo - Mhi - a; Mlo = a; Mhi = b; }
        else Dv = kap(p, q);
        Dt[(size_t)p * QC + q] = Dv;
      }
      T R[kGenMax][kGenMax];
      R[0][0] = Dv;
      int dd = 1, toff = 0;
      for (int lvl = 1; lvl < m; ++lvl) {
        const int dn = min(lvl + 1, order);
        T tot = T(0), colsum[kGenMax], rowsum[kGenMax];
        for (int a = 0; a < dd; ++a) { rowsum[a] = T(0); colsum[a] = T(0); }
        for (int a = 0; a < dd; ++a)
          for (int b = 0; b < dd; ++b) { const T v = R[a][b]; tot += v; rowsum[a] += v; colsum[b] += v; }
        const T T2 = block_excl_scan<T>(Ctot[lvl - 1], warp_tot);
        T Pr[kGenMax];
        for (int a = 0; a + 1 < dn; ++a) Pr[a] = block_excl_scan<T>(a < dd ? rowsum[a] : T(0), warp_tot);
        T F[kGenMax][kGenMax];
        F[0][0] = T2;
        for (int jj = 2; jj <= dn; ++jj) {
          F[0][jj - 1] = Ccol[lvl - 1][jj - 2] / T(jj);
          F[jj - 1][0] = Pr[jj - 2] / T(jj);
          for (int kk = 2; kk <= dn; ++kk) F[jj - 1][kk - 1] = R[jj - 2][kk - 2] / T(jj * kk);
        }
        Ctot[lvl - 1] += tot;
        for (int b = 0; b < dd; ++b) Ccol[lvl - 1][b] += colsum[b];
        for (int a = 0; a < dn; ++a)
          for (int b = 0; b < dn; ++b) {
            if (colok) Ft[(size_t)(toff + a * dn + b) * cells + (size_t)p * QC + q] = F[a][b];
            R[a][b] = Dv * F[a][b];
          }
        toff += dn * dn;
        dd = dn;
      }
    }
  }
  __syncthreads();

  // ---------------- backward: rows in reverse, levels top down ----------------
  const T* g = static_cast<const T*>(P.g);
  T gl[kGenMax + 1];
  for (int l = 1; l <= m; ++l)
    gl[l] = P.diag ? g[(size_t)l * (P.row_end - P.row_begin) + pair] : g[((size_t)l * (P.row_end - P.row_begin) + (pair / P.N2)) * P.N2 + (pair % P.N2)];
  T CtotR[kGenMax], CcolR[kGenMax][kGenMax];
  for (int l = 0; l < kGenMax; ++l) { CtotR[l] = T(0); for (int b = 0; b < kGenMax; ++b) CcolR[l][b] = T(0); }
  for (int p = PR - 1; p >= 0; --p) {
    const T Dv = colok ? Dt[(size_t)p * QC + q] : T(0);
    T Dbar = T(0);
    T Rb[kGenMax][kGenMax];
    {
      const int dtop = min(m, order);
      for (int a = 0; a < dtop; ++a)
        for (int b = 0; b < dtop; ++b) Rb[a][b] = gl[m];
    }
    int toff = NTL;
    for (int l = m; l >= 2; --l) {                  // transition level l-1 -> l; running sums indexed by lvl - 1 = l - 2
      const int dn = min(l, order), dd = min(l - 1, order);
      toff -= dn * dn;
      T F[kGenMax][kGenMax];
      for (int a = 0; a < dn; ++a)
        for (int b = 0; b < dn; ++b) {
          F[a][b] = colok ? Ft[(size_t)(toff + a * dn + b) * cells + (size_t)p * QC + q] : T(0);
          Dbar = gfma<T>(Rb[a][b], F[a][b], Dbar);
        }
      const T X00 = Rb[0][0] * Dv;
      const T totbar = block_excl_scan_rev<T>(CtotR[l - 2], warp_tot);
      T rowbar[kGenMax], colbar[kGenMax], Yj[kGenMax];
      for (int jj = 2; jj <= dn; ++jj) {
        Yj[jj - 2] = Rb[0][jj - 1] * Dv / T(jj);
        colbar[jj - 2] = CcolR[l - 2][jj - 2];
        rowbar[jj - 2] = block_excl_scan_rev<T>(Rb[jj - 1][0] * Dv / T(jj), warp_tot);
      }
      T Rn[kGenMax][kGenMax];
      for (int a = 0; a < dd; ++a)
        for (int b = 0; b < dd; ++b) {
          T v = gl[l - 1] + totbar;
          if (a + 2 <= dn) v += rowbar[a];
          if (b + 2 <= dn) v += colbar[b];
          if (a + 1 < dn && b + 1 < dn) v += Rb[a + 1][b + 1] * Dv / T((a + 2) * (b + 2));
          Rn[a][b] = v;
        }
      CtotR[l - 2] += X00;
      for (int jj = 2; jj <= dn; ++jj) CcolR[l - 2][jj - 2] += Yj[jj - 2];
      for (int a = 0; a < dd; ++a)
        for (int b = 0; b < dd; ++b) Rb[a][b] = Rn[a][b];
    }
    Dbar += m >= 2 ? Rb[0][0] : gl[1];              // level 1: R[0][0] = D
    if (colok) W[(size_t)p * QC + q] = Dbar;
  }
}

}  // namespace gpsig
