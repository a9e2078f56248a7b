// Higher-order recursions (order > 1) for sm_100a -- the general, unspecialised path.
//
//   trunc_generic_kernel     gpsig/signature_algs.py:37-74  signature_kern_higher_order
//                            (via gpsig/kernels.py:209-238 _K_seq and :190-207 _K_seq_diag)
//   tens_seq_generic_kernel  gpsig/signature_algs.py:129-160 signature_kern_tens_vs_seq_higher_order
//                            (via gpsig/kernels.py:314-341 _K_tens_vs_seq)
//
// order = 1 (the default and every BASELINE config) runs the fused kernels in
// trunc_skew_kernel.cuh / trunc_kernel.cuh / tens_kernels.cuh.  order > 1 keeps, per tile element,
// an order x order family R[a][b] per level (a/b = how often the current row / column index
// was repeated); it is what the reference's notebooks validate against signatures
// (notebooks/signature_kernel.ipynb cells 11-29).  Still fused (the lifted tensor never exists),
// but written for generality (runtime levels <= 8, order <= 8), not throughput:
// one CTA per pair, one thread per tile column sweeping the rows, the exclusive row prefixes of
// signature_algs.py:64,67 as block-wide scans.
#pragma once
#include "trunc_kernel.cuh"
#include "tens_kernels.cuh"

namespace gpsig {

constexpr int kGenMax = 8;     // max levels and max order of the generic path

struct GenericParams {
  const void* xp; const void* yp;       // packed sequences, plain layout (common.cuh)
  size_t xblk, yblk;
  int N, N2, L, L2, DP;
  int row_begin, row_end;
  int m, order, difference;
  int symmetric, diag;
  double p0, p1; int base;
  SpectralParams sp; int d;             // GPSIG_BASE_SPECTRAL: parameter block and the true state dimension
  const void* norm_x; const void* norm_x2; const void* level_scale;
  double sym_jitter; int return_levels;
  void* out;
};

// exclusive prefix sum over the threads of the block (all threads must call)
template <typename T>
__device__ __forceinline__ T block_excl_scan(T v, T* warp_tot /* [32] shared */) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  T incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const T t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();                       // previous use of warp_tot is over
  if (lane == 31) warp_tot[w] = incl;
  __syncthreads();
  T base = T(0);
  for (int k = 0; k < w && k < nw; ++k) base += warp_tot[k];
  return base + incl - v;
}

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* warp_tot) {
  v = warp_sum<T>(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = v;
  __syncthreads();
  T s = T(0);
  const int nw = (blockDim.x + 31) >> 5;
  for (int k = 0; k < nw; ++k) s += warp_tot[k];
  return s;
}

template <typename T>
__global__ void __launch_bounds__(1024) trunc_generic_kernel(const GenericParams P) {
  __shared__ T warp_tot[32];
  const int diff = P.difference ? 1 : 0;
  const int nrows = P.L - diff, ncols = P.L2 - diff;
  const int DP = P.DP, m = P.m, order = P.order;
  int i, j;
  if (P.diag) { i = j = blockIdx.x; }
  else { i = P.row_begin + blockIdx.x / P.N2; j = blockIdx.x % P.N2; }
  const T* x = static_cast<const T*>(P.xp) + (size_t)i * P.xblk;
  const T* y = static_cast<const T*>(P.yp) + (size_t)j * P.yblk;
  const T* xsq = x + (size_t)P.L * DP;
  const T* ysq = y + (size_t)P.L2 * DP;
  const int q = threadIdx.x;                       // tile column (of the differenced tile)
  const bool colok = q < ncols;
  const T p0 = T(P.p0), p1 = T(P.p1);

  auto kappa = [&](int p, int qq) {
    const T* xr = x + (size_t)p * DP; const T* yr = y + (size_t)qq * DP;
    if (P.base == GPSIG_BASE_SPECTRAL) return spectral_eval<T>(P.sp, P.d, xr, yr);
    T dot = T(0);
    for (int k = 0; k < DP; ++k) dot = gfma<T>(xr[k], yr[k], dot);
    return base_eval<T>(P.base, p0, p1, dot, xsq[p], ysq[qq]);
  };

  // running state of this column, per level i = 1..m-1 (index i-1)
  T Ctot[kGenMax], Ccol[kGenMax][kGenMax];
  T Ksum[kGenMax + 1];
  for (int l = 0; l < kGenMax; ++l) { Ctot[l] = T(0); for (int b = 0; b < kGenMax; ++b) Ccol[l][b] = T(0); }
  for (int l = 0; l <= kGenMax; ++l) Ksum[l] = T(0);

  T Mlo = T(0), Mhi = T(0);                        // M[p, q], M[p, q+1] of the previous row (difference mode)
  if (diff && colok) { Mlo = kappa(0, q); Mhi = kappa(0, q + 1); }

  for (int p = 0; p < nrows; ++p) {
    T Dv = T(0);
    if (colok) {
      if (diff) {
        const T a = kappa(p + 1, q), b = kappa(p + 1, q + 1);
        Dv = b + Mlo - Mhi - a;                    // signature_algs.py:56
        Mlo = a; Mhi = b;
      } else Dv = kappa(p, q);
    }
    // level 1: the family is the single entry R[0][0] = M (signature_algs.py:58-60)
    T R[kGenMax][kGenMax];
    R[0][0] = Dv;
    Ksum[1] += Dv;
    int d = 1;
    for (int lvl = 1; lvl < m; ++lvl) {            // level lvl -> level lvl+1 (signature_algs.py:61-72)
      const int dn = min(lvl + 1, order);
      T tot = T(0), colsum[kGenMax], rowsum[kGenMax];
      for (int a = 0; a < d; ++a) { rowsum[a] = T(0); colsum[a] = T(0); }
      for (int a = 0; a < d; ++a)
        for (int b = 0; b < d; ++b) { const T v = R[a][b]; tot += v; rowsum[a] += v; colsum[b] += v; }
      // exclusive 2-D prefix of tot: rows < p are in Ctot, columns < q by a block scan (:64)
      const T T2 = block_excl_scan<T>(Ctot[lvl - 1], warp_tot);
      T Pr[kGenMax];
      for (int a = 0; a + 1 < dn; ++a) Pr[a] = block_excl_scan<T>(a < d ? rowsum[a] : T(0), warp_tot);   // same row, columns < q (:67)
      T Rn[kGenMax][kGenMax];
      Rn[0][0] = Dv * T2;
      for (int jj = 2; jj <= dn; ++jj) {
        Rn[0][jj - 1] = Dv * Ccol[lvl - 1][jj - 2] / T(jj);                      // :66  rows < p of sum_a R[a][jj-2]
        Rn[jj - 1][0] = Dv * Pr[jj - 2] / T(jj);                                 // :67
        for (int kk = 2; kk <= dn; ++kk) Rn[jj - 1][kk - 1] = Dv * R[jj - 2][kk - 2] / T(jj * kk);   // :69
      }
      Ctot[lvl - 1] += tot;
      for (int b = 0; b < d; ++b) Ccol[lvl - 1][b] += colsum[b];
      T s = T(0);
      for (int a = 0; a < dn; ++a)
        for (int b = 0; b < dn; ++b) { R[a][b] = Rn[a][b]; s += Rn[a][b]; }
      Ksum[lvl + 1] += s;
      d = dn;
    }
  }
  // ---- totals and epilogue (kernels.py:431-477) ----
  T acc = T(0);
  for (int lvl = 0; lvl <= m; ++lvl) {
    T v = lvl == 0 ? T(1) : block_sum<T>(Ksum[lvl], warp_tot);
    if (threadIdx.x != 0) continue;
    if (P.diag) { static_cast<T*>(P.out)[(size_t)lvl * P.N + i] = v; continue; }
    if (P.symmetric && i == j) v += T(P.sym_jitter);
    if (P.norm_x) {
      const T a = gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + i]);
      const T b = gsqrt<T>(static_cast<const T*>(P.norm_x2)[(size_t)lvl * P.N2 + j]);
      v = v / (a * b);
    }
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
    const size_t rows = (size_t)(P.row_end - P.row_begin);
    const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
    if (P.return_levels) static_cast<T*>(P.out)[(size_t)lvl * rows * P.N2 + rc] = v; else acc += v;
  }
  if (threadIdx.x == 0 && !P.diag && !P.return_levels) {
    const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
    static_cast<T*>(P.out)[rc] = acc;
  }
}

template <typename T>
int launch_trunc_generic(const GenericParams& P, cudaStream_t stream) {
  const int ncols = P.L2 - (P.difference ? 1 : 0);
  if (ncols > 1024)
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "order > 1 supports at most 1024 tile columns (L2=%d)", P.L2);
  if (P.m > kGenMax || P.order > kGenMax)
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "order > 1 supports num_levels <= %d (got %d)", kGenMax, P.m);
  const size_t npairs = P.diag ? (size_t)P.N : (size_t)(P.row_end - P.row_begin) * P.N2;
  if (npairs > 0x7fffffffull) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "too many pairs in one call (%zu)", npairs);
  const int threads = (ncols + 31) / 32 * 32;
  prof_begin(stream);
  trunc_generic_kernel<T><<<(unsigned)npairs, threads, 0, stream>>>(P);
  prof_end(stream);
  GPSIG_LAUNCH_CHECK("trunc_generic_kernel");
  return GPSIG_OK;
}

// ---------------------------------------------------------------------------
// tensors vs sequences, higher order: one thread per (z, x_n) pair, sequential in time
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) tens_seq_generic_kernel(const TensSeqParams P, const int order, const int DP) {
  const int ncols = P.col_end - P.col_begin;
  const int nl = blockIdx.x * blockDim.x + threadIdx.x;
  const int z = blockIdx.y;
  if (nl >= ncols) return;
  const int n = P.col_begin + nl;
  const int zs = P.zs, m = P.m, L = P.L;
  const T* xblk = static_cast<const T*>(P.xp) + (size_t)n * P.xblk;
  const T* xsq = xblk + (size_t)L * DP;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const bool diff = P.difference != 0;

  auto kappa = [&](int t, int s, int l) {          // base kernel of tensor component (t, z, s) and x_n[l]
    const T* blk = static_cast<const T*>(P.zp) + ((size_t)t * P.Mz + z) * P.zblk;
    const T* zc = blk + s * DP; const T* xr = xblk + (size_t)l * DP;
    if (P.base == GPSIG_BASE_SPECTRAL) return spectral_eval<T>(P.sp, P.d, zc, xr);
    T dot = T(0);
    for (int k = 0; k < DP; ++k) dot = gfma<T>(zc[k], xr[k], dot);
    return base_eval<T>(P.base, p0, p1, dot, blk[zs * DP + s], xsq[l]);
  };
  auto Mval = [&](int t, int l) { return zs == 2 ? kappa(t, 1, l) - kappa(t, 0, l) : kappa(t, 0, l); };   // kernels.py:331

  T acc = T(0);
  auto emit = [&](int lvl, T v) {
    if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + n]);
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
    if (P.return_levels) static_cast<T*>(P.out)[((size_t)lvl * P.Mz + z) * ncols + nl] = v; else acc += v;
  };
  emit(0, T(1));
  int k0 = 0;
  for (int lvl = 1; lvl <= m; ++lvl) {             // signature_algs.py:140-158
    // component jj of this level: family cur[jj][0..d_jj-1] at the current time, S[jj] = sum over earlier times of its total
    T S[kGenMax], cur[kGenMax][kGenMax], Mprev[kGenMax];
    for (int jj = 0; jj < lvl; ++jj) { S[jj] = T(0); Mprev[jj] = T(0); }
    T Klvl = T(0);
    for (int l = 0; l < L; ++l) {
      const bool first = diff && l == 0;
      for (int jj = 0; jj < lvl; ++jj) {
        const T Mt = Mval(k0 + jj, l);
        T D = Mt;
        if (diff) { D = Mt - Mprev[jj]; Mprev[jj] = Mt; }     // signature_algs.py:138
        if (first) continue;
        const int d = min(jj + 1, order);
        if (jj == 0) cur[0][0] = D;
        else {
          cur[jj][0] = D * S[jj - 1];                          // :153  exclusive cumsum of sum(R)
          for (int c = 1; c < d; ++c) cur[jj][c] = D * cur[jj - 1][c - 1] / T(c + 1);   // :155
        }
      }
      if (first) continue;
      // advance the running sums AFTER the whole time step (S[jj] must exclude the current time)
      for (int jj = 0; jj < lvl; ++jj) {
        const int d = min(jj + 1, order);
        T s = T(0);
        for (int c = 0; c < d; ++c) s += cur[jj][c];
        if (jj == lvl - 1) Klvl += s; else S[jj] += s;
      }
    }
    emit(lvl, Klvl);                                           // :158
    k0 += lvl;
  }
  if (!P.return_levels) static_cast<T*>(P.out)[(size_t)z * ncols + nl] = acc;
}

template <typename T>
int launch_tens_seq_generic(const TensSeqParams& P, int order, int DP, cudaStream_t stream) {
  if (P.m > kGenMax) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "order > 1 supports num_levels <= %d (got %d)", kGenMax, P.m);
  const int ncols = P.col_end - P.col_begin;
  dim3 grid((ncols + 127) / 128, P.Mz);
  prof_begin(stream);
  tens_seq_generic_kernel<T><<<grid, 128, 0, stream>>>(P, order, DP);
  prof_end(stream);
  GPSIG_LAUNCH_CHECK("tens_seq_generic_kernel");
  return GPSIG_OK;
}

}  // namespace gpsig
