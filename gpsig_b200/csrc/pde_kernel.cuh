// Untruncated signature kernel: Goursat PDE wavefront solver for sm_100a.
//
// Replaces gpsig/sigKer_fast.pyx:15-62 (sig_kern_diag, the numerics oracle)
// and gpsig/covariance_op/untrunc_cov_op_gpu.cu:5-36 (UntruncCovKernel, which
// keeps the whole (P+1)^2 grid in global memory, reads it 3x per cell, syncs
// the block on every anti-diagonal and is limited to P+1 < 1024).
//
// Here one WARP solves one pair (x, y) as a register-systolic wavefront:
//   * the grid is cut into vertical strips of 32*W columns; lane t owns W
//     consecutive columns of the strip and keeps the previous grid row of
//     those columns in registers;
//   * at step s lane t works on grid row (s - t): the skew makes the value it
//     needs from its left neighbour (same row, computed one step earlier)
//     arrive with a single warp shuffle -- one anti-diagonal per step, no
//     barrier, no global or shared traffic for the solution itself;
//   * the right edge of a strip is parked in shared memory ((P+1) doubles) and
//     becomes the left boundary of the next strip;
//   * dyadic refinement: the increment <dx_I, dy_J>/4^o and the two stencil
//     coefficients are evaluated once per COARSE cell and reused by its
//     4^o fine cells (the reference recomputes the dot product per fine cell,
//     pyx:41-43).
// Only k(x, y) = K[P, Q] leaves the chip (kernels_pde.py:185).
#pragma once
#include "common.cuh"

namespace gpsig {

constexpr int kPdeW = 8;   // columns per lane

struct PdeParams {
  const void* xp; const void* yp;     // packed sequences (linear base kernel: raw scaled coords)
  size_t xblk, yblk;
  int N, N2, L, L2, DP;
  int order;                          // dyadic refinement
  int pair_mode;                      // GPSIG_PDE_DIAG / GPSIG_PDE_GRAM
  int row_begin, row_end;
  double sigma;
  double* out;
};

// smem per warp: dx[(L-1)][DP], dy[(L2-1)][DP], edge[P+1]
inline size_t pde_smem_bytes(int L, int L2, int DP, int order) {
  const size_t P = ((size_t)(L - 1)) << order;
  return ((size_t)(L - 1) * DP + (size_t)(L2 - 1) * DP + P + 1) * sizeof(double);
}

template <int DP, int ORDER>
__global__ void __launch_bounds__(32) pde_wavefront_kernel(const PdeParams Pm) {
  constexpr int W = kPdeW;
  constexpr int NC = (W >> ORDER) > 0 ? (W >> ORDER) : 1;   // coarse columns per lane
  extern __shared__ __align__(128) unsigned char pde_smem[];
  const int lane = threadIdx.x;
  const int L = Pm.L, L2 = Pm.L2;
  const int P = (L - 1) << ORDER, Q = (L2 - 1) << ORDER;
  double* dx = reinterpret_cast<double*>(pde_smem);
  double* dy = dx + (size_t)(L - 1) * DP;
  double* edge = dy + (size_t)(L2 - 1) * DP;

  int ix, iy;
  if (Pm.pair_mode == GPSIG_PDE_DIAG) { ix = Pm.row_begin + blockIdx.x; iy = ix; }
  else { ix = Pm.row_begin + blockIdx.x / Pm.N2; iy = blockIdx.x % Pm.N2; }
  const double* x = static_cast<const double*>(Pm.xp) + (size_t)ix * Pm.xblk;
  const double* y = static_cast<const double*>(Pm.yp) + (size_t)iy * Pm.yblk;

  // increments, scaled so that <dx, dy> is already divided by 4^order (exact: power of two)
  const double fscale = 1.0 / (double)(1 << (2 * ORDER));
  for (int e = lane; e < (L - 1) * DP; e += 32) dx[e] = (x[e + DP] - x[e]) * fscale;
  for (int e = lane; e < (L2 - 1) * DP; e += 32) dy[e] = y[e + DP] - y[e];
  for (int e = lane; e <= P; e += 32) edge[e] = 1.0;       // K[., 0] = 1 (pyx:29-31)
  __syncwarp();

  const int nstrips = (Q + 32 * W - 1) / (32 * W);
  double result = 1.0;
  for (int strip = 0; strip < nstrips; ++strip) {
    const int j0 = strip * 32 * W + lane * W;               // first fine column (cell index) of this lane
    const int ncell = max(0, min(W, Q - j0));              // valid cells of this lane
    const int last_lane = min(31, (Q - strip * 32 * W - 1) / W);
    double up[W];                                           // K[i, j0+1 .. j0+W]   (previous row)
#pragma unroll
    for (int c = 0; c < W; ++c) up[c] = 1.0;                // K[0, .] = 1
    double A[NC], B[NC];
    double diag = 1.0;                                      // K[i, j0]
    double send = 1.0;
    const int nsteps = P + last_lane + 1;
    for (int s = 0; s < nsteps; ++s) {
      const int i = s - lane;                               // fine row of this lane at this step
      double left = __shfl_up_sync(0xffffffffu, send, 1);   // K[i+1, j0] from lane-1 (computed last step)
      const bool rowok = (i >= 0) && (i < P);
      if (lane == 0) left = rowok ? edge[i + 1] : 1.0;
      if (rowok && ncell > 0) {
        if ((i & ((1 << ORDER) - 1)) == 0) {                // entering a new coarse row: refresh coefficients
          const double* dxi = dx + (size_t)(i >> ORDER) * DP;
#pragma unroll
          for (int cc = 0; cc < NC; ++cc) {
            const int J = min((j0 >> ORDER) + cc, L2 - 2);
            const double* dyj = dy + (size_t)J * DP;
            double e = 0.0;
#pragma unroll
            for (int k = 0; k < DP; ++k) e = fma(dxi[k], dyj[k], e);
            const double e2 = e * e;
            A[cc] = 1. + 0.5 * e + (1. / 12) * e2;          // pyx:48
            B[cc] = 1. - (1. / 12) * e2;
          }
        }
        const double newdiag = left;                        // becomes K[i+1, j0] = next row's diag
        double dg = diag, lf = left;
#pragma unroll
        for (int c = 0; c < W; ++c) {
          const int cc = (ORDER >= 3) ? 0 : (c >> ORDER);
          const double w = fma(up[c], A[cc], -(dg * B[cc]));  // off the dependency chain
          const double v = fma(lf, A[cc], w);                  // (k01 + k10) * A - k00 * B
          dg = up[c];
          if (c < ncell) { up[c] = v; lf = v; }
        }
        diag = newdiag;
        send = lf;
        if (lane == last_lane) {
          if (strip + 1 < nstrips) edge[i + 1] = lf;         // right edge -> next strip's left boundary
          else if (i == P - 1) result = lf;
        }
      }
    }
    __syncwarp();
  }
  // broadcast from the lane that owns K[P, Q]
  const int owner = min(31, (Q - (nstrips - 1) * 32 * W - 1) / W);
  result = __shfl_sync(0xffffffffu, result, owner);
  if (lane == 0) {
    const size_t o = Pm.pair_mode == GPSIG_PDE_DIAG ? (size_t)blockIdx.x
                                                    : (size_t)(ix - Pm.row_begin) * Pm.N2 + iy;
    Pm.out[o] = Pm.sigma * result;
  }
}

}  // namespace gpsig
