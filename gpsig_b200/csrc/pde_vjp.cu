// Reverse mode of the untruncated (Goursat PDE) kernel diagonal, as the reference defines it.
//
// The reference does NOT differentiate its stencil: its hand-written gradient (`_KdiagGrad`,
// gpsig/kernels_pde.py:465-509 for the Cython path; the same tensor algebra in
// gpsig/covariance_op/_untrunc_cov_grad.py:25-77 for the TF op) is the continuous adjoint
//     KK[p, q]     = K[p, q] * K_rev[P-1-p, P-1-q]                    (kernels_pde.py:493-496)
//     K_grad[i, :] = 1/c * sum_{p in coarse row i} sum_q KK[p, q] * dx[q >> n, :] / c      (:498-504)
//     grad_x[l, :] = -2 K_grad[l, :] + 2 K_grad[l-1, :]                                     (:508)
// with K the forward grid (2nd-order stencil, sigKer_fast.pyx:48) and K_rev the grid of the TIME-REVERSED
// path (1st-order stencil in the Cython path, pyx:50,61; the 2nd-order one in the TF op, which calls the same
// `untrunc_cov` on the reversed increments).  It needs both grids at the same cell, and they are swept from
// opposite corners.  The reference keeps both [A, P+1, P+1] grids (2 x 32 GB at C3) and forms the products
// with a chain of dense TF ops; here
//   sweep 0  solves the reversed grid and writes it once to an HBM scratch [pairs, P, P] (8 P^2 bytes/pair),
//   sweep 1  solves the forward grid on chip, reads the matching reversed value as each cell is produced,
//            reduces over q in registers and adds the two signed contributions of every fine row to grad_x.
// Both sweeps: one warp per sequence, lane = one fine row of a 32-row band, lanes skewed by one column
// (lane t works on column s - t at step s), so `up` arrives by one shuffle and `left` is the lane's own
// previous value; the last row of a band is handed to the next band through a P-entry row buffer.  The
// forward grid is never stored.  HBM traffic per pair: 16 P^2 bytes (one write + one read of the scratch).
#include "common.cuh"

namespace gpsig {

struct PdeVjpParams {
  const double* X;        // [N, L, d] coordinates (already divided by the lengthscales)
  const double* gout;     // [row_end - row_begin] upstream gradient (sigma folded in by the caller)
  double* gradX;          // [N, L, d]
  double* E;              // [chunk, L-1, L-1] coarse increments <dx_I, dx_J> / 4^n
  double* grid;           // [chunk, P, P] reversed-path grid
  double* rowbuf;         // [chunk, P]
  int L, d, order, P;
  int pair0, npairs;      // sequences [pair0, pair0 + npairs) of this launch
  int row_begin;
  int rev_solver;
};

// E[I][J] in the accumulation order of sigKer_fast.pyx:41-43 (inc += dx*dy/factor, k ascending)
__global__ void __launch_bounds__(256) pde_vjp_inc_kernel(const PdeVjpParams P) {
  const int Lm = P.L - 1;
  const size_t per = (size_t)Lm * Lm;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= per * P.npairs) return;
  const int a = (int)(idx / per);
  const int I = (int)((idx % per) / Lm), J = (int)(idx % Lm);
  const double* x = P.X + (size_t)(P.pair0 + a) * P.L * P.d;
  const double factor = (double)(1u << (2 * P.order));
  double inc = 0.;
  for (int k = 0; k < P.d; ++k)
    inc = inc + (x[(I + 1) * P.d + k] - x[I * P.d + k]) * (x[(J + 1) * P.d + k] - x[J * P.d + k]) / factor;
  P.E[idx] = inc;
}

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

// MODE 0: reversed-path grid -> HBM scratch.  MODE 1: forward grid x scratch -> gradX.
template <int MODE>
__global__ void __launch_bounds__(128) pde_vjp_sweep_kernel(const PdeVjpParams P) {
  const int lane = threadIdx.x & 31;
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (a >= P.npairs) return;
  const int n = P.order, c = 1 << n, Pf = P.P, Lm = P.L - 1, d = P.d;
  const double* E = P.E + (size_t)a * Lm * Lm;
  double* grid = P.grid + (size_t)a * Pf * Pf;
  double* rowbuf = P.rowbuf + (size_t)a * Pf;
  const double* x = P.X + (size_t)(P.pair0 + a) * P.L * d;
  const bool second = MODE == 1 || P.rev_solver == 1;

  for (int band = 0; band * 32 < Pf; ++band) {
    const int r = band * 32 + lane;
    const bool row_ok = r < Pf;
    // coarse row of the cells (r-1, .) this lane solves; MODE 0 indexes the reversed increments
    // inc_rev[I][J] = inc[L-2-I][L-2-J] (pyx:43)
    const int Ic = r > 0 ? ((min(r, Pf - 1) - 1) >> n) : 0;
    const double* Erow = E + (size_t)(MODE == 0 ? (Lm - 1 - Ic) : Ic) * Lm;
    // MODE 1: this row of the forward grid meets row P-1-r of the stored grid, walked right to left
    const double* grow = grid + (size_t)(Pf - 1 - min(r, Pf - 1)) * Pf;
    double acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = 0.;
    double sJ = 0.;
    double cur = 1., up_prev = 1.;
    double A = 1., B = 1.;
    double e_next = Erow[MODE == 0 ? (Lm - 1) : 0];
    double up0_next = 1.;                                  // lane 0: rowbuf[q] one step ahead
    if (lane == 0 && band > 0) up0_next = rowbuf[0];
    double g_next = 0.;
    if (MODE == 1 && row_ok && lane == 0) g_next = __ldg(grow + (Pf - 1));
    const int nsteps = Pf + 31;
    for (int s = 0; s < nsteps; ++s) {
      const int q = s - lane;
      double up = shfl_up1(cur);                           // (r-1, q): finished by lane-1 one step ago
      if (lane == 0) {
        up = band > 0 ? up0_next : 1.;
        if (band > 0 && s + 1 < Pf) up0_next = rowbuf[s + 1];
      }
      const bool active = row_ok && q >= 0 && q < Pf;
      double g = 0.;
      if (MODE == 1) {
        g = g_next;
        if (row_ok && q + 1 >= 0 && q + 1 < Pf) g_next = __ldg(grow + (Pf - 2 - q));
      }
      if (active) {
        if (q >= 1 && ((q - 1) & (c - 1)) == 0) {          // entering coarse column J = (q-1) >> n
          const double e = e_next;
          const int Jn = ((q - 1) >> n) + 1;
          if (Jn < Lm) e_next = Erow[MODE == 0 ? (Lm - 1 - Jn) : Jn];
          if (second) { const double e2 = e * e * (1. / 12); A = 1. + 0.5 * e + e2; B = 1. - e2; }
          else { A = 1.; B = 1. - e; }
        }
        const double v = (r == 0 || q == 0) ? 1. : (up + cur) * A - up_prev * B;
        cur = v;
        if (MODE == 0) {
          grid[(size_t)r * Pf + q] = v;
        } else {
          sJ = fma(v, g, sJ);
          if (((q + 1) & (c - 1)) == 0) {                  // last fine column of coarse column J = q >> n
            const int J = q >> n;
            const double* xa = x + (size_t)J * d;
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (k < d) acc[k] = fma(sJ, xa[d + k] - xa[k], acc[k]);
            sJ = 0.;
          }
        }
        if (lane == 31) rowbuf[q] = v;                     // top row of the next band
      }
      up_prev = up;
    }
    if (MODE == 1 && row_ok) {
      // K_grad fine row r -> coarse row i = r >> n, scaled by 1/c (row average) * 1/c (dx on the fine grid),
      // grad_x[i] -= 2 v, grad_x[i+1] += 2 v (kernels_pde.py:508), times the upstream gradient
      const double w = 2. * P.gout[P.pair0 + a - P.row_begin] / ((double)c * (double)c);
      double* gx = P.gradX + ((size_t)(P.pair0 + a) * P.L + (r >> n)) * d;
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (k < d) { const double t = w * acc[k]; atomicAdd(gx + k, -t); atomicAdd(gx + d + k, t); }
    }
    __syncwarp();
  }
}

static size_t pde_vjp_pair_bytes(int L, int order) {
  const size_t Pf = (size_t)(L - 1) << order;
  return (Pf * Pf + (size_t)(L - 1) * (L - 1) + Pf) * sizeof(double);
}

static int pde_kdiag_vjp_impl(const double* X, int N, int L, int d, int order, int rev_solver, int row_begin, int row_end,
                              const double* gout, double* gradX, void* ws, size_t ws_bytes, void* stream) {
  if (!X || !gout || !gradX) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "X/grad_out/gradX must not be NULL");
  if (d < 1 || d > 16) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "state dimension d=%d is not supported (1..16)", d);
  if (L < 2) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d)", L);
  if (order < 0 || order > 8) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "dyadic order %d out of range (0..8)", order);
  if (rev_solver != 0 && rev_solver != 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "rev_solver must be 0 (1st order) or 1 (2nd order)");
  if (row_begin < 0 || row_end > N || row_begin > row_end)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad row range [%d, %d) for N=%d", row_begin, row_end, N);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int rows = row_end - row_begin;
  if (rows == 0) return GPSIG_OK;
  GPSIG_CUDA_CHECK(cudaMemsetAsync(gradX + (size_t)row_begin * L * d, 0, (size_t)rows * L * d * sizeof(double), s));
  const size_t per = pde_vjp_pair_bytes(L, order);
  if (!ws || ws_bytes < per || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need at least %zu bytes (one sequence), got %zu", per, ws_bytes);
  const int chunk = (int)min((size_t)rows, ws_bytes / per);
  const size_t Pf = (size_t)(L - 1) << order;
  PdeVjpParams P;
  memset(&P, 0, sizeof(P));
  P.X = X; P.gout = gout; P.gradX = gradX;
  P.grid = static_cast<double*>(ws);
  P.E = P.grid + (size_t)chunk * Pf * Pf;
  P.rowbuf = P.E + (size_t)chunk * (L - 1) * (L - 1);
  P.L = L; P.d = d; P.order = order; P.P = (int)Pf; P.row_begin = row_begin; P.rev_solver = rev_solver;
  for (int p0 = row_begin; p0 < row_end; p0 += chunk) {
    P.pair0 = p0; P.npairs = min(chunk, row_end - p0);
    const size_t ninc = (size_t)P.npairs * (L - 1) * (L - 1);
    pde_vjp_inc_kernel<<<(unsigned)((ninc + 255) / 256), 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_inc_kernel");
    const unsigned grid = (unsigned)((P.npairs + 3) / 4);
    prof_begin(s);
    pde_vjp_sweep_kernel<0><<<grid, 128, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_sweep_kernel<0>");
    pde_vjp_sweep_kernel<1><<<grid, 128, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_sweep_kernel<1>");
    prof_end(s);
  }
  return GPSIG_OK;
}

}  // namespace gpsig

extern "C" {

size_t gpsig_pde_kdiag_vjp_workspace_bytes(int nseq, int L, int dyadic_order) {
  if (nseq < 1 || L < 2 || dyadic_order < 0) return 0;
  return (size_t)nseq * gpsig::pde_vjp_pair_bytes(L, dyadic_order);
}

int gpsig_pde_kdiag_vjp_f64(const double* X, int N, int L, int d, int dyadic_order, int rev_solver, int row_begin,
                            int row_end, const double* grad_out, double* gradX, void* workspace,
                            size_t workspace_bytes, void* stream) {
  return gpsig::pde_kdiag_vjp_impl(X, N, L, d, dyadic_order, rev_solver, row_begin, row_end, grad_out, gradX, workspace,
                                   workspace_bytes, stream);
}

}  // extern "C"
