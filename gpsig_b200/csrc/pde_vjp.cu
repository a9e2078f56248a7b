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
// Both sweeps: one CTA per sequence, its warps pipeline the 32-row bands of the grid (see the kernel).  The
// forward grid is never stored.  HBM traffic per sequence: 16 P^2 bytes (one coalesced write + one read).
#include "common.cuh"

namespace gpsig {

struct PdeVjpParams {
  const double* X;        // [N, L, d] coordinates (already divided by the lengthscales)
  const double* gout;     // [row_end - row_begin] upstream gradient (sigma folded in by the caller)
  double* gradX;          // [N, L, d]
  double* E;              // [chunk, L-1, L-1] coarse increments <dx_I, dx_J> / 4^n
  double* grid;           // [chunk, bands, P + 31, 32] reversed-path grid in sweep order
  double* rowbuf;         // [chunk, kVjpWarps, 32 * bands]
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

constexpr int kVjpWarps = 8;      // warps per CTA = bands of one sequence in flight
constexpr int kVjpUnroll = 4;     // steps per prefetch group of the stored grid

__host__ __device__ inline size_t vjp_bands(size_t Pf) { return (Pf + 31) / 32; }
__host__ __device__ inline size_t vjp_steps(size_t Pf) { return Pf + 31; }                 // steps of one band
__host__ __device__ inline size_t vjp_grid_elems(size_t Pf) { return vjp_bands(Pf) * vjp_steps(Pf) * 32; }
__host__ __device__ inline size_t vjp_rowbuf_elems(size_t Pf) { return (size_t)kVjpWarps * vjp_bands(Pf) * 32; }

// One CTA per sequence, kVjpWarps warps; warp w sweeps the 32-row bands w, w + W, ... of the (P x P) grid, lane =
// one fine row, lanes skewed by one column (lane t is at column s - t at step s).  `up` comes from the lane
// above by one shuffle; the last row of band b reaches band b+1 (the next warp) through a row buffer in global
// memory, published in blocks of 32 columns through a per-band progress counter in shared memory, so all bands
// of a sequence run as a pipeline.
// MODE 0 solves the grid of the time-reversed path and stores it; MODE 1 solves the forward grid, multiplies
// each value by the stored reversed value of the same cell and reduces.  The scratch is laid out in the ORDER
// THE SWEEP PRODUCES IT, [band][step][lane]: every step is one coalesced 256-byte store (MODE 0) or load
// (MODE 1).  The two sweeps meet in the same cell when MODE 1 aligns its bands to the END of the grid
// (rows r = 32 band + lane - pad): its (band, step s, lane t) reads MODE 0's (NB-1-band, P+30-s, 31-t).
template <int MODE>
__global__ void __launch_bounds__(32 * kVjpWarps) pde_vjp_sweep_kernel(const PdeVjpParams P) {
  extern __shared__ int prog[];                            // [NB] columns of band b's last row published so far
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a = blockIdx.x;
  const int n = P.order, c = 1 << n, Pf = P.P, Lm = P.L - 1, d = P.d;
  const int NB = (int)vjp_bands(Pf), NS = (int)vjp_steps(Pf);
  for (int i = threadIdx.x; i < NB; i += blockDim.x) prog[i] = 0;
  __syncthreads();
  volatile int* vprog = prog;
  const double* E = P.E + (size_t)a * Lm * Lm;
  double* S = P.grid + (size_t)a * vjp_grid_elems(Pf);
  double* rowbufs = P.rowbuf + (size_t)a * vjp_rowbuf_elems(Pf);
  const size_t Pfp = (size_t)NB * 32;
  const double* x = P.X + (size_t)(P.pair0 + a) * P.L * d;
  const bool second = MODE == 1 || P.rev_solver == 1;
  const int roff = MODE == 1 ? NB * 32 - Pf : 0;

  for (int band = warp; band < NB; band += kVjpWarps) {
    const int r = band * 32 + lane - roff;
    const bool row_ok = r >= 0 && r < Pf;
    const int rc = min(max(r, 0), Pf - 1);
    // coarse row of the cells (r-1, .) this lane solves; MODE 0 indexes the reversed increments
    // inc_rev[I][J] = inc[L-2-I][L-2-J] (pyx:43)
    const int Ic = rc > 0 ? ((rc - 1) >> n) : 0;
    const double* Erow = E + (size_t)(MODE == 0 ? (Lm - 1 - Ic) : Ic) * Lm;
    const double* rb_in = rowbufs + (size_t)((band + kVjpWarps - 1) % kVjpWarps) * Pfp;
    double* rb_out = rowbufs + (size_t)(band % kVjpWarps) * Pfp;
    double* Sw = S + (size_t)band * NS * 32 + lane;                              // MODE 0: + s * 32
    const double* Sr = S + (size_t)(NB - 1 - band) * NS * 32 + (31 - lane);      // MODE 1: + (Pf + 30 - s) * 32
    double acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = 0.;
    double sJ = 0.;
    double cur = 1., up_prev = 1.;
    double A = 1., B = 1.;
    double e_next = Erow[MODE == 0 ? (Lm - 1) : 0];
    double rbcur = 1., rbnext = 1.;
    if (band > 0) {
      while (vprog[band - 1] < min(32, Pf)) { }
      __threadfence_block();
      rbnext = __ldcg(rb_in + lane);
    }
    double gcur[kVjpUnroll], gnext[kVjpUnroll];
    if (MODE == 1) {
#pragma unroll
      for (int u = 0; u < kVjpUnroll; ++u) { const int sp = Pf + 30 - u; gnext[u] = __ldg(Sr + (size_t)sp * 32); gcur[u] = 0.; }
    }
    for (int s0 = 0; s0 < NS; s0 += kVjpUnroll) {
      if (MODE == 1) {
#pragma unroll
        for (int u = 0; u < kVjpUnroll; ++u) {
          gcur[u] = gnext[u];
          const int sp = Pf + 30 - (s0 + kVjpUnroll + u);
          if (sp >= 0) gnext[u] = __ldg(Sr + (size_t)sp * 32);
        }
      }
      if ((s0 & 31) == 0 && band > 0) {                    // next 32 columns of the row above this band
        rbcur = rbnext;
        if (s0 + 32 < Pf) {
          const int need = min(s0 + 64, Pf);
          while (vprog[band - 1] < need) { }
          __threadfence_block();
          rbnext = __ldcg(rb_in + s0 + 32 + lane);
        }
      }
#pragma unroll
      for (int u = 0; u < kVjpUnroll; ++u) {
        const int s = s0 + u;
        const int q = s - lane;
        double up = shfl_up1(cur);                         // (r-1, q): finished by lane-1 one step ago
        const double up0 = __shfl_sync(0xffffffffu, rbcur, s & 31);
        if (lane == 0) up = band > 0 ? up0 : 1.;
        const bool active = row_ok && q >= 0 && q < Pf;
        if (active) {
          if (q >= 1 && ((q - 1) & (c - 1)) == 0) {        // entering coarse column J = (q-1) >> n
            const double e = e_next;
            const int Jn = ((q - 1) >> n) + 1;
            if (Jn < Lm) e_next = Erow[MODE == 0 ? (Lm - 1 - Jn) : Jn];
            if (second) { const double e2 = e * e * (1. / 12); A = 1. + 0.5 * e + e2; B = 1. - e2; }
            else { A = 1.; B = 1. - e; }
          }
          const double v = (r == 0 || q == 0) ? 1. : (up + cur) * A - up_prev * B;
          cur = v;
          if (MODE == 0) {
            Sw[(size_t)s * 32] = v;
          } else {
            sJ = fma(v, gcur[u], sJ);
            if (((q + 1) & (c - 1)) == 0) {                // last fine column of coarse column J = q >> n
              const double* xa = x + (size_t)(q >> n) * d;
#pragma unroll
              for (int k = 0; k < 16; ++k)
                if (k < d) acc[k] = fma(sJ, __ldg(xa + d + k) - __ldg(xa + k), acc[k]);
              sJ = 0.;
            }
          }
          if (lane == 31) {                                // top row of the next band
            __stcg(rb_out + q, v);
            if ((q & 31) == 31 || q == Pf - 1) { __threadfence_block(); vprog[band] = q + 1; }
          }
        }
        up_prev = up;
      }
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
  }
}

static size_t pde_vjp_pair_bytes(int L, int order) {
  const size_t Pf = (size_t)(L - 1) << order;
  return (vjp_grid_elems(Pf) + (size_t)(L - 1) * (L - 1) + vjp_rowbuf_elems(Pf)) * sizeof(double);
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
  const size_t Pf = (size_t)(L - 1) << order;
  const size_t smem = vjp_bands(Pf) * sizeof(int);
  if (smem > 200 * 1024) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "grid too long for the PDE backward (P=%zu)", Pf);
  GPSIG_CUDA_CHECK(cudaMemsetAsync(gradX + (size_t)row_begin * L * d, 0, (size_t)rows * L * d * sizeof(double), s));
  const size_t per = pde_vjp_pair_bytes(L, order);
  if (!ws || ws_bytes < per || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need at least %zu bytes (one sequence), got %zu", per, ws_bytes);
  const int chunk = (int)min((size_t)rows, ws_bytes / per);
  PdeVjpParams P;
  memset(&P, 0, sizeof(P));
  P.X = X; P.gout = gout; P.gradX = gradX;
  P.grid = static_cast<double*>(ws);
  P.E = P.grid + (size_t)chunk * vjp_grid_elems(Pf);
  P.rowbuf = P.E + (size_t)chunk * (L - 1) * (L - 1);
  P.L = L; P.d = d; P.order = order; P.P = (int)Pf; P.row_begin = row_begin; P.rev_solver = rev_solver;
  if (smem > 48 * 1024) {
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(pde_vjp_sweep_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(pde_vjp_sweep_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  for (int p0 = row_begin; p0 < row_end; p0 += chunk) {
    P.pair0 = p0; P.npairs = min(chunk, row_end - p0);
    const size_t ninc = (size_t)P.npairs * (L - 1) * (L - 1);
    pde_vjp_inc_kernel<<<(unsigned)((ninc + 255) / 256), 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_inc_kernel");
    prof_begin(s);
    pde_vjp_sweep_kernel<0><<<(unsigned)P.npairs, 32 * kVjpWarps, smem, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_sweep_kernel<0>");
    pde_vjp_sweep_kernel<1><<<(unsigned)P.npairs, 32 * kVjpWarps, smem, s>>>(P);
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
