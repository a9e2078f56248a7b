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
  double* DX;             // [chunk, L-1, d] dx_J
  double* grid;           // [chunk, bands, P/CS + 31, CS, 32] reversed-path grid (node columns 1..P) in sweep order
  double* rowbuf;         // [chunk, kVjpWarps, 32 * bands]
  int L, d, order, P;
  int pair0, npairs;      // sequences [pair0, pair0 + npairs) of this launch
  int row_begin;
  int rev_solver;
};

// E[I][J] in the accumulation order of sigKer_fast.pyx:41-43 (inc += dx*dy/factor, k ascending); row I = L-1 of the
// launch writes DX[J][k] = x[J+1][k] - x[J][k] instead
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
  if (I == 0)
    for (int k = 0; k < P.d; ++k) P.DX[((size_t)a * Lm + J) * P.d + k] = x[(J + 1) * P.d + k] - x[J * P.d + k];
}

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

constexpr int kVjpWarps = 8;      // warps per CTA = bands of one sequence in flight
constexpr int kVjpUnroll = 2;     // steps per prefetch group of the stored grid

__host__ __device__ inline int vjp_cs(int order) { return order >= 2 ? 4 : (1 << order); }   // node columns per step
__host__ __device__ inline size_t vjp_bands(size_t Pf) { return (Pf + 31) / 32; }
__host__ __device__ inline size_t vjp_steps(size_t Pf, int cs) { return Pf / cs + 31; }      // steps of one band
__host__ __device__ inline size_t vjp_grid_elems(size_t Pf, int cs) { return vjp_bands(Pf) * vjp_steps(Pf, cs) * cs * 32; }
__host__ __device__ inline size_t vjp_rowbuf_elems(size_t Pf) { return (size_t)kVjpWarps * vjp_bands(Pf) * 32; }

// One CTA per sequence, kVjpWarps warps; warp w sweeps the 32-row bands w, w + W, ... of the grid.  Lane = one
// node row r (0 = the boundary row of ones); a STEP advances a lane by CS node columns (CS = 2^order up to 4: one
// coarse column of cells, so the stencil coefficients, the increment load and the flush of the q-reduction happen
// at the same step for every lane: no divergence), lanes skewed by one step: lane t is at column group s - t.
// `up` (the CS values of the row above) comes from the lane above by shuffles; the last row of band b reaches band
// b+1 (the next warp) through a row buffer in global memory, published in blocks of 32 columns through a per-band
// progress counter in shared memory, so all bands of a sequence run as a pipeline.
// Node columns 1..P are swept (group g = nodes g CS + 1 .. g CS + CS = cells g CS .. g CS + CS - 1); column 0 is 1.
// MODE 0 solves the grid of the time-reversed path and stores it; MODE 1 solves the forward grid, multiplies
// each value by the stored reversed value of the mirrored node and reduces.  The scratch is laid out in the
// ORDER THE SWEEP PRODUCES IT, [band][step][k][lane]: every value moves in coalesced 256-byte rows.  The two
// sweeps meet when MODE 1 aligns its bands to the END of the grid (rows r = 32 band + lane - pad): the value it
// needs at (band, step s, lane t) was written by MODE 0's (NB-1-band, ~NG+30-s, 31-t).
template <int MODE, int CS, int DA>
__global__ void __launch_bounds__(32 * kVjpWarps) pde_vjp_sweep_kernel(const PdeVjpParams P) {
  extern __shared__ int prog[];                            // [NB] row-buffer entries of band b published so far
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a = blockIdx.x;
  const int n = P.order, c = 1 << n, Pf = P.P, Lm = P.L - 1, d = P.d;
  constexpr int LCS = CS == 4 ? 2 : (CS == 2 ? 1 : 0);
  const int NB = (int)vjp_bands(Pf), NG = Pf / CS, NS = NG + 31;
  for (int i = threadIdx.x; i < NB; i += blockDim.x) prog[i] = 0;
  __syncthreads();
  volatile int* vprog = prog;
  const double* E = P.E + (size_t)a * Lm * Lm;
  const double* DX = P.DX + (size_t)a * Lm * d;
  double* S = P.grid + (size_t)a * vjp_grid_elems(Pf, CS);
  double* rowbufs = P.rowbuf + (size_t)a * vjp_rowbuf_elems(Pf);
  const size_t Pfp = (size_t)NB * 32;
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
    double* Sw = S + (size_t)band * NS * (CS * 32) + lane;                       // MODE 0: + (s CS + k) 32
    const double* Sr = S + (size_t)(NB - 1 - band) * NS * (CS * 32) + (31 - lane);
    // MODE 1: stored value of the mirrored node of node column q of this lane's row (1 on its boundary, 0 outside)
    auto rev_at = [&](int q) -> double {
      const int b = Pf - 1 - q;                            // node column of the reversed grid
      if (b <= 0) return b == 0 ? 1. : 0.;
      const int gp = (b - 1) >> LCS, kp = (b - 1) & (CS - 1);
      return __ldg(Sr + ((size_t)(gp + 31 - lane) * CS + kp) * 32);
    };
    double acc[DA];
#pragma unroll
    for (int k = 0; k < DA; ++k) acc[k] = 0.;
    double sJ = 0.;
    if (MODE == 1 && row_ok) sJ = rev_at(0);               // node column 0: K = 1
    double cur = 1., upl = 1.;                             // left neighbour (r, q-1); up-left (r-1, q-1)
    double vs[CS];
#pragma unroll
    for (int k = 0; k < CS; ++k) vs[k] = 1.;
    double A = 1., B = 1.;
    double e_next = Erow[MODE == 0 ? (Lm - 1) : 0];
    double rbcur = 1., rbnext = 1.;
    if (band > 0) {
      while (vprog[band - 1] < min(32, Pf)) { }
      __threadfence_block();
      rbnext = __ldcg(rb_in + lane);
    }
    double gcur[kVjpUnroll][CS], gnext[kVjpUnroll][CS];
    if (MODE == 1) {
#pragma unroll
      for (int u = 0; u < kVjpUnroll; ++u)
#pragma unroll
        for (int k = 0; k < CS; ++k) {
          const int g2 = u - lane;
          gnext[u][k] = (row_ok && g2 >= 0 && g2 < NG) ? rev_at(g2 * CS + 1 + k) : 0.;
          gcur[u][k] = 0.;
        }
    }
    for (int s0 = 0; s0 < NS; s0 += kVjpUnroll) {
      if (MODE == 1) {
#pragma unroll
        for (int u = 0; u < kVjpUnroll; ++u)
#pragma unroll
          for (int k = 0; k < CS; ++k) {
            gcur[u][k] = gnext[u][k];
            const int g2 = s0 + kVjpUnroll + u - lane;
            gnext[u][k] = (row_ok && g2 >= 0 && g2 < NG) ? rev_at(g2 * CS + 1 + k) : 0.;
          }
      }
#pragma unroll
      for (int u = 0; u < kVjpUnroll; ++u) {
        const int s = s0 + u;
        const int g = s - lane;                            // column group of this lane
        if (((s * CS) & 31) == 0 && band > 0) {            // next 32 entries of the row above this band
          rbcur = rbnext;
          if (s * CS + 32 < Pf) {
            const int need = min(s * CS + 64, Pf);
            while (vprog[band - 1] < need) { }
            __threadfence_block();
            rbnext = __ldcg(rb_in + s * CS + 32 + lane);
          }
        }
        double up[CS];
#pragma unroll
        for (int k = 0; k < CS; ++k) {
          up[k] = shfl_up1(vs[k]);                         // (r-1, q_k): finished by lane-1 one step ago
          const double u0 = __shfl_sync(0xffffffffu, rbcur, (s * CS + k) & 31);
          if (lane == 0) up[k] = band > 0 ? u0 : 1.;
        }
        const bool active = row_ok && g >= 0 && g < NG;
        if (active) {
          if (((g * CS) & (c - 1)) == 0) {                 // cells of a new coarse column J = (g CS) >> n
            const double e = e_next;
            const int Jn = ((g * CS) >> n) + 1;
            if (Jn < Lm) e_next = Erow[MODE == 0 ? (Lm - 1 - Jn) : Jn];
            if (second) { const double e2 = e * e * (1. / 12); A = 1. + 0.5 * e + e2; B = 1. - e2; }
            else { A = 1.; B = 1. - e; }
          }
          double ul = upl, left = cur;
#pragma unroll
          for (int k = 0; k < CS; ++k) {
            const double v = r == 0 ? 1. : fma(left, A, fma(up[k], A, -(ul * B)));   // (k01 + k10) A - k00 B
            vs[k] = v; left = v; ul = up[k];
            if (MODE == 0) Sw[(size_t)(s * CS + k) * 32] = v;
          }
          cur = left; upl = ul;
          if (MODE == 1) {
            // q-reduction: node q multiplies dx[q >> n]; the group's last node may open the next coarse column
#pragma unroll
            for (int k = 0; k < CS - 1; ++k) sJ = fma(vs[k], gcur[u][k], sJ);
            const double last = vs[CS - 1] * gcur[u][CS - 1];
            if (((g * CS + CS) & (c - 1)) == 0) {
              const double* dxa = DX + (size_t)((g * CS + CS - 1) >> n) * d;
#pragma unroll
              for (int k = 0; k < DA; ++k)
                if (k < d) acc[k] = fma(sJ, __ldg(dxa + k), acc[k]);
              sJ = last;
            } else sJ += last;
          }
          if (lane == 31) {                                // top row of the next band
#pragma unroll
            for (int k = 0; k < CS; ++k) __stcg(rb_out + g * CS + k, vs[k]);
            const int done = g * CS + CS;
            if ((done & 31) == 0 || done == Pf) { __threadfence_block(); vprog[band] = done; }
          }
        }
      }
    }
    if (MODE == 1 && row_ok) {
      // K_grad fine row r -> coarse row i = r >> n, scaled by 1/c (row average) * 1/c (dx on the fine grid),
      // grad_x[i] -= 2 v, grad_x[i+1] += 2 v (kernels_pde.py:508), times the upstream gradient
      const double w = 2. * P.gout[P.pair0 + a - P.row_begin] / ((double)c * (double)c);
      double* gx = P.gradX + ((size_t)(P.pair0 + a) * P.L + (r >> n)) * d;
#pragma unroll
      for (int k = 0; k < DA; ++k)
        if (k < d) { const double t = w * acc[k]; atomicAdd(gx + k, -t); atomicAdd(gx + d + k, t); }
    }
  }
}

static size_t pde_vjp_pair_bytes(int L, int d, int order) {
  const size_t Pf = (size_t)(L - 1) << order;
  return (vjp_grid_elems(Pf, vjp_cs(order)) + (size_t)(L - 1) * (L - 1) + (size_t)(L - 1) * d + vjp_rowbuf_elems(Pf)) * sizeof(double);
}

template <int MODE, int CS, int DA>
static int launch_vjp_sweep(const PdeVjpParams& P, size_t smem, cudaStream_t s) {
  auto kern = pde_vjp_sweep_kernel<MODE, CS, DA>;
  if (smem > 48 * 1024) GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)P.npairs, 32 * kVjpWarps, smem, s>>>(P);
  GPSIG_LAUNCH_CHECK("pde_vjp_sweep_kernel");
  return GPSIG_OK;
}
template <int MODE>
static int dispatch_vjp_sweep(const PdeVjpParams& P, size_t smem, cudaStream_t s) {
  const int cs = vjp_cs(P.order);
#define GPSIG_VJP_GO(CS_) (P.d <= 4 ? launch_vjp_sweep<MODE, CS_, 4>(P, smem, s) : P.d <= 8 ? launch_vjp_sweep<MODE, CS_, 8>(P, smem, s) \
                                                                                         : launch_vjp_sweep<MODE, CS_, 16>(P, smem, s))
  if constexpr (MODE == 0)
    return cs == 1 ? launch_vjp_sweep<0, 1, 4>(P, smem, s) : cs == 2 ? launch_vjp_sweep<0, 2, 4>(P, smem, s) : launch_vjp_sweep<0, 4, 4>(P, smem, s);
  else
    return cs == 1 ? GPSIG_VJP_GO(1) : cs == 2 ? GPSIG_VJP_GO(2) : GPSIG_VJP_GO(4);
#undef GPSIG_VJP_GO
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
  const int cs = vjp_cs(order);
  const size_t smem = vjp_bands(Pf) * sizeof(int);
  if (smem > 200 * 1024) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "grid too long for the PDE backward (P=%zu)", Pf);
  GPSIG_CUDA_CHECK(cudaMemsetAsync(gradX + (size_t)row_begin * L * d, 0, (size_t)rows * L * d * sizeof(double), s));
  const size_t per = pde_vjp_pair_bytes(L, d, order);
  if (!ws || ws_bytes < per || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need at least %zu bytes (one sequence), got %zu", per, ws_bytes);
  const int chunk = (int)min((size_t)rows, ws_bytes / per);
  PdeVjpParams P;
  memset(&P, 0, sizeof(P));
  P.X = X; P.gout = gout; P.gradX = gradX;
  P.grid = static_cast<double*>(ws);
  P.E = P.grid + (size_t)chunk * vjp_grid_elems(Pf, cs);
  P.DX = P.E + (size_t)chunk * (L - 1) * (L - 1);
  P.rowbuf = P.DX + (size_t)chunk * (L - 1) * d;
  P.L = L; P.d = d; P.order = order; P.P = (int)Pf; P.row_begin = row_begin; P.rev_solver = rev_solver;
  for (int p0 = row_begin; p0 < row_end; p0 += chunk) {
    P.pair0 = p0; P.npairs = min(chunk, row_end - p0);
    const size_t ninc = (size_t)P.npairs * (L - 1) * (L - 1);
    pde_vjp_inc_kernel<<<(unsigned)((ninc + 255) / 256), 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_inc_kernel");
    prof_begin(s);
    if (int rc = dispatch_vjp_sweep<0>(P, smem, s)) return rc;
    if (int rc = dispatch_vjp_sweep<1>(P, smem, s)) return rc;
    prof_end(s);
  }
  return GPSIG_OK;
}

}  // namespace gpsig

extern "C" {

size_t gpsig_pde_kdiag_vjp_workspace_bytes(int nseq, int L, int d, int dyadic_order) {
  if (nseq < 1 || L < 2 || d < 1 || dyadic_order < 0) return 0;
  return (size_t)nseq * gpsig::pde_vjp_pair_bytes(L, d, dyadic_order);
}

int gpsig_pde_kdiag_vjp_f64(const double* X, int N, int L, int d, int dyadic_order, int rev_solver, int row_begin,
                            int row_end, const double* grad_out, double* gradX, void* workspace,
                            size_t workspace_bytes, void* stream) {
  return gpsig::pde_kdiag_vjp_impl(X, N, L, d, dyadic_order, rev_solver, row_begin, row_end, grad_out, gradX, workspace,
                                   workspace_bytes, stream);
}

}  // extern "C"
