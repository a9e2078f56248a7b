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
  double* DX;             // [chunk, L-1, d] dx_J
  double* grid;           // [chunk, bands, P/CS + 31, CS, 32] reversed-path grid (node columns 1..P) in sweep order
  double* rowbuf;         // [chunk, kVjpWarps, 32 * bands]
  int L, d, order, P;
  int pair0, npairs;      // sequences [pair0, pair0 + npairs) of this launch
  int row_begin;
  int rev_solver;
};

// DX[a][J][k] = x[J+1][k] - x[J][k]
__global__ void __launch_bounds__(256) pde_vjp_dx_kernel(const PdeVjpParams P) {
  const int Lm = P.L - 1;
  const size_t per = (size_t)Lm * P.d;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= per * P.npairs) return;
  const int a = (int)(idx / per);
  const size_t e = idx % per;
  const double* x = P.X + (size_t)(P.pair0 + a) * P.L * P.d;
  P.DX[idx] = x[e + P.d] - x[e];
}

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

constexpr int kVjpWarps = 8;      // warps per CTA = bands of one sequence in flight

__host__ __device__ inline int vjp_cs(int order) { return order >= 2 ? 4 : (1 << order); }   // node columns per step
__host__ __device__ inline size_t vjp_bands(size_t Pf) { return (Pf + 31) / 32; }
__host__ __device__ inline size_t vjp_steps(size_t Pf, int cs) { return Pf / cs + 31; }      // steps of one band
__host__ __device__ inline size_t vjp_grid_elems(size_t Pf, int cs) { return vjp_bands(Pf) * vjp_steps(Pf, cs) * cs * 32; }
__host__ __device__ inline size_t vjp_rowbuf_elems(size_t Pf) { return (size_t)kVjpWarps * vjp_bands(Pf) * 32; }
// dynamic shared memory: progress counters [bands] (padded to 16 B) + the sequence's dx [L-1][d] when it fits
__host__ __device__ inline size_t vjp_prog_bytes(size_t Pf) { return (vjp_bands(Pf) * sizeof(int) + 15) & ~(size_t)15; }
constexpr size_t kVjpDxSmemMax = 48 * 1024;
constexpr size_t kVjpRbsBytes = (size_t)kVjpWarps * 2 * 32 * sizeof(double);   // per-warp double-buffered 32-entry window of the row above

template <int CS> __device__ __forceinline__ void vjp_load(double (&o)[CS], const double* p) {
  if constexpr (CS == 1) o[0] = __ldg(p);
  else {
#pragma unroll
    for (int h = 0; h < CS / 2; ++h) { const double2 t = __ldg(reinterpret_cast<const double2*>(p) + h); o[2 * h] = t.x; o[2 * h + 1] = t.y; }
  }
}
template <int CS> __device__ __forceinline__ void vjp_store(double* p, const double (&o)[CS]) {
  if constexpr (CS == 1) p[0] = o[0];
  else {
#pragma unroll
    for (int h = 0; h < CS / 2; ++h) reinterpret_cast<double2*>(p)[h] = make_double2(o[2 * h], o[2 * h + 1]);
  }
}

// One CTA per sequence, kVjpWarps warps; warp w sweeps the 32-row bands w, w + W, ... of the grid.  Lane = one
// node row r (0 = the boundary row of ones); a STEP advances a lane by CS node columns (CS = 2^order up to 4: one
// coarse column of cells, so the stencil coefficients and the flush of the q-reduction happen at the same step
// for every lane: no divergence), lanes skewed by one step: lane t is at column group s - t.
// `up` (the CS values of the row above) comes from the lane above by shuffles; the last row of band b reaches band
// b+1 (the next warp) through a row buffer in global memory, published in blocks of 32 columns through a per-band
// progress counter in shared memory, so all bands of a sequence run as a pipeline.  The increments
// <dx_I, dx_J> / 4^n are formed on the fly from the sequence's dx, staged once per CTA in shared memory.
// Node columns 1..P are swept (group g = nodes g CS + 1 .. g CS + CS = cells g CS .. g CS + CS - 1); column 0 is 1.
// MODE 0 solves the grid of the time-reversed path and stores it; MODE 1 solves the forward grid, multiplies
// each value by the stored reversed value of the mirrored node and reduces.  The scratch is laid out in the
// ORDER THE SWEEP PRODUCES IT, [band][step][lane][CS]: a step is one coalesced 32 CS x 8-byte store / load
// (128-bit per lane).  The two sweeps meet when MODE 1 aligns its bands to the END of the grid (rows r = 32 band
// + lane - pad): what it needs at (band, step s, lane t) was written by MODE 0's (NB-1-band, ~NG+30-s, 31-t).
template <int MODE, int CS, int DA, bool AL>
__global__ void __launch_bounds__(32 * kVjpWarps, MODE == 0 ? (DA <= 4 ? 3 : (DA <= 8 ? 2 : 1)) : (DA <= 4 ? 2 : 1))
pde_vjp_sweep_kernel(const PdeVjpParams P, const int dx_in_smem) {
  extern __shared__ __align__(16) unsigned char vjp_smem[];
  int* prog = reinterpret_cast<int*>(vjp_smem);            // [NB] row-buffer entries of band b published so far
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a = blockIdx.x;
  // AL: a step is exactly one coarse column of cells (order <= 2, CS = 2^order): n and c are compile-time
  const int n = AL ? (CS == 4 ? 2 : (CS == 2 ? 1 : 0)) : P.order, c = AL ? CS : (1 << n);
  const int Pf = P.P, Lm = P.L - 1, d = P.d;
  const int NB = (int)vjp_bands(Pf), NG = Pf / CS, NS = NG + 31;
  const double* dxs = P.DX + (size_t)a * Lm * d;
  for (int i = threadIdx.x; i < NB; i += blockDim.x) prog[i] = 0;
  double* rbs = reinterpret_cast<double*>(vjp_smem + vjp_prog_bytes(Pf)) + (size_t)warp * 64;
  if (dx_in_smem) {
    double* sdx = reinterpret_cast<double*>(vjp_smem + vjp_prog_bytes(Pf) + kVjpRbsBytes);
    for (int i = threadIdx.x; i < Lm * d; i += blockDim.x) sdx[i] = dxs[i];
    dxs = sdx;
  }
  __syncthreads();
  volatile int* vprog = prog;
  double* S = P.grid + (size_t)a * vjp_grid_elems(Pf, CS);
  double* rowbufs = P.rowbuf + (size_t)a * vjp_rowbuf_elems(Pf);
  const size_t Pfp = (size_t)NB * 32;
  const bool second = MODE == 1 || P.rev_solver == 1;
  const int roff = MODE == 1 ? NB * 32 - Pf : 0;
  const double fscale = 1.0 / (double)(1u << (2 * n));
  constexpr bool aligned = AL;
  const int jstep = (MODE == 0 ? -1 : 1) * d;              // MODE 0: inc_rev[I][J] = inc[L-2-I][L-2-J] (pyx:43)

  for (int band = warp; band < NB; band += kVjpWarps) {
    const int r = band * 32 + lane - roff;
    const bool row_ok = r >= 0 && r < Pf;
    const int rc = min(max(r, 0), Pf - 1);
    const int Ic = rc > 0 ? ((rc - 1) >> n) : 0;           // coarse row of the cells (r-1, .) this lane solves
    double dxI[DA];                                        // ... its dx, pre-scaled by 1/4^n (exact)
#pragma unroll
    for (int k = 0; k < DA; ++k) dxI[k] = k < d ? dxs[(size_t)(MODE == 0 ? Lm - 1 - Ic : Ic) * d + k] * fscale : 0.;
    const double* rb_in = rowbufs + (size_t)((band + kVjpWarps - 1) % kVjpWarps) * Pfp;
    double* rb_out = rowbufs + (size_t)(band % kVjpWarps) * Pfp;
    // scratch in units of one lane-step (CS doubles): MODE 0 writes unit (band NS + s) 32 + lane.  MODE 1 needs, for
    // node q = g CS + 1 + k of its row, the reversed grid's node b = P - 1 - q of row P - 1 - r, which MODE 0 wrote
    // from lane 31 - t of band NB-1-band as element (b-1) % CS of its group (b-1) / CS.  b - 1 = (NG-1-g) CS +
    // (CS-3-k): the reader's group g straddles writer groups h(g) = NG-1-g and h(g+1) (CS = 4), or is h(g+1)
    // (CS = 2) or h(g+2) (CS = 1), reversed inside -- a window that slides by one unit per step.
    double* pW = S + ((size_t)band * NS * 32 + lane) * CS;
    auto unit_ptr = [&](int gp) -> const double* {         // writer group gp of writer lane 31 - t
      return S + (((size_t)(NB - 1 - band) * NS + (size_t)(gp + 31 - lane)) * 32 + (31 - lane)) * CS;
    };
    auto fetch = [&](double (&o)[CS], int gg) {            // writer group h(gg) (zeros outside the grid)
      const int gp = NG - 1 - gg;
#pragma unroll
      for (int k = 0; k < CS; ++k) o[k] = 0.;
      if (row_ok && gp >= 0 && gp < NG) vjp_load<CS>(o, unit_ptr(gp));
    };
    double gq[4][CS];                                      // ring: gq[(u + j) & 3] = h(g + j) at unrolled step u
    if (MODE == 1) {
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) fetch(gq[jj], -lane + jj);
    }
    double acc[DA];
#pragma unroll
    for (int k = 0; k < DA; ++k) acc[k] = 0.;
    double sJ = 0.;
    // MODE 1, node column 0: K = 1 times the reversed grid's node column P - 1
    if (MODE == 1 && row_ok) sJ = Pf == 1 ? 1. : __ldg(unit_ptr((Pf - 2) / CS) + (Pf - 2) % CS);
    double cur = 1., upl = 1.;                             // left neighbour (r, q-1); up-left (r-1, q-1)
    double vs[CS];
#pragma unroll
    for (int k = 0; k < CS; ++k) vs[k] = 1.;
    double A = 1., B = 1.;
    const double* pJ = dxs + (size_t)(MODE == 0 ? Lm - 1 : 0) * d;    // dx of the next coarse column to enter
    double dxJ[DA], dxJn[DA];
#pragma unroll
    for (int k = 0; k < DA; ++k) { dxJ[k] = 0.; dxJn[k] = k < d ? pJ[k] : 0.; }
    double rbnext = 1.;
    if (band > 0) {
      while (vprog[band - 1] < min(32, Pf)) __nanosleep(100);
      __threadfence_block();
      rbnext = __ldcg(rb_in + lane);
    }
    const double* pRB = rb_in + 32 + lane;                 // next block of the row above
    double* pRO = rb_out - (size_t)lane * CS;              // lane 31: rb_out + g CS
    int g = -lane;                                         // column group of this lane at step s
    for (int s0 = 0; s0 < NS; s0 += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u, ++g) {
        const int s = s0 + u;
        if (((s * CS) & 31) == 0 && band > 0) {            // next 32 entries of the row above this band -> this warp's window
          __syncwarp();
          rbs[(((s * CS) >> 5) & 1) * 32 + lane] = rbnext;
          __syncwarp();
          if (s * CS + 32 < Pf) {
            const int need = min(s * CS + 64, Pf);
            while (vprog[band - 1] < need) __nanosleep(100);
            __threadfence_block();
            rbnext = __ldcg(pRB);
            pRB += 32;
          }
        }
        double gv[CS];
        if (MODE == 1) {                                   // stored values of this step; refill the ring slot 4 steps ahead
          if (CS == 4) { gv[0] = gq[u][1]; gv[1] = gq[u][0]; gv[CS > 2 ? 2 : 0] = gq[(u + 1) & 3][CS > 3 ? 3 : 0]; gv[CS > 3 ? 3 : 0] = gq[(u + 1) & 3][CS > 2 ? 2 : 0]; }
          if (CS == 2) { gv[0] = gq[(u + 1) & 3][CS > 1 ? 1 : 0]; gv[CS > 1 ? 1 : 0] = gq[(u + 1) & 3][0]; }
          if (CS == 1) gv[0] = gq[(u + 2) & 3][0];
          fetch(gq[u], g + 4);
#pragma unroll
          for (int k = 0; k < CS; ++k) {
            const int b = Pf - 2 - g * CS - k;             // = P - 1 - q: 0 = the reversed grid's boundary column, < 0 outside
            if (b == 0) gv[k] = 1.;
            if (b < 0) gv[k] = 0.;
          }
        }
        double up[CS];
#pragma unroll
        for (int k = 0; k < CS; ++k) up[k] = shfl_up1(vs[k]);   // (r-1, q_k): finished by lane-1 one step ago
        if (lane == 0) {                                   // first row of the band: from the band above (or the boundary)
          if (band > 0) {
            const double* w = rbs + ((s * CS) & 63);
            if constexpr (CS == 1) up[0] = w[0];
            else {
#pragma unroll
              for (int h = 0; h < CS / 2; ++h) { const double2 t = reinterpret_cast<const double2*>(w)[h]; up[2 * h] = t.x; up[2 * h + 1] = t.y; }
            }
          } else {
#pragma unroll
            for (int k = 0; k < CS; ++k) up[k] = 1.;
          }
        }
        const bool active = row_ok && (unsigned)g < (unsigned)NG;
        if (active) {
          if (aligned || ((g * CS) & (c - 1)) == 0) {      // cells of a new coarse column J = (g CS) >> n
            double e = 0.;
#pragma unroll
            for (int k = 0; k < DA; ++k) { dxJ[k] = dxJn[k]; e = fma(dxI[k], dxJ[k], e); }
            pJ += jstep;
            if (((g * CS) >> n) + 1 < Lm) {
#pragma unroll
              for (int k = 0; k < DA; ++k) if (k < d) dxJn[k] = pJ[k];
            }
            if (second) { const double e2 = e * e * (1. / 12); A = 1. + 0.5 * e + e2; B = 1. - e2; }
            else { A = 1.; B = 1. - e; }
            if (r == 0) { A = 0.; B = -1.; }               // boundary row: v = up-left = 1
          }
          double ul = upl, left = cur;
#pragma unroll
          for (int k = 0; k < CS; ++k) {
            const double v = fma(left, A, fma(up[k], A, -(ul * B)));   // (k01 + k10) A - k00 B
            vs[k] = v; left = v; ul = up[k];
          }
          cur = left; upl = ul;
          if (MODE == 0) vjp_store<CS>(pW, vs);
          if (MODE == 1) {
            // q-reduction: node q multiplies dx[q >> n]; the group's last node may open the next coarse column
#pragma unroll
            for (int k = 0; k < CS - 1; ++k) sJ = fma(vs[k], gv[k], sJ);
            const double last = vs[CS - 1] * gv[CS - 1];
            if (aligned || ((g * CS + CS) & (c - 1)) == 0) {
#pragma unroll
              for (int k = 0; k < DA; ++k) acc[k] = fma(sJ, dxJ[k], acc[k]);
              sJ = last;
            } else sJ += last;
          }
          if (lane == 31) {                                // top row of the next band
#pragma unroll
            for (int k = 0; k < CS; ++k) __stcg(pRO + k, vs[k]);
            const int done = g * CS + CS;
            if ((done & 31) == 0 || done == Pf) { __threadfence_block(); vprog[band] = done; }
          }
        }
        pW += CS * 32;
        pRO += CS;
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
  return (vjp_grid_elems(Pf, vjp_cs(order)) + (size_t)(L - 1) * d + vjp_rowbuf_elems(Pf)) * sizeof(double);
}

template <int MODE, int CS, int DA, bool AL>
static int launch_vjp_sweep(const PdeVjpParams& P, cudaStream_t s) {
  auto kern = pde_vjp_sweep_kernel<MODE, CS, DA, AL>;
  const size_t dxb = (size_t)(P.L - 1) * P.d * sizeof(double);
  const int dx_in_smem = dxb <= kVjpDxSmemMax;
  const size_t smem = vjp_prog_bytes((size_t)P.P) + kVjpRbsBytes + (dx_in_smem ? dxb : 0);
  if (smem > 48 * 1024) GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)P.npairs, 32 * kVjpWarps, smem, s>>>(P, dx_in_smem);
  GPSIG_LAUNCH_CHECK("pde_vjp_sweep_kernel");
  return GPSIG_OK;
}
template <int MODE>
static int dispatch_vjp_sweep(const PdeVjpParams& P, cudaStream_t s) {
  const int cs = vjp_cs(P.order);
#define GPSIG_VJP_GO(CS_, AL_) (P.d <= 4 ? launch_vjp_sweep<MODE, CS_, 4, AL_>(P, s) : P.d <= 8 ? launch_vjp_sweep<MODE, CS_, 8, AL_>(P, s) \
                                                                                             : launch_vjp_sweep<MODE, CS_, 16, AL_>(P, s))
  return cs == 1 ? GPSIG_VJP_GO(1, true) : cs == 2 ? GPSIG_VJP_GO(2, true) : (P.order == 2 ? GPSIG_VJP_GO(4, true) : GPSIG_VJP_GO(4, false));
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
  if (vjp_prog_bytes(Pf) > 128 * 1024) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "grid too long for the PDE backward (P=%zu)", Pf);
  GPSIG_CUDA_CHECK(cudaMemsetAsync(gradX + (size_t)row_begin * L * d, 0, (size_t)rows * L * d * sizeof(double), s));
  const size_t per = pde_vjp_pair_bytes(L, d, order);
  if (!ws || ws_bytes < per || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need at least %zu bytes (one sequence), got %zu", per, ws_bytes);
  const int chunk = (int)min((size_t)rows, ws_bytes / per);
  PdeVjpParams P;
  memset(&P, 0, sizeof(P));
  P.X = X; P.gout = gout; P.gradX = gradX;
  P.grid = static_cast<double*>(ws);
  P.DX = P.grid + (size_t)chunk * vjp_grid_elems(Pf, cs);
  P.rowbuf = P.DX + (size_t)chunk * (L - 1) * d;
  P.L = L; P.d = d; P.order = order; P.P = (int)Pf; P.row_begin = row_begin; P.rev_solver = rev_solver;
  for (int p0 = row_begin; p0 < row_end; p0 += chunk) {
    P.pair0 = p0; P.npairs = min(chunk, row_end - p0);
    const size_t ndx = (size_t)P.npairs * (L - 1) * d;
    pde_vjp_dx_kernel<<<(unsigned)((ndx + 255) / 256), 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("pde_vjp_dx_kernel");
    prof_begin(s);
    if (int rc = dispatch_vjp_sweep<0>(P, s)) return rc;
    if (int rc = dispatch_vjp_sweep<1>(P, s)) return rc;
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
