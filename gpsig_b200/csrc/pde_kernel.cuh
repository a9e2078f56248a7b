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


// ---------------------------------------------------------------------------
// Pipelined variant: one CTA per pair, one WARP PER STRIP.
//
// The sequential-strip kernel above leaves a pair's 2P-long dependency chain on
// a single warp (c3: 8 strips x 2028 steps, 1000 pairs = 1000 warps on 148 SMs).
// Here the strips of a pair run CONCURRENTLY as a second level of the same
// systolic skew: warp w owns strip w and trails warp w-1 by one lane-pipeline
// depth (32 steps) plus one hand-off window.  The right edge of strip w-1 flows
// to warp w through a small shared-memory ring (kPdeRing rows) with two
// monotonic progress counters per boundary (rows published / rows consumed),
// exchanged every kPdeWin steps -- no block-wide barrier anywhere.  A pair's
// critical path drops from strips*(P+32) to P + strips*(32+window) steps and
// the CTA keeps `strips` warps busy, so the FP64 pipe, not latency, bounds it.
// ---------------------------------------------------------------------------
constexpr int kPdeRing = 128;   // ring rows per strip boundary (power of two, >= 2*kPdeWin + 32)
constexpr int kPdeWin = 16;     // steps between progress exchanges
constexpr int kPdeMaxStrips = 16;

// dx rows are padded to an odd number of 16-byte units (skewed row reads of a quarter warp hit distinct
// banks).  The dy coefficients a lane needs never change (its columns are fixed): they live in registers
// when they fit (coarse columns per lane x DP <= 32 doubles), else in shared memory padded per lane.
__host__ __device__ inline int pde_coarse_cols(int order) { return (kPdeW >> order) > 0 ? (kPdeW >> order) : 1; }
__host__ __device__ inline bool pde_dy_in_regs(int DP, int order) { return pde_coarse_cols(order) * DP <= 32; }
__host__ __device__ inline size_t pde_dy_smem_elems(int L2, int DP, int order) {
  if (pde_dy_in_regs(DP, order)) return 0;
  const int nc = pde_coarse_cols(order);
  return (size_t)(L2 - 1) * DP + 2 * (size_t)((L2 - 1 + nc - 1) / nc) + 2;
}
inline size_t pde_pipe_smem_bytes(int L, int L2, int DP, int order, int nstrips) {
  return ((size_t)(L - 1) * (DP + 2) + pde_dy_smem_elems(L2, DP, order) + (size_t)nstrips * kPdeRing) * sizeof(double) +
         2 * kPdeMaxStrips * sizeof(int);
}

// keeps a loop-invariant value in a register (ptxas otherwise re-derives shared-window addresses at every use)
__device__ __forceinline__ uint32_t opaque_u32(uint32_t v) {
  uint32_t r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}
__device__ __forceinline__ int ld_volatile_shared_u(uint32_t a) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ int ld_volatile_shared(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_shared(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

template <int DP, int ORDER>
__global__ void __launch_bounds__(kPdeMaxStrips * 32) pde_pipe_kernel(const PdeParams Pm) {
  constexpr int W = kPdeW;
  constexpr int NC = (W >> ORDER) > 0 ? (W >> ORDER) : 1;   // coarse columns per lane
  constexpr int RB = kPdeRing, G = kPdeWin;
  extern __shared__ __align__(128) unsigned char pde_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, NS = blockDim.x >> 5;
  const int L = Pm.L, L2 = Pm.L2;
  const int P = (L - 1) << ORDER, Q = (L2 - 1) << ORDER;
  constexpr int XS = DP + 2;                                // padded dx row stride (doubles)
  constexpr bool DYREG = NC * DP <= 32;
  double* dx = reinterpret_cast<double*>(pde_smem);
  double* dy = dx + (size_t)(L - 1) * XS;                  // only if !DYREG: element (J, k) at J*DP + k + 2*(J/NC)
  double* ring = dy + pde_dy_smem_elems(L2, DP, ORDER);    // [NS][RB]: ring w = left boundary of strip w
  int* prod = reinterpret_cast<int*>(ring + (size_t)NS * RB);   // prod[w]: rows of boundary w published by warp w-1
  int* cons = prod + kPdeMaxStrips;                               // cons[w]: rows of boundary w consumed by warp w

  int ix, iy;
  if (Pm.pair_mode == GPSIG_PDE_DIAG) { ix = Pm.row_begin + blockIdx.x; iy = ix; }
  else { ix = Pm.row_begin + blockIdx.x / Pm.N2; iy = blockIdx.x % Pm.N2; }
  const double* x = static_cast<const double*>(Pm.xp) + (size_t)ix * Pm.xblk;
  const double* y = static_cast<const double*>(Pm.yp) + (size_t)iy * Pm.yblk;

  const double fscale = 1.0 / (double)(1 << (2 * ORDER));
  for (int e = threadIdx.x; e < (L - 1) * DP; e += blockDim.x) {
    const int I = e / DP, k = e - I * DP;
    dx[I * XS + k] = (x[e + DP] - x[e]) * fscale;
  }
  if (!DYREG)
    for (int e = threadIdx.x; e < (L2 - 1) * DP; e += blockDim.x) dy[e + 2 * ((e / DP) / NC)] = y[e + DP] - y[e];
  if (threadIdx.x < kPdeMaxStrips) { prod[threadIdx.x] = 0; cons[threadIdx.x] = 0; }
  __syncthreads();                                          // the only block-wide barrier

  const int j0 = w * 32 * W + lane * W;                     // first fine column (cell index) of this lane
  double2 dyreg[DYREG ? NC : 1][DYREG ? DP / 2 : 1];
  if (DYREG) {
#pragma unroll
    for (int cc = 0; cc < NC; ++cc) {
      const int J = min((j0 >> ORDER) + cc, L2 - 2);
#pragma unroll
      for (int k = 0; k < DP / 2; ++k)
        dyreg[DYREG ? cc : 0][DYREG ? k : 0] = make_double2(y[(J + 1) * DP + 2 * k] - y[J * DP + 2 * k],
                                                              y[(J + 1) * DP + 2 * k + 1] - y[J * DP + 2 * k + 1]);
    }
  }
  const int ncell = max(0, min(W, Q - j0));
  const int LL = min(31, (Q - w * 32 * W - 1) / W);         // last lane with cells in this strip
  const bool has_left = w > 0, has_right = w + 1 < NS;
  const uint32_t ring_in_s = opaque_u32(smem_u32(ring + (size_t)w * RB));          // shared-space addresses
  const uint32_t ring_out_s = opaque_u32(smem_u32(ring + (size_t)(w + 1) * RB));
  const uint32_t dx_s = opaque_u32(smem_u32(dx));
  const uint32_t prod_in_s = opaque_u32(smem_u32(&prod[w]));
  const uint32_t cons_out_s = opaque_u32(smem_u32(&cons[w + 1 < kPdeMaxStrips ? w + 1 : w]));

  double up[W];
#pragma unroll
  for (int c = 0; c < W; ++c) up[c] = 1.0;                  // K[0, .] = 1 (pyx:29-31)
  double An[NC], Bn[NC];
#pragma unroll
  for (int cc = 0; cc < NC; ++cc) { An[cc] = 1.0; Bn[cc] = 1.0; }
  double A[NC], B[NC];
  double diag = 1.0, send = 1.0;
  const int nsteps = P + LL + 1;
  for (int s = 0; s < nsteps; ++s) {
    if ((s & (G - 1)) == 0) {
      // ---- hand-off window: rows [s, s+G) ------------------------------------------
      if (has_left) {                                       // lane 0 will read boundary rows s+1 .. s+G
        const int need = min(s + G, P);
        while (ld_volatile_shared_u(prod_in_s) < need) __nanosleep(64);
      }
      if (has_right) {                                      // lane LL will write boundary rows up to s+G-LL
        const int top = min(s + G - LL, P);
        while (top > ld_volatile_shared_u(cons_out_s) + RB) __nanosleep(64);
      }
      __threadfence_block();
    }
    if ((s & ((1 << ORDER) - 1)) == 0 && ncell > 0) {
      // Coefficients of the coarse row this lane enters during the next 2^ORDER steps, computed at a
      // warp-uniform step (the lanes cross coarse rows at different steps because of the skew; doing
      // this where they cross would run the divergent branch on every step).  pyx:41-48.
      const int In = min(max((s - lane + (1 << ORDER) - 1) >> ORDER, 0), L - 2);
      const uint32_t dxa = dx_s + (uint32_t)In * (uint32_t)(XS * 8);
      double2 xv[DP / 2];
#pragma unroll
      for (int k = 0; k < DP / 2; ++k)
        asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(xv[k].x), "=d"(xv[k].y) : "r"(dxa + 16u * k));
#pragma unroll
      for (int cc = 0; cc < NC; ++cc) {
        double e = 0.0;
        if (DYREG) {
#pragma unroll
          for (int k = 0; k < DP / 2; ++k) {
            e = fma(xv[k].x, dyreg[DYREG ? cc : 0][DYREG ? k : 0].x, e);
            e = fma(xv[k].y, dyreg[DYREG ? cc : 0][DYREG ? k : 0].y, e);
          }
        } else {
          const int J = min((j0 >> ORDER) + cc, L2 - 2);
          const double2* dyj = reinterpret_cast<const double2*>(dy + (size_t)J * DP + 2 * (J / NC));
#pragma unroll
          for (int k = 0; k < DP / 2; ++k) { const double2 yv = dyj[k]; e = fma(xv[k].x, yv.x, e); e = fma(xv[k].y, yv.y, e); }
        }
        const double e2 = e * e;
        An[cc] = 1. + 0.5 * e + (1. / 12) * e2;             // pyx:48
        Bn[cc] = 1. - (1. / 12) * e2;
      }
    }
    const int i = s - lane;                                 // fine row of this lane at this step
    double left = __shfl_up_sync(0xffffffffu, send, 1);     // K[i+1, j0] from lane-1 (computed last step)
    const bool rowok = (i >= 0) && (i < P) && (ncell > 0);
    const uint32_t slot = (uint32_t)((i + 1) & (RB - 1)) * 8u;
    if (lane == 0) {
      left = 1.0;                                           // K[., 0] = 1 (pyx:29-31)
      if (has_left && rowok) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(left) : "r"(ring_in_s + slot) : "memory");
    }
    if (rowok) {
      if ((i & ((1 << ORDER) - 1)) == 0) {                  // entering a new coarse row
#pragma unroll
        for (int cc = 0; cc < NC; ++cc) { A[cc] = An[cc]; B[cc] = Bn[cc]; }
      }
      // All W cells are computed, also the ones past column Q of the last lane (they feed nothing:
      // the lane to their right has no cells, the result is picked from up[] at the end).
      // (k01 + k10) * A - k00 * B, split into the part that only needs the previous row (wv, no
      // dependency between cells) and the one FMA per cell on the left-to-right chain
      double wv[W];
#pragma unroll
      for (int c = 0; c < W; ++c) {
        const int cc = (ORDER >= 3) ? 0 : (c >> ORDER);
        wv[c] = fma(up[c], A[cc], -((c == 0 ? diag : up[c > 0 ? c - 1 : 0]) * B[cc]));
      }
      diag = left;
      double lf = left;
#pragma unroll
      for (int c = 0; c < W; ++c) {
        const int cc = (ORDER >= 3) ? 0 : (c >> ORDER);
        lf = fma(lf, A[cc], wv[c]);
        up[c] = lf;
      }
      send = lf;
      if (lane == 31 && has_right) asm volatile("st.shared.f64 [%0], %1;" ::"r"(ring_out_s + slot), "d"(lf) : "memory");
    }
    if (((s + 1) & (G - 1)) == 0 || s + 1 == nsteps) {
      // ---- publish progress ------------------------------------------------------------
      __syncwarp();
      if (has_right && lane == LL) {
        __threadfence_block();
        st_volatile_shared(&prod[w + 1], max(0, min(s - LL + 1, P)));
      }
      if (has_left && lane == 0) st_volatile_shared(&cons[w], min(s + 1, P));
    }
  }
  if (w == NS - 1) {
    // K[P, Q] = up[Q - 1 - j0] of the last lane with cells
    double result = up[0];
#pragma unroll
    for (int c = 1; c < W; ++c) if (c == ncell - 1) result = up[c];
    result = __shfl_sync(0xffffffffu, result, LL);
    if (lane == 0) {
      const size_t o = Pm.pair_mode == GPSIG_PDE_DIAG ? (size_t)blockIdx.x
                                                      : (size_t)(ix - Pm.row_begin) * Pm.N2 + iy;
      Pm.out[o] = Pm.sigma * result;
    }
  }
}

// ---------------------------------------------------------------------------
// Block-row, persistent variant of the pipelined kernel (the one the library launches).
//
// (1) A lane advances min(2^ORDER, 4) FINE ROWS per step instead of one.  With the lanes skewed by
//     whole coarse rows (ORDER <= 2) every lane enters a new coarse row at every step, so the increment
//     dot product and the stencil coefficients (pyx:41-48) are computed straight into A/B at a
//     warp-uniform point (no double buffer, no per-lane select), and the fixed cost of a step
//     (shuffles, ring hand-off, progress counters, loop control) is spread over 4x the cells:
//     ~4.7 issued instructions per cell instead of ~12, 3.5 of them FP64.
// (2) A CTA is PERSISTENT: it takes pair after pair from a global ticket counter, and warp w (strip w)
//     moves on to the next pair as soon as its own strip is done.  The fill/drain bubble of the strip
//     pipeline (warp w trails warp w-1 by ~36 steps: 32 % of a pair's warp-steps at c3) is paid once per
//     CTA instead of once per pair.  dx is double buffered in shared memory (filled by warp 0), the ring
//     and the progress counters run on cumulative block-row numbers, a 4-entry queue hands the ticket
//     from warp 0 to the other warps.
// Ring, counters and shared-memory layout otherwise as in pde_pipe_kernel.
// ---------------------------------------------------------------------------
// register cap: 80 = 3 CTAs of 8 strips per SM for the shapes with few coarse columns per lane; the
// low dyadic orders keep 8 / 4 coefficient pairs and dy vectors per lane and need more
__host__ __device__ constexpr int pde_block_maxreg(int DP, int ORDER) {
  return ORDER <= 1 ? 128 : (ORDER == 2 ? (DP >= 16 ? 128 : (DP >= 8 ? 96 : 80)) : (DP >= 16 ? 96 : 80));
}
#ifndef GPSIG_PDE_SLEEP
#define GPSIG_PDE_SLEEP 256
#endif
inline size_t pde_block_smem_bytes(int L, int L2, int DP, int order, int nstrips, bool persist, bool sym = false) {
  return ((size_t)(persist ? 2 : 1) * (L - 1) * (DP + 2) + pde_dy_smem_elems(L2, DP, order) + (size_t)nstrips * kPdeRing) * sizeof(double) +
         (2 * kPdeMaxStrips + 8) * sizeof(int) + (sym ? (size_t)nstrips * 264 * sizeof(double) : 0);
}

// SYM (diagonal pairs k(x, x), Y == NULL): the grid is symmetric, so strip w only solves the rows from its own diagonal
// block down (cell rows i >= 256 w), as the reference does (sigKer_fast.pyx:33-61 fills the lower triangle and doubles
// the sub-diagonal value).  The node row above its first cell row, K[256 w, 256 w + 1 ..], is by symmetry the node
// COLUMN 256 w below the diagonal, i.e. the first 256 right-edge values of strip w-1: those go to a per-boundary
// buffer (diagc) instead of the ring, strip w starts once they are all there, uses them as its initial `up` row AND as
// its left boundary for the diagonal block, and takes the ring from there on.  36 of the 64 blocks of 256 x 256 cells
// are solved at c3 (the diagonal blocks in full) instead of 64.
constexpr int kPdeDiagPad = 264;    // doubles per boundary: node column values K[256 w + r, 256 w], r = 0..256
template <int DP, int ORDER, bool SYM>
__global__ void __maxnreg__(pde_block_maxreg(DP, ORDER)) pde_block_kernel(const PdeParams Pm, const int persist,
                                                                                      unsigned* __restrict__ ticket) {
  constexpr int W = kPdeW;
  constexpr int NC = (W >> ORDER) > 0 ? (W >> ORDER) : 1;   // coarse columns per lane
  constexpr int RS = (1 << ORDER) < 4 ? (1 << ORDER) : 4;   // fine rows per step
  constexpr int CST = (1 << ORDER) / RS;                    // steps per coarse row (1 for ORDER <= 2)
  constexpr int RB = kPdeRing;                              // ring rows (fine rows)
  constexpr int RBb = RB / RS;                              // ... in block rows
  constexpr int G = (16 / RS) > 4 ? (16 / RS) : 4;          // steps between progress exchanges (2G <= RBb)
  extern __shared__ __align__(128) unsigned char pde_smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, NS = blockDim.x >> 5;
  const int L = Pm.L, L2 = Pm.L2;
  const int P = (L - 1) << ORDER, Q = (L2 - 1) << ORDER;
  const int PB = P / RS;                                    // block rows per pair
  constexpr int XS = DP + 2;                                // padded dx row stride (doubles)
  constexpr bool DYREG = NC * DP <= 32;
  const size_t dxsz = (size_t)(L - 1) * XS;
  double* dx = reinterpret_cast<double*>(pde_smem);         // [persist ? 2 : 1][L-1][XS]
  double* dy = dx + (persist ? 2 : 1) * dxsz;
  double* ring = dy + pde_dy_smem_elems(L2, DP, ORDER);
  int* prod = reinterpret_cast<int*>(ring + (size_t)NS * RB);   // cumulative block rows of boundary w published by warp w-1
  int* cons = prod + kPdeMaxStrips;                               // ... consumed by warp w
  int* pairq = cons + kPdeMaxStrips;                              // [4] ticket queue (pair index or -1)
  int* npub = pairq + 4;                                          // tickets published by warp 0
  int* fin = npub + 1;                                            // pairs finished by the last warp
  double* diagc = reinterpret_cast<double*>(fin + 3);             // SYM: [NS][kPdeDiagPad]

  if (threadIdx.x < kPdeMaxStrips) { prod[threadIdx.x] = 0; cons[threadIdx.x] = 0; }
  if (threadIdx.x == 0) { *npub = 0; *fin = 0; }
  __syncthreads();                                          // the only block-wide barrier

  const size_t npairs = Pm.pair_mode == GPSIG_PDE_DIAG ? (size_t)(Pm.row_end - Pm.row_begin) : (size_t)(Pm.row_end - Pm.row_begin) * Pm.N2;
  const int j0 = w * 32 * W + lane * W;
  const int ncell = max(0, min(W, Q - j0));
  const int LL = min(31, (Q - w * 32 * W - 1) / W);         // last lane with cells in this strip
  const bool has_left = w > 0, has_right = w + 1 < NS;
  const uint32_t ring_in_s = opaque_u32(smem_u32(ring + (size_t)w * RB));
  const uint32_t ring_out_s = opaque_u32(smem_u32(ring + (size_t)(w + 1) * RB));
  const uint32_t prod_in_s = opaque_u32(smem_u32(&prod[w]));
  const uint32_t cons_out_s = opaque_u32(smem_u32(&cons[w + 1 < kPdeMaxStrips ? w + 1 : w]));
  const uint32_t npub_s = opaque_u32(smem_u32(npub)), fin_s = opaque_u32(smem_u32(fin));
  const bool lane0 = lane == 0, lane_last = lane == 31;
  const double fscale = 1.0 / (double)(1 << (2 * ORDER));
  // SYM: first block row of this strip (its diagonal block starts at cell row 256 w) and block rows of a diagonal block
  constexpr int DBF = (32 * W) / RS;
  const int b0 = SYM ? w * DBF : 0;
  const int DBw = SYM ? min(DBF, PB - b0) : 0;               // block rows of MY diagonal block
  const int bn0 = SYM ? (w + 1) * DBF : 0;                   // first block row of the NEXT strip
  const int DBn = SYM ? min(DBF, max(PB - bn0, 0)) : 0;
  const uint32_t diag_in_s = opaque_u32(smem_u32(diagc + (size_t)w * kPdeDiagPad));
  const uint32_t diag_out_s = opaque_u32(smem_u32(diagc + (size_t)(w + 1 < NS ? w + 1 : w) * kPdeDiagPad));
  const int nsteps = PB - b0 + LL;

  for (int n = 0;; ++n) {                                    // n-th pair of this CTA
    const int base = n * PB;                                // cumulative block rows before this pair
    // ---- ticket -------------------------------------------------------------------------------
    if (w == 0) {
      // dx buffer n&1 and queue slot n&3 are free once the last warp has finished pair n-2
      while (ld_volatile_shared_u(fin_s) < n - 1) __nanosleep(GPSIG_PDE_SLEEP);
      __threadfence_block();
      long long t = 0;
      if (lane0) t = persist ? (long long)atomicAdd(ticket, 1u) : (n == 0 ? (long long)blockIdx.x : (long long)npairs);
      t = __shfl_sync(0xffffffffu, t, 0);
      const int pid = t < (long long)npairs ? (int)t : -1;
      if (pid >= 0 && (!DYREG || true)) {
        int ixn;
        if (Pm.pair_mode == GPSIG_PDE_DIAG) ixn = Pm.row_begin + pid; else ixn = Pm.row_begin + pid / Pm.N2;
        const double* xn = static_cast<const double*>(Pm.xp) + (size_t)ixn * Pm.xblk;
        double* dxb = dx + (persist ? (size_t)(n & 1) * dxsz : 0);
        for (int e = lane; e < (L - 1) * DP; e += 32) {
          const int I = e / DP, k = e - I * DP;
          dxb[I * XS + k] = (xn[e + DP] - xn[e]) * fscale;
        }
        if (!DYREG) {
          const int iyn = Pm.pair_mode == GPSIG_PDE_DIAG ? ixn : pid % Pm.N2;
          const double* yn = static_cast<const double*>(Pm.yp) + (size_t)iyn * Pm.yblk;
          for (int e = lane; e < (L2 - 1) * DP; e += 32) dy[e + 2 * ((e / DP) / NC)] = yn[e + DP] - yn[e];
        }
      }
      __syncwarp();
      if (lane0) {
        pairq[n & 3] = pid;
        __threadfence_block();
        st_volatile_shared(npub, n + 1);
      }
      __syncwarp();
    }
    while (ld_volatile_shared_u(npub_s) <= n) __nanosleep(GPSIG_PDE_SLEEP);
    __threadfence_block();
    const int pid = ld_volatile_shared(&pairq[n & 3]);
    if (pid < 0) break;
    int ix, iy;
    if (Pm.pair_mode == GPSIG_PDE_DIAG) { ix = Pm.row_begin + pid; iy = ix; }
    else { ix = Pm.row_begin + pid / Pm.N2; iy = pid % Pm.N2; }
    const double* y = static_cast<const double*>(Pm.yp) + (size_t)iy * Pm.yblk;
    const uint32_t dx_s = smem_u32(dx + (persist ? (size_t)(n & 1) * dxsz : 0));

    double2 dyreg[DYREG ? NC : 1][DYREG ? DP / 2 : 1];
    if (DYREG) {
#pragma unroll
      for (int cc = 0; cc < NC; ++cc) {
        const int J = min((j0 >> ORDER) + cc, L2 - 2);
#pragma unroll
        for (int k = 0; k < DP / 2; ++k)
          dyreg[DYREG ? cc : 0][DYREG ? k : 0] = make_double2(y[(J + 1) * DP + 2 * k] - y[J * DP + 2 * k],
                                                                y[(J + 1) * DP + 2 * k + 1] - y[J * DP + 2 * k + 1]);
      }
    }
    double up[W];
#pragma unroll
    for (int c = 0; c < W; ++c) up[c] = 1.0;                // K[0, .] = 1 (pyx:29-31)
    double A[NC], B[NC], An[NC], Bn[NC];
#pragma unroll
    for (int cc = 0; cc < NC; ++cc) { A[cc] = B[cc] = An[cc] = Bn[cc] = 1.0; }
    double send[RS];
#pragma unroll
    for (int r = 0; r < RS; ++r) send[r] = 1.0;
    double dgl = 1.0;                                       // K[i, j0]
    if (SYM && has_left) {
      // the node row above my first cell row = the node column to my left (symmetry): wait for all of it
      const int need0 = base + b0 + DBw;
      while (ld_volatile_shared_u(prod_in_s) - need0 < 0) __nanosleep(GPSIG_PDE_SLEEP);
      __threadfence_block();
      asm volatile("ld.shared.f64 %0, [%1];" : "=d"(dgl) : "r"(diag_in_s + (uint32_t)(W * lane) * 8u) : "memory");
#pragma unroll
      for (int c = 0; c < W; ++c)
        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(up[c]) : "r"(diag_in_s + (uint32_t)(W * lane + 1 + c) * 8u) : "memory");
    }
    int b = b0 - lane;                                      // block row of this lane at step s
    for (int s = 0; s < nsteps; ++s, ++b) {
      if ((s & (G - 1)) == 0) {
        // ---- hand-off window: block rows [s, s+G) of this pair ----------------------------
        if (has_left) {
          const int need = base + min(b0 + s + G, PB);
          while (ld_volatile_shared_u(prod_in_s) - need < 0) __nanosleep(GPSIG_PDE_SLEEP);
        }
        if (has_right) {
          if (!SYM) {
            const int top = base + min(s + G - LL, PB);
            while (top - ld_volatile_shared_u(cons_out_s) > RBb) __nanosleep(GPSIG_PDE_SLEEP);
          } else {
            const int topb = min(b0 + s + G - LL, PB);         // block rows (of this pair) lane LL will have written after this window
            // (a) its rows that land in the next strip's diagonal buffer: that strip must have finished reading the
            //     buffer of the PREVIOUS pair (it has, once it is past its diagonal block there)
            if (n > 0 && topb >= bn0) {
              const int freed = base - PB + bn0 + DBn;
              while (ld_volatile_shared_u(cons_out_s) - freed < 0) __nanosleep(GPSIG_PDE_SLEEP);
            }
            // (b) its rows that land in the ring (past the next strip's diagonal block): the reader has left the
            //     previous pair and is at most a ring behind, counting from where it starts to read the ring
            if (topb > bn0 + DBn) {
              for (;;) {
                const int c = ld_volatile_shared_u(cons_out_s) - base;
                if (c >= 0 && topb - max(c, bn0 + DBn) <= RBb) break;
                __nanosleep(GPSIG_PDE_SLEEP);
              }
            }
          }
        }
        __threadfence_block();
      }
      if ((CST == 1 || (s & (CST - 1)) == 0) && ncell > 0) {
        // coefficients of the coarse row this lane enters within the next CST steps (pyx:41-48)
        const int In = min(max(b + CST - 1, 0) / CST, L - 2);
        const uint32_t dxa = dx_s + (uint32_t)In * (uint32_t)(XS * 8);
        double2 xv[DP / 2];
#pragma unroll
        for (int k = 0; k < DP / 2; ++k)
          asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(xv[k].x), "=d"(xv[k].y) : "r"(dxa + 16u * k));
#pragma unroll
        for (int cc = 0; cc < NC; ++cc) {
          double e = 0.0;
          if (DYREG) {
#pragma unroll
            for (int k = 0; k < DP / 2; ++k) {
              e = fma(xv[k].x, dyreg[DYREG ? cc : 0][DYREG ? k : 0].x, e);
              e = fma(xv[k].y, dyreg[DYREG ? cc : 0][DYREG ? k : 0].y, e);
            }
          } else {
            const int J = min((j0 >> ORDER) + cc, L2 - 2);
            const double2* dyj = reinterpret_cast<const double2*>(dy + (size_t)J * DP + 2 * (J / NC));
#pragma unroll
            for (int k = 0; k < DP / 2; ++k) { const double2 yv = dyj[k]; e = fma(xv[k].x, yv.x, e); e = fma(xv[k].y, yv.y, e); }
          }
          const double e2 = e * e;
          An[cc] = 1. + 0.5 * e + (1. / 12) * e2;           // pyx:48
          Bn[cc] = 1. - (1. / 12) * e2;
        }
      }
      // K[i+1, j0] of this block row's RS fine rows, from the lane to the left (computed one step ago)
      double recv[RS];
#pragma unroll
      for (int r = 0; r < RS; ++r) recv[r] = __shfl_up_sync(0xffffffffu, send[r], 1);
      const bool rowok = (b >= b0) && (b < PB) && (ncell > 0);
      const uint32_t slot = (uint32_t)(((base + b) * RS) & (RB - 1)) * 8u;
      if (lane0) {
#pragma unroll
        for (int r = 0; r < RS; ++r) recv[r] = 1.0;         // K[., 0] = 1 (pyx:29-31)
        if (SYM && has_left && rowok && b < b0 + DBw) {
          // inside my diagonal block the left boundary comes from the diagonal buffer (node rows RS b + 1 ..)
#pragma unroll
          for (int r = 0; r < RS; ++r)
            asm volatile("ld.shared.f64 %0, [%1];" : "=d"(recv[r]) : "r"(diag_in_s + (uint32_t)((b - b0) * RS + 1 + r) * 8u) : "memory");
        } else if (has_left && rowok) {
          if (RS == 4) {
            asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(recv[0]), "=d"(recv[RS > 1 ? 1 : 0]) : "r"(ring_in_s + slot) : "memory");
            asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(recv[RS > 2 ? 2 : 0]), "=d"(recv[RS > 3 ? 3 : 0]) : "r"(ring_in_s + slot + 16u) : "memory");
          } else if (RS == 2) {
            asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(recv[0]), "=d"(recv[RS > 1 ? 1 : 0]) : "r"(ring_in_s + slot) : "memory");
          } else {
            asm volatile("ld.shared.f64 %0, [%1];" : "=d"(recv[0]) : "r"(ring_in_s + slot) : "memory");
          }
        }
      }
      if (rowok) {
        if (CST == 1 || (b & (CST - 1)) == 0) {             // entering a new coarse row
#pragma unroll
          for (int cc = 0; cc < NC; ++cc) { A[cc] = An[cc]; B[cc] = Bn[cc]; }
        }
        // (k01 + k10) * A - k00 * B per cell (pyx:48).  The RS x W block of this step is swept by
        // ANTI-DIAGONALS, written out in that order: the up-to-RS cells of a diagonal are independent, each
        // is DMUL -> DFMA (previous row only) -> DFMA (left neighbour, the chain), and a value is needed
        // one diagonal (~12 FP64 instructions) after it was produced: no issue stalls on the 8.5-cycle
        // DFMA latency.  (Row-by-row order left the warp waiting on every chain FMA.)  All W cells are
        // computed, also those past column Q of the last lane.
        double Kv[RS + 1][W + 1];                           // Kv[r+1][c+1] = cell (r, c); row 0 / column 0 = boundary
        Kv[0][0] = dgl;
#pragma unroll
        for (int c = 0; c < W; ++c) Kv[0][c + 1] = up[c];
#pragma unroll
        for (int r = 0; r < RS; ++r) Kv[r + 1][0] = recv[r];
#pragma unroll
        for (int d = 0; d < RS + W - 1; ++d) {
          double t0[RS], t1[RS];
#pragma unroll
          for (int r = 0; r < RS; ++r) {
            const int c = d - r;
            if (c >= 0 && c < W) t0[r] = Kv[r][c] * B[(ORDER >= 3) ? 0 : (c >> ORDER)];
          }
#pragma unroll
          for (int r = 0; r < RS; ++r) {
            const int c = d - r;
            if (c >= 0 && c < W) t1[r] = fma(Kv[r][c + 1], A[(ORDER >= 3) ? 0 : (c >> ORDER)], -t0[r]);
          }
#pragma unroll
          for (int r = 0; r < RS; ++r) {
            const int c = d - r;
            if (c >= 0 && c < W) Kv[r + 1][c + 1] = fma(Kv[r + 1][c], A[(ORDER >= 3) ? 0 : (c >> ORDER)], t1[r]);
          }
        }
        dgl = recv[RS - 1];
#pragma unroll
        for (int c = 0; c < W; ++c) up[c] = Kv[RS][c + 1];
#pragma unroll
        for (int r = 0; r < RS; ++r) send[r] = Kv[r + 1][W];
        if (SYM && lane_last && has_right && b < bn0 + DBn) {
          // node rows RS b + 1 .. RS b + RS of my right edge = entries RS (b - bn0) + 1 + r of the next strip's diagonal
          // buffer (entry 0, the corner K[256 (w+1), 256 (w+1)], is the last value of block row bn0 - 1)
#pragma unroll
          for (int r = 0; r < RS; ++r) {
            const int e = (b - bn0) * RS + 1 + r;
            if (e >= 0) asm volatile("st.shared.f64 [%0], %1;" ::"r"(diag_out_s + (uint32_t)e * 8u), "d"(send[r]) : "memory");
          }
        } else if (lane_last && has_right) {
          if (RS == 4) {
            asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(ring_out_s + slot), "d"(send[0]), "d"(send[RS > 1 ? 1 : 0]) : "memory");
            asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(ring_out_s + slot + 16u), "d"(send[RS > 2 ? 2 : 0]), "d"(send[RS > 3 ? 3 : 0]) : "memory");
          } else if (RS == 2) {
            asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(ring_out_s + slot), "d"(send[0]), "d"(send[RS > 1 ? 1 : 0]) : "memory");
          } else {
            asm volatile("st.shared.f64 [%0], %1;" ::"r"(ring_out_s + slot), "d"(send[0]) : "memory");
          }
        }
      }
      if (((s + 1) & (G - 1)) == 0 || s + 1 == nsteps) {
        // ---- publish progress (cumulative block rows) -------------------------------------
        __syncwarp();
        if (has_right && lane == LL) {
          __threadfence_block();
          st_volatile_shared(&prod[w + 1], base + max(0, min(b0 + s - LL + 1, PB)));
        }
        if (has_left && lane0) st_volatile_shared(&cons[w], base + min(b0 + s + 1, PB));
      }
    }
    if (w == NS - 1) {
      double result = up[0];
#pragma unroll
      for (int c = 1; c < W; ++c) if (c == ncell - 1) result = up[c];
      result = __shfl_sync(0xffffffffu, result, LL);
      if (lane0) {
        const size_t o = Pm.pair_mode == GPSIG_PDE_DIAG ? (size_t)pid : (size_t)(ix - Pm.row_begin) * Pm.N2 + iy;
        Pm.out[o] = Pm.sigma * result;
        __threadfence_block();
        st_volatile_shared(fin, n + 1);
      }
      __syncwarp();
    }
  }
}

}  // namespace gpsig
