// Vector-Jacobian products (reverse mode) of the first-order truncated signature kernels for sm_100a.
//
// The reference gets these from TensorFlow's autodiff of gpsig/signature_algs.py:8-35, 76-127 (every
// tf.cumsum / tf.reduce_sum / base-kernel op of the lifted tensor is differentiated and the lifted
// tensor is kept alive for the backward pass).  Here each op has a hand-written adjoint:
//
//   trunc_vjp_*      raw levels K_l(x_i, y_j), l = 1..m  (Gram rows or diagonal pairs)   -> dL/dX, dL/dX2
//   tens_seq_vjp_*   raw levels of inducing tensors vs sequences (signature_algs.py:101-127) -> dL/dZ, dL/dX
//   tens_tens_vjp_*  raw levels of inducing tensors vs themselves (signature_algs.py:76-99)  -> dL/dZ
//
// Gradients are taken w.r.t. the coordinates the kernels are CALLED with (the host layer scales by the
// lengthscales / applies lags with differentiable torch ops and chains the rest).
//
// Adjoint of the level recursion for one pair.  With D[p,q] the (differenced) base-kernel tile,
//   F_a[p,q] = sum over a-tuples strictly inside [0,p) x [0,q)   (F_0 = 1; the forward 2-D exclusive prefix sums)
//   G_b[p,q] = sum over b-tuples strictly inside (p,P) x (q,Q)   (G_0 = 1; the same recursion run backwards)
//   K_l = sum_{p,q} D F_{l-1}           dK_l/dD[p,q] = sum_{a+b=l-1} F_a[p,q] G_b[p,q]
// so  W[p,q] = dL/dD[p,q] = sum_a F_a[p,q] * (sum_b g_{a+b+1} G_b[p,q]).
// One THREAD owns one pair: a forward sweep stores F_1..F_{m-1} in an HBM scratch table laid out with the
// pair index fastest (coalesced across the threads of a warp), a backward sweep forms G on the fly and
// stores W; a second kernel (one thread per (pair, point)) turns W into dL/dM by the adjoint of the 2-D
// difference (signature_algs.py:26) and chains it through the base kernel's partial derivatives into the
// coordinates, accumulating with atomics (several pairs share a sequence).
// First version: parallel over pairs only, sized for the training shapes (minibatch x inducing points);
// the large-Gram backward is chunked over rows by the host layer.  order == 1 only.
#pragma once
#include "common.cuh"

namespace gpsig {

// partial derivatives of the base kernel f(dot, xs, ys) (kernels.py:946-1173) w.r.t. x.y, |x|^2, |y|^2
template <typename T>
__device__ __forceinline__ void base_partials(int base, T p0, T p1, T dot, T xs, T ys, T& fd, T& fx, T& fy) {
  fd = fx = fy = T(0);
  switch (base) {
    case GPSIG_BASE_LINEAR: fd = T(1); break;
    case GPSIG_BASE_RBF: { const T f = gexp<T>(-((T(-2) * dot + xs) + ys) / T(2)); fd = f; fx = fy = T(-0.5) * f; break; }
    case GPSIG_BASE_COSINE: {
      const T sx = gsqrt<T>(xs), sy = gsqrt<T>(ys);
      const T f = dot / (sx * sy);
      fd = T(1) / (sx * sy); fx = T(-0.5) * f / xs; fy = T(-0.5) * f / ys; break;
    }
    case GPSIG_BASE_POLY: fd = p1 * gpow<T>(dot + p0, p1 - T(1)); break;
    case GPSIG_BASE_MIX: {
      const T e = gexp<T>(-(xs + ys - T(2) * dot) / T(2));
      fd = p0 * e + (T(1) - p0); fx = fy = T(-0.5) * p0 * e; break;
    }
    default: {
      const T r2 = (T(-2) * dot + xs) + ys;
      if (!(r2 > T(1e-40))) break;                        // clamped: zero gradient, as tf.maximum in _euclid_dist (kernels.py:959-961)
      const T r = gsqrt<T>(r2);
      T dr2;                                              // df/d(r^2)
      if (base == GPSIG_BASE_MATERN12) dr2 = -gexp<T>(-r) / (T(2) * r);
      else if (base == GPSIG_BASE_MATERN32) dr2 = T(-1.5) * gexp<T>(-T(1.7320508075688772) * r);
      else { const T s5 = T(2.23606797749979); dr2 = T(-5.0 / 6.0) * (T(1) + s5 * r) * gexp<T>(-s5 * r); }
      fd = T(-2) * dr2; fx = fy = dr2;
    }
  }
}

template <typename T>
__device__ __forceinline__ void dot3(const T* __restrict__ x, const T* __restrict__ y, int d, T& dot, T& xs, T& ys) {
  dot = xs = ys = T(0);
  for (int k = 0; k < d; ++k) { const T u = x[k], v = y[k]; dot = gfma<T>(u, v, dot); xs = gfma<T>(u, u, xs); ys = gfma<T>(v, v, ys); }
}
template <typename T>
__device__ __forceinline__ T kappa_raw(const T* x, const T* y, int d, int base, T p0, T p1) {
  T dot, xs, ys;
  dot3(x, y, d, dot, xs, ys);
  return base_eval<T>(base, p0, p1, dot, xs, ys);
}

// out-of-line copies: the base-kernel switch is large (f64 exp / pow / sqrt), and kernels that run one warp per
// SM partition stall on instruction fetch when it is inlined at every unrolled site
template <typename T>
__device__ __noinline__ T kappa_raw_call(const T* x, const T* y, int d, int base, T p0, T p1) {
  return kappa_raw<T>(x, y, d, base, p0, p1);
}
template <typename T>
__device__ __noinline__ void base_partials_call(int base, T p0, T p1, T dot, T xs, T ys, T& fd, T& fx, T& fy) {
  base_partials<T>(base, p0, p1, dot, xs, ys, fd, fx, fy);
}

template <typename T> __device__ __forceinline__ void atomic_add(T* p, T v) { atomicAdd(p, v); }

struct TruncVjpParams {
  const void* X; const void* Y;      // dense [N, L, d] / [N2, L2, d] (Y == X for the symmetric / diagonal case)
  int N, N2, L, L2, d, m;
  int diag;                          // pairs (i, i) instead of rows x N2
  int row_begin, row_end;
  int difference;
  double p0, p1; int base;
  const void* g;                     // upstream dL/dK_l: [m+1, R, N2] or [m+1, R] (diag)
  void* gradX; void* gradY;          // [N, L, d] / [N2, L2, d], accumulated
  void* F; void* C; void* W;         // scratch: F [(m-1), P, Q, npairs], C [(m-1), Q, npairs], W [P, Q, npairs]
  size_t npairs;
  int w_cell_major;                  // W is [npairs, P, Q] (written by the CTA-per-pair sweeps) instead of [P, Q, npairs]
};

// sweep 1+2: one thread per pair
template <typename T>
__global__ void trunc_vjp_tables_kernel(const TruncVjpParams P) {
  const size_t pair = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= P.npairs) return;
  const size_t np = P.npairs;
  const int m = P.m, d = P.d, diff = P.difference ? 1 : 0;
  const int PR = P.L - diff, QC = P.L2 - diff;                        // rows / columns of D
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * P.L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * P.L2 * d;
  const T p0 = T(P.p0), p1 = T(P.p1);
  T* F = static_cast<T*>(P.F);
  T* C = static_cast<T*>(P.C);
  T* W = static_cast<T*>(P.W);
  auto Fat = [&](int a, int p, int q) -> T& { return F[(((size_t)(a - 1) * PR + p) * QC + q) * np + pair]; };   // a = 1..m-1
  auto Cat = [&](int a, int q) -> T& { return C[((size_t)(a - 1) * QC + q) * np + pair]; };
  // upstream gradients of this pair
  T g[9];
  const size_t gstride = P.diag ? (size_t)(P.row_end - P.row_begin) : (size_t)(P.row_end - P.row_begin) * P.N2;
  for (int l = 0; l <= m; ++l) g[l] = static_cast<const T*>(P.g)[(size_t)l * gstride + pair];

  // D[p, q] from rolling base-kernel values: m00 = M[p, q], m01 = M[p, q+1], m10 = M[p+1, q], m11 = M[p+1, q+1]
  auto Dcell = [&](int p, int q, T& m00, T& m10, bool fresh) -> T {
    if (!diff) return kappa_raw<T>(x + (size_t)p * d, y + (size_t)q * d, d, P.base, p0, p1);
    if (fresh) {
      m00 = kappa_raw<T>(x + (size_t)p * d, y + (size_t)q * d, d, P.base, p0, p1);
      m10 = kappa_raw<T>(x + (size_t)(p + 1) * d, y + (size_t)q * d, d, P.base, p0, p1);
    }
    const T m01 = kappa_raw<T>(x + (size_t)p * d, y + (size_t)(q + 1) * d, d, P.base, p0, p1);
    const T m11 = kappa_raw<T>(x + (size_t)(p + 1) * d, y + (size_t)(q + 1) * d, d, P.base, p0, p1);
    const T D = (m11 - m10) - (m01 - m00);                           // signature_algs.py:26
    m00 = m01; m10 = m11;
    return D;
  };

  // ---- forward sweep: F_a[p, q], a = 1..m-1 ----------------------------------------------------
  for (int a = 1; a < m; ++a)
    for (int q = 0; q < QC; ++q) Cat(a, q) = T(0);
  for (int p = 0; p < PR; ++p) {
    T E[8];
    for (int a = 1; a < m; ++a) E[a] = T(0);
    T m00 = T(0), m10 = T(0);
    for (int q = 0; q < QC; ++q) {
      const T D = Dcell(p, q, m00, m10, q == 0);
      T Fprev = T(1);                                                // F_0
      for (int a = 1; a < m; ++a) {
        const T c = Cat(a, q);
        const T Fa = E[a];                                           // rows < p, columns < q
        Fat(a, p, q) = Fa;
        E[a] += c;
        Cat(a, q) = gfma<T>(D, Fprev, c);                            // column sum now includes row p
        Fprev = Fa;
      }
    }
  }
  // ---- backward sweep: G_b on the fly, W = sum_a F_a H_a ------------------------------------------
  for (int b = 1; b < m; ++b)
    for (int q = 0; q < QC; ++q) Cat(b, q) = T(0);
  for (int p = PR - 1; p >= 0; --p) {
    T E[8];
    for (int b = 1; b < m; ++b) E[b] = T(0);
    // walk the columns right to left; rolling values mirrored: keep M[p, q+1], M[p+1, q+1] from the previous cell
    T r01 = T(0), r11 = T(0);
    for (int q = QC - 1; q >= 0; --q) {
      T D;
      if (!diff) D = kappa_raw<T>(x + (size_t)p * d, y + (size_t)q * d, d, P.base, p0, p1);
      else {
        if (q == QC - 1) {
          r01 = kappa_raw<T>(x + (size_t)p * d, y + (size_t)(q + 1) * d, d, P.base, p0, p1);
          r11 = kappa_raw<T>(x + (size_t)(p + 1) * d, y + (size_t)(q + 1) * d, d, P.base, p0, p1);
        }
        const T m00 = kappa_raw<T>(x + (size_t)p * d, y + (size_t)q * d, d, P.base, p0, p1);
        const T m10 = kappa_raw<T>(x + (size_t)(p + 1) * d, y + (size_t)q * d, d, P.base, p0, p1);
        D = (r11 - m10) - (r01 - m00);
        r01 = m00; r11 = m10;
      }
      T G[8];
      G[0] = T(1);
      T Gprev = T(1);
      for (int b = 1; b < m; ++b) {
        const T c = Cat(b, q);
        const T Gb = E[b];                                           // rows > p, columns > q
        G[b] = Gb;
        E[b] += c;
        Cat(b, q) = gfma<T>(D, Gprev, c);
        Gprev = Gb;
      }
      T w = T(0);
      for (int a = 0; a < m; ++a) {
        T H = T(0);
        for (int b = 0; a + b + 1 <= m; ++b) H = gfma<T>(g[a + b + 1], G[b], H);
        const T Fa = a == 0 ? T(1) : Fat(a, p, q);
        w = gfma<T>(Fa, H, w);
      }
      W[((size_t)p * QC + q) * np + pair] = w;
    }
  }
}

// sweep 1+2, one WARP per pair (training shapes: few pairs, e.g. the 50 diagonal pairs of a minibatch -- with one
// thread per pair the sweeps are a single serial chain per pair: 60 ms for 50 pairs of L = 100).  Lane t owns the
// CW consecutive columns [t CW, (t+1) CW) of D; rows are swept one at a time.  The column sums C_a[q] live in
// registers; the exclusive prefix over q that turns them into F_a[p, .] (G_b[p, .] on the way back: suffix) is a
// lane-local scan plus one warp scan of the lane totals per level (the levels of a row are independent).  Same
// scratch tables and layout as the thread-per-pair kernel (the grad kernel reads W from it).
template <typename T, int CW, int MLV>
__global__ void __launch_bounds__(128) trunc_vjp_tables_warp_kernel(const TruncVjpParams P) {
  const size_t pair = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pair >= P.npairs) return;
  const int lane = threadIdx.x & 31;
  const size_t np = P.npairs;
  const int m = P.m, d = P.d, diff = P.difference ? 1 : 0;
  const int PR = P.L - diff, QC = P.L2 - diff;
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * P.L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * P.L2 * d;
  const T p0 = T(P.p0), p1 = T(P.p1);
  T* F = static_cast<T*>(P.F);
  T* W = static_cast<T*>(P.W);
  auto Fat = [&](int a, int p, int q) -> T& { return F[(((size_t)(a - 1) * PR + p) * QC + q) * np + pair]; };   // a = 1..m-1
  T g[9];
  const size_t gstride = P.diag ? (size_t)(P.row_end - P.row_begin) : (size_t)(P.row_end - P.row_begin) * P.N2;
  for (int l = 0; l <= m; ++l) g[l] = static_cast<const T*>(P.g)[(size_t)l * gstride + pair];
  const int q0 = lane * CW;
  // base-kernel row r on this lane's columns q0 .. q0 + CW (one more than it owns: the difference needs q + 1)
  auto Mrow = [&](int r, T (&o)[CW + 1]) {
#pragma unroll
    for (int c = 0; c <= CW; ++c) {
      const int q = min(q0 + c, P.L2 - 1);
      o[c] = kappa_raw_call<T>(x + (size_t)r * d, y + (size_t)q * d, d, P.base, p0, p1);
    }
  };
  auto Drow = [&](const T (&lo)[CW + 1], const T (&hi)[CW + 1], T (&D)[CW]) {   // rows p (lo) and p+1 (hi)
#pragma unroll
    for (int c = 0; c < CW; ++c) {
      const T v = diff ? (hi[c + 1] - hi[c]) - (lo[c + 1] - lo[c]) : lo[c];     // signature_algs.py:26
      D[c] = (q0 + c < QC) ? v : T(0);
    }
  };
  // exclusive prefix (DIR = +1) / suffix (DIR = -1) over all columns of the warp
  auto scan = [&](const T (&v)[CW], T (&o)[CW], int dir) {
    T tot = T(0);
#pragma unroll
    for (int c = 0; c < CW; ++c) tot += v[c];
    T incl = tot;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const T t = dir > 0 ? __shfl_up_sync(0xffffffffu, incl, s) : __shfl_down_sync(0xffffffffu, incl, s);
      if (dir > 0 ? lane >= s : lane + s < 32) incl += t;
    }
    T run = incl - tot;                                                         // columns of the lanes before / after
    if (dir > 0) {
#pragma unroll
      for (int c = 0; c < CW; ++c) { o[c] = run; run += v[c]; }
    } else {
#pragma unroll
      for (int c = CW - 1; c >= 0; --c) { o[c] = run; run += v[c]; }
    }
  };

  T C[MLV][CW];
  T lo[CW + 1], hi[CW + 1], D[CW];
  // ---- forward sweep: F_a[p, q], a = 1..m-1 ----------------------------------------------------
#pragma unroll
  for (int a = 0; a < MLV; ++a)
#pragma unroll
    for (int c = 0; c < CW; ++c) C[a][c] = T(0);
  Mrow(0, lo);
  for (int p = 0; p < PR; ++p) {
    if (diff) Mrow(p + 1, hi);
    Drow(lo, hi, D);
    T Fprev[CW];
#pragma unroll
    for (int c = 0; c < CW; ++c) Fprev[c] = T(1);                               // F_0
#pragma unroll
    for (int a = 0; a < MLV; ++a) {
      if (a + 1 < m) {
        T Fa[CW];
        scan(C[a], Fa, +1);                                                     // rows < p, columns < q
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          if (q0 + c < QC) Fat(a + 1, p, q0 + c) = Fa[c];
          C[a][c] = gfma<T>(D[c], Fprev[c], C[a][c]);                           // column sum now includes row p
          Fprev[c] = Fa[c];
        }
      }
    }
    if (diff) {
#pragma unroll
      for (int c = 0; c <= CW; ++c) lo[c] = hi[c];
    } else if (p + 1 < PR) Mrow(p + 1, lo);
  }
  // ---- backward sweep: G_b on the fly, W = sum_a F_a H_a ------------------------------------------
#pragma unroll
  for (int a = 0; a < MLV; ++a)
#pragma unroll
    for (int c = 0; c < CW; ++c) C[a][c] = T(0);
  if (diff) Mrow(PR, hi); else Mrow(PR - 1, lo);
  for (int p = PR - 1; p >= 0; --p) {
    if (diff) Mrow(p, lo);
    Drow(lo, hi, D);
    T G[MLV + 1][CW];
#pragma unroll
    for (int c = 0; c < CW; ++c) G[0][c] = T(1);
#pragma unroll
    for (int b = 0; b < MLV; ++b) {
      if (b + 1 < m) {
        scan(C[b], G[b + 1], -1);                                               // rows > p, columns > q
#pragma unroll
        for (int c = 0; c < CW; ++c) C[b][c] = gfma<T>(D[c], G[b][c], C[b][c]);
      }
    }
#pragma unroll
    for (int c = 0; c < CW; ++c) {
      if (q0 + c < QC) {
        T w = T(0);
#pragma unroll
        for (int a = 0; a <= MLV; ++a) {
          if (a < m) {
            T H = T(0);
#pragma unroll
            for (int b = 0; b <= MLV; ++b)
              if (a + b + 1 <= m) H = gfma<T>(g[a + b + 1], G[b][c], H);
            const T Fa = a == 0 ? T(1) : Fat(a, p, q0 + c);
            w = gfma<T>(Fa, H, w);
          }
        }
        W[((size_t)p * QC + q0 + c) * np + pair] = w;
      }
    }
    if (diff) {
#pragma unroll
      for (int c = 0; c <= CW; ++c) hi[c] = lo[c];
    } else if (p > 0) Mrow(p - 1, lo);
  }
}

// ---------------------------------------------------------------------------
// sweep 1+2 with one CTA per pair: D and the current table (F_{a-1} or G_{b-1}) live in shared memory, the 2-D
// exclusive prefix sum of each level is two parallel passes (down the rows: one thread per column; along the rows:
// one warp per row, warp scans), and W accumulates in registers (cell e = tid + k * threads).  F_1..F_{m-1} go
// through the HBM scratch in a private [pair][a][cell] layout (L2-resident, coalesced); W is written cell-fastest
// per pair (`w_cell_major`), which the warp-per-point gradient kernel reads back coalesced.  Needs L * L2 <= 12288 cells.
// ---------------------------------------------------------------------------
constexpr int kVjpCtaThreads = 256;
constexpr int kVjpCtaCells = 48;                                     // cells per thread at most

template <typename T>
__device__ __forceinline__ T warp_excl_prefix1(T v, int lane, T& total) {
  T inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
  total = __shfl_sync(0xffffffffu, inc, 31);
  const T ex = __shfl_up_sync(0xffffffffu, inc, 1);
  return lane == 0 ? T(0) : ex;
}
template <typename T>
__device__ __forceinline__ T warp_excl_suffix1(T v, int lane, T& total) {
  T inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_down_sync(0xffffffffu, inc, o); if (lane + o < 32) inc += t; }
  total = __shfl_sync(0xffffffffu, inc, 0);
  const T ex = __shfl_down_sync(0xffffffffu, inc, 1);
  return lane == 31 ? T(0) : ex;
}

template <typename T>
__global__ void __launch_bounds__(kVjpCtaThreads) trunc_vjp_tables_cta_kernel(const TruncVjpParams P) {
  extern __shared__ __align__(16) unsigned char vjp_cta_smem[];
  constexpr int NT = kVjpCtaThreads;
  const size_t pair = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t np = P.npairs;
  const int m = P.m, d = P.d, diff = P.difference ? 1 : 0;
  const int L = P.L, L2 = P.L2, PR = L - diff, QC = L2 - diff, cells = PR * QC;
  T* Dt = reinterpret_cast<T*>(vjp_cta_smem);                        // [PR][QC]
  T* Tc = Dt + (size_t)L * L2;                                       // [L][L2] base kernel, then [PR][QC] current table
  T* gs = Tc + (size_t)L * L2;                                       // [m + 1]
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * L2 * d;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const size_t gstride = P.diag ? (size_t)(P.row_end - P.row_begin) : (size_t)(P.row_end - P.row_begin) * P.N2;
  if (tid <= m) gs[tid] = static_cast<const T*>(P.g)[(size_t)tid * gstride + pair];
  for (int e = tid; e < L * L2; e += NT) {
    const int p = e / L2, q = e - p * L2;
    Tc[e] = kappa_raw_call<T>(x + (size_t)p * d, y + (size_t)q * d, d, P.base, p0, p1);
  }
  __syncthreads();
  for (int e = tid; e < cells; e += NT) {
    const int p = e / QC, q = e - p * QC;
    const T* M0 = Tc + (size_t)p * L2 + q;
    Dt[e] = diff ? (M0[L2 + 1] - M0[L2]) - (M0[1] - M0[0]) : M0[0];  // signature_algs.py:26
  }
  __syncthreads();
  // Tc <- 2-D exclusive prefix (dir = +1) or suffix (dir = -1) sums of D * Tc
  auto step_level = [&](int dir) {
    for (int q = tid; q < QC; q += NT) {                             // down the rows
      T run = T(0);
      // the next row's operands are read before this row's result is stored (the compiler cannot hoist them itself)
      if (dir > 0) {
        T dn = Dt[q], tn = Tc[q];
        for (int p = 0; p < PR; ++p) {
          const int e = p * QC + q;
          const T v = dn * tn;
          if (p + 1 < PR) { dn = Dt[e + QC]; tn = Tc[e + QC]; }
          Tc[e] = run; run += v;
        }
      } else {
        T dn = Dt[(PR - 1) * QC + q], tn = Tc[(PR - 1) * QC + q];
        for (int p = PR - 1; p >= 0; --p) {
          const int e = p * QC + q;
          const T v = dn * tn;
          if (p > 0) { dn = Dt[e - QC]; tn = Tc[e - QC]; }
          Tc[e] = run; run += v;
        }
      }
    }
    __syncthreads();
    if (QC & 1) {                                                    // along the rows, one thread per row: an odd row
      for (int p = tid; p < PR; p += NT) {                           // stride keeps the 32 rows of a warp on distinct banks
        T* row = Tc + (size_t)p * QC;
        T run = T(0);
        if (dir > 0) { T nxt = row[0]; for (int q = 0; q < QC; ++q) { const T v = nxt; if (q + 1 < QC) nxt = row[q + 1]; row[q] = run; run += v; } }
        else { T nxt = row[QC - 1]; for (int q = QC - 1; q >= 0; --q) { const T v = nxt; if (q > 0) nxt = row[q - 1]; row[q] = run; run += v; } }
      }
      __syncthreads();
      return;
    }
    const int nch = (QC + 31) >> 5;
    for (int p = warp; p < PR; p += NT / 32) {                       // even stride: one warp per row, warp scans
      T carry = T(0), tot;
      T* row = Tc + (size_t)p * QC;
      if (dir > 0) for (int c = 0; c < nch; ++c) {
        const int q = c * 32 + lane;
        const T ex = warp_excl_prefix1<T>(q < QC ? row[q] : T(0), lane, tot);
        if (q < QC) row[q] = carry + ex;
        carry += tot;
      } else for (int c = nch - 1; c >= 0; --c) {
        const int q = c * 32 + lane;
        const T ex = warp_excl_suffix1<T>(q < QC ? row[q] : T(0), lane, tot);
        if (q < QC) row[q] = carry + ex;
        carry += tot;
      }
    }
    __syncthreads();
  };
  // ---- forward: F_a, a = 1..m-1 ---------------------------------------------------------------
  T* Fg = static_cast<T*>(P.F) + pair * (size_t)(m - 1) * cells;
  for (int e = tid; e < cells; e += NT) Tc[e] = T(1);
  __syncthreads();
  for (int a = 1; a < m; ++a) {
    step_level(+1);
    for (int e = tid; e < cells; e += NT) Fg[(size_t)(a - 1) * cells + e] = Tc[e];
    __syncthreads();
  }
  // ---- backward: G_b on the fly, W = sum_{a + b + 1 <= m} g[a + b + 1] F_a G_b -------------------
  T w[kVjpCtaCells];
#pragma unroll
  for (int k = 0; k < kVjpCtaCells; ++k) w[k] = T(0);
  for (int e = tid; e < cells; e += NT) Tc[e] = T(1);
  __syncthreads();
  for (int b = 0; b < m; ++b) {
#pragma unroll
    for (int k = 0; k < kVjpCtaCells; ++k) {
      const int e = tid + k * NT;
      if (e < cells) {
        T h = gs[b + 1];                                             // a = 0: F_0 = 1
#pragma unroll
        for (int a = 1; a < 8; ++a)                                  // m <= 8; branch-free so the loads of all cells overlap
          if (a + b + 1 <= m) h = gfma<T>(gs[a + b + 1], Fg[(size_t)(a - 1) * cells + e], h);
        w[k] = gfma<T>(Tc[e], h, w[k]);
      }
    }
    if (b + 1 < m) { __syncthreads(); step_level(-1); }
  }
  T* W = static_cast<T*>(P.W);
#pragma unroll
  for (int k = 0; k < kVjpCtaCells; ++k) {
    const int e = tid + k * NT;
    if (e < cells) W[pair * (size_t)cells + e] = w[k];                 // cell-fastest: coalesced, read back the same way
  }
}

template <typename T>
static bool launch_trunc_vjp_tables_cta(const TruncVjpParams& P, cudaStream_t s) {
  const size_t table = (size_t)P.L * P.L2;
  if (table > (size_t)kVjpCtaThreads * kVjpCtaCells || P.m > 8) return false;
  const size_t smem = (2 * table + 16) * sizeof(T);
  if (smem > 220 * 1024) return false;
  auto* kern = trunc_vjp_tables_cta_kernel<T>;
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
  kern<<<(unsigned)P.npairs, kVjpCtaThreads, smem, s>>>(P);
  return true;
}

template <typename T>
static bool launch_trunc_vjp_tables_warp(const TruncVjpParams& P, cudaStream_t s) {
  const int QC = P.L2 - (P.difference ? 1 : 0);
  const int cw = (QC + 31) / 32, mlv = P.m - 1;
  if (cw > 8 || mlv > 7) return false;
  const unsigned grid = (unsigned)((P.npairs + 3) / 4);
#define GPSIG_VW(CW_, MLV_) trunc_vjp_tables_warp_kernel<T, CW_, MLV_><<<grid, 128, 0, s>>>(P)
  if (mlv <= 3) { if (cw <= 2) GPSIG_VW(2, 3); else if (cw <= 4) GPSIG_VW(4, 3); else GPSIG_VW(8, 3); }
  else { if (cw <= 2) GPSIG_VW(2, 7); else if (cw <= 4) GPSIG_VW(4, 7); else GPSIG_VW(8, 7); }
#undef GPSIG_VW
  return true;
}

// sweep 3: one thread per (pair, point): r < L -> point x_r of the pair, else point y_{r-L}
template <typename T>
__global__ void trunc_vjp_grad_kernel(const TruncVjpParams P) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t np = P.npairs;
  const int LL = P.L + P.L2;
  if (t >= np * LL) return;
  const int r = (int)(t / np);
  const size_t pair = t % np;
  const int d = P.d, diff = P.difference ? 1 : 0;
  const int PR = P.L - diff, QC = P.L2 - diff;
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * P.L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * P.L2 * d;
  const T* W = static_cast<const T*>(P.W);
  const T p0 = T(P.p0), p1 = T(P.p1);
  auto Wat = [&](int p, int q) -> T { return (p < 0 || q < 0 || p >= PR || q >= QC) ? T(0) : W[((size_t)p * QC + q) * np + pair]; };
  // dL/dM[pp, qq]: adjoint of the 2-D difference
  auto dM = [&](int pp, int qq) -> T {
    if (!diff) return Wat(pp, qq);
    return (Wat(pp - 1, qq - 1) - Wat(pp - 1, qq)) - (Wat(pp, qq - 1) - Wat(pp, qq));
  };
  // d > 16: the coordinates are taken 16 at a time (the partials are re-evaluated per chunk; the large-d training
  // shapes are small, the forward Gram at large d runs the lifted path)
  for (int k0 = 0; k0 < d; k0 += 16) {
  const int kd = min(16, d - k0);
  T acc[16];
  for (int k = 0; k < kd; ++k) acc[k] = T(0);
  if (r < P.L) {
    const int pp = r;
    const T* xp = x + (size_t)pp * d;
    for (int qq = 0; qq < P.L2; ++qq) {
      const T a = dM(pp, qq);
      if (a == T(0)) continue;
      const T* yq = y + (size_t)qq * d;
      T dot, xs, ys, fd, fx, fy;
      dot3(xp, yq, d, dot, xs, ys);
      base_partials<T>(P.base, p0, p1, dot, xs, ys, fd, fx, fy);
      for (int k = 0; k < kd; ++k) acc[k] = gfma<T>(a, fd * yq[k0 + k] + T(2) * fx * xp[k0 + k], acc[k]);
    }
    T* gx = static_cast<T*>(P.gradX) + ((size_t)i * P.L + pp) * d + k0;
    for (int k = 0; k < kd; ++k) atomic_add<T>(gx + k, acc[k]);
  } else {
    const int qq = r - P.L;
    const T* yq = y + (size_t)qq * d;
    for (int pp = 0; pp < P.L; ++pp) {
      const T a = dM(pp, qq);
      if (a == T(0)) continue;
      const T* xp = x + (size_t)pp * d;
      T dot, xs, ys, fd, fx, fy;
      dot3(xp, yq, d, dot, xs, ys);
      base_partials<T>(P.base, p0, p1, dot, xs, ys, fd, fx, fy);
      for (int k = 0; k < kd; ++k) acc[k] = gfma<T>(a, fd * xp[k0 + k] + T(2) * fy * yq[k0 + k], acc[k]);
    }
    T* gy = static_cast<T*>(P.gradY) + ((size_t)j * P.L2 + qq) * d + k0;
    for (int k = 0; k < kd; ++k) atomic_add<T>(gy + k, acc[k]);
  }
  }
}

// the same with one warp per (pair, point): for few pairs (training shapes), where a thread per point leaves the GPU idle
template <typename T>
__global__ void __launch_bounds__(256) trunc_vjp_grad_warp_kernel(const TruncVjpParams P) {
  const size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;   // one warp per (pair, point); lanes stride over the other sequence
  const int lane = threadIdx.x & 31;
  const size_t np = P.npairs;
  const int LL = P.L + P.L2;
  if (t >= np * LL) return;
  // cell-major W: the 8 warps of a CTA take 8 consecutive points of one pair (its W rows are shared through L1 / L2)
  const int r = P.w_cell_major ? (int)(t % LL) : (int)(t / np);
  const size_t pair = P.w_cell_major ? t / LL : t % np;
  const int d = P.d, diff = P.difference ? 1 : 0;
  const int PR = P.L - diff, QC = P.L2 - diff;
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * P.L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * P.L2 * d;
  const T* W = static_cast<const T*>(P.W);
  const T p0 = T(P.p0), p1 = T(P.p1);
  auto Wat = [&](int p, int q) -> T { return (p < 0 || q < 0 || p >= PR || q >= QC) ? T(0) : (P.w_cell_major ? W[pair * ((size_t)PR * QC) + (size_t)p * QC + q] : W[((size_t)p * QC + q) * np + pair]); };
  // dL/dM[pp, qq]: adjoint of the 2-D difference
  auto dM = [&](int pp, int qq) -> T {
    if (!diff) return Wat(pp, qq);
    return (Wat(pp - 1, qq - 1) - Wat(pp - 1, qq)) - (Wat(pp, qq - 1) - Wat(pp, qq));
  };
  for (int k0 = 0; k0 < d; k0 += 16) {                               // d > 16: 16 coordinates at a time
  T acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = T(0);
  if (r < P.L) {
    const int pp = r;
    const T* xp = x + (size_t)pp * d;
    for (int qq = lane; qq < P.L2; qq += 32) {
      const T a = dM(pp, qq);
      if (a == T(0)) continue;
      const T* yq = y + (size_t)qq * d;
      T dot, xs, ys, fd, fx, fy;
      dot3(xp, yq, d, dot, xs, ys);
      base_partials_call<T>(P.base, p0, p1, dot, xs, ys, fd, fx, fy);
#pragma unroll
      for (int k = 0; k < 16; ++k) if (k0 + k < d) acc[k] = gfma<T>(a, fd * yq[k0 + k] + T(2) * fx * xp[k0 + k], acc[k]);
    }
    T* gx = static_cast<T*>(P.gradX) + ((size_t)i * P.L + pp) * d + k0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (k0 + k >= d) continue;                                          // warp-uniform
      T v = acc[k];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v != T(0)) atomic_add<T>(gx + k, v);
    }
  } else {
    const int qq = r - P.L;
    const T* yq = y + (size_t)qq * d;
    for (int pp = lane; pp < P.L; pp += 32) {
      const T a = dM(pp, qq);
      if (a == T(0)) continue;
      const T* xp = x + (size_t)pp * d;
      T dot, xs, ys, fd, fx, fy;
      dot3(xp, yq, d, dot, xs, ys);
      base_partials_call<T>(P.base, p0, p1, dot, xs, ys, fd, fx, fy);
#pragma unroll
      for (int k = 0; k < 16; ++k) if (k0 + k < d) acc[k] = gfma<T>(a, fd * xp[k0 + k] + T(2) * fy * yq[k0 + k], acc[k]);
    }
    T* gy = static_cast<T*>(P.gradY) + ((size_t)j * P.L2 + qq) * d + k0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (k0 + k >= d) continue;                                          // warp-uniform
      T v = acc[k];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v != T(0)) atomic_add<T>(gy + k, v);
    }
  }
}
}

// ---------------------------------------------------------------------------
// inducing tensors vs sequences
// ---------------------------------------------------------------------------
struct TensSeqVjpParams {
  const void* Z; const void* X;      // dense [T, Mz, zs, d], [N, L, d]
  int Mz, zs, N, L, d, m;
  int col_begin, col_end;
  int difference;
  double p0, p1; int base;
  const void* g;                     // [m+1, Mz, C]
  void* gradZ; void* gradX;
  void* S;                           // scratch [T, L, npairs]: running sums S_k(l) of the forward recursion
  size_t npairs;
  int kc0, kcn;                      // coordinates [kc0, kc0 + kcn) of this launch (d > 16 runs the reverse sweep in chunks; kcn = 0: all)
  int order;                         // > 1: higher-order recursion (signature_algs.py:129-160), thread-per-pair kernel only
};

// Forward (tens_level in tens_kernels.cuh): within level I, components jj = 0..I-1 (global index k),
//   S_jj(l) = S_jj(l-1) + D_jj(l) * S_{jj-1}(l-1)   (S_{-1} = 1),  K_I = S_{I-1}(L-1).
// Reverse: lam_jj(l-1) += lam_jj(l);  aD_jj(l) = lam_jj(l) * S_{jj-1}(l-1);  lam_{jj-1}(l-1) += lam_jj(l) * D_jj(l).
constexpr int kTsvChunk = 64;                         // coordinates per launch at d > 16
template <typename T, int GZ>                         // GZ >= T * zs * (coordinates of this launch): per-thread accumulators of dL/dZ
__global__ void tens_seq_vjp_kernel(const TensSeqVjpParams P) {
  const size_t pair = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= P.npairs) return;
  const size_t np = P.npairs;
  const int ncols = P.col_end - P.col_begin;
  const int z = (int)(pair / ncols), nl = (int)(pair % ncols), n = P.col_begin + nl;
  const int m = P.m, d = P.d, zs = P.zs, L = P.L, TT = m * (m + 1) / 2;
  const bool diff = P.difference != 0;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* Z = static_cast<const T*>(P.Z);
  const T* x = static_cast<const T*>(P.X) + (size_t)n * L * d;
  T* S = static_cast<T*>(P.S);
  auto zpt = [&](int k, int s) { return Z + (((size_t)k * P.Mz + z) * zs + s) * d; };
  auto Mt = [&](int k, int l) -> T {
    const T* xl = x + (size_t)l * d;
    if (zs == 2) return kappa_raw<T>(zpt(k, 1), xl, d, P.base, p0, p1) - kappa_raw<T>(zpt(k, 0), xl, d, P.base, p0, p1);   // kernels.py:331
    return kappa_raw<T>(zpt(k, 0), xl, d, P.base, p0, p1);
  };
  auto Sat = [&](int k, int l) -> T& { return S[((size_t)k * L + l) * np + pair]; };
  const int l0 = diff ? 1 : 0;                                       // first time index that carries a D
  const int kc0 = P.kcn > 0 ? P.kc0 : 0, kcn = P.kcn > 0 ? P.kcn : d;
  // ---- forward: store S_k(l) (first chunk only: the scratch persists across the launches of one call) ----------
  const int order = P.order > 1 ? P.order : 1;
  if (kc0 == 0 && order > 1) {
    // higher order (signature_algs.py:140-158): component jj of a level carries a family cur[jj][0..min(jj+1,order)-1];
    // stored: S_k(l) = sum over EARLIER times of the component's family total (what cur[jj+1][0] multiplies, :153)
    T Sv[36], Mprev[36];
    for (int k = 0; k < TT; ++k) { Sv[k] = T(0); Mprev[k] = T(0); }
    for (int l = 0; l < L; ++l) {
      for (int k = 0; k < TT; ++k) Sat(k, l) = Sv[k];
      int k0 = 0;
      for (int I = 1; I <= m; ++I) {
        T cur[8][8];
        for (int jj = 0; jj < I; ++jj) {
          const int k = k0 + jj;
          const T mt = Mt(k, l);
          T D = mt;
          if (diff) { D = mt - Mprev[k]; Mprev[k] = mt; }
          if (l < l0) continue;
          const int dj = min(jj + 1, order);
          cur[jj][0] = jj == 0 ? D : D * Sat(k - 1, l);
          for (int c = 1; c < dj; ++c) cur[jj][c] = D * cur[jj - 1][c - 1] / T(c + 1);              // :155
        }
        if (l >= l0)
          for (int jj = 0; jj + 1 < I; ++jj) {
            T sm = T(0);
            for (int c = 0; c < min(jj + 1, order); ++c) sm += cur[jj][c];
            Sv[k0 + jj] += sm;
          }
        k0 += I;
      }
    }
  } else if (kc0 == 0) {
    T Sv[36], Mprev[36];
    for (int k = 0; k < TT; ++k) { Sv[k] = T(0); Mprev[k] = T(0); }
    for (int l = 0; l < L; ++l) {
      int k0 = 0;
      for (int I = 1; I <= m; ++I) {
        for (int jj = I - 1; jj >= 0; --jj) {
          const int k = k0 + jj;
          const T mt = Mt(k, l);
          T D = mt;
          if (diff) { D = mt - Mprev[k]; Mprev[k] = mt; }
          if (l >= l0) Sv[k] = jj > 0 ? gfma<T>(D, Sv[k - 1], Sv[k]) : Sv[k] + D;
        }
        k0 += I;
      }
      for (int k = 0; k < TT; ++k) Sat(k, l) = Sv[k];
    }
  }
  // ---- reverse -----------------------------------------------------------------------------------
  T lam[36], aDnext[36], gz[GZ];
  for (int k = 0; k < TT; ++k) { lam[k] = T(0); aDnext[k] = T(0); }
  for (int e = 0; e < TT * zs * kcn; ++e) gz[e] = T(0);
  {
    int k0 = 0;
    for (int I = 1; I <= m; ++I) { lam[k0 + I - 1] = static_cast<const T*>(P.g)[((size_t)I * P.Mz + z) * ncols + nl]; k0 += I; }
  }
  T Aacc[36];                                                       // order > 1: adjoint of S_k, summed over LATER times
  for (int k = 0; k < TT; ++k) Aacc[k] = T(0);
  for (int l = L - 1; l >= 0; --l) {
    const T* xl = x + (size_t)l * d;
    T gx[kTsvChunk];
    for (int q = 0; q < kcn; ++q) gx[q] = T(0);
    T aDk[36];
    if (order > 1) {
      int kq = 0;
      for (int I = 1; I <= m; ++I) {
        if (l < l0) { for (int jj = 0; jj < I; ++jj) aDk[kq + jj] = T(0); kq += I; continue; }
        const T gI = static_cast<const T*>(P.g)[((size_t)I * P.Mz + z) * ncols + nl];
        T Dl[8], cur[8][8], cb[8][8];
        for (int jj = 0; jj < I; ++jj) Dl[jj] = diff ? Mt(kq + jj, l) - Mt(kq + jj, l - 1) : Mt(kq + jj, l);
        for (int jj = 0; jj < I; ++jj) {                              // the forward families at this time, from the stored sums
          const int dj = min(jj + 1, order);
          cur[jj][0] = jj == 0 ? Dl[0] : Dl[jj] * Sat(kq + jj - 1, l);
          for (int c = 1; c < dj; ++c) cur[jj][c] = Dl[jj] * cur[jj - 1][c - 1] / T(c + 1);
        }
        for (int jj = I - 1; jj >= 0; --jj) {
          const int dj = min(jj + 1, order);
          const T sbar = jj == I - 1 ? gI : Aacc[kq + jj];
          for (int c = 0; c < dj; ++c) {
            T v = sbar;
            if (jj + 1 < I && c + 1 < min(jj + 2, order)) v += cb[jj + 1][c + 1] * Dl[jj + 1] / T(c + 2);
            cb[jj][c] = v;
          }
          T aD = jj == 0 ? cb[0][0] : cb[jj][0] * Sat(kq + jj - 1, l);
          for (int c = 1; c < dj; ++c) aD += cb[jj][c] * cur[jj - 1][c - 1] / T(c + 1);
          aDk[kq + jj] = aD;
        }
        for (int jj = 1; jj < I; ++jj) Aacc[kq + jj - 1] += cb[jj][0] * Dl[jj];
        kq += I;
      }
    }
    int k0 = 0;
    for (int I = 1; I <= m; ++I) {
      for (int jj = 0; jj < I; ++jj) {
        const int k = k0 + jj;
        // adjoint of D_k(l) (zero where the forward skipped the update)
        T aD = T(0);
        if (order > 1) aD = aDk[k];
        else if (l >= l0) {
          const T Sprev = jj > 0 ? (l > 0 ? Sat(k - 1, l - 1) : T(0)) : T(1);
          aD = lam[k] * Sprev;
        }
        // adjoint of Mt_k(l): D(l) = Mt(l) - Mt(l-1), D(l+1) = Mt(l+1) - Mt(l)
        const T aM = diff ? aD - aDnext[k] : aD;
        aDnext[k] = aD;
        if (aM != T(0)) {
          for (int s = 0; s < zs; ++s) {
            const T sgn = (zs == 2 && s == 0) ? T(-1) : T(1);
            const T* zp = zpt(k, s);
            T dot, zsq, xsq, fd, fz, fx;
            dot3(zp, xl, d, dot, zsq, xsq);
            base_partials<T>(P.base, p0, p1, dot, zsq, xsq, fd, fz, fx);
            const T w = sgn * aM;
            for (int q = 0; q < kcn; ++q) {
              gz[(k * zs + s) * kcn + q] = gfma<T>(w, fd * xl[kc0 + q] + T(2) * fz * zp[kc0 + q], gz[(k * zs + s) * kcn + q]);
              gx[q] = gfma<T>(w, fd * zp[kc0 + q] + T(2) * fx * xl[kc0 + q], gx[q]);
            }
          }
        }
      }
      k0 += I;
    }
    // propagate lam to time l-1 (needs D_k(l), highest component first within a level)
    if (order == 1 && l >= l0) {
      k0 = 0;
      for (int I = 1; I <= m; ++I) {
        for (int jj = 1; jj < I; ++jj) {
          const int k = k0 + jj;
          const T D = diff ? Mt(k, l) - Mt(k, l - 1) : Mt(k, l);
          lam[k - 1] = gfma<T>(lam[k], D, lam[k - 1]);
        }
        k0 += I;
      }
    }
    T* gxo = static_cast<T*>(P.gradX) + ((size_t)n * L + l) * d + kc0;
    for (int q = 0; q < kcn; ++q) if (gx[q] != T(0)) atomic_add<T>(gxo + q, gx[q]);
  }
  for (int k = 0; k < TT; ++k)
    for (int s = 0; s < zs; ++s) {
      T* gzo = static_cast<T*>(P.gradZ) + (((size_t)k * P.Mz + z) * zs + s) * d + kc0;
      for (int q = 0; q < kcn; ++q) atomic_add<T>(gzo + q, gz[(k * zs + s) * kcn + q]);
    }
}

// ---------------------------------------------------------------------------
// inducing tensors vs sequences, training shapes: one warp per (sequence n, chunk of inducing tensors).
// Lane owns CW consecutive time steps of x_n (registers); the recursion over time becomes warp scans:
//   E_jj(l)   = S_{jj-1}(l-1) = sum_{l' < l} D_{jj-1}(l') E_{jj-1}(l')      (E_0 = 1)
//   lam_jj(l) = sum_{l' > l} lam_{jj+1}(l') D_{jj+1}(l')                      (lam_{I-1} = g_I)
//   aD_jj(l)  = lam_jj(l) E_jj(l)
// so no scratch in HBM is needed; D / E and the base-kernel values of the current level live in shared memory as
// [jj][c][lane] (for the RBF kernel the partial derivatives follow from the stored value: one exp per element, not two).
// dL/dx accumulates in registers over the warp's tensors, dL/dz is reduced over lanes per (component, sign).
// ---------------------------------------------------------------------------
constexpr int kTsvWarps = 4;

template <typename T, int CW>
__device__ __forceinline__ void warp_excl_prefix(T (&v)[CW], int lane) {
  T tot = T(0);
#pragma unroll
  for (int c = 0; c < CW; ++c) { const T t = v[c]; v[c] = tot; tot += t; }
  T inc = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
  T exc = __shfl_up_sync(0xffffffffu, inc, 1);
  if (lane == 0) exc = T(0);
#pragma unroll
  for (int c = 0; c < CW; ++c) v[c] += exc;
}
template <typename T, int CW>
__device__ __forceinline__ void warp_excl_suffix(T (&v)[CW], int lane) {
  T tot = T(0);
#pragma unroll
  for (int c = CW - 1; c >= 0; --c) { const T t = v[c]; v[c] = tot; tot += t; }
  T inc = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_down_sync(0xffffffffu, inc, o); if (lane + o < 32) inc += t; }
  T exc = __shfl_down_sync(0xffffffffu, inc, 1);
  if (lane == 31) exc = T(0);
#pragma unroll
  for (int c = 0; c < CW; ++c) v[c] += exc;
}
// sum of v[i] over the 32 lanes for all i < NV with NV-1 + (5 - log2 NV) shuffles; the lane ends up with entry `idx`
template <typename T, int NV>
__device__ __forceinline__ T warp_transpose_reduce(T (&v)[NV], int lane, int& idx) {
  int off = 16;
  idx = 0;
#pragma unroll
  for (int n = NV; n > 1; n >>= 1) {
    const bool hi = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < n / 2; ++i) {
      const T send = hi ? v[i] : v[i + n / 2];
      const T keep = hi ? v[i + n / 2] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
    if (hi) idx += n / 2;
    off >>= 1;
  }
  for (; off > 0; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
  return v[0];
}
// ---------------------------------------------------------------------------
// gradient kernel, one CTA per pair (after the CTA-per-pair sweeps, cell-major W, L2 <= 128): warp = a point of x,
// lane = up to CQ points of y held in registers.  The base-kernel partials are evaluated ONCE per lifted element and
// feed both sides: dL/dx_p is reduced over the lanes per point, dL/dy_q accumulates in the lane's registers over the
// warp's points and is reduced over the 8 warps through shared memory.
// ---------------------------------------------------------------------------
template <typename T, int DP, int CQ>
__global__ void __launch_bounds__(256) trunc_vjp_grad_cta_kernel(const TruncVjpParams P) {
  extern __shared__ __align__(16) unsigned char vjp_grad_smem[];
  T* red = reinterpret_cast<T*>(vjp_grad_smem);                      // [8 warps][CQ * 32][DP]
  const size_t pair = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int d = P.d, diff = P.difference ? 1 : 0, L = P.L, L2 = P.L2;
  const int PR = L - diff, QC = L2 - diff;
  int i, j;
  if (P.diag) { i = P.row_begin + (int)pair; j = i; }
  else { i = P.row_begin + (int)(pair / P.N2); j = (int)(pair % P.N2); }
  const T* x = static_cast<const T*>(P.X) + (size_t)i * L * d;
  const T* y = static_cast<const T*>(P.Y) + (size_t)j * L2 * d;
  const T* W = static_cast<const T*>(P.W) + pair * ((size_t)PR * QC);
  const T p0 = T(P.p0), p1 = T(P.p1);
  auto Wat = [&](int p, int q) -> T { return (p < 0 || q < 0 || p >= PR || q >= QC) ? T(0) : W[(size_t)p * QC + q]; };
  T yr[CQ][DP], ysq[CQ], gy[CQ][DP], by[CQ];
#pragma unroll
  for (int c = 0; c < CQ; ++c) {
    const int q = lane + 32 * c;
    ysq[c] = by[c] = T(0);
#pragma unroll
    for (int k = 0; k < DP; ++k) {
      yr[c][k] = (q < L2 && k < d) ? y[(size_t)q * d + k] : T(0);
      ysq[c] = gfma<T>(yr[c][k], yr[c][k], ysq[c]);
      gy[c][k] = T(0);
    }
  }
  for (int p = warp; p < L; p += 8) {
    T xr[DP], gx[DP], xsq = T(0), bx = T(0);
#pragma unroll
    for (int k = 0; k < DP; ++k) { xr[k] = k < d ? x[(size_t)p * d + k] : T(0); xsq = gfma<T>(xr[k], xr[k], xsq); gx[k] = T(0); }
#pragma unroll
    for (int c = 0; c < CQ; ++c) {
      const int q = lane + 32 * c;
      if (q >= L2) continue;
      const T a = diff ? (Wat(p - 1, q - 1) - Wat(p - 1, q)) - (Wat(p, q - 1) - Wat(p, q)) : Wat(p, q);   // adjoint of the 2-D difference
      if (a == T(0)) continue;
      T dot = T(0), fd, fx, fy;
#pragma unroll
      for (int k = 0; k < DP; ++k) dot = gfma<T>(xr[k], yr[c][k], dot);
      base_partials_call<T>(P.base, p0, p1, dot, xsq, ysq[c], fd, fx, fy);
      const T A = a * fd;
      bx = gfma<T>(a, fx, bx);
      by[c] = gfma<T>(a, fy, by[c]);
#pragma unroll
      for (int k = 0; k < DP; ++k) { gx[k] = gfma<T>(A, yr[c][k], gx[k]); gy[c][k] = gfma<T>(A, xr[k], gy[c][k]); }
    }
#pragma unroll
    for (int k = 0; k < DP; ++k) gx[k] = gfma<T>(T(2) * bx, xr[k], gx[k]);
    int idx;
    const T tot = warp_transpose_reduce<T, DP>(gx, lane, idx);
    if ((lane & (32 / DP - 1)) == 0 && idx < d && tot != T(0))
      atomic_add<T>(static_cast<T*>(P.gradX) + ((size_t)i * L + p) * d + idx, tot);
  }
#pragma unroll
  for (int c = 0; c < CQ; ++c)
#pragma unroll
    for (int k = 0; k < DP; ++k)
      red[((size_t)warp * (CQ * 32) + c * 32 + lane) * DP + k] = gfma<T>(T(2) * by[c], yr[c][k], gy[c][k]);
  __syncthreads();
  for (int e = tid; e < CQ * 32 * DP; e += 256) {
    const int q = e / DP, k = e - q * DP;
    if (q >= L2 || k >= d) continue;
    T v = T(0);
#pragma unroll
    for (int w = 0; w < 8; ++w) v += red[((size_t)w * (CQ * 32) + q) * DP + k];
    if (v != T(0)) atomic_add<T>(static_cast<T*>(P.gradY) + ((size_t)j * L2 + q) * d + k, v);
  }
}

template <typename T, int DP, int CQ>
static bool launch_vjp_grad_cta_inst(const TruncVjpParams& P, cudaStream_t s) {
  const size_t smem = (size_t)8 * CQ * 32 * DP * sizeof(T);
  auto* kern = trunc_vjp_grad_cta_kernel<T, DP, CQ>;
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
  kern<<<(unsigned)P.npairs, 256, smem, s>>>(P);
  return true;
}
// needs the cell-major W of the CTA-per-pair sweeps; false when y is too long for the lanes' registers
template <typename T>
static bool launch_trunc_vjp_grad_cta(const TruncVjpParams& P, cudaStream_t s) {
  if (!P.w_cell_major || P.L2 > 128 || P.d > 16) return false;
  const int cq = P.L2 <= 32 ? 1 : (P.L2 <= 64 ? 2 : 4);
  if (P.d <= 8) {
    if (cq == 1) return launch_vjp_grad_cta_inst<T, 8, 1>(P, s);
    if (cq == 2) return launch_vjp_grad_cta_inst<T, 8, 2>(P, s);
    return launch_vjp_grad_cta_inst<T, 8, 4>(P, s);
  }
  if (cq == 1) return launch_vjp_grad_cta_inst<T, 16, 1>(P, s);
  if (cq == 2) return launch_vjp_grad_cta_inst<T, 16, 2>(P, s);
  return false;
}

template <typename T, int CW, int DP>
__global__ void __launch_bounds__(kTsvWarps * 32, 2) tens_seq_vjp_warp_kernel(const TensSeqVjpParams P, int zchunks, int zper) {
  extern __shared__ __align__(16) unsigned char tsv_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m = P.m, d = P.d, zs = P.zs, L = P.L, Mz = P.Mz;
  const int ncols = P.col_end - P.col_begin;
  const long wid = (long)blockIdx.x * kTsvWarps + warp;
  if (wid >= (long)ncols * zchunks) return;                         // warp-uniform; the kernel has no block barrier
  const int nl = (int)(wid / zchunks), zc = (int)(wid % zchunks), n = P.col_begin + nl;
  const int z_lo = zc * zper, z_hi = min(Mz, z_lo + zper);
  const bool diff = P.difference != 0;
  const int l0 = diff ? 1 : 0;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* Z = static_cast<const T*>(P.Z);
  const T* g = static_cast<const T*>(P.g);
  T* Ds = reinterpret_cast<T*>(tsv_smem) + (size_t)warp * 4 * m * CW * 32;
  T* Es = Ds + (size_t)m * CW * 32;
  T* Ks = Es + (size_t)m * CW * 32;                                  // forward base-kernel values [jj][s][c][lane]
  const bool from_value = P.base == GPSIG_BASE_RBF;                  // partials follow from the value: no second exp
  const double tab = exp2((double)lane * (1.0 / 32.0));             // exp2_tab32's table, one entry per lane
  (void)tab;
  auto at = [&](T* base, int jj, int c) -> T& { return base[(jj * CW + c) * 32 + lane]; };

  T xr[CW][DP], xsq[CW], gx[CW][DP], gxb[CW];
  {
    const T* x = static_cast<const T*>(P.X) + (size_t)n * L * d;
#pragma unroll
    for (int c = 0; c < CW; ++c) {
      const int l = lane * CW + c;
      xsq[c] = T(0); gxb[c] = T(0);
#pragma unroll
      for (int q = 0; q < DP; ++q) {
        xr[c][q] = (l < L && q < d) ? x[(size_t)l * d + q] : T(0);
        xsq[c] = gfma<T>(xr[c][q], xr[c][q], xsq[c]);
        gx[c][q] = T(0);
      }
    }
  }
  for (int z = z_lo; z < z_hi; ++z) {
    int k0 = 0;
    for (int I = 1; I <= m; k0 += I, ++I) {
      const T gI = g[((size_t)I * Mz + z) * ncols + nl];
      if (gI == T(0)) continue;
      // ---- forward: D_jj(l) of this level's components ------------------------------------------
      for (int jj = 0; jj < I; ++jj) {
        T mt[CW];
#pragma unroll
        for (int c = 0; c < CW; ++c) mt[c] = T(0);
        for (int s = 0; s < zs; ++s) {
          const T* zp = Z + (((size_t)(k0 + jj) * Mz + z) * zs + s) * d;
          T zr[DP], zsq = T(0);
#pragma unroll
          for (int q = 0; q < DP; ++q) { zr[q] = q < d ? zp[q] : T(0); zsq = gfma<T>(zr[q], zr[q], zsq); }
          const bool neg = zs == 2 && s == 0;                        // kernels.py:331
#pragma unroll
          for (int c = 0; c < CW; ++c) {
            T dot = T(0);
#pragma unroll
            for (int q = 0; q < DP; ++q) dot = gfma<T>(zr[q], xr[c][q], dot);
            T kv;
            if constexpr (sizeof(T) == 8) {                          // float64 RBF: the table exp2 of the forward kernels (9 FP64 instructions)
              if (from_value) kv = exp2_tab32(((T(-2) * dot + zsq) + xsq[c]) * T(-0.5 * 1.4426950408889634), tab);   // exp(-r^2 / 2) = 2^(-r^2 log2(e) / 2)
              else kv = base_eval_call<T>(P.base, p0, p1, dot, zsq, xsq[c]);
            } else {
              kv = base_eval_call<T>(P.base, p0, p1, dot, zsq, xsq[c]);
            }
            Ks[((jj * 2 + s) * CW + c) * 32 + lane] = kv;
            mt[c] += neg ? -kv : kv;
          }
        }
        const T prev = __shfl_up_sync(0xffffffffu, mt[CW - 1], 1);
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          const int l = lane * CW + c;
          const T pm = c > 0 ? mt[c > 0 ? c - 1 : 0] : prev;
          at(Ds, jj, c) = (l >= l0 && l < L) ? (diff ? mt[c] - pm : mt[c]) : T(0);
        }
      }
      // ---- forward: E_jj(l) = S_{jj-1}(l-1) -----------------------------------------------------
      {
        T e[CW];
#pragma unroll
        for (int c = 0; c < CW; ++c) { e[c] = T(1); at(Es, 0, c) = T(1); }
        for (int jj = 1; jj < I; ++jj) {
#pragma unroll
          for (int c = 0; c < CW; ++c) e[c] *= at(Ds, jj - 1, c);
          warp_excl_prefix<T, CW>(e, lane);
#pragma unroll
          for (int c = 0; c < CW; ++c) at(Es, jj, c) = e[c];
        }
      }
      // ---- reverse: lam, aD, aM, gradients ------------------------------------------------------
      T lam[CW];
#pragma unroll
      for (int c = 0; c < CW; ++c) lam[c] = gI;
      for (int jj = I - 1; jj >= 0; --jj) {
        T aM[CW];
#pragma unroll
        for (int c = 0; c < CW; ++c) {
          const int l = lane * CW + c;
          aM[c] = (l >= l0 && l < L) ? lam[c] * at(Es, jj, c) : T(0);
        }
        if (jj > 0) {
#pragma unroll
          for (int c = 0; c < CW; ++c) lam[c] *= at(Ds, jj, c);
          warp_excl_suffix<T, CW>(lam, lane);
        }
        if (diff) {                                                  // D(l) = Mt(l) - Mt(l-1)
          T nxt = __shfl_down_sync(0xffffffffu, aM[0], 1);
          if (lane == 31) nxt = T(0);
#pragma unroll
          for (int c = 0; c < CW; ++c) aM[c] -= (c + 1 < CW) ? aM[c + 1 < CW ? c + 1 : c] : nxt;
        }
        for (int s = 0; s < zs; ++s) {
          const T* zp = Z + (((size_t)(k0 + jj) * Mz + z) * zs + s) * d;
          T zr[DP], gz[DP], zsq = T(0), gzb = T(0);
#pragma unroll
          for (int q = 0; q < DP; ++q) { zr[q] = q < d ? zp[q] : T(0); zsq = gfma<T>(zr[q], zr[q], zsq); gz[q] = T(0); }
          const bool neg = zs == 2 && s == 0;
#pragma unroll
          for (int c = 0; c < CW; ++c) {
            T fd, fz, fx;
            if (from_value) {                                        // RBF: f_dot = f, f_|z|^2 = f_|x|^2 = -f / 2
              fd = Ks[((jj * 2 + s) * CW + c) * 32 + lane];
              fz = fx = T(-0.5) * fd;
            } else {
              T dot = T(0);
#pragma unroll
              for (int q = 0; q < DP; ++q) dot = gfma<T>(zr[q], xr[c][q], dot);
              base_partials_call<T>(P.base, p0, p1, dot, zsq, xsq[c], fd, fz, fx);
            }
            const T w = neg ? -aM[c] : aM[c];
            const T A = w * fd;
            gzb = gfma<T>(w, fz, gzb);
            gxb[c] = gfma<T>(w, fx, gxb[c]);
#pragma unroll
            for (int q = 0; q < DP; ++q) { gz[q] = gfma<T>(A, xr[c][q], gz[q]); gx[c][q] = gfma<T>(A, zr[q], gx[c][q]); }
          }
#pragma unroll
          for (int q = 0; q < DP; ++q) gz[q] = gfma<T>(T(2) * gzb, zr[q], gz[q]);
          int idx;
          const T tot = warp_transpose_reduce<T, DP>(gz, lane, idx);
          if ((lane & (32 / DP - 1)) == 0 && idx < d)
            atomic_add<T>(static_cast<T*>(P.gradZ) + (((size_t)(k0 + jj) * Mz + z) * zs + s) * d + idx, tot);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CW; ++c) {
    const int l = lane * CW + c;
    if (l >= L) continue;
    T* gxo = static_cast<T*>(P.gradX) + ((size_t)n * L + l) * d;
#pragma unroll
    for (int q = 0; q < DP; ++q) {
      const T v = gfma<T>(T(2) * gxb[c], xr[c][q], gx[c][q]);
      if (q < d && v != T(0)) atomic_add<T>(gxo + q, v);
    }
  }
}

template <typename T, int CW, int DP>
static bool launch_tsv_warp_inst(const TensSeqVjpParams& P, cudaStream_t s) {
  const int ncols = P.col_end - P.col_begin;
  int zchunks = (148 * 16 + ncols - 1) / ncols;
  zchunks = zchunks < 1 ? 1 : (zchunks > P.Mz ? P.Mz : zchunks);
  const int zper = (P.Mz + zchunks - 1) / zchunks;
  zchunks = (P.Mz + zper - 1) / zper;
  const size_t smem = (size_t)kTsvWarps * 4 * P.m * CW * 32 * sizeof(T);
  auto* kern = tens_seq_vjp_warp_kernel<T, CW, DP>;
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
  const long warps = (long)ncols * zchunks;
  kern<<<(unsigned)((warps + kTsvWarps - 1) / kTsvWarps), kTsvWarps * 32, smem, s>>>(P, zchunks, zper);
  return true;
}
// false when the shape does not fit the register budget (the caller falls back to one thread per pair)
template <typename T>
static bool launch_tens_seq_vjp_warp(const TensSeqVjpParams& P, cudaStream_t s) {
  const int dp = P.d <= 8 ? 8 : 16;
  const int cw = P.L <= 64 ? 2 : (P.L <= 128 ? 4 : (P.L <= 256 ? 8 : 0));
  if (cw == 0 || P.d > 16) return false;
  if (cw * dp * (int)(sizeof(T) / 4) > 64) return false;
  if (dp == 8) {
    if (cw == 2) return launch_tsv_warp_inst<T, 2, 8>(P, s);
    if (cw == 4) return launch_tsv_warp_inst<T, 4, 8>(P, s);
    if constexpr (sizeof(T) == 4) { if (cw == 8) return launch_tsv_warp_inst<T, 8, 8>(P, s); }
  } else {
    if (cw == 2) return launch_tsv_warp_inst<T, 2, 16>(P, s);
    if constexpr (sizeof(T) == 4) { if (cw == 4) return launch_tsv_warp_inst<T, 4, 16>(P, s); }
  }
  return false;
}

// ---------------------------------------------------------------------------
// inducing tensors vs themselves: one warp per (component k, row a), lanes stride over b; no atomics.
//   R_i(a, b) = prod_{jj < i} Mt_{k_i + jj}(a, b),  Mt symmetric in (a, b)
//   dL/dZ[k, a] = sum_b (g_i[a, b] + g_i[b, a]) * prod_{others} Mt * dMt_k(a, b)/dz_a
// ---------------------------------------------------------------------------
struct TensTensVjpParams {
  const void* Z; int Mz, zs, m, d;
  double p0, p1; int base;
  const void* g;                     // [m+1, Mz, Mz]
  void* gradZ;                       // [T, Mz, zs, d], overwritten
};

template <typename T, int DP>
__global__ void __launch_bounds__(128) tens_tens_vjp_kernel(const TensTensVjpParams P) {
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int k = gridDim.y - 1 - blockIdx.y;                          // global component index; the top level's components cost the most: first
  if (a >= P.Mz) return;
  const int d = P.d, zs = P.zs, Mz = P.Mz;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* Z = static_cast<const T*>(P.Z);
  int I = 1, k0 = 0;
  while (k0 + I <= k) { k0 += I; ++I; }                               // level of component k, first component k0
  // coordinates of tensor zz, component kk, into registers (zero-padded) with their squared norms
  auto ld = [&](int kk, int zz, T (&r)[2][DP], T (&rq)[2]) {
    const T* zp = Z + ((size_t)kk * Mz + zz) * zs * d;
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      rq[s] = T(0);
#pragma unroll
      for (int q = 0; q < DP; ++q) { r[s][q] = (s < zs && q < d) ? zp[s * d + q] : T(0); rq[s] = gfma<T>(r[s][q], r[s][q], rq[s]); }
    }
  };
  auto dotp = [&](const T (&u)[DP], const T (&v)[DP]) {
    T dot = T(0);
#pragma unroll
    for (int q = 0; q < DP; ++q) dot = gfma<T>(u[q], v[q], dot);
    return dot;
  };
  const T* g = static_cast<const T*>(P.g) + (size_t)I * Mz * Mz;
  T za[2][DP], zaq[2], acc[2][DP];
  ld(k, a, za, zaq);
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int q = 0; q < DP; ++q) acc[s][q] = T(0);
  for (int b = lane; b < Mz; b += 32) {
    T w = g[(size_t)a * Mz + b] + g[(size_t)b * Mz + a];
    if (w == T(0)) continue;
    for (int jj = 0; jj < I; ++jj) {
      if (k0 + jj == k) continue;
      T ua[2][DP], ub[2][DP], uaq[2], ubq[2];
      ld(k0 + jj, a, ua, uaq);
      ld(k0 + jj, b, ub, ubq);
      auto ev = [&](int s, int s2) { return base_eval_call<T>(P.base, p0, p1, dotp(ua[s], ub[s2]), uaq[s], ubq[s2]); };
      w *= zs == 2 ? ev(1, 1) + ev(0, 0) - ev(1, 0) - ev(0, 1) : ev(0, 0);                           // kernels.py:278
    }
    // d Mt_k(a, b) / d z_a^s: sum over the two kappas that contain z_a^s
    T zb[2][DP], zbq[2];
    ld(k, b, zb, zbq);
#pragma unroll
    for (int s = 0; s < 2; ++s) {
#pragma unroll
      for (int s2 = 0; s2 < 2; ++s2) {
        if (s >= zs || s2 >= zs) continue;
        const T sgn = (zs == 2 && s != s2) ? T(-1) : T(1);
        T fd, fx, fy;
        base_partials_call<T>(P.base, p0, p1, dotp(za[s], zb[s2]), zaq[s], zbq[s2], fd, fx, fy);
        const T wf = sgn * w * fd, wx = T(2) * sgn * w * fx;
#pragma unroll
        for (int q = 0; q < DP; ++q) acc[s][q] = gfma<T>(wf, zb[s2][q], gfma<T>(wx, za[s][q], acc[s][q]));
      }
    }
  }
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int q = 0; q < DP; ++q) {
      if (s >= zs || q >= d) continue;                               // warp-uniform
      T v = acc[s][q];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) static_cast<T*>(P.gradZ)[(((size_t)k * Mz + a) * zs + s) * d + q] = v;
    }
}

// d > 16: the same adjoint with one warp per (component, row a) working on ONE column b at a time: the lanes stride over
// the coordinates, dot products are warp reductions, lane t accumulates dL/dz_a[q] for q = t, t + 32, ... (256 per pass)
template <typename T>
__device__ __forceinline__ void warp_dot3(const T* __restrict__ u, const T* __restrict__ v, int d, int lane, T& dot, T& us, T& vs) {
  dot = us = vs = T(0);
  for (int q = lane; q < d; q += 32) { const T a = u[q], b = v[q]; dot = gfma<T>(a, b, dot); us = gfma<T>(a, a, us); vs = gfma<T>(b, b, vs); }
  dot = warp_sum<T>(dot); us = warp_sum<T>(us); vs = warp_sum<T>(vs);
}

template <typename T>
__global__ void __launch_bounds__(128) tens_tens_vjp_big_kernel(const TensTensVjpParams P) {
  const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int k = gridDim.y - 1 - blockIdx.y;
  if (a >= P.Mz) return;
  const int d = P.d, zs = P.zs, Mz = P.Mz;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* Z = static_cast<const T*>(P.Z);
  int I = 1, k0 = 0;
  while (k0 + I <= k) { k0 += I; ++I; }
  const T* g = static_cast<const T*>(P.g) + (size_t)I * Mz * Mz;
  auto zpt = [&](int kk, int zz, int sidx) { return Z + (((size_t)kk * Mz + zz) * zs + sidx) * d; };
  auto ev = [&](int kk, int b, int sa, int sb) {
    T dot, us, vs;
    warp_dot3<T>(zpt(kk, a, sa), zpt(kk, b, sb), d, lane, dot, us, vs);
    return base_eval_call<T>(P.base, p0, p1, dot, us, vs);
  };
  for (int q0 = 0; q0 < d; q0 += 256) {
    T acc[2][8];
#pragma unroll
    for (int sidx = 0; sidx < 2; ++sidx)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[sidx][c] = T(0);
    for (int b = 0; b < Mz; ++b) {
      T w = g[(size_t)a * Mz + b] + g[(size_t)b * Mz + a];
      if (w == T(0)) continue;                                         // warp-uniform
      for (int jj = 0; jj < I; ++jj) {
        if (k0 + jj == k) continue;
        w *= zs == 2 ? ev(k0 + jj, b, 1, 1) + ev(k0 + jj, b, 0, 0) - ev(k0 + jj, b, 1, 0) - ev(k0 + jj, b, 0, 1) : ev(k0 + jj, b, 0, 0);   // kernels.py:278
      }
      for (int sa = 0; sa < zs; ++sa)
        for (int sb = 0; sb < zs; ++sb) {
          const T sgn = (zs == 2 && sa != sb) ? T(-1) : T(1);
          const T* za = zpt(k, a, sa); const T* zb = zpt(k, b, sb);
          T dot, us, vs, fd, fx, fy;
          warp_dot3<T>(za, zb, d, lane, dot, us, vs);
          base_partials_call<T>(P.base, p0, p1, dot, us, vs, fd, fx, fy);
          const T wf = sgn * w * fd, wx = T(2) * sgn * w * fx;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int q = q0 + lane + 32 * c;
            if (q < d) acc[sa][c] = gfma<T>(wf, zb[q], gfma<T>(wx, za[q], acc[sa][c]));
          }
        }
    }
    for (int sa = 0; sa < zs; ++sa)
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int q = q0 + lane + 32 * c;
        if (q < d) static_cast<T*>(P.gradZ)[(((size_t)k * Mz + a) * zs + sa) * d + q] = acc[sa][c];
      }
  }
}

}  // namespace gpsig
