// Kzz (inducing tensors against themselves) and its adjoint on 16 x 16 tiles of tensor pairs, any state dimension.
//
// Reference: gpsig/kernels.py:264-284 (_K_tens: the four base-kernel values of an increment pair, kernels.py:278) and
// gpsig/signature_algs.py:76-99 (signature_kern_tens_vs_tens: R_i = prod of the level's components); the adjoint is what
// TF autodiff derives from those.
//
// The first kernels (tens_kernels.cuh: one thread per entry, coordinates in registers up to d = 16, a runtime loop from global
// memory above; vjp_kernels.cuh: one warp per (component, row), above d = 16 one column at a time with warp reductions) are
// latency bound: Kzz forward + backward of 500 tensors took 0.86 ms at d = 8 and 10.3 ms at d = 28, where the arithmetic is
// 0.3 GFMA.  Here a CTA of 16 x 16 threads owns a tile of tensor pairs (thread = one pair (a, b)), the coordinates of the 16 + 16
// tensors of one component are staged in shared memory ([s][tensor][d + 1 padding]: a row of threads reads one tensor as a
// broadcast and 16 consecutive tensors conflict-free), and
//   forward: the CTA walks the components, each thread multiplies up its levels;
//   adjoint: one CTA per (component k, tile of rows a) walks the tiles of columns b; per tile it rebuilds the other components'
//            factors, puts the 16 x 32 weights of d kappa / d(dot) in shared memory and accumulates dL/dz_a as a small
//            matrix product against the staged z_b (thread = (a, coordinates tx, tx + 16, ...)): no atomics, no warp reductions.
#pragma once
#include "common.cuh"
#include "trunc_kernel.cuh"
#include "tens_kernels.cuh"
#include "vjp_kernels.cuh"

namespace gpsig {

constexpr int kTT = 16;       // tensors per tile side

// stage component k's tensors t0 .. t0 + 15 (both increment ends) into dst[s][t][dp], zero past Mz; optional lengthscales
template <typename T>
__device__ __forceinline__ void tt_stage(T* dst, const T* Z, const T* len, int k, int t0, int Mz, int zs, int d, int dp, int tid) {
  // 16 threads per tensor walk its zs * d coordinates (consecutive addresses; zs <= 2: no division)
  const int t = tid >> 4, per = zs * d;
  const bool live = t0 + t < Mz;
  const T* src = Z + ((size_t)k * Mz + (live ? t0 + t : 0)) * per;
  for (int r = tid & 15; r < per; r += 16) {
    const int s = r >= d ? 1 : 0, q = r - s * d;
    T v = T(0);
    if (live) { v = src[r]; if (len) v = v / len[q]; }
    dst[(s * kTT + t) * dp + q] = v;
  }
}

// the (up to) four base-kernel values of a tensor pair from staged coordinates, combined as in kernels.py:278
template <typename T>
__device__ __forceinline__ T tt_rbf64(T dot, T xs, T ys, double tab) {      // exp(-(|x|^2 + |y|^2 - 2 x.y) / 2) by the table exp2 (convergent warp)
  constexpr double kLog2e = 1.4426950408889634;
  return T(exp2_tab32(((double)dot - 0.5 * ((double)xs + (double)ys)) * kLog2e, tab));
}

template <typename T>
__device__ __forceinline__ T tt_entry(const T* A, const T* B, int ty, int tx, int zs, int d, int dp, int base, T p0, T p1, double tab) {
  const T* a0 = A + ty * dp; const T* a1 = A + (kTT + ty) * dp;
  const T* b0 = B + tx * dp; const T* b1 = B + (kTT + tx) * dp;
  if (zs == 2) {
    T d00 = T(0), d01 = T(0), d10 = T(0), d11 = T(0), ua0 = T(0), ua1 = T(0), vb0 = T(0), vb1 = T(0);
#pragma unroll 4
    for (int q = 0; q < d; ++q) {
      const T x0 = a0[q], x1 = a1[q], y0 = b0[q], y1 = b1[q];
      d00 = gfma<T>(x0, y0, d00); d01 = gfma<T>(x0, y1, d01); d10 = gfma<T>(x1, y0, d10); d11 = gfma<T>(x1, y1, d11);
      ua0 = gfma<T>(x0, x0, ua0); ua1 = gfma<T>(x1, x1, ua1); vb0 = gfma<T>(y0, y0, vb0); vb1 = gfma<T>(y1, y1, vb1);
    }
    if (sizeof(T) == 8 && base == GPSIG_BASE_RBF)      // 9 FP64 instructions per value against ~35 of exp() behind a call
      return tt_rbf64<T>(d11, ua1, vb1, tab) + tt_rbf64<T>(d00, ua0, vb0, tab) - tt_rbf64<T>(d10, ua1, vb0, tab) - tt_rbf64<T>(d01, ua0, vb1, tab);
    return base_eval_call<T>(base, p0, p1, d11, ua1, vb1) + base_eval_call<T>(base, p0, p1, d00, ua0, vb0)
         - base_eval_call<T>(base, p0, p1, d10, ua1, vb0) - base_eval_call<T>(base, p0, p1, d01, ua0, vb1);
  }
  T d00 = T(0), ua0 = T(0), vb0 = T(0);
#pragma unroll 4
  for (int q = 0; q < d; ++q) { const T x0 = a0[q], y0 = b0[q]; d00 = gfma<T>(x0, y0, d00); ua0 = gfma<T>(x0, x0, ua0); vb0 = gfma<T>(y0, y0, vb0); }
  if (sizeof(T) == 8 && base == GPSIG_BASE_RBF) return tt_rbf64<T>(d00, ua0, vb0, tab);
  return base_eval_call<T>(base, p0, p1, d00, ua0, vb0);
}

// slots: staged components per side (forward 1; adjoint 2 = component k and one other at a time, or all of a level)
static inline size_t tt_tile_smem(int d, int zs, size_t elem, int slots, bool vjp) {
  const size_t tile = (size_t)zs * kTT * (d + 1) * elem;
  return 2 * slots * tile + (vjp ? (size_t)2 * kTT * 2 * kTT * 2 * elem : 0);     // adjoint: + the two weight tiles
}

template <typename T>
__global__ void __launch_bounds__(256) tens_tens_tile_kernel(const TensTensParams P) {
  extern __shared__ __align__(16) unsigned char tt_smem[];
  const int d = P.d, zs = P.zs, dp = d + 1, Mz = P.Mz;
  T* A = reinterpret_cast<T*>(tt_smem);
  T* B = A + (size_t)zs * kTT * dp;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int a0 = blockIdx.y * kTT, b0 = blockIdx.x * kTT, a = a0 + ty, b = b0 + tx;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* ls = static_cast<const T*>(P.level_scale);
  const T* len = static_cast<const T*>(P.ls);
  const T* Z = static_cast<const T*>(P.Z);
  T* out = static_cast<T*>(P.out);
  const bool live = a < Mz && b < Mz;
  const double tab = sizeof(T) == 8 ? exp2((double)(tid & 31) * (1.0 / 32.0)) : 0.0;
  const size_t MM = (size_t)Mz * Mz, ab = (size_t)a * Mz + b;
  T acc = ls ? ls[0] : T(1);
  if (P.return_levels && live) out[ab] = acc;
  int k = 0;
  for (int i = 1; i <= P.m; ++i) {
    T R = T(1);
    for (int jj = 0; jj < i; ++jj, ++k) {
      __syncthreads();
      tt_stage<T>(A, Z, len, k, a0, Mz, zs, d, dp, tid);
      tt_stage<T>(B, Z, len, k, b0, Mz, zs, d, dp, tid);
      __syncthreads();
      const T Mt = tt_entry<T>(A, B, ty, tx, zs, d, dp, P.base, p0, p1, tab);
      R = jj == 0 ? Mt : Mt * R;                                                                     // signature_algs.py:91-96
    }
    if (ls) R *= ls[i];
    if (P.return_levels) { if (live) out[(size_t)i * MM + ab] = R; } else acc += R;
  }
  if (!P.return_levels && live) out[ab] = acc;
}

// NQ: coordinates per thread in the accumulation (d <= 16 NQ).  ALL: the tiles of every component of the level are staged at
// once (slots = m per side: the rows' tiles once, the columns' tiles once per column tile: three barriers and one round of
// global loads per column tile instead of 2 I + 1 and I); otherwise component k's tiles and one other component at a time.
template <typename T, int NQ, bool ALL>
__global__ void __launch_bounds__(256, NQ <= 2 ? 3 : 2) tens_tens_tile_vjp_kernel(const TensTensVjpParams P, const int slots) {
  extern __shared__ __align__(16) unsigned char tt_smem[];
  const int d = P.d, zs = P.zs, dp = d + 1, Mz = P.Mz;
  const size_t tile = (size_t)zs * kTT * dp;
  T* A = reinterpret_cast<T*>(tt_smem);
  T* B = A + slots * tile;
  T* Wf = B + slots * tile;                   // [a][s][b][s']: sgn * w * d kappa / d dot
  T* Wx = Wf + kTT * 2 * kTT * 2;             // [a][s][b][s']: 2 sgn * w * d kappa / d |z_a|^2
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int k = gridDim.y - 1 - blockIdx.y;                           // the top level's components cost the most: first
  const int a0 = blockIdx.x * kTT, a = a0 + ty;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* Z = static_cast<const T*>(P.Z);
  int I = 1, k0 = 0;
  while (k0 + I <= k) { k0 += I; ++I; }                               // level of component k, first component k0
  const T* g = static_cast<const T*>(P.g) + (size_t)I * Mz * Mz;
  T acc[2][NQ], axs[2] = {T(0), T(0)};
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int c = 0; c < NQ; ++c) acc[s][c] = T(0);
  const double tab = sizeof(T) == 8 ? exp2((double)(tid & 31) * (1.0 / 32.0)) : 0.0;
  const bool rbf64 = sizeof(T) == 8 && P.base == GPSIG_BASE_RBF;
  const int kk = ALL ? k - k0 : 0;
  T* Ak = A + kk * tile; T* Bk = B + kk * tile;
  if (ALL) { for (int jj = 0; jj < I; ++jj) tt_stage<T>(A + jj * tile, Z, nullptr, k0 + jj, a0, Mz, zs, d, dp, tid); }
  else tt_stage<T>(Ak, Z, nullptr, k, a0, Mz, zs, d, dp, tid);
  for (int b0 = 0; b0 < Mz; b0 += kTT) {
    const int b = b0 + tx;
    T w = T(0);
    if (a < Mz && b < Mz) w = g[(size_t)a * Mz + b] + g[(size_t)b * Mz + a];
    if (ALL) {
      __syncthreads();                                                 // the previous tile's accumulation is done with B, Wf, Wx
      for (int jj = 0; jj < I; ++jj) tt_stage<T>(B + jj * tile, Z, nullptr, k0 + jj, b0, Mz, zs, d, dp, tid);
      __syncthreads();
      for (int jj = 0; jj < I; ++jj)
        if (jj != kk) w *= tt_entry<T>(A + jj * tile, B + jj * tile, ty, tx, zs, d, dp, P.base, p0, p1, tab);
    } else {
      T* Aj = A + tile; T* Bj = B + tile;
      for (int jj = 0; jj < I; ++jj) {
        if (k0 + jj == k) continue;
        __syncthreads();
        tt_stage<T>(Aj, Z, nullptr, k0 + jj, a0, Mz, zs, d, dp, tid);
        tt_stage<T>(Bj, Z, nullptr, k0 + jj, b0, Mz, zs, d, dp, tid);
        __syncthreads();
        w *= tt_entry<T>(Aj, Bj, ty, tx, zs, d, dp, P.base, p0, p1, tab);
      }
      __syncthreads();                                                 // the previous tile's accumulation is done with Bk, Wf, Wx
      tt_stage<T>(Bk, Z, nullptr, k, b0, Mz, zs, d, dp, tid);
      __syncthreads();
    }
    // weights of this pair: d Mt_k(a, b) / d z_a^s = sum_s' sgn (fd z_b^s' + 2 fx z_a^s)
    {
      T dt[2][2] = {{T(0), T(0)}, {T(0), T(0)}}, uq[2] = {T(0), T(0)}, vq[2] = {T(0), T(0)};
      const T* x0p = Ak + ty * dp; const T* x1p = Ak + (kTT + ty) * dp;
      const T* y0p = Bk + tx * dp; const T* y1p = Bk + (kTT + tx) * dp;
      if (zs == 2) {
#pragma unroll 4
        for (int q = 0; q < d; ++q) {
          const T x0 = x0p[q], x1 = x1p[q], y0 = y0p[q], y1 = y1p[q];
          dt[0][0] = gfma<T>(x0, y0, dt[0][0]); dt[0][1] = gfma<T>(x0, y1, dt[0][1]);
          dt[1][0] = gfma<T>(x1, y0, dt[1][0]); dt[1][1] = gfma<T>(x1, y1, dt[1][1]);
          uq[0] = gfma<T>(x0, x0, uq[0]); uq[1] = gfma<T>(x1, x1, uq[1]); vq[0] = gfma<T>(y0, y0, vq[0]); vq[1] = gfma<T>(y1, y1, vq[1]);
        }
      } else {
        for (int q = 0; q < d; ++q) { const T x0 = x0p[q], y0 = y0p[q]; dt[0][0] = gfma<T>(x0, y0, dt[0][0]); uq[0] = gfma<T>(x0, x0, uq[0]); vq[0] = gfma<T>(y0, y0, vq[0]); }
      }
#pragma unroll
      for (int s = 0; s < 2; ++s)
#pragma unroll
        for (int s2 = 0; s2 < 2; ++s2) {
          T wf = T(0), wx = T(0);
          if (rbf64) {                                   // every lane evaluates (table shuffle), f_dot = f, f_|x|^2 = -f / 2
            const T f = tt_rbf64<T>(dt[s][s2], uq[s], vq[s2], tab);
            if (s < zs && s2 < zs) { wf = ((zs == 2 && s != s2) ? -w : w) * f; wx = -wf; }
          } else if (s < zs && s2 < zs && w != T(0)) {
            const T sgn = (zs == 2 && s != s2) ? T(-1) : T(1);
            T fd, fx, fy;
            base_partials_call<T>(P.base, p0, p1, dt[s][s2], uq[s], vq[s2], fd, fx, fy);
            wf = sgn * w * fd; wx = T(2) * sgn * w * fx;
          }
          Wf[((ty * 2 + s) * kTT + tx) * 2 + s2] = wf;
          Wx[((ty * 2 + s) * kTT + tx) * 2 + s2] = wx;
        }
    }
    __syncthreads();
    // dL/dz_a^s[q] += sum_{b, s'} Wf[a][s][b][s'] z_b^s'[q]  (thread: row a = ty, coordinates q = tx + 16 c)
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (s >= zs) continue;
      const T* wf = Wf + (ty * 2 + s) * kTT * 2;
      const T* wx = Wx + (ty * 2 + s) * kTT * 2;
      T sx = T(0);
      if (zs == 2) {
#pragma unroll
        for (int bb = 0; bb < kTT; ++bb) {
          const T f0 = wf[bb * 2], f1 = wf[bb * 2 + 1];
          sx += wx[bb * 2] + wx[bb * 2 + 1];
          const T* z0 = Bk + bb * dp; const T* z1 = Bk + (kTT + bb) * dp;
#pragma unroll
          for (int c = 0; c < NQ; ++c) { const int q = tx + 16 * c; if (q < d) acc[s][c] = gfma<T>(f1, z1[q], gfma<T>(f0, z0[q], acc[s][c])); }
        }
      } else {
#pragma unroll
        for (int bb = 0; bb < kTT; ++bb) {
          const T f0 = wf[bb * 2];
          sx += wx[bb * 2];
          const T* z0 = Bk + bb * dp;
#pragma unroll
          for (int c = 0; c < NQ; ++c) { const int q = tx + 16 * c; if (q < d) acc[s][c] = gfma<T>(f0, z0[q], acc[s][c]); }
        }
      }
      axs[s] += sx;
    }
  }
  if (a < Mz) {
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (s >= zs) continue;
#pragma unroll
      for (int c = 0; c < NQ; ++c) {
        const int q = tx + 16 * c;
        if (q < d) static_cast<T*>(P.gradZ)[(((size_t)k * Mz + a) * zs + s) * d + q] = gfma<T>(axs[s], Ak[(s * kTT + ty) * dp + q], acc[s][c]);
      }
    }
  }
}

// false: shape outside the tiled kernels (the caller keeps the first kernels)
template <typename T>
bool launch_tens_tens_tile(const TensTensParams& P, cudaStream_t s) {
  const size_t smem = tt_tile_smem(P.d, P.zs, sizeof(T), 1, false);
  if (P.base == GPSIG_BASE_SPECTRAL || smem > kMaxSmem) return false;
  auto kern = tens_tens_tile_kernel<T>;
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
  const int nt = (P.Mz + kTT - 1) / kTT;
  kern<<<dim3(nt, nt), 256, smem, s>>>(P);
  return true;
}

template <typename T>
bool launch_tens_tens_tile_vjp(const TensTensVjpParams& P, cudaStream_t s) {
  const bool all = tt_tile_smem(P.d, P.zs, sizeof(T), P.m, true) <= 100 * 1024;      // two CTAs per SM
  const int slots = all ? P.m : 2;
  const size_t smem = tt_tile_smem(P.d, P.zs, sizeof(T), slots, true);
  if (P.d > 128 || smem > kMaxSmem) return false;
  if (sizeof(T) == 4 && P.d <= 4) return false;      // float32, d <= 4: the register kernel is faster (500 tensors, m = 4, d = 4: 0.16 against 0.24 ms; d = 8: 0.30 against 0.26)
  const int TT = P.m * (P.m + 1) / 2;
  const dim3 grid((P.Mz + kTT - 1) / kTT, TT);
#define GPSIG_TTV2(NQ_, ALL_)                                                                                                      \
  do {                                                                                                                               \
    auto kern = tens_tens_tile_vjp_kernel<T, NQ_, ALL_>;                                                                             \
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false; \
    kern<<<grid, 256, smem, s>>>(P, slots);                                                                                          \
  } while (0)
#define GPSIG_TTV(NQ_) do { if (all) GPSIG_TTV2(NQ_, true); else GPSIG_TTV2(NQ_, false); } while (0)
  if (P.d <= 16) GPSIG_TTV(1); else if (P.d <= 32) GPSIG_TTV(2); else if (P.d <= 64) GPSIG_TTV(4); else GPSIG_TTV(8);
#undef GPSIG_TTV
#undef GPSIG_TTV2
  return true;
}

}  // namespace gpsig
