// Inducing-tensor covariances for sm_100a.
//
//   tens_vs_seq_kernel  Kzx: gpsig/kernels.py:314-341 (_K_tens_vs_seq) +
//                       gpsig/signature_algs.py:101-127, epilogue kernels.py:604-620.
//   tens_tens_kernel    Kzz: gpsig/kernels.py:264-284 (_K_tens) +
//                       gpsig/signature_algs.py:76-99 (tensor_kern).
//
// The reference materialises M[T, Mz, (2,) N, L] (80 GB at N=10k) and scans it
// along time with tf.cumsum.  Here one THREAD owns one (z, x_n) pair and walks
// the time axis once: the exclusive cumulative sums of signature_algs.py:123
// become T = m(m+1)/2 running sums held in registers,
//     S[k+j] += D[k+j](l) * S[k+j-1](<l),     K_i = S[k_i + i - 1] at the end,
// so nothing is scanned across threads and nothing is written but the result.
// Lanes of a warp are 32 consecutive sequences (coalesced output); the 8 warps
// of a CTA are 8 inducing tensors sharing those sequences' rows through L1.
#pragma once
#include "common.cuh"

namespace gpsig {

constexpr int kTensWarps = 8;

struct TensSeqParams {
  const void* zp;      // packed tensors: block (t*Mz + z) = coords[zs][DP] + sqn[LPz]
  const void* xp;      // packed sequences
  size_t zblk, xblk;
  int Mz, zs;          // zs = 2 with increments else 1
  int N, L, m;
  int col_begin, col_end;
  int difference;
  double p0, p1; int base;
  SpectralParams sp; int d;   // GPSIG_BASE_SPECTRAL (general path only)
  const void* norm_x;  // [m+1, N] (already jittered) or NULL
  const void* level_scale;
  int return_levels;
  void* out;           // [(m+1,)? Mz, col_end-col_begin]
};

// Level I of signature_algs.py:101-127 for one (z, x_n) pair: I running sums.
// zc/zq point at this level's I (x zs) tensor components.
// RBF: the packed coordinates are pre-scaled by sqrt(log2 e) (abi.cu), so the exponent is in base-2 units:
// float64 exp2_tab32 (9 FP64 instructions, table lookup by warp shuffle -> the warp must be convergent),
// float32 one MUFU.EX2.
template <typename T> struct TensExp2;
template <> struct TensExp2<double> { static __device__ __forceinline__ double ex(double t, double tab) { return exp2_tab32(t, tab); } };
template <> struct TensExp2<float> { static __device__ __forceinline__ float ex(float t, double) { return exp2_approx(t); } };

template <typename T, int DP, int OP, int I>
__device__ __forceinline__ T tens_level(const T* zc, const T* zq, int zs, const T* xblk, const T* xsq, int L,
                                        bool difference, int base, T p0, T p1, double tab) {
  using Op = BaseOp<T, OP>;
  T S[I], Mprev[I];
#pragma unroll
  for (int t = 0; t < I; ++t) { S[t] = T(0); Mprev[t] = T(0); }
  for (int l = 0; l < L; ++l) {
    T xv[DP];
#pragma unroll
    for (int k = 0; k < DP; ++k) xv[k] = xblk[(size_t)l * DP + k];
    const T xc = Op::col_const(xsq[l]);
    auto kappa = [&](int e) {
      T acc = Op::init(zq[e], xc);
#pragma unroll
      for (int k = 0; k < DP; ++k) acc = gfma<T>(zc[e * DP + k], xv[k], acc);
      if (OP == OP_GENERIC) return base_eval_call<T>(base, p0, p1, acc, zq[e], xc);
      if (OP == OP_RBF) return TensExp2<T>::ex(acc, tab);
      return Op::finish(acc, zq[e], xc, base, p0, p1);
    };
    const bool first = difference && l == 0;
#pragma unroll
    for (int jj = I - 1; jj >= 0; --jj) {
      T Mt = zs == 2 ? kappa(2 * jj + 1) - kappa(2 * jj) : kappa(jj);       // kernels.py:331
      T D = Mt;
      if (difference) { D = Mt - Mprev[jj]; Mprev[jj] = Mt; }               // signature_algs.py:114
      if (!first) {
        if (jj > 0) S[jj] = gfma<T>(D, S[jj - 1], S[jj]);                     // signature_algs.py:123
        else S[jj] += D;
      }
    }
  }
  return S[I - 1];                                                           // signature_algs.py:125
}

template <typename T, int DP, int OP>
__global__ void __launch_bounds__(kTensWarps * 32) tens_vs_seq_kernel(const TensSeqParams P) {
  using Op = BaseOp<T, OP>;
  extern __shared__ __align__(128) unsigned char tens_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int zs = P.zs, m = P.m;
  const int TT = m * (m + 1) / 2;
  // per-warp z tile: coords [TT*zs][DP], then row constants [TT*zs]
  T* zc = reinterpret_cast<T*>(tens_smem) + (size_t)warp * (TT * 2 * (DP + 1));
  T* zq = zc + TT * 2 * DP;
  const int z = blockIdx.y * kTensWarps + warp;
  if (z >= P.Mz) return;
  for (int e = lane; e < TT * zs; e += 32) {
    const int t = e / zs, s = e % zs;
    const T* blk = static_cast<const T*>(P.zp) + ((size_t)t * P.Mz + z) * P.zblk;
    for (int k = 0; k < DP; ++k) zc[e * DP + k] = blk[s * DP + k];
    zq[e] = Op::row_const(blk[zs * DP + s]);
  }
  __syncwarp();
  const int ncols = P.col_end - P.col_begin;
  const int nl = blockIdx.x * 32 + lane;
  const bool live = nl < ncols;                           // lanes past the last column compute on a clamped column (the warp stays convergent for the table shuffle)
  const int n = P.col_begin + (live ? nl : ncols - 1);
  const double tab = exp2((double)lane * (1.0 / 32.0));
  const T* xblk = static_cast<const T*>(P.xp) + (size_t)n * P.xblk;
  const T* xsq = xblk + (size_t)P.L * DP;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const bool diff = P.difference != 0;

  // epilogue (kernels.py:604-620) applied level by level
  T acc = T(0);
  auto emit = [&](int lvl, T v) {
    if (!live) return;
    if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + n]);
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
    if (P.return_levels) static_cast<T*>(P.out)[((size_t)lvl * P.Mz + z) * ncols + nl] = v; else acc += v;
  };
  emit(0, T(1));
#define GPSIG_LEVEL(I_) if (m >= I_) { constexpr int k = I_ * (I_ - 1) / 2; \
    const T v_ = tens_level<T, DP, OP, I_>(zc + (size_t)k * zs * DP, zq + k * zs, zs, xblk, xsq, P.L, diff, P.base, p0, p1, tab); emit(I_, v_); }
  GPSIG_LEVEL(1) GPSIG_LEVEL(2) GPSIG_LEVEL(3) GPSIG_LEVEL(4)
  GPSIG_LEVEL(5) GPSIG_LEVEL(6) GPSIG_LEVEL(7) GPSIG_LEVEL(8)
#undef GPSIG_LEVEL
  if (!P.return_levels && live) static_cast<T*>(P.out)[(size_t)z * ncols + nl] = acc;
}

template <typename T, int OP, int DP>
int launch_tens_vs_seq(const TensSeqParams& P, int m, cudaStream_t stream) {
  if (m < 1 || m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", m);
  const int TT = m * (m + 1) / 2;
  const size_t smem = (size_t)kTensWarps * TT * 2 * (DP + 1) * sizeof(T);
  auto kern = tens_vs_seq_kernel<T, DP, OP>;
  GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int ncols = P.col_end - P.col_begin;
  dim3 grid((ncols + 31) / 32, (P.Mz + kTensWarps - 1) / kTensWarps);
  prof_begin(stream);
  kern<<<grid, kTensWarps * 32, smem, stream>>>(P);
  prof_end(stream);
  GPSIG_LAUNCH_CHECK("tens_vs_seq_kernel");
  return GPSIG_OK;
}

// ---------------------------------------------------------------------------
// Kzz: one thread per (z, z') entry; tiny (Mz^2 * T * 4 base-kernel values)
// ---------------------------------------------------------------------------
struct TensTensParams {
  const void* Z;       // raw [T, Mz, zs, d]
  const void* ls;      // lengthscales [d] or NULL (kernels.py:367-399)
  int Mz, zs, m, d;
  double p0, p1; int base;
  SpectralParams sp;
  const void* level_scale; int return_levels;
  void* out;           // [(m+1,)? Mz, Mz]
};

template <typename T>
__global__ void tens_tens_kernel(const TensTensParams P) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;   // column z'
  const int a = blockIdx.y;                              // row z
  if (b >= P.Mz) return;
  const int d = P.d, zs = P.zs;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* ls = static_cast<const T*>(P.level_scale);
  const T* len = static_cast<const T*>(P.ls);
  T* out = static_cast<T*>(P.out);
  const size_t MM = (size_t)P.Mz * P.Mz, ab = (size_t)a * P.Mz + b;
  auto kappa = [&](const T* za, const T* zb) {
    if (P.base == GPSIG_BASE_SPECTRAL) {
      T u[16], v[16];
      for (int k = 0; k < d; ++k) { u[k] = za[k]; v[k] = zb[k]; if (len) { u[k] = u[k] / len[k]; v[k] = v[k] / len[k]; } }
      return spectral_eval<T>(P.sp, d, u, v);
    }
    T dot = T(0), xs = T(0), ys = T(0);
    for (int k = 0; k < d; ++k) {
      T u = za[k], v = zb[k];
      if (len) { u = u / len[k]; v = v / len[k]; }
      dot = gfma<T>(u, v, dot); xs = gfma<T>(u, u, xs); ys = gfma<T>(v, v, ys);
    }
    return base_eval<T>(P.base, p0, p1, dot, xs, ys);
  };
  T acc = ls ? ls[0] : T(1);
  if (P.return_levels) out[ab] = acc;
  int k = 0;
  for (int i = 1; i <= P.m; ++i) {
    T R = T(1);
    for (int jj = 0; jj < i; ++jj, ++k) {
      const T* za = static_cast<const T*>(P.Z) + ((size_t)k * P.Mz + a) * zs * d;
      const T* zb = static_cast<const T*>(P.Z) + ((size_t)k * P.Mz + b) * zs * d;
      T Mt;
      if (zs == 2) Mt = kappa(za + d, zb + d) + kappa(za, zb) - kappa(za + d, zb) - kappa(za, zb + d);  // kernels.py:278
      else Mt = kappa(za, zb);
      R = jj == 0 ? Mt : Mt * R;                                                                     // signature_algs.py:91-96
    }
    if (ls) R *= ls[i];
    if (P.return_levels) out[(size_t)i * MM + ab] = R; else acc += R;
  }
  if (!P.return_levels) out[ab] = acc;
}

}  // namespace gpsig
