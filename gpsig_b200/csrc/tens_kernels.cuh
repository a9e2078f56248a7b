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
#include <type_traits>

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
// register-resident kernel (tens_vs_seq_kernel2): float64 exp2 from an 8192-entry shared-memory table (6 FP64 instructions,
// exp2_tab8k; 2048 entries and 7 instructions with GPSIG_TENS_EXP2_8K=0: c4 float64 14.65 -> see DESIGN 4.3)
// instead of 9; `tab` carries the table's shared-memory address in its low word)
template <typename T> struct TensExp2T;
#ifndef GPSIG_TENS_EXP2_8K
#define GPSIG_TENS_EXP2_8K 1
#endif
template <> struct TensExp2T<double> {
  template <int EBITS> static __device__ __forceinline__ double ex(double t, double tab) {
    if constexpr (EBITS == 13) return exp2_tab8k(t, (uint32_t)__double2loint(tab));
    else return exp2_tab2048(t, (uint32_t)__double2loint(tab));
  }
};
template <> struct TensExp2T<float> { template <int EBITS> static __device__ __forceinline__ float ex(float t, double) { return exp2_approx(t); } };

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

// ---------------------------------------------------------------------------
// Kzx, register-resident tensors (the variant the library launches for the common shapes).
//
// tens_vs_seq_kernel above keeps z in shared memory and gives every lane its own sequence: every FMA of the
// contraction reads a z operand through the shared-memory pipe (1 LDS per 2-4 FMAs) and every lane streams its own
// x rows through L1 -- ncu: issue slots 26-31 % busy, stalled on the MIO queue / short scoreboard.  Here the roles
// are swapped: LANE = inducing tensor (its I x zs x DP coordinates of the current level live in REGISTERS), the
// CTA's warps share one sequence x_n at a time, brought by one TMA bulk copy into a double-buffered shared-memory
// slot and read as warp-uniform broadcasts (DP/2 + 1 loads per time step against I zs (DP + exp) FMAs).
// Levels are the outer loop (z registers are loaded once per level and CTA), the CH sequences of the CTA's chunk
// the inner loop; results are collected in a shared-memory tile [CH][ZB+1] and written out transposed, so the
// stores are coalesced along the sequence index.
// ---------------------------------------------------------------------------
template <typename T> struct TensVec;   // 16-byte vector of T
template <> struct TensVec<double> { static constexpr int V = 2; using type = double2;
  static __device__ __forceinline__ double get(const double2& v, int e) { return e == 0 ? v.x : v.y; } };
template <> struct TensVec<float> { static constexpr int V = 4; using type = float4;
  static __device__ __forceinline__ float get(const float4& v, int e) { return e == 0 ? v.x : (e == 1 ? v.y : (e == 2 ? v.z : v.w)); } };
constexpr size_t kTens2MaxSmem = 200 * 1024;
constexpr int kTens2Warps = 4;      // 128 tensors per CTA
constexpr int kTens2Chunk = 16;     // sequences per CTA at most (fewer when the grid would not fill the GPU)

template <typename T, int DP, int OP, int ZS, int I, int EBITS>
__device__ __forceinline__ T tens2_level(const T (&zr)[I * ZS * DP], const T (&zq)[I * ZS], const T* xs, const T* xsq, int L,
                                         bool difference, int base, T p0, T p1, double tab) {
  using Op = BaseOp<T, OP>;
  T S[I], Mprev[I];
#pragma unroll
  for (int t = 0; t < I; ++t) { S[t] = T(0); Mprev[t] = T(0); }
  for (int l = 0; l < L; ++l) {
    T xv[DP];
    using V16 = typename TensVec<T>::type;
    constexpr int V = TensVec<T>::V;
#pragma unroll
    for (int kv = 0; kv < DP / V; ++kv) {
      const V16 v = *reinterpret_cast<const V16*>(xs + (size_t)l * DP + kv * V);      // warp-uniform: broadcast
#pragma unroll
      for (int e = 0; e < V; ++e) xv[kv * V + e] = TensVec<T>::get(v, e);
    }
    const T xc = Op::col_const(xsq[l]);
    auto kappa = [&](int e) {
      T acc = Op::init(zq[e], xc);
#pragma unroll
      for (int k = 0; k < DP; ++k) acc = gfma<T>(zr[e * DP + k], xv[k], acc);
      if (OP == OP_GENERIC) return base_eval_call<T>(base, p0, p1, acc, zq[e], xc);
      if (OP == OP_RBF) return TensExp2T<T>::template ex<EBITS>(acc, tab);
      return Op::finish(acc, zq[e], xc, base, p0, p1);
    };
    const bool first = difference && l == 0;
#pragma unroll
    for (int jj = I - 1; jj >= 0; --jj) {
      T Mt = ZS == 2 ? kappa(2 * jj + 1) - kappa(2 * jj) : kappa(jj);       // kernels.py:331
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

// registers the z tile of the top level needs (32-bit units); the kernel is used when this fits
template <typename T> constexpr int tens2_zregs(int m, int zs, int DP) { return m * zs * (DP + 1) * (int)(sizeof(T) / 4); }

template <typename T, int DP, int OP, int ZS, int MMAX, int EBITS = 11>
__global__ void __launch_bounds__(kTens2Warps * 32) tens_vs_seq_kernel2(const TensSeqParams P, const int CH) {
  using Op = BaseOp<T, OP>;
  constexpr int ZB = kTens2Warps * 32;
  extern __shared__ __align__(128) unsigned char tens_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(tens_smem);                     // [2]
  const size_t xbytes = P.xblk * sizeof(T);
  const size_t xslot = (xbytes + 127) / 128 * 128;
  unsigned char* xbuf = tens_smem + 128;                                       // [2][xslot]
  T* tile = reinterpret_cast<T*>(xbuf + 2 * xslot);                            // [CH][ZB + 1]
  const int lane = threadIdx.x & 31;
  const int m = P.m;
  constexpr bool ETAB = sizeof(T) == 8 && OP == OP_RBF;                        // 2^(j/2048) table behind the tile
  const int ncols = P.col_end - P.col_begin;
  const int c0 = blockIdx.x * CH, nseq = min(CH, ncols - c0);                  // this CTA's sequences (local column indices)
  const int zraw = blockIdx.y * ZB + threadIdx.x;
  const bool zlive = zraw < P.Mz;
  const int z = zlive ? zraw : P.Mz - 1;                                       // clamped: the warp stays convergent (table shuffle)
  double tab = exp2((double)lane * (1.0 / 32.0));
  const T p0 = T(P.p0), p1 = T(P.p1);
  const bool diff = P.difference != 0;
  const T* xg = static_cast<const T*>(P.xp) + (size_t)(P.col_begin + c0) * P.xblk;

  if (threadIdx.x == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); fence_mbar_init(); }
  if constexpr (ETAB) {
    double* etab = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(tile) + (((size_t)CH * (ZB + 1) * sizeof(T) + 15) & ~(size_t)15));
    exp2_table_fill<EBITS>(etab, etab + (1 << EBITS), threadIdx.x, ZB);
    tab = __hiloint2double(0, (int)smem_u32(etab));
  }
  __syncthreads();
  const int nitems = m * nseq;                                                 // (level, sequence) items, level-major
  if (threadIdx.x == 0 && nitems > 0) { mbar_expect_tx(&full[0], (uint32_t)xbytes); tma_load_1d(xbuf, xg, (uint32_t)xbytes, &full[0]); }
  // level 0 is the constant 1 (signature_algs.py:110)
  T l0 = T(1);
  if (P.level_scale) l0 *= static_cast<const T*>(P.level_scale)[0];
  for (int j = 0; j < nseq; ++j) {
    T v = T(1);
    if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)P.col_begin + c0 + j]);
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[0];
    tile[j * (ZB + 1) + threadIdx.x] = v;
  }
  (void)l0;
  auto flush = [&](int lvl) {                                                  // tile -> out[(lvl,) z, c0 + j], coalesced along j
    __syncthreads();
    for (int e = threadIdx.x; e < ZB * CH; e += ZB) {
      const int zz = e / CH, j = e % CH;
      const int zo = blockIdx.y * ZB + zz;
      if (zo < P.Mz && j < nseq) {
        T* o = static_cast<T*>(P.out) + (P.return_levels ? (size_t)lvl * P.Mz * ncols : 0) + (size_t)zo * ncols + c0 + j;
        *o = tile[j * (ZB + 1) + zz];
      }
    }
    __syncthreads();
  };
  if (P.return_levels) flush(0);

  int it = 0;
  auto run_level = [&](auto IC) {
    constexpr int I = decltype(IC)::value;
    if (I > m) return;
    constexpr int k0 = I * (I - 1) / 2;
    T zr[I * ZS * DP], zq[I * ZS];
#pragma unroll
    for (int t = 0; t < I; ++t) {
      const T* blk = static_cast<const T*>(P.zp) + ((size_t)(k0 + t) * P.Mz + z) * P.zblk;
#pragma unroll
      for (int s = 0; s < ZS; ++s) {
#pragma unroll
        for (int k = 0; k < DP; ++k) zr[(t * ZS + s) * DP + k] = blk[s * DP + k];
        zq[t * ZS + s] = Op::row_const(blk[ZS * DP + s]);
      }
    }
    T ls = T(1);
    if (P.level_scale) ls = static_cast<const T*>(P.level_scale)[I];
    for (int j = 0; j < nseq; ++j, ++it) {
      const int b = it & 1;
      // the other slot was released by the barrier that ended the previous item
      if (threadIdx.x == 0 && it + 1 < nitems) {
        const int jn = (j + 1 < nseq) ? j + 1 : 0;
        mbar_expect_tx(&full[b ^ 1], (uint32_t)xbytes);
        tma_load_1d(xbuf + (size_t)(b ^ 1) * xslot, xg + (size_t)jn * P.xblk, (uint32_t)xbytes, &full[b ^ 1]);
      }
      mbar_wait(&full[b], (it >> 1) & 1);
      const T* xs = reinterpret_cast<const T*>(xbuf + (size_t)b * xslot);
      T v = tens2_level<T, DP, OP, ZS, I, EBITS>(zr, zq, xs, xs + (size_t)P.L * DP, P.L, diff, P.base, p0, p1, tab);
      if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)I * P.N + P.col_begin + c0 + j]);   // kernels.py:610-614
      v *= ls;
      T* tl = tile + j * (ZB + 1) + threadIdx.x;
      if (P.return_levels) *tl = v; else *tl += v;
      __syncthreads();
    }
    if (P.return_levels) flush(I);
  };
#define GPSIG_L2(I_) if constexpr (MMAX >= I_) run_level(std::integral_constant<int, I_>{});
  GPSIG_L2(1) GPSIG_L2(2) GPSIG_L2(3) GPSIG_L2(4) GPSIG_L2(5) GPSIG_L2(6) GPSIG_L2(7) GPSIG_L2(8)
#undef GPSIG_L2
  if (!P.return_levels) flush(0);
}

template <typename T, int OP, int DP, int ZS, int MMAX, int EBITS = 11>
int launch_tens2(const TensSeqParams& P, cudaStream_t stream) {
  const size_t xslot = (P.xblk * sizeof(T) + 127) / 128 * 128;
  const int ncols = P.col_end - P.col_begin;
  const int zblocks = (P.Mz + kTens2Warps * 32 - 1) / (kTens2Warps * 32);
  int ch = (int)((long)ncols * zblocks / (2 * 148));                              // at least two CTAs per SM before a CTA takes a second sequence
  ch = ch < 1 ? 1 : (ch > kTens2Chunk ? kTens2Chunk : ch);
  const size_t etab = (sizeof(T) == 8 && OP == OP_RBF) ? ((size_t)(1 << EBITS) + 256) * sizeof(double) + 16 : 0;   // table + fill scratch
  if constexpr (EBITS == 11 && GPSIG_TENS_EXP2_8K != 0 && sizeof(T) == 8 && OP == OP_RBF) {
    // the 8192-entry table whenever two CTAs per SM still fit with it
    const size_t big = 128 + 2 * xslot + (size_t)kTens2Chunk * (kTens2Warps * 32 + 1) * sizeof(T) + ((size_t)(1 << 13) + 256) * sizeof(double) + 16;
    if (2 * (big + 1024) <= 227 * 1024) return launch_tens2<T, OP, DP, ZS, MMAX, 13>(P, stream);
  }
  const size_t smem = 128 + 2 * xslot + (size_t)ch * (kTens2Warps * 32 + 1) * sizeof(T) + etab;
  if (128 + 2 * xslot + (size_t)kTens2Chunk * (kTens2Warps * 32 + 1) * sizeof(T) + etab > kTens2MaxSmem) return -100;                                           // caller falls back to tens_vs_seq_kernel
  auto kern = tens_vs_seq_kernel2<T, DP, OP, ZS, MMAX, EBITS>;
  GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((ncols + ch - 1) / ch, zblocks);
  prof_begin(stream);
  kern<<<grid, kTens2Warps * 32, smem, stream>>>(P, ch);
  prof_end(stream);
  GPSIG_LAUNCH_CHECK("tens_vs_seq_kernel2");
  return GPSIG_OK;
}

template <typename T, int OP, int DP>
int launch_tens_vs_seq(const TensSeqParams& P, int m, cudaStream_t stream) {
  if (m < 1 || m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", m);
  // register-resident tensors when the top level's z tile fits in <= 144 registers (C4: m=4, zs=2, DP=8: 72 f32 / 144 f64)
  if (getenv("GPSIG_TENS_OLD") == nullptr) {
    const int mm = m <= 2 ? 2 : (m <= 4 ? 4 : 8);
    if (tens2_zregs<T>(mm, P.zs, DP) <= 144) {
      int rc = -100;
#define GPSIG_T2(ZS_, MM_) rc = launch_tens2<T, OP, DP, ZS_, MM_>(P, stream)
      if (P.zs == 2) { if (mm == 2) GPSIG_T2(2, 2); else if (mm == 4) GPSIG_T2(2, 4); else GPSIG_T2(2, 8); }
      else { if (mm == 2) GPSIG_T2(1, 2); else if (mm == 4) GPSIG_T2(1, 4); else GPSIG_T2(1, 8); }
#undef GPSIG_T2
      if (rc != -100) return rc;
    }
  }
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

// One thread per entry (a, b).  Each component's coordinates (both tensors, both increment ends, scaled by the
// lengthscales) are loaded into registers once and the four base-kernel values of kernels.py:278 come from those:
// the loads are issued back to back instead of one dependent L1 round trip per term.
template <typename T, int DP>
__global__ void __launch_bounds__(128) tens_tens_kernel(const TensTensParams P) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;   // column z'
  const int a = blockIdx.y;                              // row z
  if (b >= P.Mz) return;
  const int d = P.d, zs = P.zs;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* ls = static_cast<const T*>(P.level_scale);
  const T* len = static_cast<const T*>(P.ls);
  T* out = static_cast<T*>(P.out);
  const size_t MM = (size_t)P.Mz * P.Mz, ab = (size_t)a * P.Mz + b;
  auto kappa_spectral = [&](const T* za, const T* zb) {
    T u[16], v[16];
    for (int k = 0; k < d; ++k) { u[k] = za[k]; v[k] = zb[k]; if (len) { u[k] = u[k] / len[k]; v[k] = v[k] / len[k]; } }
    return spectral_eval<T>(P.sp, d, u, v);
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
      if (P.base == GPSIG_BASE_SPECTRAL) {
        if (zs == 2) Mt = kappa_spectral(za + d, zb + d) + kappa_spectral(za, zb) - kappa_spectral(za + d, zb) - kappa_spectral(za, zb + d);
        else Mt = kappa_spectral(za, zb);
      } else {
        T u[2][DP], v[2][DP], uq[2], vq[2];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          uq[s] = vq[s] = T(0);
#pragma unroll
          for (int q = 0; q < DP; ++q) {
            const bool live = s < zs && q < d;
            T uu = live ? za[s * d + q] : T(0), vv = live ? zb[s * d + q] : T(0);
            if (len && live) { uu = uu / len[q]; vv = vv / len[q]; }
            u[s][q] = uu; v[s][q] = vv;
            uq[s] = gfma<T>(uu, uu, uq[s]); vq[s] = gfma<T>(vv, vv, vq[s]);
          }
        }
        auto ev = [&](int s, int s2) {
          T dot = T(0);
#pragma unroll
          for (int q = 0; q < DP; ++q) dot = gfma<T>(u[s][q], v[s2][q], dot);
          return base_eval_call<T>(P.base, p0, p1, dot, uq[s], vq[s2]);
        };
        if (zs == 2) Mt = ev(1, 1) + ev(0, 0) - ev(1, 0) - ev(0, 1);                                  // kernels.py:278
        else Mt = ev(0, 0);
      }
      R = jj == 0 ? Mt : Mt * R;                                                                     // signature_algs.py:91-96
    }
    if (ls) R *= ls[i];
    if (P.return_levels) out[(size_t)i * MM + ab] = R; else acc += R;
  }
  if (!P.return_levels) out[ab] = acc;
}

// d > 16: the same entry by entry with the contraction as a runtime loop over d (Kzz is tiny: Mz^2 * T * 4 values)
template <typename T>
__global__ void __launch_bounds__(128) tens_tens_big_kernel(const TensTensParams P) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int a = blockIdx.y;
  if (b >= P.Mz) return;
  const int d = P.d, zs = P.zs;
  const T p0 = T(P.p0), p1 = T(P.p1);
  const T* ls = static_cast<const T*>(P.level_scale);
  const T* len = static_cast<const T*>(P.ls);
  T* out = static_cast<T*>(P.out);
  const size_t MM = (size_t)P.Mz * P.Mz, ab = (size_t)a * P.Mz + b;
  auto ev = [&](const T* u, const T* v) {
    T dot = T(0), uq = T(0), vq = T(0);
    for (int q = 0; q < d; ++q) {
      T uu = u[q], vv = v[q];
      if (len) { uu = uu / len[q]; vv = vv / len[q]; }
      dot = gfma<T>(uu, vv, dot); uq = gfma<T>(uu, uu, uq); vq = gfma<T>(vv, vv, vq);
    }
    return base_eval_call<T>(P.base, p0, p1, dot, uq, vq);
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
      if (zs == 2) Mt = ev(za + d, zb + d) + ev(za, zb) - ev(za + d, zb) - ev(za, zb + d);             // kernels.py:278
      else Mt = ev(za, zb);
      R = jj == 0 ? Mt : Mt * R;                                                                     // signature_algs.py:91-96
    }
    if (ls) R *= ls[i];
    if (P.return_levels) out[(size_t)i * MM + ab] = R; else acc += R;
  }
  if (!P.return_levels) out[ab] = acc;
}

}  // namespace gpsig
