// Kzx (inducing tensors vs sequences, gpsig/kernels.py:322-341 + signature_algs.py:101-127) at state dimensions above 16:
// the lifted form.
//
// At d <= 16 tens_vs_seq_kernel2 keeps a tensor's coordinates in registers and evaluates the base kernel inside the time
// recursion.  Above that the coordinates do not fit and the general kernel (a runtime loop over d per base-kernel value,
// one thread per (z, x) pair) took 785 ms at d = 28 where d = 8 takes 14 ms (c4 shape).  At large d the base-kernel
// matrix is what costs -- 2 d flops per entry against ~3 for the recursion -- and it is a plain GEMM:
//   (A) tens_lift_gemm_kernel: C[r, e] = kappa(zpoint_r, xpoint_e) for ALL tensor points r = ((c zs + s) Mz + z) against
//       the points e = l nb + n of a block of nb sequences (time-major, so that what the recursion reads per step is
//       contiguous along the sequence index): the register-tiled, double-buffered, persistent DFMA / FFMA GEMM of
//       lift_tile_kernel2 (lifted_kernels.cuh) with the base-kernel epilogue fused;
//   (B) tens_lift_recur_kernel: one thread per (tensor z, sequence n), lanes along n: one pass over time advances the
//       running sums of all levels (signature_algs.py:119-125) from coalesced reads of C, then the epilogue of
//       kernels.py:604-620.
// C (8 bytes x Mz m(m+1)/2 zs x L per sequence) lives in the caller's workspace, sized for blocks of sequences.
#pragma once
#include <stdlib.h>

#include "tens_kernels.cuh"
#include "trunc_kernel.cuh"

namespace gpsig {

constexpr size_t kTensLiftCap = (size_t)1 << 30;       // bytes of lifted values in flight
// state dimensions that take the lifted form: above 16, and float64 from 9 (the register tile of tens_vs_seq_kernel2 holds
// m zs (DP + 1) doubles: it stops fitting at DP = 16)
inline bool tens_lift_dims(int d, int elem) { return d > 16 || (d > 8 && elem == 8); }
// sequences per block (a multiple of 32 when there are that many) and the scratch it takes
inline int tens_lift_block(int Mz, int zs, int m, int L, int ncols, int elem) {
  const size_t per_seq = (size_t)(m * (m + 1) / 2) * zs * Mz * L * elem;
  size_t cap = kTensLiftCap;
  if (const char* e = getenv("GPSIG_TENS_LIFT_CAP_KB")) { const long v = atol(e); if (v > 0) cap = (size_t)v << 10; }   // debug switch (tests: many blocks)
  long nb = (long)(cap / (per_seq ? per_seq : 1));
  if (nb >= 32) nb = nb / 32 * 32;
  if (nb < 1) nb = 1;
  if (nb > ncols) nb = ncols;
  return (int)nb;
}
inline size_t tens_lift_bytes(int Mz, int zs, int m, int L, int ncols, int elem) {
  const int nb = tens_lift_block(Mz, zs, m, L, ncols, elem);
  return ((size_t)(m * (m + 1) / 2) * zs * Mz * L * nb * elem + 255) & ~(size_t)255;
}

template <typename T, int TM, int TN>
__global__ void __launch_bounds__(256) tens_lift_gemm_kernel(const TensSeqParams P, const int DP, const int n0, const int nb,
                                                              T* __restrict__ C, const int nrb, const int ncb) {
  constexpr int BM = 16 * TM, BN = 16 * TN, BK = 16;
  constexpr int SA = BM + 1, SB = BN + 1;
  constexpr int VEC = 16 / sizeof(T), UPR = BK / VEC;                 // elements per 16-byte load, loads per row and slab
  constexpr int UA = (BM * UPR + 255) / 256, UB = (BN * UPR + 255) / 256;
  using V16 = typename Vec16<T>::type;
  extern __shared__ __align__(16) unsigned char tl_smem[];
  T* As = reinterpret_cast<T*>(tl_smem);
  T* Bs = As + 2 * BK * SA;
  T* Qs = Bs + 2 * BK * SB;                                           // [BM + BN] squared norms
  static_assert(BM + BN <= 256, "one squared norm per thread");
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int L = P.L, Mz = P.Mz, zs = P.zs;
  const int R = P.m * (P.m + 1) / 2 * zs * Mz, E = L * nb;
  const T* zp = static_cast<const T*>(P.zp);
  const T* xp = static_cast<const T*>(P.xp);
  auto zrow = [&](int r, const T*& blk, int& sidx) {                  // r = (c zs + s) Mz + z
    r = min(r, R - 1);
    const int cs = r / Mz, z = r - cs * Mz, c = cs / zs;
    sidx = cs - c * zs;
    blk = zp + (size_t)(c * Mz + z) * P.zblk;
  };
  auto xrow = [&](int e, const T*& blk, int& l) {                     // e = l nb + n
    e = min(e, E - 1);
    l = e / nb;
    blk = xp + (size_t)(P.col_begin + n0 + (e - l * nb)) * P.xblk;
  };
  V16 ra[UA], rb[UB];
  auto fetch = [&](int row0, int col0, int k0) {
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int idx = tid + 256 * u, r = idx / UPR, kk = (idx - r * UPR) * VEC;
      if (idx < BM * UPR) { const T* blk; int s; zrow(row0 + r, blk, s); ra[u] = *reinterpret_cast<const V16*>(blk + s * DP + k0 + kk); }
    }
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int idx = tid + 256 * u, r = idx / UPR, kk = (idx - r * UPR) * VEC;
      if (idx < BN * UPR) { const T* blk; int l; xrow(col0 + r, blk, l); rb[u] = *reinterpret_cast<const V16*>(blk + (size_t)l * DP + k0 + kk); }
    }
  };
  auto stash = [&](int buf) {
    T* A = As + buf * BK * SA;
    T* B = Bs + buf * BK * SB;
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int idx = tid + 256 * u, r = idx / UPR, kk = (idx - r * UPR) * VEC;
      if (idx < BM * UPR) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) A[(kk + v) * SA + r] = vget(ra[u], v);
      }
    }
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int idx = tid + 256 * u, r = idx / UPR, kk = (idx - r * UPR) * VEC;
      if (idx < BN * UPR) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) B[(kk + v) * SB + r] = vget(rb[u], v);
      }
    }
  };
  const T p0 = T(P.p0), p1 = T(P.p1);
  const int base = P.base;
  const double tab = exp2((double)(tid & 31) * (1.0 / 32.0));
  const int nitems = nrb * ncb;
  int it = blockIdx.x;
  if (it >= nitems) return;
  fetch((it / ncb) * BM, (it % ncb) * BN, 0);
  for (;;) {
    const int row0 = (it / ncb) * BM, col0 = (it % ncb) * BN;
    stash(0);
    __syncthreads();
    const int nx = it + (int)gridDim.x;
    T qv = T(0);
    if (tid < BM) { const T* blk; int s; zrow(row0 + tid, blk, s); qv = blk[zs * DP + s]; }
    else if (tid < BM + BN) { const T* blk; int l; xrow(col0 + tid - BM, blk, l); qv = blk[(size_t)L * DP + l]; }
    T acc[TM][TN];
#pragma unroll
    for (int a = 0; a < TM; ++a)
#pragma unroll
      for (int b = 0; b < TN; ++b) acc[a][b] = T(0);
    int buf = 0;
    for (int k0 = 0; k0 < DP; k0 += BK) {
      const bool more = k0 + BK < DP;
      if (more) fetch(row0, col0, k0 + BK);
      else if (nx < nitems) fetch((nx / ncb) * BM, (nx % ncb) * BN, 0);
      const T* A = As + buf * BK * SA + ty;
      const T* B = Bs + buf * BK * SB + tx;
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        T av[TM], bv[TN];
#pragma unroll
        for (int a = 0; a < TM; ++a) av[a] = A[k * SA + 16 * a];
#pragma unroll
        for (int b = 0; b < TN; ++b) bv[b] = B[k * SB + 16 * b];
#pragma unroll
        for (int a = 0; a < TM; ++a)
#pragma unroll
          for (int b = 0; b < TN; ++b) acc[a][b] = gfma<T>(av[a], bv[b], acc[a][b]);
      }
      if (more) stash(buf ^ 1);
      __syncthreads();
      buf ^= 1;
    }
    if (tid < BM + BN) Qs[tid] = qv;
    __syncthreads();
    const T* zsq = Qs + ty;
    const T* xsq = Qs + BM + tx;
    if (base == GPSIG_BASE_RBF) {
      // 2^((z.x - |z|^2/2 - |x|^2/2) log2 e): exp2_tab32 (float64; the warp stays convergent, stores predicated) / MUFU.EX2
      constexpr double kLog2e = 1.4426950408889634;
      T xc[TN];
#pragma unroll
      for (int b = 0; b < TN; ++b) xc[b] = T(-0.5 * kLog2e) * xsq[16 * b];
#pragma unroll
      for (int a = 0; a < TM; ++a) {
        const int r = row0 + ty + 16 * a;
        const T zc = T(-0.5 * kLog2e) * zsq[16 * a];
#pragma unroll
        for (int b = 0; b < TN; ++b) {
          const int e = col0 + tx + 16 * b;
          const T t = gfma<T>(acc[a][b], T(kLog2e), zc + xc[b]);
          T v;
          if constexpr (sizeof(T) == 8) v = T(exp2_tab32((double)t, tab)); else v = T(exp2_approx((float)t));
          if (r < R && e < E) C[(size_t)r * E + e] = v;
        }
      }
    } else {
#pragma unroll
      for (int a = 0; a < TM; ++a) {
        const int r = row0 + ty + 16 * a;
        if (r >= R) continue;
        const T zq = zsq[16 * a];
#pragma unroll
        for (int b = 0; b < TN; ++b) {
          const int e = col0 + tx + 16 * b;
          if (e >= E) continue;
          C[(size_t)r * E + e] = base == GPSIG_BASE_LINEAR ? acc[a][b] : base_eval_call<T>(base, p0, p1, acc[a][b], zq, xsq[16 * b]);
        }
      }
    }
    if (nx >= nitems) break;
    it = nx;
  }
}

// (B) one thread per (tensor z, sequence n of the block); MMAX >= m levels in registers
template <typename T, int MMAX>
__global__ void __launch_bounds__(128) tens_lift_recur_kernel(const TensSeqParams P, const int n0, const int nb, const T* __restrict__ C) {
  const int n = blockIdx.x * 128 + threadIdx.x, z = blockIdx.y;
  if (n >= nb) return;
  const int L = P.L, Mz = P.Mz, zs = P.zs, m = P.m;
  const size_t E = (size_t)L * nb;
  constexpr int TT = MMAX * (MMAX + 1) / 2;
  T S[TT], Mprev[TT];
#pragma unroll
  for (int k = 0; k < TT; ++k) { S[k] = T(0); Mprev[k] = T(0); }
  const bool diff = P.difference != 0;
  for (int l = 0; l < L; ++l) {
    const T* col = C + (size_t)l * nb + n;
    const bool first = diff && l == 0;
#pragma unroll
    for (int I = 1; I <= MMAX; ++I) {
      if (I > m) break;
      constexpr int dummy = 0; (void)dummy;
      const int k0 = I * (I - 1) / 2;
#pragma unroll
      for (int jj = I - 1; jj >= 0; --jj) {
        const int c = k0 + jj;
        T Mt;
        if (zs == 2) Mt = col[((size_t)(c * 2 + 1) * Mz + z) * E] - col[((size_t)(c * 2) * Mz + z) * E];      // kernels.py:331
        else Mt = col[((size_t)c * Mz + z) * E];
        T D = Mt;
        if (diff) { D = Mt - Mprev[c]; Mprev[c] = Mt; }                                                          // signature_algs.py:114
        if (!first) {
          if (jj > 0) S[c] = gfma<T>(D, S[c - 1], S[c]);                                                        // signature_algs.py:123
          else S[c] += D;
        }
      }
    }
  }
  // epilogue (kernels.py:604-620), level by level
  const int ncols = P.col_end - P.col_begin;
  const int nl = n0 + n, ng = P.col_begin + nl;
  T acc = T(0);
  auto emit = [&](int lvl, T v) {
    if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + ng]);
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
    if (P.return_levels) static_cast<T*>(P.out)[((size_t)lvl * Mz + z) * ncols + nl] = v; else acc += v;
  };
  emit(0, T(1));
#pragma unroll
  for (int I = 1; I <= MMAX; ++I) if (I <= m) emit(I, S[I * (I - 1) / 2 + I - 1]);                              // signature_algs.py:125
  if (!P.return_levels) static_cast<T*>(P.out)[(size_t)z * ncols + nl] = acc;
}

// (B') per-level output (the differentiable path asks for raw levels): one thread per (tensor z, sequence n, LEVEL I), the
// levels being independent chains - m times the threads of (B) for the small blocks of a training minibatch
template <typename T, int MMAX>
__global__ void __launch_bounds__(128) tens_lift_recur_level_kernel(const TensSeqParams P, const int n0, const int nb, const T* __restrict__ C) {
  const int n = blockIdx.x * 128 + threadIdx.x, z = blockIdx.y;
  const int I = P.m - (int)blockIdx.z;
  if (n >= nb) return;
  const int L = P.L, Mz = P.Mz, zs = P.zs;
  const size_t E = (size_t)L * nb, rs = (size_t)Mz * E;
  const T* base = C + n + ((size_t)(I * (I - 1) / 2) * zs * Mz + z) * E;
  const bool diff = P.difference != 0;
  T S[MMAX], Mp[MMAX];
#pragma unroll
  for (int k = 0; k < MMAX; ++k) { S[k] = T(0); Mp[k] = T(0); }
  for (int l = 0; l < L; ++l) {
    const T* col = base + (size_t)l * nb;
    const bool first = diff && l == 0;
#pragma unroll
    for (int jj = MMAX - 1; jj >= 0; --jj) {
      if (jj >= I) continue;
      const T Mt = zs == 2 ? col[(size_t)(jj * 2 + 1) * rs] - col[(size_t)(jj * 2) * rs] : col[(size_t)jj * rs];      // kernels.py:331
      T D = Mt;
      if (diff) { D = Mt - Mp[jj]; Mp[jj] = Mt; }                                                                    // signature_algs.py:114
      if (!first) { if (jj > 0) S[jj] = gfma<T>(D, S[jj - 1], S[jj]); else S[jj] += D; }                             // signature_algs.py:123
    }
  }
  const int ncols = P.col_end - P.col_begin;
  const int nl = n0 + n, ng = P.col_begin + nl;
  auto emit = [&](int lvl, T v) {                                                                                    // kernels.py:604-620
    if (P.norm_x) v = v / gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + ng]);
    if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
    static_cast<T*>(P.out)[((size_t)lvl * Mz + z) * ncols + nl] = v;
  };
  if (I == 1) emit(0, T(1));
  T top = T(0);
#pragma unroll
  for (int jj = 0; jj < MMAX; ++jj) if (jj == I - 1) top = S[jj];
  emit(I, top);                                                                                                      // signature_algs.py:125
}

// (C) reverse mode on the lifted values: one thread per (tensor z, sequence n, LEVEL I) - the levels' recursions are
// independent of each other (level I only touches its own I components), so a thread keeps I <= MMAX values per array instead
// of m (m + 1) / 2 (the first version, one thread per (z, n), spilled 700 bytes at m = 4 in float64 and had a quarter of the
// threads: 1.10 ms at minibatch 50, d = 28).  C holds kappa (the forward GEMM) and is overwritten IN PLACE by
// Gd[r, e] = dL/dkappa(r, e) (times kappa for the RBF kernel, whose partials are kappa (x - z)), from which the caller forms
// dL/dZ = Gd X - rowsum(Gd) Z and dL/dX = Gd^T Z - colsum(Gd) X (linear: without the sums) as two plain GEMMs.  Forward
// recursion S_jj(l) = S_jj(l-1) + D_jj(l) S_{jj-1}(l-1); the states are rebuilt backwards from the final ones
// (S_jj(l-1) = S_jj(l) - D_jj(l) S_{jj-1}(l-1)), so nothing but C is stored; adjoints lam_jj as in vjp_kernels.cuh.
template <typename T, int MMAX>
__global__ void __launch_bounds__(128) tens_lift_reverse_kernel(const TensSeqParams P, const int nb, T* __restrict__ C,
                                                                 const T* __restrict__ g, const int scale_by_value) {
  const int n = blockIdx.x * 128 + threadIdx.x, z = blockIdx.y;
  const int I = P.m - (int)blockIdx.z;                  // the top level (the longest chain) first
  if (n >= nb) return;
  const int L = P.L, Mz = P.Mz, zs = P.zs;
  const size_t E = (size_t)L * nb;
  const bool diff = P.difference != 0;
  const int kb = I * (I - 1) / 2;
  T* base = C + n + ((size_t)kb * zs * Mz + z) * E;     // component kb + jj, end s: + (jj zs + s) Mz E
  const size_t rs = (size_t)Mz * E;
  T S[MMAX], Mp[MMAX];
#pragma unroll
  for (int k = 0; k < MMAX; ++k) { S[k] = T(0); Mp[k] = T(0); }
  // pass 1: the final running sums
  for (int l = 0; l < L; ++l) {
    const T* col = base + (size_t)l * nb;
    const bool first = diff && l == 0;
#pragma unroll
    for (int jj = MMAX - 1; jj >= 0; --jj) {
      if (jj >= I) continue;
      const T Mt = zs == 2 ? col[(size_t)(jj * 2 + 1) * rs] - col[(size_t)(jj * 2) * rs] : col[(size_t)jj * rs];
      T D = Mt;
      if (diff) { D = Mt - Mp[jj]; Mp[jj] = Mt; }
      if (!first) { if (jj > 0) S[jj] = gfma<T>(D, S[jj - 1], S[jj]); else S[jj] += D; }
    }
  }
  // pass 2: backwards
  T lam[MMAX], aDn[MMAX], k1[MMAX], k0v[MMAX];
#pragma unroll
  for (int k = 0; k < MMAX; ++k) { lam[k] = T(0); aDn[k] = T(0); k0v[k] = T(0); k1[k] = T(0); }
  const int ncols = P.col_end - P.col_begin;
  const T gI = g[((size_t)I * Mz + z) * ncols + n];
#pragma unroll
  for (int jj = 0; jj < MMAX; ++jj) if (jj == I - 1) lam[jj] = gI;
  {                                                     // kappa values of the last time step
    const T* col = base + (size_t)(L - 1) * nb;
#pragma unroll
    for (int jj = 0; jj < MMAX; ++jj)
      if (jj < I) { k0v[jj] = col[(size_t)(jj * zs) * rs]; if (zs == 2) k1[jj] = col[(size_t)(jj * zs + 1) * rs]; }
  }
  for (int l = L - 1; l >= 0; --l) {
    T* col = base + (size_t)l * nb;
    const bool first = diff && l == 0;
    T p1[MMAX], p0[MMAX];                               // kappa values of time l - 1, in flight while time l is worked on
#pragma unroll
    for (int jj = 0; jj < MMAX; ++jj) { p0[jj] = T(0); p1[jj] = T(0); }
    if (l > 0) {
      const T* colp = col - nb;
#pragma unroll
      for (int jj = 0; jj < MMAX; ++jj)
        if (jj < I) { p0[jj] = colp[(size_t)(jj * zs) * rs]; if (zs == 2) p1[jj] = colp[(size_t)(jj * zs + 1) * rs]; }
    }
    T Sprev_below = T(1);                               // S_{jj-1}(l-1), S_{-1} = 1
#pragma unroll
    for (int jj = 0; jj < MMAX; ++jj) {                 // ascending: lam[jj-1] is used before it is updated
      if (jj >= I) continue;
      const T Mt = zs == 2 ? k1[jj] - k0v[jj] : k0v[jj];
      const T Mtp = zs == 2 ? p1[jj] - p0[jj] : p0[jj];
      const T D = diff ? Mt - Mtp : Mt;
      T aD = T(0);
      T Sprev = S[jj];
      if (!first) {
        Sprev = jj > 0 ? gfma<T>(-D, Sprev_below, S[jj]) : S[jj] - D;          // S_jj(l-1)
        aD = jj > 0 ? lam[jj] * Sprev_below : lam[jj];
        if (jj > 0) lam[jj - 1] = gfma<T>(lam[jj], D, lam[jj - 1]);
      }
      // dL/dMt_jj(l) = aD(l) - aD(l+1) in difference mode, aD(l) otherwise
      const T gM = diff ? aD - aDn[jj] : aD;
      aDn[jj] = aD;
      if (zs == 2) {
        col[(size_t)(jj * 2 + 1) * rs] = scale_by_value ? gM * k1[jj] : gM;
        col[(size_t)(jj * 2) * rs] = scale_by_value ? -gM * k0v[jj] : -gM;
      } else col[(size_t)jj * rs] = scale_by_value ? gM * k0v[jj] : gM;
      Sprev_below = Sprev;                              // becomes S_jj(l-1) for the component above
      S[jj] = Sprev;
    }
#pragma unroll
    for (int jj = 0; jj < MMAX; ++jj) { k0v[jj] = p0[jj]; k1[jj] = p1[jj]; }
  }
}

template <typename T>
int launch_tens_seq_lifted(const TensSeqParams& P, int DP, void* scratch, size_t scratch_bytes, cudaStream_t s);
// kappa of all tensor points against the columns [col_begin, col_end) into C[R, L * ncols], then (g != NULL) the reverse pass in place
template <typename T>
int launch_tens_seq_lifted_gd(const TensSeqParams& P, int DP, T* C, const T* g, cudaStream_t s);

}  // namespace gpsig
