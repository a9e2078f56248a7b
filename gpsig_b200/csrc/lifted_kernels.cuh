// Truncated signature-kernel Gram for LARGE state dimension (d > 16) on sm_100a.
//
// Same contract as trunc_skew_kernel.cuh / trunc_kernel.cuh (gpsig/kernels.py:209-238 _K_seq, :190-207
// _K_seq_diag, gpsig/signature_algs.py:8-35, epilogue kernels.py:431-477).  The reference has no limit on the
// state dimension (kernels.py:343-365 reshapes to num_features * (num_lags + 1)); its own experiment driver runs
// d_eff = 26 ... 1928 (benchmarks/datasets.json with num_lags = 1 and a time channel).
//
// At d <= 16 the K = d contraction is a quarter of the per-element work and lives in the lanes' FMA streams of
// the systolic kernel.  At large d it is most of the work (2 d flops against ~20 for the recursion) and it is a
// plain GEMM, so the path is split where the work changes character:
//
//   (A) LIFT: for a block of sequence pairs, the base-kernel tile M[p, q] = kappa(x_i[p], y_j[q])
//       (kernels.py:225-231) is produced as a GEMM with the base-kernel epilogue fused
//         * float32: tcgen05.mma kind::tf32 with a 3-term split (x = hi + lo; hi.hi + hi.lo + lo.hi), operands
//           pre-packed in the canonical K-major core-matrix layout and brought in by 1-D TMA bulk copies,
//           accumulator in tensor memory, epilogue warps read it back with tcgen05.ld (lift_mma_kernel);
//         * float64 (no FP64 MMA kind exists): register-tiled DFMA GEMM (lift_tile_kernel; also float32 for
//           shapes the MMA kernel does not take)
//       and written to a scratch in the SKEWED order the recursion consumes it: element (p, q) of the n-th pair
//       of a stream, owned by lane t = q / COLS, goes to row n*L + p + t of a [rows][32][COLS] array, so that
//       the warp that runs the recursion reads ONE contiguous 32*COLS-element line per step.
//   (B) RECURSION: one warp per stream of pairs, the lane-skewed systolic recursion of trunc_skew_kernel.cuh
//       (lane t one row behind lane t-1, chain values by one shuffle, level totals carried to lane 31 at the
//       first row of the next pair) with the base-kernel evaluation replaced by that one coalesced load.
//
// The lifted tile takes a round trip through HBM/L2 here (4-8 bytes per element against >= 2 d flops); the
// reference round-trips the WHOLE lifted tensor 3 + 4 (m-1) times.
#pragma once
#include <stdlib.h>

#include "trunc_kernel.cuh"

namespace gpsig {

struct LiftParams {
  const void* xp; const void* yp;        // packed sequences, plain layout (common.cuh), row stride DP
  size_t xblk, yblk;
  int DP;
  int N, N2, L, L2;
  int diag;                              // pairs (i, i)
  int row_begin, row_end;                // rows of the whole call (output indexing)
  int ib, ie, jb, je;                    // this block: rows [ib, ie) x columns [jb, je)   (diag: pairs ib..ie)
  int SP;                                // pairs per stream (consecutive rows i, same column j)
  int nci;                               // streams per column: ceil((ie - ib) / SP)
  int nstreams;
  size_t stream_elems;                   // (SP * L + 32) * 32 * COLS
  int m, difference, symmetric;
  double p0, p1; int base;
  const void* norm_x; const void* norm_x2; const void* level_scale;
  double sym_jitter; int return_levels;
  void* scratch;
  void* out;
};

// pair n of stream sg -> (i, j); false if the stream has no such pair
__device__ __forceinline__ bool lift_pair(const LiftParams& P, int sg, int n, int& i, int& j) {
  if (P.diag) { i = P.ib + sg * P.SP + n; j = i; }
  else { const int jj = sg / P.nci, ci = sg - jj * P.nci; i = P.ib + ci * P.SP + n; j = P.jb + jj; }
  return i < P.ie;
}
__device__ __forceinline__ int lift_stream_pairs(const LiftParams& P, int sg) {
  const int ci = P.diag ? sg : sg % P.nci;
  return min(P.SP, P.ie - P.ib - ci * P.SP);
}

// ---------------------------------------------------------------------------
// (A) register-tiled GEMM with the base-kernel epilogue: one CTA per (pair, 64-row block)
// ---------------------------------------------------------------------------
template <typename T, int COLS>
__global__ void __launch_bounds__(256) lift_tile_kernel(const LiftParams P) {
  constexpr int BM = 64, BN = 64, BK = 16, PAD = 4;
  __shared__ __align__(16) T As[BK][BM + PAD];
  __shared__ __align__(16) T Bs[BK][BN + PAD];
  const int pi = blockIdx.x;                       // pair of the block
  const int sg = pi / P.SP, n = pi - sg * P.SP;
  int i, j;
  if (sg >= P.nstreams || !lift_pair(P, sg, n, i, j)) return;
  const int L = P.L, L2 = P.L2, DP = P.DP;
  const int row0 = blockIdx.y * BM;
  if (row0 >= L) return;
  const T* x = static_cast<const T*>(P.xp) + (size_t)i * P.xblk;
  const T* y = static_cast<const T*>(P.yp) + (size_t)j * P.yblk;
  const T* xsq = x + (size_t)L * DP;
  const T* ysq = y + (size_t)L2 * DP;
  T* S = static_cast<T*>(P.scratch) + (size_t)sg * P.stream_elems;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int lr = tid >> 2, lk = (tid & 3) * 4;     // loader: row of the tile, first k of 4
  const T p0 = T(P.p0), p1 = T(P.p1);
  constexpr int LP2 = 32 * COLS;
  for (int col0 = 0; col0 < LP2; col0 += BN) {
    T acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = T(0);
    const T* xr = x + (size_t)min(row0 + lr, L - 1) * DP;
    const T* yr = y + (size_t)min(col0 + lr, L2 - 1) * DP;       // columns past L2 repeat the last point: zero increments
    for (int k0 = 0; k0 < DP; k0 += BK) {
      T xa[4], ya[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) { xa[e] = xr[k0 + lk + e]; ya[e] = yr[k0 + lk + e]; }
      __syncthreads();
#pragma unroll
      for (int e = 0; e < 4; ++e) { As[lk + e][lr] = xa[e]; Bs[lk + e][lr] = ya[e]; }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        T av[4], bv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) av[a] = As[k][ty * 4 + a];
#pragma unroll
        for (int b = 0; b < 4; ++b) bv[b] = Bs[k][tx * 4 + b];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = gfma<T>(av[a], bv[b], acc[a][b]);
      }
    }
    // base kernel (kernels.py:946-1173) + store in the skewed order of the recursion
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      const int p = row0 + ty * 4 + a;
      if (p >= L) continue;
      const T xs = xsq[p];
      T v[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int q = min(col0 + tx * 4 + b, L2 - 1);
        v[b] = base_eval<T>(P.base, p0, p1, acc[a][b], xs, ysq[q]);
      }
      const int q = col0 + tx * 4;
      if constexpr (COLS >= 4) {
        const int t = q / COLS, c = q - t * COLS;
        T* dst = S + ((size_t)(n * L + p + t) * 32 + t) * COLS + c;
#pragma unroll
        for (int b = 0; b < 4; ++b) dst[b] = v[b];
      } else {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int t = (q + b) / COLS, c = (q + b) - t * COLS;
          S[((size_t)(n * L + p + t) * 32 + t) * COLS + c] = v[b];
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// (A') float64 lift, second version: one CTA per (pair, BM x BN block) with BM = 16 TM, BN = 16 TN and a TM x TN register
// tile per thread whose rows / columns are STRIDED by 16 (thread (ty, tx) owns rows ty + 16 a, columns tx + 16 b), so the
// block can be 112 x 112 (TM = TN = 7: a 100-point pair wastes 20 % of the tile, not the 39 % of 64-blocks), the
// shared-memory reads are a broadcast (A) and 16 consecutive doubles (B), 49 DFMA per 14 loads.  k-slabs of 16 are
// double buffered: the next slab is fetched into registers while the current one is multiplied, one barrier per slab.
// Only the L2 real columns go through the GEMM; the slots past the last column (the recursion wants 32 COLS of them)
// repeat the last column's value, written by the thread that owns it.
// c_large_d float64 (N = N2 = 1000, L = 100, d = 126): lift 430 -> see DESIGN 4.8.
// ---------------------------------------------------------------------------
template <typename T, int COLS, int TM, int TN>
__global__ void __launch_bounds__(256) lift_tile_kernel2(const LiftParams P, const int ncb, const int nblk, const int nitems) {
  constexpr int BM = 16 * TM, BN = 16 * TN, BK = 16;
  constexpr int SA = BM + 1, SB = BN + 1;            // odd strides: the transposing stores of a k-slab are conflict free
  extern __shared__ __align__(16) unsigned char lift2_smem[];
  T* As = reinterpret_cast<T*>(lift2_smem);          // [2][BK][SA]
  T* Bs = As + 2 * BK * SA;                          // [2][BK][SB]
  T* Qs = Bs + 2 * BK * SB;                          // [BM + BN] squared norms of the block's rows / columns
  T* VL = Qs + BM + BN;                              // [BM] the last real column's value per row (for the padded slots)
  static_assert(sizeof(T) == 8, "lift_tile_kernel2 is the float64 lift");
  static_assert(BM + BN <= 256, "one squared norm per thread");
  const int L = P.L, L2 = P.L2, DP = P.DP;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  // Work items (pair, block of the pair) are taken round robin by a PERSISTENT grid: with 200 registers only one CTA
  // fits an SM, so the first k-slab of the NEXT item is fetched (into registers) before the epilogue of the current
  // one - the global-load latency and the epilogue would otherwise be exposed once per item.
  struct Item { const T* x; const T* y; T* S; int n, row0, col0; };
  auto setup = [&](int it, Item& I) -> bool {
    const int pi = it / nblk, blk = it - pi * nblk;
    const int sg = pi / P.SP;
    I.n = pi - sg * P.SP;
    int i, j;
    if (sg >= P.nstreams || !lift_pair(P, sg, I.n, i, j)) return false;
    I.row0 = (blk / ncb) * BM; I.col0 = (blk % ncb) * BN;
    I.x = static_cast<const T*>(P.xp) + (size_t)i * P.xblk;
    I.y = static_cast<const T*>(P.yp) + (size_t)j * P.yblk;
    I.S = static_cast<T*>(P.scratch) + (size_t)sg * P.stream_elems;
    return true;
  };
  auto next_valid = [&](int it, Item& I) -> int {
    while (it < nitems && !setup(it, I)) it += gridDim.x;
    return it;
  };
  // loader: unit u = (row, pair of k); a slab of A has BM * 8 units of two doubles
  constexpr int UA = (BM * 8 + 255) / 256, UB = (BN * 8 + 255) / 256;
  using V2 = typename Vec16<T>::type;                // two doubles
  V2 ra[UA], rb[UB];
  auto fetch = [&](const Item& I, int k0) {
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int idx = tid + 256 * u, r = idx >> 3, kk = (idx & 7) * 2;
      if (idx < BM * 8) ra[u] = *reinterpret_cast<const V2*>(I.x + (size_t)min(I.row0 + r, L - 1) * DP + k0 + kk);
    }
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int idx = tid + 256 * u, r = idx >> 3, kk = (idx & 7) * 2;
      if (idx < BN * 8) rb[u] = *reinterpret_cast<const V2*>(I.y + (size_t)min(I.col0 + r, L2 - 1) * DP + k0 + kk);
    }
  };
  auto stash = [&](int buf) {
    T* A = As + buf * BK * SA;
    T* B = Bs + buf * BK * SB;
#pragma unroll
    for (int u = 0; u < UA; ++u) {
      const int idx = tid + 256 * u, r = idx >> 3, kk = (idx & 7) * 2;
      if (idx < BM * 8) { A[kk * SA + r] = ra[u].x; A[(kk + 1) * SA + r] = ra[u].y; }
    }
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const int idx = tid + 256 * u, r = idx >> 3, kk = (idx & 7) * 2;
      if (idx < BN * 8) { B[kk * SB + r] = rb[u].x; B[(kk + 1) * SB + r] = rb[u].y; }
    }
  };
  const T p0 = T(P.p0), p1 = T(P.p1);
  const int base = P.base;
  constexpr int LP2 = 32 * COLS;
  const double tab = exp2((double)(tid & 31) * (1.0 / 32.0));

  Item cur;
  int it = next_valid((int)blockIdx.x, cur);
  if (it >= nitems) return;
  fetch(cur, 0);
  for (;;) {
    stash(0);
    __syncthreads();
    Item nxt;
    const int nx = next_valid(it + (int)gridDim.x, nxt);
    // squared norms of this block's points: requested now, parked in shared memory before the epilogue
    T qv = T(0);
    if (tid < BM) qv = cur.x[(size_t)L * DP + min(cur.row0 + tid, L - 1)];
    else if (tid < BM + BN) qv = cur.y[(size_t)L2 * DP + min(cur.col0 + tid - BM, L2 - 1)];
    T acc[TM][TN];
#pragma unroll
    for (int a = 0; a < TM; ++a)
#pragma unroll
      for (int b = 0; b < TN; ++b) acc[a][b] = T(0);
    int buf = 0;
    for (int k0 = 0; k0 < DP; k0 += BK) {
      const bool more = k0 + BK < DP;
      if (more) fetch(cur, k0 + BK);
      else if (nx < nitems) fetch(nxt, 0);
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
    // base kernel (kernels.py:946-1173) + store in the skewed order of the recursion
    const int row0 = cur.row0, col0 = cur.col0, n = cur.n;
    T* S = cur.S;
    if (tid < BM + BN) Qs[tid] = qv;
    __syncthreads();
    const T* xsq = Qs + ty;                           // [16 a]
    const T* ysq = Qs + BM + tx;                      // [16 b]
    auto put = [&](int p, int q, T v) {
      const int t = q / COLS, c = q - t * COLS;
      S[((size_t)(n * L + p + t) * 32 + t) * COLS + c] = v;
    };
    const bool last_cols = col0 <= L2 - 1 && L2 - 1 < col0 + BN;      // this block holds the last real column
    if (base == GPSIG_BASE_RBF) {
      // exp(-(|x|^2 + |y|^2 - 2 x.y) / 2) = 2^((x.y - |x|^2/2 - |y|^2/2) log2 e) by the table exp2 of the fused kernels
      // (exp2_tab32: 9 FP64 instructions + a table shuffle against ~35 for exp(); the warp stays convergent: every lane
      // evaluates all its TM x TN slots on clamped indices, only the stores are predicated)
      constexpr double kLog2e = 1.4426950408889634;
      T yc[TN];
#pragma unroll
      for (int b = 0; b < TN; ++b) yc[b] = T(-0.5 * kLog2e) * ysq[16 * b];
#pragma unroll
      for (int a = 0; a < TM; ++a) {
        const int p = row0 + ty + 16 * a;
        const T xc = T(-0.5 * kLog2e) * xsq[16 * a];
#pragma unroll
        for (int b = 0; b < TN; ++b) {
          const int q = col0 + tx + 16 * b;
          const T v = T(exp2_tab32((double)gfma<T>(acc[a][b], T(kLog2e), xc + yc[b]), tab));
          if (p < L && q < L2) put(p, q, v);
          if (q == L2 - 1) VL[ty + 16 * a] = v;
        }
      }
    } else {
#pragma unroll
      for (int a = 0; a < TM; ++a) {
        const int p = row0 + ty + 16 * a;
        if (p >= L) continue;
        const T xs = xsq[16 * a];
#pragma unroll
        for (int b = 0; b < TN; ++b) {
          const int q = col0 + tx + 16 * b;
          if (q >= L2) continue;
          // one out-of-line copy of the base-kernel switch: inlined at all TM x TN sites it made 34 k instructions of
          // epilogue, which ran out of the instruction cache (55 % of the samples)
          const T v = base == GPSIG_BASE_LINEAR ? acc[a][b] : base_eval_call<T>(base, p0, p1, acc[a][b], xs, ysq[16 * b]);
          put(p, q, v);
          if (q == L2 - 1) VL[ty + 16 * a] = v;
        }
      }
    }
    // the slots past the last column repeat it (zero increments): all threads share the rows x (32 COLS - L2) stores
    if (last_cols && LP2 > L2) {                      // uniform over the CTA
      __syncthreads();
      const int npad = LP2 - L2, nrow = min(BM, L - row0);
      for (int e = tid; e < nrow * npad; e += 256) {
        const int r = e / npad, qq = L2 + (e - r * npad);
        put(row0 + r, qq, VL[r]);
      }
    }
    if (nx >= nitems) break;
    cur = nxt;
    it = nx;
  }
}

template <typename T, int COLS, int TM, int TN>
int launch_lift_tile2(const LiftParams& P, int npair_slots, cudaStream_t s) {
  constexpr int BM = 16 * TM, BN = 16 * TN, BK = 16;
  const int nrb = (P.L + BM - 1) / BM, ncb = (P.L2 + BN - 1) / BN;
  const size_t smem = ((size_t)2 * BK * ((BM + 1) + (BN + 1)) + 2 * BM + BN) * sizeof(T);
  auto kern = lift_tile_kernel2<T, COLS, TM, TN>;
  if (smem > 48 * 1024) GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem) != cudaSuccess || occ < 1) occ = 1;
  const long nitems = (long)npair_slots * nrb * ncb;
  if (nitems > 0x7fffffffL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "too many lift tiles in one block (%ld)", nitems);
  const long want = (long)occ * device_sm_count();
  const unsigned grid = (unsigned)(nitems < want ? nitems : want);
  kern<<<grid, 256, smem, s>>>(P, ncb, nrb * ncb, (int)nitems);
  GPSIG_LAUNCH_CHECK("lift_tile_kernel2");
  return GPSIG_OK;
}
// block shape by the waste of the pair's L x L2 points: 112 x 112 unless 64-blocks tile them tighter
inline bool lift_tile2_wide(int L, int L2) {
  auto cover = [](int n, int b) { return (long)((n + b - 1) / b) * b; };
  return cover(L, 112) * cover(L2, 112) <= cover(L, 64) * cover(L2, 64);
}

// ---------------------------------------------------------------------------
// (B) the lane-skewed recursion on a stream of lifted tiles: one warp per stream
// ---------------------------------------------------------------------------
template <typename T, int COLS> struct LiftVec;
template <> struct LiftVec<float, 2> { using type = float2; };
template <> struct LiftVec<float, 4> { using type = float4; };
template <> struct LiftVec<float, 8> { using type = float4; };
template <> struct LiftVec<double, 2> { using type = double2; };
template <> struct LiftVec<double, 4> { using type = double2; };
template <> struct LiftVec<double, 8> { using type = double2; };

template <typename T, int COLS>
__device__ __forceinline__ void lift_load_row(T (&dst)[COLS], const T* src) {
  using VT = typename LiftVec<T, COLS>::type;
  constexpr int NV = COLS * sizeof(T) / sizeof(VT);
  union { VT v[NV]; T e[COLS]; } u;
#pragma unroll
  for (int k = 0; k < NV; ++k) u.v[k] = __ldcs(reinterpret_cast<const VT*>(src) + k);    // streamed once: evict first
#pragma unroll
  for (int c = 0; c < COLS; ++c) dst[c] = u.e[c];
}

constexpr int kLiftWarps = 4;       // streams per CTA of the recursion kernel
#ifndef GPSIG_LIFT_BATCH
#define GPSIG_LIFT_BATCH 4
#endif
constexpr int kLiftBatch = GPSIG_LIFT_BATCH;   // lines per request burst

template <typename T, int COLS, int MLEV>
__global__ void __launch_bounds__(kLiftWarps * 32) lift_recur_kernel(const LiftParams P) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sg = blockIdx.x * kLiftWarps + warp;
  if (sg >= P.nstreams) return;
  const int npairs = lift_stream_pairs(P, sg);
  if (npairs <= 0) return;
  const int L = P.L, L2 = P.L2;
  const bool diff = P.difference != 0;
  const T* S = static_cast<const T*>(P.scratch) + (size_t)sg * P.stream_elems + (size_t)lane * COLS;
  constexpr size_t LINE = 32 * COLS;
  const int q0 = lane * COLS;

  T C[MLEV - 1][COLS], Ct[COLS], Mp[COLS];
#pragma unroll
  for (int c = 0; c < COLS; ++c) {
    Ct[c] = T(0); Mp[c] = T(0);
#pragma unroll
    for (int l = 0; l < MLEV - 1; ++l) C[l][c] = T(0);
  }
  T send[MLEV], sendU = T(0);
#pragma unroll
  for (int l = 0; l < MLEV; ++l) send[l] = T(0);

  const int nsteps = npairs * L + 32;              // + the step that carries the last totals to lane 31
  int r = -lane, seg = 0;                          // this lane's row within its pair, and its pair
  // lines are requested kLiftBatch at a time (one contiguous kLiftBatch * 32 * COLS-element run per request burst:
  // DRAM pages are opened for 2-4 KB instead of one 512-byte line), one batch ahead of the batch being consumed
  T bufA[kLiftBatch][COLS], bufB[kLiftBatch][COLS];
  auto load_batch = [&](T (&dst)[kLiftBatch][COLS], int s0) {
#pragma unroll
    for (int a = 0; a < kLiftBatch; ++a) lift_load_row<T, COLS>(dst[a], S + (size_t)min(s0 + a, nsteps - 1) * LINE);
  };
  load_batch(bufA, 0);
  const size_t rows_out = (size_t)(P.row_end - P.row_begin);

  auto step = [&](const T (&cur)[COLS]) {
    // chain values of the lane to the left (the same row, one step ago)
    T recv[MLEV];
    T recvU = __shfl_up_sync(0xffffffffu, sendU, 1);
#pragma unroll
    for (int l = 0; l < MLEV; ++l) {
      recv[l] = __shfl_up_sync(0xffffffffu, send[l], 1);
      if (lane == 0) recv[l] = T(0);
    }
    const bool started = r >= 0;
    T D[COLS];
    if (diff) {
      // 2-D increment (signature_algs.py:26): row 0 of a pair is consumed by the first-row branch,
      // padded columns repeat the last one (increment exactly 0)
      T U[COLS];
#pragma unroll
      for (int c = 0; c < COLS; ++c) { U[c] = cur[c] - Mp[c]; Mp[c] = cur[c]; }
      const T left = lane == 0 ? U[0] : recvU;
#pragma unroll
      for (int c = 0; c < COLS; ++c) D[c] = U[c] - (c > 0 ? U[c - 1] : left);
      sendU = U[COLS - 1];
    } else {
#pragma unroll
      for (int c = 0; c < COLS; ++c) D[c] = (q0 + c < L2) ? cur[c] : T(0);
    }
    if (started && r == 0) {
      // first row of a pair: the chain carries the level totals of the pair just finished
      T tot[MLEV];
#pragma unroll
      for (int l = 0; l < MLEV; ++l) {
        T e = recv[l];
        if (l < MLEV - 1) {
#pragma unroll
          for (int c = 0; c < COLS; ++c) e += C[l < MLEV - 1 ? l : 0][c];
        } else {
#pragma unroll
          for (int c = 0; c < COLS; ++c) e += Ct[c];
        }
        tot[l] = e;
        send[l] = e;
      }
      if (lane == 31 && seg > 0) {
        int i, j;
        lift_pair(P, sg, seg - 1, i, j);
        // ---- epilogue (kernels.py:431-477; raw levels for the diagonal, kernels.py:190-207) ----
        T sum = T(0);
#pragma unroll
        for (int lvl = 0; lvl <= MLEV; ++lvl) {
          if (lvl <= P.m) {
            T v = lvl == 0 ? T(1) : tot[lvl > 0 ? lvl - 1 : 0];
            if (P.diag) { static_cast<T*>(P.out)[(size_t)lvl * P.N + i] = v; continue; }
            if (P.symmetric && i == j) v += T(P.sym_jitter);
            if (P.norm_x) {
              const T a = gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + i]);
              const T b = gsqrt<T>(static_cast<const T*>(P.norm_x2)[(size_t)lvl * P.N2 + j]);
              v = v / (a * b);
            }
            if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
            const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
            if (P.return_levels) static_cast<T*>(P.out)[(size_t)lvl * rows_out * P.N2 + rc] = v;
            else sum += v;
          }
        }
        if (!P.diag && !P.return_levels) static_cast<T*>(P.out)[(size_t)(i - P.row_begin) * P.N2 + j] = sum;
      }
#pragma unroll
      for (int c = 0; c < COLS; ++c) {
        Ct[c] = T(0);
#pragma unroll
        for (int l = 1; l < MLEV - 1; ++l) C[l][c] = T(0);
        C[0][c] = diff ? T(0) : D[c];
      }
    } else {
      // C_{l+1}[q] += D[p, q] * sum_{q' < q} C_l[q'], highest level first (signature_algs.py:32 re-associated)
#pragma unroll
      for (int l = MLEV - 2; l >= 0; --l) {
        T e = recv[l];
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          const T t = e;
          e += C[l][c];
          if (l == MLEV - 2) Ct[c] = gfma<T>(D[c], t, Ct[c]);
          else C[l + 1 < MLEV - 1 ? l + 1 : 0][c] = gfma<T>(D[c], t, C[l + 1 < MLEV - 1 ? l + 1 : 0][c]);
        }
        send[l] = e;
      }
#pragma unroll
      for (int c = 0; c < COLS; ++c) C[0][c] += D[c];
    }
    if (++r == L) { r = 0; ++seg; }
  };

  for (int s = 0; s < nsteps; s += 2 * kLiftBatch) {
    load_batch(bufB, s + kLiftBatch);
#pragma unroll
    for (int a = 0; a < kLiftBatch; ++a) if (s + a < nsteps) step(bufA[a]);
    load_batch(bufA, s + 2 * kLiftBatch);
#pragma unroll
    for (int a = 0; a < kLiftBatch; ++a) if (s + kLiftBatch + a < nsteps) step(bufB[a]);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
inline int lift_cols(int L2) { return L2 <= 64 ? 2 : (L2 <= 128 ? 4 : 8); }
inline bool lift_eligible(int L, int L2) { return L2 <= 256 && L >= 1; }
constexpr size_t kLiftScratchCap = (size_t)2 << 30;      // bytes of lifted tiles in flight
constexpr int kLiftWantStreams = 148 * 32;               // streams (= warps of the recursion kernel) wanted per block

struct LiftPlan { int COLS, SP, max_streams; size_t stream_elems, scratch_bytes; };
// npairs: pairs of the whole call (sizes the streams); deterministic in (npairs, L, L2, elem)
inline LiftPlan lift_plan(size_t npairs, int L, int L2, int elem) {
  LiftPlan pl;
  pl.COLS = lift_cols(L2);
  size_t capb = kLiftScratchCap;
  if (const char* e = getenv("GPSIG_LIFT_CAP_MB")) { const long v = atol(e); if (v > 0) capb = (size_t)v << 20; }   // debug switch (tests: many blocks)
  // pairs per stream: long streams amortise the 31-step pipeline fill, short ones give the recursion kernel enough
  // warps per block of scratch: as long as kLiftWantStreams streams fit the cap, and the call has that many
  const size_t line = (size_t)32 * pl.COLS * elem;
  long long rows_fit = (long long)(capb / ((size_t)kLiftWantStreams * line)) - 32;
  long long sp = rows_fit / L;
  const long long sp_call = (long long)(npairs / (148 * 24));
  if (sp > sp_call) sp = sp_call;
  pl.SP = sp < 1 ? 1 : (sp > 8 ? 8 : (int)sp);
  pl.stream_elems = ((size_t)pl.SP * L + 32) * 32 * pl.COLS;
  const size_t per = pl.stream_elems * elem;
  size_t want = (npairs + pl.SP - 1) / pl.SP;
  size_t cap = capb / per;
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  pl.max_streams = (int)(want > 0x3fffffff ? 0x3fffffff : want);
  pl.scratch_bytes = (size_t)pl.max_streams * per;
  return pl;
}

struct LiftMmaOperands;
// P: everything but the block fields; scratch >= lift_plan(...).scratch_bytes; mma: the pre-split tensor-core
// operands of the float32 path or NULL (lifted.cu)
template <typename T>
int launch_lift(LiftParams P, void* scratch, size_t scratch_bytes, const LiftMmaOperands* mma, cudaStream_t s);

}  // namespace gpsig
