// Fused truncated signature-kernel Gram for sm_100a.
//
// Replaces gpsig/kernels.py:209-238 (_K_seq) + :190-207 (_K_seq_diag) +
// gpsig/signature_algs.py:8-35 (signature_kern_first_order) and the
// normalisation / variance epilogue of kernels.py:431-477.  The reference
// materialises the lifted tensor M[N, L, N2, L2] in HBM and streams it through
// ~3+4(m-1) TensorFlow kernels; here one WARP owns one pair (x_i, y_j) and the
// L x L2 tile never leaves registers:
//
//   lane t owns COLS consecutive columns q of the tile and sweeps the rows p;
//   per row it evaluates the base kernel M[p+1, q] (x row broadcast from
//   shared memory, y columns resident in registers or lane-interleaved shared
//   memory), forms the 2-D increment D[p, q] (signature_algs.py:26) from the
//   previous row and the neighbour lane, and advances ALL levels at once:
//       C_i[q]   = sum_{p' < p} R_i[p', q]            (column prefix, registers)
//       E_i[q]   = sum_{q' < q} C_i[q']                (warp-shuffle prefix scan)
//       R_{i+1}[p, q] = D[p, q] * E_i[q]               (signature_algs.py:32)
//       K_i      = sum_q C_i[q] after the last row     (signature_algs.py:33)
//   which is the reference's  R <- M * cumsum_excl(cumsum_excl(R, rows), cols)
//   re-associated so that the row-cumsum is the loop-carried state.
//
// A CTA is 8 warps = 8 columns sequences y_j; it walks a chunk of rows x_i,
// each x_i arriving in shared memory as one TMA bulk copy (double buffered,
// mbarrier completion), shared by the 8 warps.
#pragma once
#include "common.cuh"

namespace gpsig {

constexpr int kTruncWarps = 8;

struct TruncParams {
  const void* xp;          // packed X  (see common.cuh)
  const void* yp;          // packed Y  (== xp for K(X, X) and for the diagonal)
  size_t xblk, yblk;       // packed block sizes in elements
  int xrs, yrs;            // row strides of the packed blocks in elements (DP, or the padded stride of the systolic kernel)
  int N, N2, L, L2;
  int row_begin, row_end;  // rows of X this call covers
  int ichunk;              // rows per CTA
  int m;                   // levels wanted (<= MLEV)
  int difference;
  int nchunks;             // column chunks of 32*COLS-1
  int x_in_smem;           // stage x_i with TMA (else broadcast loads from L1/L2)
  int symmetric;           // X2 == NULL
  int tri;                 // symmetric Gram over all rows: CTAs strictly below the diagonal exit, mirror_lower_kernel fills them in
  int diag;                // pairs (i, i) only: one pair per warp, x rows read through L1
  double p0, p1; int base; // base-kernel parameters
  const void* norm_x; const void* norm_x2; const void* level_scale;
  double sym_jitter; int return_levels;
  void* out;
  int dry_run;             // systolic launcher only: report whether the shape fits its on-chip plan, launch nothing
};

template <typename T> struct Vec16;   // 16-byte vector of T
template <> struct Vec16<double> { static constexpr int V = 2; using type = double2; };
template <> struct Vec16<float> { static constexpr int V = 4; using type = float4; };
__device__ __forceinline__ double vget(const double2& v, int e) { return e == 0 ? v.x : v.y; }
__device__ __forceinline__ float vget(const float4& v, int e) { return e == 0 ? v.x : (e == 1 ? v.y : (e == 2 ? v.z : v.w)); }

// y operand placement: registers (fp32: COLS*DP floats fit) or, for fp64, a
// lane-interleaved per-warp shared-memory tile read with conflict-free LDS.128.
// In registers when the operand is at most 64 32-bit registers per lane.
template <typename T, int DP, int COLS> struct YPolicy { static constexpr bool kInRegs = COLS * DP * sizeof(T) <= 256; };

template <typename T, int DP, int COLS>
__host__ __device__ constexpr size_t trunc_ywarp_bytes() { return YPolicy<T, DP, COLS>::kInRegs ? 0 : (size_t)COLS * 32 * DP * sizeof(T); }

__host__ __device__ inline size_t align_up_dev(size_t x) { return (x + 127) / 128 * 128; }

template <typename T, int DP, int COLS, int MLEV, int OP>
__global__ void __launch_bounds__(kTruncWarps * 32, 1) trunc_gram_kernel(const TruncParams P) {
  using V16 = typename Vec16<T>::type;
  constexpr int V = Vec16<T>::V;
  constexpr int KV = DP / V;
  constexpr int CS = 32 * COLS - 1;           // D columns per chunk (one M column of overlap)
  constexpr bool YREG = YPolicy<T, DP, COLS>::kInRegs;
  using Op = BaseOp<T, OP>;

  extern __shared__ __align__(128) unsigned char trunc_smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(trunc_smem);          // [2]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const bool DIAG = P.diag != 0;
  const int diff = P.difference ? 1 : 0;
  const int nrows = P.L - diff, ncols = P.L2 - diff;
  const size_t xbytes = P.xblk * sizeof(T);
  unsigned char* sp = trunc_smem + 128;
  T* xbuf0 = reinterpret_cast<T*>(sp);
  T* xbuf1 = reinterpret_cast<T*>(sp + (P.x_in_smem && !DIAG ? align_up_dev(xbytes) : 0));
  sp += (P.x_in_smem && !DIAG) ? 2 * align_up_dev(xbytes) : 0;
  T* ysm = reinterpret_cast<T*>(sp) + (size_t)warp * (COLS * 32 * DP);   // only if !YREG
  sp += YREG ? 0 : nwarps * trunc_ywarp_bytes<T, DP, COLS>();
  // row carries of the column-chunked scan: [warp][MLEV-1][nrows]
  T* carry = reinterpret_cast<T*>(sp) + (size_t)warp * (MLEV - 1) * nrows;

  const int j = blockIdx.x * nwarps + warp;
  const bool active = j < P.N2;
  const T* yblk = static_cast<const T*>(P.yp) + (size_t)(active ? j : 0) * P.yblk;
  const T* ycoord = yblk;
  const T* ysqn = yblk + (size_t)P.L2 * DP;

  int i_lo, i_hi;
  if (DIAG) { i_lo = j; i_hi = j + 1; }
  else {
    i_lo = P.row_begin + blockIdx.y * P.ichunk;
    i_hi = min(P.row_end, i_lo + P.ichunk);
  }
  const bool use_tma = P.x_in_smem && !DIAG;
  if (use_tma && threadIdx.x == 0) {
    mbar_init(&full_bar[0], 1); mbar_init(&full_bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (use_tma && threadIdx.x == 0 && i_lo < i_hi) {
    mbar_expect_tx(&full_bar[0], (uint32_t)xbytes);
    tma_load_1d(xbuf0, static_cast<const T*>(P.xp) + (size_t)i_lo * P.xblk, (uint32_t)xbytes, &full_bar[0]);
  }

  for (int i = i_lo, it = 0; i < i_hi; ++i, ++it) {
    const T* xrows; const T* xsqn;
    if (use_tma) {
      const int buf = it & 1;
      if (threadIdx.x == 0 && i + 1 < i_hi) {      // prefetch x_{i+1}; buffer buf^1 was released by the barrier below
        mbar_expect_tx(&full_bar[buf ^ 1], (uint32_t)xbytes);
        tma_load_1d(buf ? xbuf0 : xbuf1, static_cast<const T*>(P.xp) + (size_t)(i + 1) * P.xblk, (uint32_t)xbytes, &full_bar[buf ^ 1]);
      }
      mbar_wait(&full_bar[buf], (it >> 1) & 1);
      xrows = buf ? xbuf1 : xbuf0;
    } else {
      xrows = static_cast<const T*>(P.xp) + (size_t)(DIAG ? (active ? i : 0) : i) * P.xblk;
    }
    xsqn = xrows + (size_t)P.L * DP;

    if (active) {
      T ktot[MLEV];
#pragma unroll
      for (int l = 0; l < MLEV; ++l) ktot[l] = T(0);

      for (int chunk = 0; chunk < P.nchunks; ++chunk) {
        const int q0 = chunk * CS + lane * COLS;       // first tile column of this lane
        // ---- y operand for this chunk --------------------------------------
        T yc[COLS];
        V16 yreg[YREG ? COLS : 1][YREG ? KV : 1];
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          const int q = min(q0 + c, P.L2 - 1);
          yc[c] = Op::col_const(ysqn[q]);
#pragma unroll
          for (int kv = 0; kv < KV; ++kv) {
            V16 v = *reinterpret_cast<const V16*>(ycoord + (size_t)q * DP + kv * V);
            if (YREG) yreg[YREG ? c : 0][YREG ? kv : 0] = v;
            else *reinterpret_cast<V16*>(ysm + ((size_t)(c * KV + kv) * 32 + lane) * V) = v;
          }
        }
        if (!YREG) __syncwarp();

        auto eval_row = [&](int prow, T (&Mrow)[COLS]) {
          const T xc = Op::row_const(xsqn[prow]);
          T acc[COLS];
#pragma unroll
          for (int c = 0; c < COLS; ++c) acc[c] = Op::init(xc, yc[c]);
#pragma unroll
          for (int kv = 0; kv < KV; ++kv) {
            const V16 xv = *reinterpret_cast<const V16*>(xrows + (size_t)prow * DP + kv * V);
#pragma unroll
            for (int c = 0; c < COLS; ++c) {
              V16 yv;
              if (YREG) yv = yreg[YREG ? c : 0][YREG ? kv : 0];
              else yv = *reinterpret_cast<const V16*>(ysm + ((size_t)(c * KV + kv) * 32 + lane) * V);
#pragma unroll
              for (int e = 0; e < V; ++e) acc[c] = gfma<T>(vget(xv, e), vget(yv, e), acc[c]);
            }
          }
#pragma unroll
          for (int c = 0; c < COLS; ++c) Mrow[c] = Op::finish(acc[c], xc, yc[c], P.base, T(P.p0), T(P.p1));
        };

        // validity of this lane's D columns
        bool valid[COLS];
#pragma unroll
        for (int c = 0; c < COLS; ++c) valid[c] = (q0 + c < ncols) && (lane * COLS + c < CS);

        T C[MLEV][COLS];
#pragma unroll
        for (int l = 0; l < MLEV; ++l)
#pragma unroll
          for (int c = 0; c < COLS; ++c) C[l][c] = T(0);

        T Mprev[COLS];
        if (diff) eval_row(0, Mprev);

        for (int r = 0; r < nrows; ++r) {
          T D[COLS];
          {
            T Mcur[COLS];
            eval_row(r + diff, Mcur);
            if (diff) {
              T U[COLS];
#pragma unroll
              for (int c = 0; c < COLS; ++c) { U[c] = Mcur[c] - Mprev[c]; Mprev[c] = Mcur[c]; }
              const T Unext = __shfl_down_sync(0xffffffffu, U[0], 1);
#pragma unroll
              for (int c = 0; c < COLS; ++c) D[c] = (c + 1 < COLS ? U[c + 1] : Unext) - U[c];
            } else {
#pragma unroll
              for (int c = 0; c < COLS; ++c) D[c] = Mcur[c];
            }
#pragma unroll
            for (int c = 0; c < COLS; ++c) D[c] = valid[c] ? D[c] : T(0);
          }
          // ---- all levels, highest first so each C_i is read before row r updates it
#pragma unroll
          for (int l = MLEV - 2; l >= 0; --l) {
            T loc[COLS];
            T run = T(0);
#pragma unroll
            for (int c = 0; c < COLS; ++c) { loc[c] = run; run += C[l][c]; }
            T incl = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const T t = __shfl_up_sync(0xffffffffu, incl, o);
              if (lane >= o) incl += t;
            }
            T off = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) off = T(0);
            if (P.nchunks > 1) {
              T* cw = carry + (size_t)l * nrows + r;
              const T base = chunk > 0 ? *cw : T(0);
              off += base;
              __syncwarp();
              if (lane == 31 && chunk + 1 < P.nchunks) *cw = base + incl;
            }
#pragma unroll
            for (int c = 0; c < COLS; ++c) C[l + 1][c] = gfma<T>(D[c], off + loc[c], C[l + 1][c]);
          }
#pragma unroll
          for (int c = 0; c < COLS; ++c) C[0][c] += D[c];
        }
#pragma unroll
        for (int l = 0; l < MLEV; ++l) {
          T s = T(0);
#pragma unroll
          for (int c = 0; c < COLS; ++c) s += C[l][c];
          ktot[l] += s;
        }
        if (P.nchunks > 1) __syncwarp();
      }

      // ---- epilogue: reduce, normalise, scale, store (kernels.py:431-477) ----
      T mine = T(1);                                   // level 0 is identically 1 (signature_algs.py:20)
#pragma unroll
      for (int l = 0; l < MLEV; ++l) {
        const T k = warp_sum<T>(ktot[l]);
        if (lane == l + 1) mine = k;
      }
      const int lvl = lane;
      if (DIAG) {
        if (lvl <= P.m) static_cast<T*>(P.out)[(size_t)lvl * P.N + i] = mine;
      } else {
        T v = T(0);
        if (lvl <= P.m) {
          v = mine;
          if (P.symmetric && i == j) v += T(P.sym_jitter);
          if (P.norm_x) {
            const T a = gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + i]);
            const T b = gsqrt<T>(static_cast<const T*>(P.norm_x2)[(size_t)lvl * P.N2 + j]);
            v = v / (a * b);
          }
          if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
        }
        const size_t rows = (size_t)(P.row_end - P.row_begin);
        const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
        if (P.return_levels) {
          if (lvl <= P.m) static_cast<T*>(P.out)[(size_t)lvl * rows * P.N2 + rc] = v;
        } else {
          v = warp_sum<T>(v);
          if (lane == 0) static_cast<T*>(P.out)[rc] = v;
        }
      }
    }
    if (use_tma) __syncthreads();     // everyone is done with x_i before its buffer is refilled
  }
}

// ---------------------------------------------------------------------------
// host-side launcher: picks the instantiation and the shared-memory plan
// ---------------------------------------------------------------------------
constexpr size_t kMaxSmem = 227 * 1024;

template <typename T, int DP, int COLS, int MLEV, int OP>
int launch_trunc_inst(TruncParams P, cudaStream_t stream) {
  const bool DIAG = P.diag != 0;
  const int diff = P.difference ? 1 : 0;
  const int nrows = P.L - diff, ncols = P.L2 - diff;
  constexpr int CS = 32 * COLS - 1;
  P.nchunks = (ncols + CS - 1) / CS;
  const size_t xbytes = align_up_dev(P.xblk * sizeof(T));
  int nwarps = kTruncWarps;
  size_t fixed = 0;
  for (; nwarps >= 1; nwarps >>= 1) {   // fewer warps per CTA when the per-warp tiles are large
    fixed = 128 + nwarps * trunc_ywarp_bytes<T, DP, COLS>() +
            (P.nchunks > 1 ? (size_t)nwarps * (MLEV - 1) * nrows * sizeof(T) : 0);
    if (fixed <= kMaxSmem) break;
  }
  if (nwarps < 1) nwarps = 1;
  P.x_in_smem = (!DIAG && fixed + 2 * xbytes <= kMaxSmem) ? 1 : 0;
  const size_t smem = fixed + (P.x_in_smem ? 2 * xbytes : 0);
  if (smem > kMaxSmem)
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "sequence too long for the on-chip plan: needs %zu bytes of shared memory (L=%d, L2=%d, levels=%d)",
               smem, P.L, P.L2, MLEV);
  auto kern = trunc_gram_kernel<T, DP, COLS, MLEV, OP>;
  GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid;
  grid.x = (P.N2 + nwarps - 1) / nwarps;
  if (DIAG) { grid.y = 1; P.ichunk = 1; }
  else {
    // >= ~24 waves of CTAs over 148 SMs (1 CTA/SM) while keeping >= 8 rows per CTA to amortise the y load
    const int rows = P.row_end - P.row_begin;
    int want_y = (24 * 148 + (int)grid.x - 1) / (int)grid.x;
    int ichunk = (rows + want_y - 1) / want_y;
    if (ichunk < 8) ichunk = 8;
    if (ichunk > rows) ichunk = rows;
    P.ichunk = ichunk;
    grid.y = (rows + ichunk - 1) / ichunk;
  }
  if (!DIAG) prof_begin(stream);
  kern<<<grid, nwarps * 32, smem, stream>>>(P);
  if (!DIAG) prof_end(stream);
  GPSIG_LAUNCH_CHECK("trunc_gram_kernel");
  return GPSIG_OK;
}

// Instantiation tiers.  The hot specialisations (linear / RBF base kernel, d <= 8)
// get every (columns-per-lane, levels) shape; the rest get a sparse grid and run
// the next larger shape (extra levels are computed and dropped, extra columns idle).
#ifdef GPSIG_FAST_BUILD   // developer builds: only the shapes of the BASELINE configs
#define GPSIG_COLS_FULL(X) X(2) X(4) X(7)
#define GPSIG_MLEV_FULL(X) X(4) X(5)
#define GPSIG_COLS_SPARSE(X) X(4) X(8)
#define GPSIG_MLEV_SPARSE(X) X(5)
#else
#define GPSIG_COLS_FULL(X) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8)
#define GPSIG_MLEV_FULL(X) X(3) X(4) X(5) X(6) X(8)
#define GPSIG_COLS_SPARSE(X) X(2) X(4) X(8)
#define GPSIG_MLEV_SPARSE(X) X(4) X(6) X(8)
#endif

template <int OP, int DP> struct TruncTier { static constexpr bool kFull = (OP != OP_GENERIC) && DP <= 8; };

template <typename T, int OP, int DP, int COLS>
int launch_trunc_mlev(const TruncParams& P, cudaStream_t stream) {
#define GPSIG_CASE(M_) if (P.m <= M_) return launch_trunc_inst<T, DP, COLS, M_, OP>(P, stream);
  if constexpr (TruncTier<OP, DP>::kFull) { GPSIG_MLEV_FULL(GPSIG_CASE) } else { GPSIG_MLEV_SPARSE(GPSIG_CASE) }
#undef GPSIG_CASE
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported by this build", P.m);
}

// One translation unit per (T, OP, DP) instantiates this.
// COLS = columns per lane: the smallest built value with 32*COLS-1 >= ncols, else the
// largest one with the row column-chunked.
template <typename T, int OP, int DP>
int launch_trunc(const TruncParams& P, cudaStream_t stream) {
  if (P.m < 1 || P.m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", P.m);
  const int ncols = P.L2 - (P.difference ? 1 : 0);
  const int need = (ncols + 1 + 31) / 32;
  int last = 0;
#define GPSIG_CASE(C_) last = C_; if (need <= C_) return launch_trunc_mlev<T, OP, DP, C_>(P, stream);
  if constexpr (TruncTier<OP, DP>::kFull) { GPSIG_COLS_FULL(GPSIG_CASE) } else { GPSIG_COLS_SPARSE(GPSIG_CASE) }
#undef GPSIG_CASE
  // longer than one warp pass: column-chunk with the widest shape
#define GPSIG_CASE(C_) if (last == C_) return launch_trunc_mlev<T, OP, DP, C_>(P, stream);
  if constexpr (TruncTier<OP, DP>::kFull) { GPSIG_COLS_FULL(GPSIG_CASE) } else { GPSIG_COLS_SPARSE(GPSIG_CASE) }
#undef GPSIG_CASE
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "no kernel shape for %d columns", ncols);
}

}  // namespace gpsig
