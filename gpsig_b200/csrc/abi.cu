// extern "C" entry points of libgpsig_b200.so (see include/gpsig_b200.h).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "pde_kernel.cuh"
#include "vjp_kernels.cuh"
#include "tens_kernels.cuh"
#include "trunc_kernel.cuh"
#include "trunc_skew_kernel.cuh"
#include "generic_kernels.cuh"
#include "lift_mma_kernel.cuh"
#include "generic_vjp_kernels.cuh"
#include "tens_lifted_kernels.cuh"

namespace gpsig {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ---- diagnostics ------------------------------------------------------------
static unsigned long long g_launches = 0;
void count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }

static thread_local int g_prof_on = 0;
static thread_local cudaEvent_t g_prof_ev[2] = {nullptr, nullptr};
static thread_local int g_prof_valid = 0;
void prof_begin(cudaStream_t s) {
  if (!g_prof_on) return;
  if (!g_prof_ev[0]) { cudaEventCreate(&g_prof_ev[0]); cudaEventCreate(&g_prof_ev[1]); }
  cudaEventRecord(g_prof_ev[0], s);
}
void prof_end(cudaStream_t s) {
  if (!g_prof_on) return;
  cudaEventRecord(g_prof_ev[1], s);
  g_prof_valid = 1;
}

// FMA-chain microbenchmark: 8 independent chains per thread, 2 flops per FMA.
template <typename T>
__global__ void __launch_bounds__(256) fma_peak_kernel(T* out, int iters, T a, T b) {
  T v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = T(threadIdx.x + k);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = gfma<T>(v[k], a, b);
  }
  T s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == T(123.456)) out[0] = s;   // never true; keeps the chains alive
}

// explicit instantiations live in trunc_inst.cu / tens_inst.cu (one object per T, OP, DP)
#define GPSIG_DECL(T, OP, DP)                                                             \
  extern template int launch_trunc<T, OP, DP>(const TruncParams&, cudaStream_t);          \
  extern template int launch_trunc_skew<T, OP, DP>(const TruncParams&, cudaStream_t);     \
  extern template int launch_tens_vs_seq<T, OP, DP>(const TensSeqParams&, int, cudaStream_t);
#define GPSIG_DECL_T(T) \
  GPSIG_DECL(T, 0, 4) GPSIG_DECL(T, 0, 8) GPSIG_DECL(T, 0, 16) \
  GPSIG_DECL(T, 1, 4) GPSIG_DECL(T, 1, 8) GPSIG_DECL(T, 1, 16) \
  GPSIG_DECL(T, 2, 4) GPSIG_DECL(T, 2, 8) GPSIG_DECL(T, 2, 16)
GPSIG_DECL_T(float)
GPSIG_DECL_T(double)

extern template int launch_lift<float>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);
extern template int launch_lift<double>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);
// tens_tens.cu (tens_tens_tile_kernels.cuh): Kzz and its adjoint on 16 x 16 tiles of tensor pairs; false = shape not covered
template <typename T> bool launch_tens_tens_tile(const TensTensParams&, cudaStream_t);
template <typename T> bool launch_tens_tens_tile_vjp(const TensTensVjpParams&, cudaStream_t);
extern template bool launch_tens_tens_tile<float>(const TensTensParams&, cudaStream_t);
extern template bool launch_tens_tens_tile<double>(const TensTensParams&, cudaStream_t);
extern template bool launch_tens_tens_tile_vjp<float>(const TensTensVjpParams&, cudaStream_t);
extern template bool launch_tens_tens_tile_vjp<double>(const TensTensVjpParams&, cudaStream_t);
extern template int launch_tens_seq_lifted<float>(const TensSeqParams&, int, void*, size_t, cudaStream_t);
extern template int launch_tens_seq_lifted<double>(const TensSeqParams&, int, void*, size_t, cudaStream_t);
extern template int launch_tens_seq_lifted_gd<float>(const TensSeqParams&, int, float*, const float*, cudaStream_t);
extern template int launch_tens_seq_lifted_gd<double>(const TensSeqParams&, int, double*, const double*, cudaStream_t);

// K(X, X) computed on and above the diagonal only: out[l, i, j] = out[l, j, i] for i > j (32 x 32 tiles through
// shared memory, both sides coalesced).  Every i > j entry is overwritten, so the result is exactly symmetric.
template <typename T>
__global__ void __launch_bounds__(256) mirror_lower_kernel(T* out, int N, int nplanes) {
  __shared__ T tile[32][33];
  const int tj = blockIdx.x, ti = blockIdx.y;      // destination tile (rows ti, columns tj), ti >= tj
  if (ti < tj) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int pl = blockIdx.z; pl < nplanes; pl += gridDim.z) {
    T* o = out + (size_t)pl * N * N;
    for (int r = ty; r < 32; r += 8) {             // source tile (rows tj, columns ti)
      const int sr = tj * 32 + r, sc = ti * 32 + tx;
      if (sr < N && sc < N) tile[r][tx] = o[(size_t)sr * N + sc];
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
      const int dr = ti * 32 + r, dc = tj * 32 + tx;
      if (dr < N && dc < N && dr > dc) o[(size_t)dr * N + dc] = tile[tx][r];
    }
    __syncthreads();
  }
}

template <typename T>
static int mirror_lower_impl(T* out, int N, int nplanes, void* stream) {
  if (!out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "out must not be NULL");
  if (N < 0 || nplanes < 0) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad sizes N=%d nplanes=%d", N, nplanes);
  if (N == 0 || nplanes == 0) return GPSIG_OK;
  const int nt = (N + 31) / 32;
  if (nt > 65535) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "matrix too large to mirror in one launch (N=%d)", N);
  mirror_lower_kernel<T><<<dim3(nt, nt, nplanes < 64 ? nplanes : 64), 256, 0, static_cast<cudaStream_t>(stream)>>>(out, N, nplanes);
  GPSIG_LAUNCH_CHECK("mirror_lower_kernel");
  return GPSIG_OK;
}

static int op_of_base(int base) {
  return base == GPSIG_BASE_LINEAR ? OP_LINEAR : (base == GPSIG_BASE_RBF ? OP_RBF : OP_GENERIC);
}

template <typename T>
static int dispatch_trunc(const TruncParams& P, int base, int DP, cudaStream_t s) {
  const int op = op_of_base(base);
#define GPSIG_GO(OP) (DP == 4 ? launch_trunc<T, OP, 4>(P, s) : DP == 8 ? launch_trunc<T, OP, 8>(P, s) : launch_trunc<T, OP, 16>(P, s))
  return op == OP_LINEAR ? GPSIG_GO(0) : (op == OP_RBF ? GPSIG_GO(1) : GPSIG_GO(2));
#undef GPSIG_GO
}
template <typename T>
static int dispatch_skew(const TruncParams& P, int base, int DP, cudaStream_t s) {
  const int op = op_of_base(base);
#define GPSIG_GO(OP) (DP == 4 ? launch_trunc_skew<T, OP, 4>(P, s) : DP == 8 ? launch_trunc_skew<T, OP, 8>(P, s) : launch_trunc_skew<T, OP, 16>(P, s))
  return op == OP_LINEAR ? GPSIG_GO(0) : (op == OP_RBF ? GPSIG_GO(1) : GPSIG_GO(2));
#undef GPSIG_GO
}
template <typename T>
static int dispatch_tens_seq(const TensSeqParams& P, int base, int DP, int m, cudaStream_t s) {
  const int op = op_of_base(base);
#define GPSIG_GO(OP) (DP == 4 ? launch_tens_vs_seq<T, OP, 4>(P, m, s) : DP == 8 ? launch_tens_vs_seq<T, OP, 8>(P, m, s) : launch_tens_vs_seq<T, OP, 16>(P, m, s))
  return op == OP_LINEAR ? GPSIG_GO(0) : (op == OP_RBF ? GPSIG_GO(1) : GPSIG_GO(2));
#undef GPSIG_GO
}

static int check_desc(const gpsig_trunc_desc* d) {
  if (!d) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "desc is NULL");
  if (d->base < GPSIG_BASE_LINEAR || d->base > GPSIG_BASE_SPECTRAL)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "unknown base kernel id %d", d->base);
  if (d->base == GPSIG_BASE_SPECTRAL) {
    if (!d->spectral || d->spectral_q < 1 || d->spectral_q > 64)
      GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "spectral base kernel needs its parameter block and 1 <= Q <= 64 (Q=%d)", d->spectral_q);
    if (d->spectral_family != GPSIG_SPECTRAL_RBF && d->spectral_family != GPSIG_SPECTRAL_EXP)
      GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "unknown spectral family %d", d->spectral_family);
  }
  if (d->num_levels < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "num_levels must be >= 1 (got %d)", d->num_levels);
  if (d->num_levels > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d > 8 is not supported", d->num_levels);
  if (d->order < 1 || d->order > d->num_levels)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "order must be in 1..num_levels (got %d, num_levels=%d)", d->order, d->num_levels);
  return GPSIG_OK;
}

template <typename T>
static int pack(const T* X, int N, int L, int d, const T* ls, T* dst, cudaStream_t s, int RS = 0, double scale = 1.0, int dup = 0) {
  const size_t tot = (size_t)N * L;
  if (tot == 0) return GPSIG_OK;
  if (RS == 0) RS = padded_dim(d);
  const int threads = 256;
  const size_t blocks = (tot + threads - 1) / threads;
  pack_sequences_kernel<T><<<(unsigned)blocks, threads, 0, s>>>(X, N, L, d, padded_dim(d), RS, ls, dst, seq_block_elems_rs(L, RS, sizeof(T)), T(scale), dup);
  GPSIG_LAUNCH_CHECK("pack_sequences_kernel");
  return GPSIG_OK;
}

static size_t packed_bytes(int N, int L, int d, int elem) {
  return align_up((size_t)N * seq_block_elems(L, d, elem) * elem, 128);
}
// with the padded row stride of the systolic Gram kernel (an upper bound of the plain layout)
static size_t packed_bytes_padded(int N, int L, int d, int elem) {
  return align_up((size_t)N * seq_block_elems_rs(L, skew_row_stride(d, elem, GPSIG_SKEW_DUP_F32 && elem == 4), elem) * elem, 128);
}

// float32 tensor-core lift (lift_mma_kernel.cuh): pre-split, pre-tiled operands behind the lifted-tile scratch
static size_t lift_mma_bytes(int N, int L, int N2, int L2, int d) {
  return align_up((size_t)N * lift_mma_a_elems(L, d) * 4, 128) + align_up((size_t)N2 * lift_mma_b_elems(L2, d) * 4, 128);
}
template <typename T>
static int lift_mma_prepare(const T*, int, int, const T*, int, int, int, char*, LiftMmaOperands*, cudaStream_t) { return GPSIG_OK; }
template <>
int lift_mma_prepare<float>(const float* xp, int N, int L, const float* yp, int N2, int L2, int d, char* dst,
                            LiftMmaOperands* op, cudaStream_t s) {
  const int KP = lift_mma_kp(d), DP = padded_dim(d), R2 = 32 * lift_cols(L2), MT = (L + 127) / 128;
  op->KP = KP; op->xop_blk = lift_mma_a_elems(L, d); op->yop_blk = lift_mma_b_elems(L2, d);
  float* a = reinterpret_cast<float*>(dst);
  float* b = reinterpret_cast<float*>(dst + align_up((size_t)N * op->xop_blk * 4, 128));
  const size_t ta = (size_t)N * MT * 128 * (KP / 4), tb = (size_t)N2 * R2 * (KP / 4);
  lift_pack_mma_kernel<<<(unsigned)((ta + 255) / 256), 256, 0, s>>>(xp, seq_block_elems(L, d, 4), DP, N, L, d, 128, MT, KP, 0, a, op->xop_blk);
  GPSIG_LAUNCH_CHECK("lift_pack_mma_kernel");
  lift_pack_mma_kernel<<<(unsigned)((tb + 255) / 256), 256, 0, s>>>(yp, seq_block_elems(L2, d, 4), DP, N2, L2, d, R2, 1, KP, 1, b, op->yop_blk);
  GPSIG_LAUNCH_CHECK("lift_pack_mma_kernel");
  op->xop = a; op->yop = b;
  return GPSIG_OK;
}

template <typename T>
static int trunc_gram_impl(const gpsig_trunc_desc* desc, const T* X, int N, int L, const T* X2, int N2, int L2, int d,
                           const T* ls, int row_begin, int row_end, const gpsig_epilogue* epi, T* out,
                           void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  const bool sym = X2 == nullptr;
  if (sym) { N2 = N; L2 = L; }
  if (N < 0 || N2 < 0 || d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad sizes N=%d N2=%d d=%d", N, N2, d);
  const int minlen = desc->difference ? 2 : 1;
  if (L < minlen || L2 < minlen) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d, L2=%d)", L, L2);
  if (row_begin < 0 || row_end > N || row_begin > row_end)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad row range [%d, %d) for N=%d", row_begin, row_end, N);
  if (epi && ((epi->norm_x == nullptr) != (epi->norm_x2 == nullptr)))
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "norm_x and norm_x2 must be given together");
  if (row_begin == row_end || N2 == 0) return GPSIG_OK;
  if (!X || !out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "X/out must not be NULL");
  const size_t need = gpsig_trunc_gram_workspace_bytes(N, L, sym ? 0 : N2, L2, d, sizeof(T));
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes (16-byte aligned), got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // order > 1: the general higher-order kernel; else the systolic kernel (padded rows) takes the
  // throughput shapes and trunc_gram_kernel the rest
  // d > 16: the lifted path (GEMM-shaped lift + recursion on the lifted tiles, lifted_kernels.cuh); sequences it does
  // not take (L2 > 256) run the general kernel, whose contraction is a runtime loop over d
  const bool big = d > 16;
  const bool lifted = big && desc->order == 1 && desc->base != GPSIG_BASE_SPECTRAL && lift_eligible(L, L2);
  const bool generic = desc->order > 1 || desc->base == GPSIG_BASE_SPECTRAL || (big && !lifted);
  bool skew = !big && !generic && desc->difference && skew_eligible(L, L2, sizeof(T)) && getenv("GPSIG_NO_SKEW") == nullptr;
  bool dup = skew && skew_uses_dup(d, L2, sizeof(T), desc->num_levels, desc->base);      // float32: coordinates stored twice for the packed-pair contraction
  if (skew) {
    // rows too long for the three-slot x ring of the systolic kernel (about L > 800 at d = 8 in float64): the general
    // kernel streams them instead (asked of the launcher itself, which knows the instantiation's shared-memory plan)
    TruncParams Q;
    memset(&Q, 0, sizeof(Q));
    const int rs0 = skew_row_stride(d, sizeof(T), dup);
    Q.xblk = seq_block_elems_rs(L, rs0, sizeof(T)); Q.yblk = seq_block_elems_rs(L2, rs0, sizeof(T));
    Q.xrs = Q.yrs = rs0; Q.N = N; Q.N2 = N2; Q.L = L; Q.L2 = L2; Q.row_begin = row_begin; Q.row_end = row_end;
    Q.m = desc->num_levels; Q.difference = 1; Q.base = desc->base; Q.dry_run = 1;
    if (dispatch_skew<T>(Q, desc->base, padded_dim(d), s) != GPSIG_OK) { skew = false; dup = false; }
  }
  const int RS = skew ? skew_row_stride(d, sizeof(T), dup) : padded_dim(d);
  T* xp = static_cast<T*>(ws);
  T* yp = sym ? xp : reinterpret_cast<T*>(static_cast<char*>(ws) + packed_bytes_padded(N, L, d, sizeof(T)));
  // float32 RBF on the systolic kernel: coordinates pre-scaled by sqrt(log2 e) so that the exponent
  // x.y - |x|^2/2 - |y|^2/2 comes out in base-2 units and exp() is one MUFU.EX2 (trunc_skew_kernel.cuh)
  // (float64: same pre-scaling, the exponent feeds exp2_tab32)
  // (the tensor-core lift of the large-d path reads its exponent the same way, lift_mma_kernel.cuh)
  const bool lift_mma = lifted && sizeof(T) == 4 && getenv("GPSIG_NO_MMA") == nullptr;
  const double cscale = (((skew && (sizeof(T) == 4 || GPSIG_SKEW_EXP2)) || lift_mma) && desc->base == GPSIG_BASE_RBF) ? 1.2011224087864498 : 1.0;
  if (int rc = pack<T>(X, N, L, d, ls, xp, s, RS, cscale, dup)) return rc;
  if (!sym) if (int rc = pack<T>(X2, N2, L2, d, ls, yp, s, RS, cscale, dup)) return rc;

  TruncParams P;
  memset(&P, 0, sizeof(P));
  P.xp = xp; P.yp = yp;
  P.xblk = seq_block_elems_rs(L, RS, sizeof(T)); P.yblk = seq_block_elems_rs(L2, RS, sizeof(T));
  P.xrs = P.yrs = RS;
  P.N = N; P.N2 = N2; P.L = L; P.L2 = L2;
  P.row_begin = row_begin; P.row_end = row_end;
  P.m = desc->num_levels; P.difference = desc->difference;
  P.symmetric = sym; P.diag = 0;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  if (epi) {
    P.norm_x = epi->norm_x; P.norm_x2 = epi->norm_x2; P.level_scale = epi->level_scale;
    P.sym_jitter = epi->sym_jitter; P.return_levels = epi->return_levels;
  } else P.return_levels = 1;
  P.out = out;
  const bool upper_only = sym && epi && epi->upper_only;
  const bool whole_sym = sym && !upper_only && row_begin == 0 && row_end == N && N >= 2;
  if (generic) {
    GenericParams G;
    memset(&G, 0, sizeof(G));
    G.xp = xp; G.yp = yp; G.xblk = P.xblk; G.yblk = P.yblk; G.N = N; G.N2 = N2; G.L = L; G.L2 = L2; G.DP = padded_dim(d);
    G.row_begin = row_begin; G.row_end = row_end; G.m = desc->num_levels; G.order = desc->order; G.difference = desc->difference;
    G.symmetric = sym; G.diag = 0; G.p0 = P.p0; G.p1 = P.p1; G.base = P.base;
    G.sp.p = desc->spectral; G.sp.q = desc->spectral_q; G.sp.family = desc->spectral_family; G.d = d;
    G.norm_x = P.norm_x; G.norm_x2 = P.norm_x2; G.level_scale = P.level_scale; G.sym_jitter = P.sym_jitter;
    G.return_levels = P.return_levels; G.out = out;
    if (int rc = launch_trunc_generic<T>(G, s)) return rc;
  } else if (lifted) {
    LiftParams Q;
    memset(&Q, 0, sizeof(Q));
    Q.xp = xp; Q.yp = yp; Q.xblk = P.xblk; Q.yblk = P.yblk; Q.DP = padded_dim(d);
    Q.N = N; Q.N2 = N2; Q.L = L; Q.L2 = L2; Q.diag = 0; Q.row_begin = row_begin; Q.row_end = row_end;
    Q.m = desc->num_levels; Q.difference = desc->difference; Q.symmetric = sym;
    Q.p0 = P.p0; Q.p1 = P.p1; Q.base = P.base;
    Q.norm_x = P.norm_x; Q.norm_x2 = P.norm_x2; Q.level_scale = P.level_scale; Q.sym_jitter = P.sym_jitter;
    Q.return_levels = P.return_levels; Q.out = out;
    const size_t off = packed_bytes_padded(N, L, d, sizeof(T)) + (sym ? 0 : packed_bytes_padded(N2, L2, d, sizeof(T)));
    const size_t sbytes = lift_plan((size_t)N * (size_t)N2, L, L2, sizeof(T)).scratch_bytes;
    LiftMmaOperands mma;
    memset(&mma, 0, sizeof(mma));
    if (lift_mma)
      if (int rc = lift_mma_prepare<T>(xp, N, L, yp, N2, L2, d, static_cast<char*>(ws) + off + sbytes, &mma, s)) return rc;
    if (int rc = launch_lift<T>(Q, static_cast<char*>(ws) + off, sbytes, mma.xop ? &mma : nullptr, s)) return rc;
  } else if (skew) {
    // the whole symmetric Gram in one call: compute only the CTAs that touch i <= j
    P.tri = upper_only || (whole_sym && N >= 64 && getenv("GPSIG_NO_TRI") == nullptr);
    if (int rc = dispatch_skew<T>(P, desc->base, padded_dim(d), s)) return rc;
  } else {
    if (int rc = dispatch_trunc<T>(P, desc->base, padded_dim(d), s)) return rc;
  }
  // K(X, X) over all rows is returned EXACTLY symmetric whichever kernel ran: the upper triangle is mirrored down
  if (whole_sym) return mirror_lower_impl<T>(out, N, P.return_levels ? desc->num_levels + 1 : 1, stream);
  return GPSIG_OK;
}

template <typename T>
static int trunc_diag_impl(const gpsig_trunc_desc* desc, const T* X, int N, int L, int d, const T* ls, T* out,
                           void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (!X || !out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "X/out must not be NULL");
  if (d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad state dimension d=%d", d);
  if (L < (desc->difference ? 2 : 1)) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d)", L);
  if (N == 0) return GPSIG_OK;
  const size_t need = gpsig_trunc_diag_workspace_bytes(N, L, d, sizeof(T));
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  T* xp = static_cast<T*>(ws);
  const bool lifted = d > 16 && desc->order == 1 && desc->base != GPSIG_BASE_SPECTRAL && lift_eligible(L, L);
  const bool lift_mma = lifted && sizeof(T) == 4 && getenv("GPSIG_NO_MMA") == nullptr;
  // tensor-core lift + RBF: coordinates pre-scaled by sqrt(log2 e), the epilogue takes exp2 (lift_mma_kernel.cuh)
  if (int rc = pack<T>(X, N, L, d, ls, xp, s, 0, (lift_mma && desc->base == GPSIG_BASE_RBF) ? 1.2011224087864498 : 1.0)) return rc;
  TruncParams P;
  memset(&P, 0, sizeof(P));
  P.xp = xp; P.yp = xp;
  P.xblk = P.yblk = seq_block_elems(L, d, sizeof(T));
  P.xrs = P.yrs = padded_dim(d);
  P.N = N; P.N2 = N; P.L = L; P.L2 = L;
  P.row_begin = 0; P.row_end = N;
  P.m = desc->num_levels; P.difference = desc->difference;
  P.symmetric = 1; P.diag = 1; P.return_levels = 1;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  P.out = out;
  if (lifted) {
    LiftParams Q;
    memset(&Q, 0, sizeof(Q));
    Q.xp = xp; Q.yp = xp; Q.xblk = Q.yblk = P.xblk; Q.DP = padded_dim(d);
    Q.N = N; Q.N2 = N; Q.L = L; Q.L2 = L; Q.diag = 1; Q.row_begin = 0; Q.row_end = N;
    Q.m = desc->num_levels; Q.difference = desc->difference; Q.symmetric = 1;
    Q.p0 = P.p0; Q.p1 = P.p1; Q.base = P.base; Q.return_levels = 1; Q.out = out;
    const size_t off = packed_bytes(N, L, d, sizeof(T));
    const size_t sbytes = lift_plan((size_t)N, L, L, sizeof(T)).scratch_bytes;
    LiftMmaOperands mma;
    memset(&mma, 0, sizeof(mma));
    if (lift_mma)
      if (int rc = lift_mma_prepare<T>(xp, N, L, xp, N, L, d, static_cast<char*>(ws) + off + sbytes, &mma, s)) return rc;
    return launch_lift<T>(Q, static_cast<char*>(ws) + off, sbytes, mma.xop ? &mma : nullptr, s);
  }
  if (desc->order > 1 || desc->base == GPSIG_BASE_SPECTRAL || d > 16) {
    GenericParams G;
    memset(&G, 0, sizeof(G));
    G.xp = xp; G.yp = xp; G.xblk = G.yblk = P.xblk; G.N = N; G.N2 = N; G.L = L; G.L2 = L; G.DP = padded_dim(d);
    G.row_begin = 0; G.row_end = N; G.m = desc->num_levels; G.order = desc->order; G.difference = desc->difference;
    G.symmetric = 1; G.diag = 1; G.p0 = P.p0; G.p1 = P.p1; G.base = P.base; G.return_levels = 1; G.out = out;
    G.sp.p = desc->spectral; G.sp.q = desc->spectral_q; G.sp.family = desc->spectral_family; G.d = d;
    return launch_trunc_generic<T>(G, s);
  }
  return dispatch_trunc<T>(P, desc->base, padded_dim(d), s);
}

template <typename T>
static int tens_vs_seq_impl(const gpsig_trunc_desc* desc, const T* Z, int Mz, int increments, const T* X, int N, int L,
                            int d, const T* ls, int col_begin, int col_end, const gpsig_epilogue* epi, T* out,
                            void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (!Z || !X || !out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "Z/X/out must not be NULL");
  if (d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad state dimension d=%d", d);
  if (L < (desc->difference ? 2 : 1)) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d)", L);
  if (col_begin < 0 || col_end > N || col_begin > col_end)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad column range [%d, %d) for N=%d", col_begin, col_end, N);
  if (col_begin == col_end || Mz == 0) return GPSIG_OK;
  const int m = desc->num_levels, TT = m * (m + 1) / 2, zs = increments ? 2 : 1;
  const size_t need = gpsig_tens_vs_seq_workspace_bytes(Mz, increments, N, L, d, m, sizeof(T));
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  T* zp = static_cast<T*>(ws);
  T* xp = reinterpret_cast<T*>(static_cast<char*>(ws) + packed_bytes(TT * Mz, zs, d, sizeof(T)));
  // first-order RBF: coordinates pre-scaled by sqrt(log2 e), the exponent feeds exp2 (tens_kernels.cuh)
  // d > 16 runs the general kernel (runtime loop over d), natural-base exponent
  const double cscale = (d <= 16 && !tens_lift_dims(d, sizeof(T)) && desc->order == 1 && desc->base == GPSIG_BASE_RBF) ? 1.2011224087864498 : 1.0;
  if (int rc = pack<T>(Z, TT * Mz, zs, d, ls, zp, s, 0, cscale)) return rc;
  if (int rc = pack<T>(X, N, L, d, ls, xp, s, 0, cscale)) return rc;
  TensSeqParams P;
  memset(&P, 0, sizeof(P));
  P.zp = zp; P.xp = xp;
  P.zblk = seq_block_elems(zs, d, sizeof(T)); P.xblk = seq_block_elems(L, d, sizeof(T));
  P.Mz = Mz; P.zs = zs; P.N = N; P.L = L; P.m = m; P.col_begin = col_begin; P.col_end = col_end;
  P.difference = desc->difference;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  if (epi) { P.norm_x = epi->norm_x2; P.level_scale = epi->level_scale; P.return_levels = epi->return_levels; }
  else P.return_levels = 1;
  P.out = out;
  P.sp.p = desc->spectral; P.sp.q = desc->spectral_q; P.sp.family = desc->spectral_family; P.d = d;
  // d > 16, first order: the lifted form (GEMM of all tensor points against blocks of sequences + time recursion,
  // tens_lifted_kernels.cuh); higher order / spectral: the general kernel
  // (also float64 at d = 9..16, where the tensor's coordinates no longer fit the register tile of tens_vs_seq_kernel2)
  if (tens_lift_dims(d, sizeof(T)) && desc->order == 1 && desc->base != GPSIG_BASE_SPECTRAL && getenv("GPSIG_TENS_NO_LIFT") == nullptr) {
    const size_t off = packed_bytes(TT * Mz, zs, d, sizeof(T)) + packed_bytes(N, L, d, sizeof(T));
    return launch_tens_seq_lifted<T>(P, padded_dim(d), static_cast<char*>(ws) + off, ws_bytes - off, s);
  }
  if (desc->order > 1 || desc->base == GPSIG_BASE_SPECTRAL || d > 16) return launch_tens_seq_generic<T>(P, desc->order, padded_dim(d), s);
  return dispatch_tens_seq<T>(P, desc->base, padded_dim(d), m, s);
}

template <typename T>
static int tens_tens_impl(const gpsig_trunc_desc* desc, const T* Z, int Mz, int increments, int d, const T* ls,
                          const gpsig_epilogue* epi, T* out, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (!Z || !out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "Z/out must not be NULL");
  if (Mz == 0) return GPSIG_OK;
  TensTensParams P;
  memset(&P, 0, sizeof(P));
  P.Z = Z; P.ls = ls; P.Mz = Mz; P.zs = increments ? 2 : 1; P.m = desc->num_levels; P.d = d;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  P.sp.p = desc->spectral; P.sp.q = desc->spectral_q; P.sp.family = desc->spectral_family;
  if (d > 16 && desc->base == GPSIG_BASE_SPECTRAL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "spectral Kzz supports d <= 16 (got %d)", d);
  if (epi) { P.level_scale = epi->level_scale; P.return_levels = epi->return_levels; } else P.return_levels = 1;
  P.out = out;
  if (!getenv("GPSIG_TENS_TENS_OLD") && launch_tens_tens_tile<T>(P, static_cast<cudaStream_t>(stream))) {
    GPSIG_LAUNCH_CHECK("tens_tens_tile_kernel");
    return GPSIG_OK;
  }
  dim3 grid((Mz + 127) / 128, Mz);
  if (P.d > 16) tens_tens_big_kernel<T><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  else if (P.d <= 8) tens_tens_kernel<T, 8><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  else tens_tens_kernel<T, 16><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  GPSIG_LAUNCH_CHECK("tens_tens_kernel");
  return GPSIG_OK;
}

static int pde_impl(const double* X, int N, int L, const double* Y, int N2, int L2, int d, const double* ls,
                    int order, int pair_mode, int row_begin, int row_end, double sigma, double* out,
                    void* ws, size_t ws_bytes, void* stream) {
  if (!X || !out) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "X/out must not be NULL");
  const bool same = Y == nullptr;
  if (same) { N2 = N; L2 = L; }
  if (d < 1 || d > 16) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "state dimension d=%d is not supported (1..16)", d);
  if (L < 2 || L2 < 2) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d, L2=%d)", L, L2);
  if (order < 0 || order > 6) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "dyadic order %d is not supported (0..6)", order);
  if (pair_mode != GPSIG_PDE_DIAG && pair_mode != GPSIG_PDE_GRAM) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad pair_mode %d", pair_mode);
  if (pair_mode == GPSIG_PDE_DIAG && N2 != N) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "diagonal pairing needs N == N2 (%d vs %d)", N, N2);
  if (row_begin < 0 || row_end > N || row_begin > row_end)
    GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad row range [%d, %d) for N=%d", row_begin, row_end, N);
  if (row_begin == row_end || N2 == 0) return GPSIG_OK;
  const size_t need = gpsig_pde_workspace_bytes(N, L, same ? 0 : N2, L2, d, order, pair_mode);
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  double* xp = static_cast<double*>(ws);
  double* yp = same ? xp : reinterpret_cast<double*>(static_cast<char*>(ws) + packed_bytes(N, L, d, 8));
  if (int rc = pack<double>(X, N, L, d, ls, xp, s)) return rc;
  if (!same) if (int rc = pack<double>(Y, N2, L2, d, ls, yp, s)) return rc;
  PdeParams P;
  memset(&P, 0, sizeof(P));
  P.xp = xp; P.yp = yp; P.xblk = seq_block_elems(L, d, 8); P.yblk = seq_block_elems(L2, d, 8);
  P.N = N; P.N2 = N2; P.L = L; P.L2 = L2; P.DP = padded_dim(d); P.order = order; P.pair_mode = pair_mode;
  P.row_begin = row_begin; P.row_end = row_end; P.sigma = sigma; P.out = out;
  const size_t npairs = pair_mode == GPSIG_PDE_DIAG ? (size_t)(row_end - row_begin) : (size_t)(row_end - row_begin) * N2;
  if (npairs > 0x7fffffffull) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "too many pairs in one call (%zu)", npairs);
  // one warp per 32*kPdeW-column strip, all strips of a pair in flight, persistent CTAs taking pairs from a
  // ticket counter (pde_block_kernel); pairs with more than kPdeMaxStrips strips run the strips one after
  // the other on one warp (pde_wavefront_kernel)
  const int Qf = (L2 - 1) << order;
  const int nstrips = (Qf + 32 * kPdeW - 1) / (32 * kPdeW);
  const bool pipe = nstrips <= kPdeMaxStrips && getenv("GPSIG_PDE_SEQ") == nullptr;
  // dyadic order 0 has one fine row per coarse row: the block-row kernel degenerates to the one-row-per-step
  // kernel plus ticket overhead, so order 0 keeps pde_pipe_kernel (measured: 0.41 ms against 0.53 ms at N=1000, L=500)
  const bool rowstep = order == 0 || getenv("GPSIG_PDE_ROWSTEP") != nullptr;
  // persistent mode needs the dy coefficients in registers (they are per pair) and room for two dx buffers
  // diagonal pairs k(x, x): the grid is symmetric and pde_block_kernel<.., SYM> lets the strips solve from their diagonal
  // block down (36 of 64 blocks at c3).  OPT-IN (GPSIG_PDE_SYM=1), because it is SLOWER as measured (c3: 1.89 ms against
  // 1.49 ms): with warp = strip the CTA's period per pair is set by strip 0, which still sweeps all rows, and the
  // strips that finish early only leave their warps idle
  bool sym = pipe && !rowstep && same && pair_mode == GPSIG_PDE_DIAG && nstrips >= 2 && getenv("GPSIG_PDE_SYM") != nullptr;
  // persistent mode needs the dy coefficients in registers (they are per pair) and room for two dx buffers
  bool persist = pipe && !rowstep && pde_dy_in_regs(P.DP, order) && getenv("GPSIG_PDE_NOPERSIST") == nullptr &&
                 pde_block_smem_bytes(L, L2, P.DP, order, nstrips, true, sym) <= kMaxSmem;
  if (sym && pde_block_smem_bytes(L, L2, P.DP, order, nstrips, persist, true) > kMaxSmem) sym = false;
  const size_t smem = !pipe ? pde_smem_bytes(L, L2, P.DP, order)
                            : rowstep ? pde_pipe_smem_bytes(L, L2, P.DP, order, nstrips)
                                      : pde_block_smem_bytes(L, L2, P.DP, order, nstrips, persist, sym);
  if (smem > kMaxSmem)
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "PDE grid too large for the on-chip plan: %zu bytes of shared memory (L=%d, L2=%d, order=%d)", smem, L, L2, order);
  unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(ws) + need - 128);
  if (persist) GPSIG_CUDA_CHECK(cudaMemsetAsync(ticket, 0, sizeof(unsigned), s));
  int sms = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
#define GPSIG_PDE_BLOCK(DP_, O_, SYM_) do { auto kern = pde_block_kernel<DP_, O_, SYM_>; \
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    unsigned grid = (unsigned)npairs; \
    if (persist) { int occ = 1; GPSIG_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * nstrips, smem)); \
      const size_t slots = (size_t)(occ > 0 ? occ : 1) * sms; if (npairs > slots) grid = (unsigned)slots; else persist = false; } \
    prof_begin(s); kern<<<grid, 32 * nstrips, smem, s>>>(P, persist ? 1 : 0, ticket); prof_end(s); } while (0)
#define GPSIG_PDE_GO(DP_, O_) do { if (pipe && !rowstep) { if (sym) GPSIG_PDE_BLOCK(DP_, O_, true); else GPSIG_PDE_BLOCK(DP_, O_, false); } else if (pipe) { auto kern = pde_pipe_kernel<DP_, O_>; \
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    prof_begin(s); kern<<<(unsigned)npairs, 32 * nstrips, smem, s>>>(P); prof_end(s); } else { auto kern = pde_wavefront_kernel<DP_, O_>; \
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    prof_begin(s); kern<<<(unsigned)npairs, 32, smem, s>>>(P); prof_end(s); } } while (0)
#define GPSIG_PDE_DP(O_) do { if (P.DP == 4) GPSIG_PDE_GO(4, O_); else if (P.DP == 8) GPSIG_PDE_GO(8, O_); else GPSIG_PDE_GO(16, O_); } while (0)
  switch (order) {
    case 0: GPSIG_PDE_DP(0); break; case 1: GPSIG_PDE_DP(1); break; case 2: GPSIG_PDE_DP(2); break;
    case 3: GPSIG_PDE_DP(3); break; case 4: GPSIG_PDE_DP(4); break; case 5: GPSIG_PDE_DP(5); break;
    default: GPSIG_PDE_DP(6); break;
  }
  GPSIG_LAUNCH_CHECK("pde_wavefront_kernel");
  return GPSIG_OK;
}


// ---- reverse mode (vjp_kernels.cuh) ---------------------------------------------------------------
extern "C" size_t gpsig_tens_vs_seq_vjp_workspace_bytes(int Mz, int ncols, int L, int m, int elem);
static size_t trunc_vjp_ws(size_t npairs, int L, int L2, int m, int difference, int elem, int order = 1) {
  const size_t PR = L - (difference ? 1 : 0), QC = L2 - (difference ? 1 : 0);
  const size_t mm = m > 1 ? m - 1 : 0;
  // order > 1 (generic_vjp_kernels.cuh): D, the factor tiles of every level, W
  if (order > 1) return align_up(npairs * (size_t)(generic_vjp_tiles(m, order) + 2) * PR * QC * elem, 128);
  return align_up(npairs * (mm * PR * QC + mm * QC + PR * QC) * elem, 128);
}

template <typename T>
static int trunc_vjp_impl(const gpsig_trunc_desc* desc, const T* X, int N, int L, const T* X2, int N2, int L2, int d,
                          int pair_mode, int row_begin, int row_end, const T* grad_levels, T* gradX, T* gradX2,
                          void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (desc->base == GPSIG_BASE_SPECTRAL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode is not implemented for the spectral base kernel");
  if (desc->num_levels > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode supports num_levels <= 8");
  const bool sym = X2 == nullptr;
  if (sym) { X2 = X; N2 = N; L2 = L; gradX2 = gradX; }
  if (!X || !grad_levels || !gradX || !gradX2) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "X/grad_levels/gradX must not be NULL");
  if (d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad state dimension d=%d", d);
  const int minlen = desc->difference ? 2 : 1;
  if (L < minlen || L2 < minlen) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d, L2=%d)", L, L2);
  if (row_begin < 0 || row_end > N || row_begin > row_end) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad row range [%d, %d) for N=%d", row_begin, row_end, N);
  const bool diag = pair_mode == GPSIG_PDE_DIAG;
  if (diag && (N2 != N || L2 != L)) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "diagonal pairing needs matching X and X2");
  const size_t npairs = diag ? (size_t)(row_end - row_begin) : (size_t)(row_end - row_begin) * N2;
  if (npairs == 0) return GPSIG_OK;
  const size_t need = trunc_vjp_ws(npairs, L, L2, desc->num_levels, desc->difference, sizeof(T), desc->order);
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TruncVjpParams P;
  memset(&P, 0, sizeof(P));
  P.X = X; P.Y = X2; P.N = N; P.N2 = N2; P.L = L; P.L2 = L2; P.d = d; P.m = desc->num_levels;
  P.diag = diag; P.row_begin = row_begin; P.row_end = row_end; P.difference = desc->difference;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  P.g = grad_levels; P.gradX = gradX; P.gradY = gradX2; P.npairs = npairs;
  const size_t PR = L - (desc->difference ? 1 : 0), QC = L2 - (desc->difference ? 1 : 0), mm = desc->num_levels - 1;
  T* w = static_cast<T*>(ws);
  if (desc->order > 1) {
    // higher-order recursion: one CTA per pair stores D and the factor tiles, sweeps back and leaves W = dL/dD
    // (cell-major); the chain into the coordinates is the first-order path's gradient kernel
    if (QC > 1024) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "order > 1 supports at most 1024 tile columns (L2=%d)", L2);
    if (npairs > 0x7fffffffull) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "too many pairs in one call (%zu)", npairs);
    GenericVjpParams G;
    memset(&G, 0, sizeof(G));
    G.order = desc->order;
    G.Dt = w; w += npairs * PR * QC;
    G.Ft = w; w += npairs * (size_t)generic_vjp_tiles(desc->num_levels, desc->order) * PR * QC;
    P.W = w; P.w_cell_major = 1;
    G.V = P;
    prof_begin(s);
    trunc_generic_vjp_kernel<T><<<(unsigned)npairs, (unsigned)((QC + 31) / 32 * 32), 0, s>>>(G);
    prof_end(s);
    GPSIG_LAUNCH_CHECK("trunc_generic_vjp_kernel");
    const size_t tot = npairs * (size_t)(L + L2);
    if (getenv("GPSIG_VJP_NO_GRAD_CTA") == nullptr && launch_trunc_vjp_grad_cta<T>(P, s)) {}
    else trunc_vjp_grad_warp_kernel<T><<<(unsigned)((tot + 7) / 8), 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("trunc_vjp_grad_kernel");
    return GPSIG_OK;
  }
  P.F = w; w += npairs * mm * PR * QC;
  P.C = w; w += npairs * mm * QC;
  P.W = w;
  const int threads = 128;
  prof_begin(s);
  // few pairs (training shapes): one warp per pair; many pairs: one thread per pair fills the machine
  // sequences short enough for shared memory: one CTA per pair (any number of pairs)
  P.w_cell_major = 1;
  const bool cta_per_pair = npairs < (1u << 30) && getenv("GPSIG_VJP_THREAD") == nullptr && getenv("GPSIG_VJP_NO_CTA") == nullptr &&
                            launch_trunc_vjp_tables_cta<T>(P, s);
  if (!cta_per_pair) P.w_cell_major = 0;
  const bool warp_per_pair = !cta_per_pair && npairs < 32768 && getenv("GPSIG_VJP_THREAD") == nullptr && launch_trunc_vjp_tables_warp<T>(P, s);
  if (!cta_per_pair && !warp_per_pair) trunc_vjp_tables_kernel<T><<<(unsigned)((npairs + threads - 1) / threads), threads, 0, s>>>(P);
  prof_end(s);
  GPSIG_LAUNCH_CHECK("trunc_vjp_tables_kernel");
  const size_t tot = npairs * (size_t)(L + L2);
  if (cta_per_pair && getenv("GPSIG_VJP_NO_GRAD_CTA") == nullptr && launch_trunc_vjp_grad_cta<T>(P, s)) {}
  else if (cta_per_pair || (npairs < 32768 && getenv("GPSIG_VJP_THREAD") == nullptr)) trunc_vjp_grad_warp_kernel<T><<<(unsigned)((tot + 7) / 8), 256, 0, s>>>(P);
  else trunc_vjp_grad_kernel<T><<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(P);
  GPSIG_LAUNCH_CHECK("trunc_vjp_grad_kernel");
  return GPSIG_OK;
}

template <typename T>
static int tens_seq_vjp_impl(const gpsig_trunc_desc* desc, const T* Z, int Mz, int increments, const T* X, int N, int L, int d,
                             int col_begin, int col_end, const T* grad_levels, T* gradZ, T* gradX,
                             void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (desc->base == GPSIG_BASE_SPECTRAL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode is not implemented for the spectral base kernel");
  if (desc->num_levels > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode supports num_levels <= 8");
  if (!Z || !X || !grad_levels || !gradZ || !gradX) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "Z/X/grad_levels/gradZ/gradX must not be NULL");
  if (d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad state dimension d=%d", d);
  if (L < (desc->difference ? 2 : 1)) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d)", L);
  if (col_begin < 0 || col_end > N || col_begin > col_end) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad column range [%d, %d) for N=%d", col_begin, col_end, N);
  const size_t npairs = (size_t)Mz * (col_end - col_begin);
  if (npairs == 0) return GPSIG_OK;
  const int m = desc->num_levels, TT = m * (m + 1) / 2, zs = increments ? 2 : 1;
  const size_t need = gpsig_tens_vs_seq_vjp_workspace_bytes(Mz, col_end - col_begin, L, m, sizeof(T));
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  TensSeqVjpParams P;
  memset(&P, 0, sizeof(P));
  P.Z = Z; P.X = X; P.Mz = Mz; P.zs = zs; P.N = N; P.L = L; P.d = d; P.m = m;
  P.col_begin = col_begin; P.col_end = col_end; P.difference = desc->difference;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  P.g = grad_levels; P.gradZ = gradZ; P.gradX = gradX; P.S = ws; P.npairs = npairs; P.order = desc->order;
  const int threads = 64;
  const unsigned blocks = (unsigned)((npairs + threads - 1) / threads);
  prof_begin(s);
  const bool warp_per_seq = desc->order == 1 && getenv("GPSIG_VJP_THREAD") == nullptr && launch_tens_seq_vjp_warp<T>(P, s);
  if (warp_per_seq) {}
  else {
    // d > 16: the reverse sweep runs once per chunk of coordinates (per-thread accumulators of dL/dZ: T * zs * chunk <= 1152)
    int chunk = d;
    if (d > 16 || TT * zs * d > 1152) { chunk = 1152 / (TT * zs); if (chunk > kTsvChunk) chunk = kTsvChunk; if (chunk > d) chunk = d; }
    if (chunk < 1) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode: too many tensor components for d=%d (num_levels=%d)", d, m);
    for (int k0 = 0; k0 < d; k0 += chunk) {
      P.kc0 = k0; P.kcn = d - k0 < chunk ? d - k0 : chunk;
      const int gz = TT * zs * P.kcn;
      if (gz <= 160) tens_seq_vjp_kernel<T, 160><<<blocks, threads, 0, s>>>(P);
      else if (gz <= 480) tens_seq_vjp_kernel<T, 480><<<blocks, threads, 0, s>>>(P);
      else tens_seq_vjp_kernel<T, 1152><<<blocks, threads, 0, s>>>(P);
    }
  }
  prof_end(s);
  GPSIG_LAUNCH_CHECK("tens_seq_vjp_kernel");
  return GPSIG_OK;
}

// Reverse mode of Kzx on the lifted form (tens_lifted_kernels.cuh): Gd[r, e] = dL/dkappa (x kappa for RBF) for ALL tensor points
// r = (c zs + s) Mz + z against the points e = l C + n of the columns [col_begin, col_end), C = col_end - col_begin.
template <typename T>
static int tens_seq_lifted_gd_impl(const gpsig_trunc_desc* desc, const T* Z, int Mz, int increments, const T* X, int N, int L, int d,
                                   int col_begin, int col_end, const T* grad_levels, T* Gd, void* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (desc->order != 1 || (desc->base != GPSIG_BASE_RBF && desc->base != GPSIG_BASE_LINEAR))
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "the lifted reverse mode takes the first-order linear and RBF kernels");
  if (!Z || !X || !grad_levels || !Gd) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "Z/X/grad_levels/Gd must not be NULL");
  if (d < 9) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "the lifted reverse mode is for state dimensions above 8 (d=%d)", d);
  if (L < (desc->difference ? 2 : 1)) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "sequences too short (L=%d)", L);
  if (col_begin < 0 || col_end > N || col_begin > col_end) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad column range [%d, %d) for N=%d", col_begin, col_end, N);
  if (col_begin == col_end || Mz == 0) return GPSIG_OK;
  const int m = desc->num_levels, TT = m * (m + 1) / 2, zs = increments ? 2 : 1;
  if (m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode supports num_levels <= 8");
  const size_t need = packed_bytes(TT * Mz, zs, d, sizeof(T)) + packed_bytes(N, L, d, sizeof(T));
  if (!ws || ws_bytes < need || (reinterpret_cast<uintptr_t>(ws) & 15))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  T* zp = static_cast<T*>(ws);
  T* xp = reinterpret_cast<T*>(static_cast<char*>(ws) + packed_bytes(TT * Mz, zs, d, sizeof(T)));
  if (int rc = pack<T>(Z, TT * Mz, zs, d, nullptr, zp, s, 0, 1.0)) return rc;
  if (int rc = pack<T>(X, N, L, d, nullptr, xp, s, 0, 1.0)) return rc;
  TensSeqParams P;
  memset(&P, 0, sizeof(P));
  P.zp = zp; P.xp = xp;
  P.zblk = seq_block_elems(zs, d, sizeof(T)); P.xblk = seq_block_elems(L, d, sizeof(T));
  P.Mz = Mz; P.zs = zs; P.N = N; P.L = L; P.m = m; P.col_begin = col_begin; P.col_end = col_end;
  P.difference = desc->difference;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base; P.d = d; P.return_levels = 1;
  return launch_tens_seq_lifted_gd<T>(P, padded_dim(d), Gd, grad_levels, s);
}

template <typename T>
static int tens_tens_vjp_impl(const gpsig_trunc_desc* desc, const T* Z, int Mz, int increments, int d, const T* grad_levels,
                              T* gradZ, void* stream) {
  if (int rc = check_desc(desc)) return rc;
  if (desc->base == GPSIG_BASE_SPECTRAL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "reverse mode is not implemented for the spectral base kernel");
  if (!Z || !grad_levels || !gradZ) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "Z/grad_levels/gradZ must not be NULL");
  if (d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad state dimension d=%d", d);
  if (Mz == 0) return GPSIG_OK;
  TensTensVjpParams P;
  memset(&P, 0, sizeof(P));
  P.Z = Z; P.Mz = Mz; P.zs = increments ? 2 : 1; P.m = desc->num_levels; P.d = d;
  P.p0 = desc->base_params[0]; P.p1 = desc->base_params[1]; P.base = desc->base;
  P.g = grad_levels; P.gradZ = gradZ;
  const int TT = P.m * (P.m + 1) / 2;
  if (!getenv("GPSIG_TENS_TENS_OLD") && launch_tens_tens_tile_vjp<T>(P, static_cast<cudaStream_t>(stream))) {
    GPSIG_LAUNCH_CHECK("tens_tens_tile_vjp_kernel");
    return GPSIG_OK;
  }
  dim3 grid((Mz + 3) / 4, TT);                                       // 4 warps per CTA, one warp per (component, row)
  if (d > 16) tens_tens_vjp_big_kernel<T><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  else if (d <= 8) tens_tens_vjp_kernel<T, 8><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  else tens_tens_vjp_kernel<T, 16><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(P);
  GPSIG_LAUNCH_CHECK("tens_tens_vjp_kernel");
  return GPSIG_OK;
}

}  // namespace gpsig

using namespace gpsig;

extern "C" {

const char* gpsig_last_error_string(void) { return g_err; }
int gpsig_version(void) { return GPSIG_VERSION; }

unsigned long long gpsig_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }
void gpsig_profile_enable(int on) { g_prof_on = on; g_prof_valid = 0; }
double gpsig_profile_last_kernel_ms(void) {
  if (!g_prof_valid) return -1.0;
  if (cudaEventSynchronize(g_prof_ev[1]) != cudaSuccess) return -1.0;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, g_prof_ev[0], g_prof_ev[1]) != cudaSuccess) return -1.0;
  return (double)ms;
}
int gpsig_fma_peak(int elem_size, int blocks, int iters, void* scratch, void* stream) {
  if (!scratch || blocks < 1 || iters < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad arguments to gpsig_fma_peak");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (elem_size == 8) fma_peak_kernel<double><<<blocks, 256, 0, s>>>(static_cast<double*>(scratch), iters, 0.999999, 1e-7);
  else if (elem_size == 4) fma_peak_kernel<float><<<blocks, 256, 0, s>>>(static_cast<float*>(scratch), iters, 0.999999f, 1e-7f);
  else GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "elem_size must be 4 or 8");
  GPSIG_LAUNCH_CHECK("fma_peak_kernel");
  return GPSIG_OK;
}

size_t gpsig_trunc_gram_workspace_bytes(int N, int L, int N2, int L2, int d, int elem) {
  size_t b = packed_bytes_padded(N, L, d, elem) + (N2 > 0 ? packed_bytes_padded(N2, L2, d, elem) : 0);
  if (d > 16 && lift_eligible(L, N2 > 0 ? L2 : L))       // the lifted tiles in flight (lifted_kernels.cuh)
    b += lift_plan((size_t)N * (size_t)(N2 > 0 ? N2 : N), L, N2 > 0 ? L2 : L, elem).scratch_bytes +
         (elem == 4 ? lift_mma_bytes(N, L, N2 > 0 ? N2 : N, N2 > 0 ? L2 : L, d) : 0);
  return b;
}
size_t gpsig_trunc_diag_workspace_bytes(int N, int L, int d, int elem) {
  size_t b = packed_bytes(N, L, d, elem);
  if (d > 16 && lift_eligible(L, L)) b += lift_plan((size_t)N, L, L, elem).scratch_bytes + (elem == 4 ? lift_mma_bytes(N, L, N, L, d) : 0);
  return b;
}
size_t gpsig_tens_vs_seq_workspace_bytes(int Mz, int increments, int N, int L, int d, int m, int elem) {
  size_t b = packed_bytes(m * (m + 1) / 2 * Mz, increments ? 2 : 1, d, elem) + packed_bytes(N, L, d, elem);
  if (gpsig::tens_lift_dims(d, elem)) b += gpsig::tens_lift_bytes(Mz, increments ? 2 : 1, m, L, N, elem);      // lifted Kzx values of a block of sequences
  return b;
}

int gpsig_trunc_gram_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, const double* X2, int N2, int L2,
                         int d, const double* ls, int rb, int re, const gpsig_epilogue* epi, double* out,
                         void* ws, size_t wsb, void* stream) {
  return trunc_gram_impl<double>(desc, X, N, L, X2, N2, L2, d, ls, rb, re, epi, out, ws, wsb, stream);
}
int gpsig_trunc_gram_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L, const float* X2, int N2, int L2,
                         int d, const float* ls, int rb, int re, const gpsig_epilogue* epi, float* out,
                         void* ws, size_t wsb, void* stream) {
  return trunc_gram_impl<float>(desc, X, N, L, X2, N2, L2, d, ls, rb, re, epi, out, ws, wsb, stream);
}
int gpsig_trunc_diag_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, int d, const double* ls,
                         double* out, void* ws, size_t wsb, void* stream) {
  return trunc_diag_impl<double>(desc, X, N, L, d, ls, out, ws, wsb, stream);
}
int gpsig_trunc_diag_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L, int d, const float* ls,
                         float* out, void* ws, size_t wsb, void* stream) {
  return trunc_diag_impl<float>(desc, X, N, L, d, ls, out, ws, wsb, stream);
}
int gpsig_tens_vs_seq_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int inc, const double* X, int N, int L,
                          int d, const double* ls, int cb, int ce, const gpsig_epilogue* epi, double* out,
                          void* ws, size_t wsb, void* stream) {
  return tens_vs_seq_impl<double>(desc, Z, Mz, inc, X, N, L, d, ls, cb, ce, epi, out, ws, wsb, stream);
}
int gpsig_tens_vs_seq_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int inc, const float* X, int N, int L,
                          int d, const float* ls, int cb, int ce, const gpsig_epilogue* epi, float* out,
                          void* ws, size_t wsb, void* stream) {
  return tens_vs_seq_impl<float>(desc, Z, Mz, inc, X, N, L, d, ls, cb, ce, epi, out, ws, wsb, stream);
}

size_t gpsig_trunc_vjp_workspace_bytes(int npairs, int L, int L2, int m, int difference, int elem) {
  return trunc_vjp_ws((size_t)npairs, L, L2, m, difference, elem);
}
size_t gpsig_trunc_vjp_workspace_bytes_order(int npairs, int L, int L2, int m, int order, int difference, int elem) {
  return trunc_vjp_ws((size_t)npairs, L, L2, m, difference, elem, order);
}
size_t gpsig_tens_vs_seq_lifted_gd_workspace_bytes(int Mz, int increments, int N, int L, int d, int m, int elem) {
  return packed_bytes(m * (m + 1) / 2 * Mz, increments ? 2 : 1, d, elem) + packed_bytes(N, L, d, elem);
}
int gpsig_tens_vs_seq_lifted_gd_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int inc, const double* X, int N, int L, int d,
                                    int cb, int ce, const double* g, double* Gd, void* ws, size_t wsb, void* stream) {
  return gpsig::tens_seq_lifted_gd_impl<double>(desc, Z, Mz, inc, X, N, L, d, cb, ce, g, Gd, ws, wsb, stream);
}
int gpsig_tens_vs_seq_lifted_gd_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int inc, const float* X, int N, int L, int d,
                                    int cb, int ce, const float* g, float* Gd, void* ws, size_t wsb, void* stream) {
  return gpsig::tens_seq_lifted_gd_impl<float>(desc, Z, Mz, inc, X, N, L, d, cb, ce, g, Gd, ws, wsb, stream);
}
size_t gpsig_tens_vs_seq_vjp_workspace_bytes(int Mz, int ncols, int L, int m, int elem) {
  return align_up((size_t)Mz * ncols * (m * (m + 1) / 2) * L * elem, 128);
}
int gpsig_trunc_vjp_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, const double* X2, int N2, int L2, int d,
                        int pair_mode, int rb, int re, const double* grad_levels, double* gradX, double* gradX2,
                        void* ws, size_t wsb, void* stream) {
  return trunc_vjp_impl<double>(desc, X, N, L, X2, N2, L2, d, pair_mode, rb, re, grad_levels, gradX, gradX2, ws, wsb, stream);
}
int gpsig_trunc_vjp_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L, const float* X2, int N2, int L2, int d,
                        int pair_mode, int rb, int re, const float* grad_levels, float* gradX, float* gradX2,
                        void* ws, size_t wsb, void* stream) {
  return trunc_vjp_impl<float>(desc, X, N, L, X2, N2, L2, d, pair_mode, rb, re, grad_levels, gradX, gradX2, ws, wsb, stream);
}
int gpsig_tens_vs_seq_vjp_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int inc, const double* X, int N, int L, int d,
                              int cb, int ce, const double* grad_levels, double* gradZ, double* gradX, void* ws, size_t wsb, void* stream) {
  return tens_seq_vjp_impl<double>(desc, Z, Mz, inc, X, N, L, d, cb, ce, grad_levels, gradZ, gradX, ws, wsb, stream);
}
int gpsig_tens_vs_seq_vjp_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int inc, const float* X, int N, int L, int d,
                              int cb, int ce, const float* grad_levels, float* gradZ, float* gradX, void* ws, size_t wsb, void* stream) {
  return tens_seq_vjp_impl<float>(desc, Z, Mz, inc, X, N, L, d, cb, ce, grad_levels, gradZ, gradX, ws, wsb, stream);
}
int gpsig_tens_tens_vjp_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int inc, int d, const double* grad_levels,
                            double* gradZ, void* stream) {
  return tens_tens_vjp_impl<double>(desc, Z, Mz, inc, d, grad_levels, gradZ, stream);
}
int gpsig_tens_tens_vjp_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int inc, int d, const float* grad_levels,
                            float* gradZ, void* stream) {
  return tens_tens_vjp_impl<float>(desc, Z, Mz, inc, d, grad_levels, gradZ, stream);
}

int gpsig_tens_tens_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int inc, int d, const double* ls,
                        const gpsig_epilogue* epi, double* out, void* stream) {
  return tens_tens_impl<double>(desc, Z, Mz, inc, d, ls, epi, out, stream);
}
int gpsig_tens_tens_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int inc, int d, const float* ls,
                        const gpsig_epilogue* epi, float* out, void* stream) {
  return tens_tens_impl<float>(desc, Z, Mz, inc, d, ls, epi, out, stream);
}

size_t gpsig_pde_workspace_bytes(int N, int L, int N2, int L2, int d, int order, int pair_mode) {
  (void)order; (void)pair_mode;
  return packed_bytes(N, L, d, 8) + (N2 > 0 ? packed_bytes(N2, L2, d, 8) : 0) + 128;   // + ticket counter of the persistent kernel
}
int gpsig_mirror_lower_f64(double* out, int N, int nplanes, void* stream) { return mirror_lower_impl<double>(out, N, nplanes, stream); }
int gpsig_mirror_lower_f32(float* out, int N, int nplanes, void* stream) { return mirror_lower_impl<float>(out, N, nplanes, stream); }
int gpsig_pde_f64(const double* X, int N, int L, const double* Y, int N2, int L2, int d, const double* ls,
                  int order, int pair_mode, int rb, int re, double sigma, double* out, void* ws, size_t wsb, void* stream) {
  return pde_impl(X, N, L, Y, N2, L2, d, ls, order, pair_mode, rb, re, sigma, out, ws, wsb, stream);
}

}  // extern "C"
