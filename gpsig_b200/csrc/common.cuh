// Shared device/host helpers of the gpsig_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

#include "../../include/gpsig_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "gpsig_b200 is written for sm_100a (B200) only"
#endif

namespace gpsig {

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
#define GPSIG_FAIL(code, ...) do { ::gpsig::set_error(__VA_ARGS__); return (code); } while (0)
#define GPSIG_CUDA_CHECK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) \
    GPSIG_FAIL(GPSIG_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define GPSIG_LAUNCH_CHECK(name) do { ::gpsig::count_launch(); cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) \
    GPSIG_FAIL(GPSIG_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e_)); } while (0)

// ---------------------------------------------------------------------------
// diagnostics (gpsig_profile_* / gpsig_launch_count in the C ABI)
// ---------------------------------------------------------------------------
void count_launch();                       // one per kernel launch of this library
// When profiling is enabled, the launcher of each MAIN kernel (the fused Gram,
// tensors-vs-sequences, PDE wavefront) brackets it with CUDA events on the
// caller's stream; gpsig_profile_last_kernel_ms() reads them back.
void prof_begin(cudaStream_t s);
void prof_end(cudaStream_t s);

// ---------------------------------------------------------------------------
// packed sequence layout in HBM (produced by pack_sequences_kernel)
//
// One block per sequence, contiguous, a multiple of 16 bytes so that a whole
// sequence is ONE 1-D TMA bulk copy (cp.async.bulk) into shared memory:
//     T coords[L][DP]   scaled by 1/lengthscales (kernels.py:358-359), zero
//                       padded from d to DP (zeros do not change dot products)
//     T sqn[LP]         |x_p|^2 of the scaled point (kernels.py:948), LP = L
//                       rounded up to 16 bytes
// ---------------------------------------------------------------------------
// d > 16 (the large-state-dimension path, lifted_kernels.cuh, and the general kernels): a multiple of 16
__host__ __device__ inline int padded_dim(int d) { return d <= 4 ? 4 : (d <= 8 ? 8 : (d + 15) / 16 * 16); }
__host__ __device__ inline int padded_len(int L, int elem) { int q = 16 / elem; return (L + q - 1) / q * q; }
__host__ __device__ inline size_t seq_block_elems(int L, int d, int elem) {
  return (size_t)L * padded_dim(d) + padded_len(L, elem);
}
__host__ __device__ inline size_t seq_block_elems_rs(int L, int RS, int elem) {
  return (size_t)L * RS + padded_len(L, elem);
}
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// RS = row stride of the packed rows in elements (>= DP; the pad is never read).  dup: every coordinate is written
// twice (x0, x0, x1, x1, ...; RS >= 2 DP) for the packed-pair float32 contraction of the systolic kernel.
template <typename T>
__global__ void pack_sequences_kernel(const T* __restrict__ X, int N, int L, int d, int DP, int RS,
                                      const T* __restrict__ lengthscales, T* __restrict__ packed,
                                      size_t block_elems, T scale, int dup = 0) {
  // one thread per point (n, p): reads d contiguous values, writes DP + 1
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)N * L) return;
  int n = (int)(idx / L), p = (int)(idx % L);
  const T* src = X + idx * d;
  T* blk = packed + (size_t)n * block_elems;
  T* dst = blk + (size_t)p * RS;
  T s = 0;
  for (int k = 0; k < DP; ++k) {
    T v = 0;
    if (k < d) { v = src[k]; if (lengthscales) v = v / lengthscales[k]; if (scale != T(1)) v *= scale; }
    if (dup) { dst[2 * k] = v; dst[2 * k + 1] = v; } else dst[k] = v;
    s += v * v;              // tf.reduce_sum(tf.square(X), axis=-1), kernels.py:948
  }
  blk[(size_t)L * RS + p] = s;
}

// ---------------------------------------------------------------------------
// base kernels as functors of (x.y, |x|^2, |y|^2) -- gpsig/kernels.py:946-1173
// ---------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ T gexp(T x);
template <> __device__ __forceinline__ double gexp<double>(double x) { return exp(x); }
template <> __device__ __forceinline__ float gexp<float>(float x) { return expf(x); }
template <typename T> __device__ __forceinline__ T gsqrt(T x);
template <> __device__ __forceinline__ double gsqrt<double>(double x) { return sqrt(x); }
template <> __device__ __forceinline__ float gsqrt<float>(float x) { return sqrtf(x); }
template <typename T> __device__ __forceinline__ T gpow(T x, T y);
template <> __device__ __forceinline__ double gpow<double>(double x, double y) { return pow(x, y); }
template <> __device__ __forceinline__ float gpow<float>(float x, float y) { return powf(x, y); }
template <typename T> __device__ __forceinline__ T gfma(T a, T b, T c);
template <> __device__ __forceinline__ double gfma<double>(double a, double b, double c) { return fma(a, b, c); }
template <> __device__ __forceinline__ float gfma<float>(float a, float b, float c) { return fmaf(a, b, c); }

// SMs of the current device (grids are sized in whole waves of it)
inline int device_sm_count() {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  return sms;
}

struct BaseParams { double p0, p1; int base; };

// Generic evaluation (runtime switch; used for everything but the two hot
// specialisations in the fused kernels and by the small tensor kernels).
template <typename T>
__device__ __forceinline__ T base_eval(int base, T p0, T p1, T dot, T xs, T ys) {
  switch (base) {
    case GPSIG_BASE_LINEAR: return dot;
    case GPSIG_BASE_RBF: return gexp<T>(-((T(-2) * dot + xs) + ys) / T(2));
    case GPSIG_BASE_COSINE: return dot / (gsqrt<T>(xs) * gsqrt<T>(ys));
    case GPSIG_BASE_POLY: return gpow<T>(dot + p0, p1);
    case GPSIG_BASE_MIX: { T ds = xs + ys - T(2) * dot; return p0 * gexp<T>(-ds / T(2)) + (T(1) - p0) * dot; }
    default: {
      T r2 = (T(-2) * dot + xs) + ys;
      T r = gsqrt<T>(r2 > T(1e-40) ? r2 : T(1e-40));
      if (base == GPSIG_BASE_MATERN12) return gexp<T>(-r);
      if (base == GPSIG_BASE_MATERN32) { T s3 = T(1.7320508075688772); return (T(1) + s3 * r) * gexp<T>(-s3 * r); }
      T s5 = T(2.23606797749979);
      return (T(1) + s5 * r + T(5.0 / 3.0) * (r * r)) * gexp<T>(-s5 * r);
    }
  }
}

// Spectral mixture kernel on raw coordinates (kernels.py:1101-1122); x, y hold >= d values, parameters are doubles.
struct SpectralParams { const double* p; int q, family; };
template <typename T>
__device__ __forceinline__ T spectral_eval(const SpectralParams& S, int d, const T* x, const T* y) {
  const double* alpha = S.p; const double* omega = S.p + S.q; const double* gamma = omega + (size_t)S.q * d;
  T tot = T(0);
  for (int q = 0; q < S.q; ++q) {
    T quad = T(0), ph = T(0);
    for (int f = 0; f < d; ++f) {
      const T df = x[f] - y[f];
      const T t = df * T(gamma[q * d + f]);
      quad += t * t;                                // reduce_sum(square(diff * gamma))
      ph += df * T(omega[q * d + f]);               // reduce_sum(diff * omega)
    }
    const T kern = S.family == GPSIG_SPECTRAL_EXP ? gexp<T>(-gsqrt<T>(quad) / T(2)) : gexp<T>(-quad / T(2));
    const T c = sizeof(T) == 8 ? T(cos(2. * 3.141592653589793 * (double)ph)) : T(cosf(2.f * 3.14159265358979f * (float)ph));
    tot += kern * c * T(alpha[q]);
  }
  return tot;
}

// float64, base-2 exponent t (coordinates pre-scaled by sqrt(log2 e)): 9 FP64-pipe instructions.
//   t = n/32 + f, n = round(32 t) by the magic-number add (exact, |f| <= 1/64),
//   2^t = 2^(n>>5) * 2^((n&31)/32) * 2^f, table as above, 2^f - 1 = f*(c1 + ... + c5 f^4) minimax
//   (1.4e-16 relative).  t is clamped to >= -1021 by an integer min on its sign/exponent word, so the
//   exponent-field add never leaves the normal range (the result there is ~1e-307, i.e. 0 at any tolerance).
__device__ __forceinline__ double exp2_tab32(double t, double tab_lane) {
  const unsigned hi = min((unsigned)__double2hiint(t), 0xC08FE800u);
  t = __hiloint2double((int)hi, __double2loint(t));
  const double s = t + 0x1.8p+47;
  const int n = __double2loint(s);
  const double f = t - (s - 0x1.8p+47);
  const double tj = __shfl_sync(0xffffffffu, tab_lane, n);          // source lane = n mod 32
  double q = fma(0x1.5d886e6d047f4p-10, f, 0x1.3b2b1bee89b8dp-7);
  q = fma(q, f, 0x1.c6b08d70300ecp-5);
  q = fma(q, f, 0x1.ebfbdff8131c3p-3);
  q = fma(q, f, 0x1.62e42fefa39efp-1);
  const double w = q * f;
  const double res = fma(tj, w, tj);
  return __hiloint2double(__double2hiint(res) + (n & ~31) * 32768, __double2loint(res));   // + (n >> 5) << 20
}

// The same with a 2048-entry table 2^(j/2048) in SHARED memory (16 KB per CTA, one LDS.64 per call instead of two
// shuffles) and a degree-3 polynomial (|f| <= 2^-12: truncation 3.4e-17): 7 FP64-pipe instructions.
// in-place update of the high word of a double (sign / exponent / top mantissa bits): written so that ptxas keeps the
// register pair where it is instead of assembling a new one with two moves
__device__ __forceinline__ void hi_min_u32(double& v, unsigned cap) {
  asm("{\n\t.reg .b32 lo, hi;\n\tmov.b64 {lo, hi}, %0;\n\tmin.u32 hi, hi, %1;\n\tmov.b64 %0, {lo, hi};\n\t}" : "+d"(v) : "r"(cap));
}
__device__ __forceinline__ void hi_mad(double& v, int k, int scale) {      // hi += k * scale
  asm("{\n\t.reg .b32 lo, hi;\n\tmov.b64 {lo, hi}, %0;\n\tmad.lo.s32 hi, %1, %2, hi;\n\tmov.b64 %0, {lo, hi};\n\t}" : "+d"(v) : "r"(k), "r"(scale));
}
__device__ __forceinline__ double exp2_tab2048(double t, uint32_t table_smem) {
  hi_min_u32(t, 0xC08FE800u);                                       // t >= -1021 (integer min on the sign/exponent word)
  const double s = t + 0x1.8p+41;                                   // ulp = 2^-11
  const int n = __double2loint(s);
  const double f = t - (s - 0x1.8p+41);
  double tj;
  asm("ld.shared.f64 %0, [%1];" : "=d"(tj) : "r"(table_smem + ((uint32_t)(n & 2047) << 3)));
  double q = fma(0x1.c6b08d704a0c0p-5, f, 0x1.ebfbdff82c58fp-3);     // (ln 2)^3 / 6, (ln 2)^2 / 2
  q = fma(q, f, 0x1.62e42fefa39efp-1);                              // ln 2
  const double w = q * f;
  double res = fma(tj, w, tj);
  hi_mad(res, n >> 11, 1 << 20);                                    // * 2^(n >> 11): exponent-field add
  return res;
}

// The same with a 16384-entry table 2^(j/16384) (128 KB of shared memory) and a degree-2 minimax polynomial
// (|f| <= 2^-15: 3.9e-16 relative): 6 FP64-pipe instructions.
__device__ __forceinline__ double exp2_tab16k(double t, uint32_t table_smem) {
  hi_min_u32(t, 0xC08FE800u);                                       // t >= -1021
  const double s = t + 0x1.8p+38;                                   // ulp = 2^-14
  const int n = __double2loint(s);
  const double f = t - (s - 0x1.8p+38);
  double tj;
  asm("ld.shared.f64 %0, [%1];" : "=d"(tj) : "r"(table_smem + ((uint32_t)(n & 16383) << 3)));
  const double q = fma(0x1.ebfbdff82c58fp-3, f, 0x1.62e42feff8e01p-1);   // (ln 2)^2 / 2, ln 2 + (ln 2)^3 2^-30 / 8
  const double w = q * f;
  double res = fma(tj, w, tj);
  hi_mad(res, n >> 14, 1 << 20);
  return res;
}

// ... and with an 8192-entry table (64 KB: two CTAs of the Kzx kernel per SM keep theirs) and the degree-2 polynomial
// on |f| <= 2^-14: 3.2e-15 relative, 6 FP64-pipe instructions.
__device__ __forceinline__ double exp2_tab8k(double t, uint32_t table_smem) {
  hi_min_u32(t, 0xC08FE800u);                                       // t >= -1021
  const double s = t + 0x1.8p+39;                                   // ulp = 2^-13
  const int n = __double2loint(s);
  const double f = t - (s - 0x1.8p+39);
  double tj;
  asm("ld.shared.f64 %0, [%1];" : "=d"(tj) : "r"(table_smem + ((uint32_t)(n & 8191) << 3)));
  const double q = fma(0x1.ebfbdff82c58fp-3, f, 0x1.62e42ff0f8a36p-1);   // (ln 2)^2 / 2, ln 2 + (ln 2)^3 2^-28 / 8
  const double w = q * f;
  double res = fma(tj, w, tj);
  hi_mad(res, n >> 13, 1 << 20);
  return res;
}
// fills table[0 .. 2^BITS) with 2^(j / 2^BITS) from 2^(BITS/2 rounded up) + 2^(BITS/2) exp2 calls and one product per entry
// (<= 1.5 ulp), all threads of the CTA; `scratch` holds 2^ceil(BITS/2) + 2^floor(BITS/2) doubles; ends with a block barrier
template <int BITS>
__device__ __forceinline__ void exp2_table_fill(double* table, double* scratch, int tid, int nthreads) {
  constexpr int HB = (BITS + 1) / 2, LB = BITS / 2;                 // j = (hi << LB) | lo
  for (int e = tid; e < (1 << HB) + (1 << LB); e += nthreads)
    scratch[e] = e < (1 << HB) ? exp2((double)e * (1.0 / (1 << HB))) : exp2((double)(e - (1 << HB)) * (1.0 / (1 << BITS)));
  __syncthreads();
  for (int j = tid; j < (1 << BITS); j += nthreads) table[j] = scratch[j >> LB] * scratch[(1 << HB) + (j & ((1 << LB) - 1))];
  __syncthreads();
}

__device__ __forceinline__ float exp2_approx(float t) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t));
  return r;
}

// ---- base-kernel specialisations on the fused path -------------------------
enum { OP_LINEAR = 0, OP_RBF = 1, OP_GENERIC = 2 };

template <typename T, int OP> struct BaseOp;
template <typename T> struct BaseOp<T, OP_LINEAR> {
  static __device__ __forceinline__ T col_const(T ysq) { return T(0); }
  static __device__ __forceinline__ T row_const(T xsq) { return T(0); }
  static __device__ __forceinline__ T init(T xc, T yc) { return T(0); }
  static __device__ __forceinline__ T finish(T acc, T xc, T yc, int base, T p0, T p1) { return acc; }
};
template <typename T> struct BaseOp<T, OP_RBF> {
  // exp(-(|x|^2 + |y|^2 - 2 x.y)/2) = exp(x.y - |x|^2/2 - |y|^2/2)   (kernels.py:1042-1044)
  static __device__ __forceinline__ T col_const(T ysq) { return T(-0.5) * ysq; }
  static __device__ __forceinline__ T row_const(T xsq) { return T(-0.5) * xsq; }
  static __device__ __forceinline__ T init(T xc, T yc) { return xc + yc; }
  static __device__ __forceinline__ T finish(T acc, T xc, T yc, int base, T p0, T p1) { return gexp<T>(acc); }
};
template <typename T> struct BaseOp<T, OP_GENERIC> {
  static __device__ __forceinline__ T col_const(T ysq) { return ysq; }
  static __device__ __forceinline__ T row_const(T xsq) { return xsq; }
  static __device__ __forceinline__ T init(T xc, T yc) { return T(0); }
  static __device__ __forceinline__ T finish(T acc, T xc, T yc, int base, T p0, T p1) {
    return base_eval<T>(base, p0, p1, acc, xc, yc);
  }
};
// out-of-line copy for kernels that evaluate the base kernel at many unrolled sites
template <typename T>
__device__ __noinline__ T base_eval_call(int base, T p0, T p1, T dot, T xs, T ys) {
  return base_eval<T>(base, p0, p1, dot, xs, ys);
}


// ---------------------------------------------------------------------------
// PTX: mbarrier + 1-D TMA bulk copy (SASS: UBLKCP / SYNCS)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy, completion signalled on an mbarrier (bytes % 16 == 0)
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace gpsig
