// Fused truncated signature-kernel Gram for sm_100a -- systolic ("skewed") variant.
//
// Same contract as trunc_kernel.cuh (gpsig/kernels.py:209-238 + signature_algs.py:8-35
// + epilogue kernels.py:431-477) for the shapes that carry the throughput:
// K(X, X2) with 32 <= L and L2 <= 32*COLS (one warp pass over the columns).
//
// One WARP owns one column sequence y_j and streams the row sequences x_i of its
// CTA's row chunk.  Lane t owns COLS consecutive columns q of the L x L2 base-kernel
// tile and sweeps the rows; lane t runs ONE ROW-STEP BEHIND lane t-1:
//
//     step s:   lane t works on stream row (s - t)
//
// so everything lane t needs from the columns to its left at that row
//   - the last first-difference U[p, q0-1] (for the 2-D increment D[p, q0-1]),
//   - the row prefixes  e_l[p] = sum_{q' < q0} C_l[p, q'],  l = 0..m-2
//     (C_l = running column sums of level l; signature_algs.py:32 re-associated),
// was finished by lane t-1 one step earlier and arrives with ONE warp shuffle per
// value -- there is no log-step prefix scan and no barrier in the recursion.  The
// stream of rows continues across pairs: at the first row of the next x_i the same
// shuffle chain carries the level totals K_l = sum_q C_l[L-1, q] to lane 31, which
// applies the epilogue and stores.  The pipeline is filled once per CTA.
//
// x_i arrives in shared memory as one TMA bulk copy into a 3-deep ring shared by
// the CTA's warps (full/empty mbarriers; lane 0 of warp 0 is the producer).  Rows
// are padded to an odd number of 16-byte units so that the skewed row reads of a
// quarter warp fall in distinct banks.
#pragma once
#include <type_traits>

#include "trunc_kernel.cuh"

namespace gpsig {

constexpr int kSkewBufs = 3;

// float32: the streamed x rows are stored with every coordinate DUPLICATED (x0, x0, x1, x1, ...), so that one 16-byte
// shared-memory load yields two ready-made (x_k, x_k) pairs for the packed FFMA2 of the contraction: two adjacent
// columns per instruction, no register moves to assemble the pairs (the y pairs are loop invariant).  Round 1 measured
// it at c2 (N = N2 = 1000, 12 warps at 168 registers) 4-5 % faster where registers allow (RBF m = 4: 43.96 against
// 45.67 ms, linear m = 5: 42.8 against 45.1 ms) and 4 % SLOWER at the headline shape (RBF m = 5: 50.6 against 48.4 ms,
// 16 bytes of spills) -- the contraction drops from 8 to 4.6 instructions per element; it also moves 67 % more x bytes.
// Round 2 (after the ring bookkeeping moved behind one test per step the headline instantiation no longer spills): ON for the
// instantiations that hold it without spills (SkewDup below): c2 float32 45.64 -> 44.67 ms.
#ifndef GPSIG_SKEW_DUP_F32
#define GPSIG_SKEW_DUP_F32 1
#endif
// row stride (in elements) of the padded packed layout: an odd number of 16-byte units
__host__ __device__ inline int skew_row_stride(int d, int elem, bool dup = false) {
  const int dp = padded_dim(d) * (dup ? 2 : 1);
  int u = dp * elem / 16;
  if ((u & 1) == 0) ++u;
  return u * 16 / elem;
}

// float64: the lane's column tile of y (COLS x DP doubles) does not fit in registers next to the
// recursion state.  It lives in TENSOR MEMORY instead: every lane owns one TMEM lane, writes its tile
// once with tcgen05.st and re-reads it every step with tcgen05.ld (SASS STTM / LDTM).  TMEM is a
// datapath of its own (measured 385 B/clk/SM at 8 warps, tools/microbench2.cu, against 127 B/clk/SM
// for shared memory), so the tile no longer competes with the x-row loads and the warp shuffles of
// the chain for the shared-memory pipe (which the lane-interleaved shared-memory tile kept 45 % busy).
#ifndef GPSIG_SKEW_TMEM
#define GPSIG_SKEW_TMEM 1
#endif
// float64 RBF: packed coordinates pre-scaled by sqrt(log2 e) (abi.cu) and exp2_tab32 below
#ifndef GPSIG_SKEW_TMEM_F32
#define GPSIG_SKEW_TMEM_F32 0     // float32 tile in tensor memory as well (measured: no gain over registers at 12 warps, 20.4 vs 20.2 M pairs/s)
#endif
#ifndef GPSIG_SKEW_GS_F64
#define GPSIG_SKEW_GS_F64 2
#endif
#ifndef GPSIG_SKEW_GS_F32
#define GPSIG_SKEW_GS_F32 2
#endif
#ifndef GPSIG_TMEM_TIE
#define GPSIG_TMEM_TIE 1
#endif
#ifndef GPSIG_SKEW_EXP2
#define GPSIG_SKEW_EXP2 1
#endif
// float64 RBF: exp2 with a 2048-entry shared-memory table and a degree-3 polynomial (7 FP64 instructions, common.cuh)
// instead of the 32-entry shuffle table and degree 5 (9)
#ifndef GPSIG_SKEW_EXP2_T2048
#define GPSIG_SKEW_EXP2_T2048 1
#endif
// ... or with a 16384-entry table (128 KB per CTA) and a degree-2 polynomial (6 FP64 instructions; 3.9e-16 relative),
// taken whenever the table fits next to the x ring (launch_skew_inst); 98.96 -> 96.69 ms at c2
#ifndef GPSIG_SKEW_EXP2_T16K
#define GPSIG_SKEW_EXP2_T16K 1
#endif
// first row of a pair on the common recursion path (see the recursion below): 1 = on.  Measured at c2, N = N2 = 1000:
// float64 98.96 -> 97.43 ms; float32 46.17 -> 46.78 ms with the plain contraction (its top level keeps one partial sum per
// column: 6 more adds per step), but 177.3 -> 165.5 ms at N = N2 = 2000 next to the packed FFMA2 contraction (GPSIG_SKEW_DUP_F32)
#ifndef GPSIG_SKEW_FRF_F64
#define GPSIG_SKEW_FRF_F64 1
#endif
#ifndef GPSIG_SKEW_FRF_F32
#define GPSIG_SKEW_FRF_F32 1
#endif
// lane 0 as a dummy lane (see the kernel): no per-step selects for the start of the chains; taken when L2 <= 31 COLS.
// Measured at c2: float64 95.5 -> 93.3 ms (the selects sat right behind the chain shuffles and stalled the in-order
// warp on their latency); float32 46.17 -> 46.34 ms
#ifndef GPSIG_SKEW_L0D_F64
#define GPSIG_SKEW_L0D_F64 1
#endif
#ifndef GPSIG_SKEW_L0D_F32
#define GPSIG_SKEW_L0D_F32 0
#endif
// x rows read through a 32-bit shared-memory address held in a register (ld.shared by inline PTX) instead of a generic
// pointer: at the float32 kernel's 168 registers the compiler otherwise rebuilds the shared window base from
// SR_CgaCtaId (S2R, ~20 cycles on the short scoreboard) in every step
// (measured at c2: float64 93.29 -> 92.90 ms; float32, where the S2R was seen, 45.63 -> 45.89 ms: off there)
#ifndef GPSIG_SKEW_SADDR_F32
#define GPSIG_SKEW_SADDR_F32 0
#endif
#ifndef GPSIG_SKEW_SADDR_F64
#define GPSIG_SKEW_SADDR_F64 1
#endif
__device__ __forceinline__ float4 lds16(uint32_t a, float) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ double2 lds16(uint32_t a, double) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds1(uint32_t a, float) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ double lds1(uint32_t a, double) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
// ring bookkeeping behind one test per step instead of two (see the step prologue): float32 46.17 -> 45.63 ms at c2
// (there the compiler keeps the step counters in vector registers and every test costs a divergence check), float64 unchanged
#ifndef GPSIG_SKEW_RING1
#define GPSIG_SKEW_RING1 1
#endif
static_assert(kSkewBufs == 3 || GPSIG_SKEW_RING1 == 0, "the merged ring bookkeeping assumes three slots");
// partial sums of the top level (float64; 4 measured slower: 93.3 -> 94.9 ms)
#ifndef GPSIG_SKEW_NT_F64
#define GPSIG_SKEW_NT_F64 2
#endif
// float64, DP = 8: the y tile column-major in tensor memory, one tcgen05.ld.x16 per column instead of two x8
// (7 instead of 14 loads and address moves per step; 98.96 -> 98.32 ms at c2)
#ifndef GPSIG_SKEW_TM16
#define GPSIG_SKEW_TM16 1
#endif
constexpr size_t kSkewExpTableBytes = 2048 * sizeof(double);
constexpr size_t kSkewExpTableBytesBig = 16384 * sizeof(double);
template <typename T, int OP> struct SkewExpTable { static constexpr bool value = GPSIG_SKEW_EXP2 && GPSIG_SKEW_EXP2_T2048 && sizeof(T) == 8 && OP == OP_RBF; };

// ---- tensor memory as per-lane storage (PTX tcgen05.alloc / st / ld) ---------
__device__ __forceinline__ uint32_t tmem_alloc(uint32_t* slot, int ncols) {    // whole CTA; warp 0 allocates
  if ((threadIdx.x >> 5) == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *reinterpret_cast<volatile uint32_t*>(slot);
}
__device__ __forceinline__ void tmem_free(uint32_t base, int ncols) {           // whole CTA
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if ((threadIdx.x >> 5) == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(ncols) : "memory");
}
// 4 doubles (8 words) per lane
__device__ __forceinline__ void tmem_st4(uint32_t addr, double a, double b, double c, double d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(addr), "r"(__double2loint(a)), "r"(__double2hiint(a)), "r"(__double2loint(b)), "r"(__double2hiint(b)),
                 "r"(__double2loint(c)), "r"(__double2hiint(c)), "r"(__double2loint(d)), "r"(__double2hiint(d)) : "memory");
}
__device__ __forceinline__ void tmem_st8w(uint32_t addr, const uint32_t (&w)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
struct TmemQuad { uint32_t w[8]; };
__device__ __forceinline__ void tmem_ld4(TmemQuad& r, uint32_t addr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
               : "r"(addr));
}
struct TmemHex { uint32_t w[16]; };
__device__ __forceinline__ void tmem_ld8d(TmemHex& r, uint32_t addr) {      // 8 doubles (16 words) per lane
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7]),
                 "=r"(r.w[8]), "=r"(r.w[9]), "=r"(r.w[10]), "=r"(r.w[11]), "=r"(r.w[12]), "=r"(r.w[13]), "=r"(r.w[14]), "=r"(r.w[15])
               : "r"(addr));
}
__device__ __forceinline__ void tmem_tie(TmemHex& r) {
  asm volatile("" : "+r"(r.w[0]), "+r"(r.w[1]), "+r"(r.w[2]), "+r"(r.w[3]), "+r"(r.w[4]), "+r"(r.w[5]), "+r"(r.w[6]), "+r"(r.w[7]),
                    "+r"(r.w[8]), "+r"(r.w[9]), "+r"(r.w[10]), "+r"(r.w[11]), "+r"(r.w[12]), "+r"(r.w[13]), "+r"(r.w[14]), "+r"(r.w[15]));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// empty volatile asm that redefines the loaded registers after the wait: no use of them can be scheduled above it
__device__ __forceinline__ void tmem_tie(TmemQuad& r) {
  asm volatile("" : "+r"(r.w[0]), "+r"(r.w[1]), "+r"(r.w[2]), "+r"(r.w[3]), "+r"(r.w[4]), "+r"(r.w[5]), "+r"(r.w[6]), "+r"(r.w[7]));
}
// shuffle-up by one lane as a volatile asm: stays where it is written relative to the tensor-memory
// loads / waits, i.e. in between the FMA groups of the contraction instead of in one burst at the step start
__device__ __forceinline__ double shfl_up1_pinned(double v) {
  int lo = __double2loint(v), hi = __double2hiint(v);
  asm volatile("shfl.sync.up.b32 %0, %0, 1, 0, 0xffffffff;" : "+r"(lo));
  asm volatile("shfl.sync.up.b32 %0, %0, 1, 0, 0xffffffff;" : "+r"(hi));
  return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double tmem_quad_get(const TmemQuad& r, int e) { return __hiloint2double((int)r.w[2 * e + 1], (int)r.w[2 * e]); }
template <typename T> __device__ __forceinline__ T tmem_get(const TmemQuad& r, int e);
template <> __device__ __forceinline__ double tmem_get<double>(const TmemQuad& r, int e) { return tmem_quad_get(r, e); }
template <> __device__ __forceinline__ float tmem_get<float>(const TmemQuad& r, int e) { return __uint_as_float(r.w[e]); }
__device__ __forceinline__ float shfl_up1_pinned(float v) {
  int x = __float_as_int(v);
  asm volatile("shfl.sync.up.b32 %0, %0, 1, 0, 0xffffffff;" : "+r"(x));
  return __int_as_float(x);
}
__host__ __device__ constexpr int tmem_cols_pow2(int n) { return n <= 32 ? 32 : n <= 64 ? 64 : n <= 128 ? 128 : n <= 256 ? 256 : 512; }

// ---- exp for the RBF base kernel --------------------------------------------
// float64: table-driven, 12 FP64-pipe instructions instead of ~22 for exp():
//   x = (32 k + j) ln2/32 + r,  exp(x) = 2^k * 2^(j/32) * exp(r),  |r| <= ln2/64,
// the 32-entry table 2^(j/32) lives one entry per lane and is read with a shuffle
// (so the call must be warp-convergent), exp(r) is a degree-6 Taylor polynomial
// (truncation 3.5e-18).  Accurate to ~1 ulp for x <= 0 (for x < -693 it returns ~1e-301, i.e. 0 at any tolerance).
__device__ __forceinline__ double exp_tab32(double x, double tab_lane) {
  const double s = fma(x, 46.16624130844683, 6755399441055744.0);
  const int n = __double2loint(s);
  const double nf = s - 6755399441055744.0;
  double r = fma(nf, -0x1.62e42fee00000p-6, x);
  r = fma(nf, -0x1.a39ef35793c76p-38, r);
  const double tj = __shfl_sync(0xffffffffu, tab_lane, n & 31);
  double q = fma(0.001388888888888889, r, 0.008333333333333333);
  q = fma(q, r, 0.041666666666666664);
  q = fma(q, r, 0.16666666666666666);
  q = fma(q, r, 0.5);
  q = fma(q, r, 1.0);
  const double w = q * r;
  const double res = fma(tj, w, tj);
  // 2^k by an exponent-field add.  k is clamped at -1000 (the result is ~1e-301 instead of a
  // denormal) and forced there for x <= -800, where n may have wrapped (integer test on the sign/exponent word).
  int k = max(n >> 5, -1000);
  if ((unsigned)__double2hiint(x) >= 0xC0890000u) k = -1000;
  return __hiloint2double(__double2hiint(res) + (k << 20), __double2loint(res));
}

// exp2_tab32 (float64, base-2 exponent, 9 FP64-pipe instructions) lives in common.cuh

// ---- packed FP32 pairs (Blackwell FFMA2 / FADD2: two FP32 operations per lane and issue slot) ---------
// The float32 kernel is bound by issue slots (75 % busy, profiles/r01f_ncu_trunc_skew_f32.txt); its FP32
// work on ADJACENT COLUMNS (contraction, first differences, level updates) can be issued as packed pairs:
// same flops, half the instructions (measured: FFMA2 at 1.9 warp-instr/clk/SM = the FFMA flop rate).
// OFF by default: with the operands where they are today the pairs have to be assembled with MOVs every
// step (x duplicated into a register pair, the y columns re-paired out of their float4 registers), which
// costs more than the packing saves: 19.0 M pairs/s against 20.1 M at c2.  It needs the x rows stored
// duplicated in shared memory and the y tile loaded as column pairs (next round).
#ifndef GPSIG_SKEW_PACK_F32
#define GPSIG_SKEW_PACK_F32 0
#endif
using u64 = unsigned long long;
__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpk2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
template <typename T> struct PairOps {          // generic: two scalar operations
  static constexpr bool kPacked = false;
  static __device__ __forceinline__ void fma(T& a0, T& a1, T x0, T x1, T y0, T y1) { a0 = gfma<T>(x0, y0, a0); a1 = gfma<T>(x1, y1, a1); }
  static __device__ __forceinline__ void add(T& r0, T& r1, T a0, T a1, T b0, T b1) { r0 = a0 + b0; r1 = a1 + b1; }
  static __device__ __forceinline__ void sub(T& r0, T& r1, T a0, T a1, T b0, T b1) { r0 = a0 - b0; r1 = a1 - b1; }
};
#if GPSIG_SKEW_PACK_F32
template <> struct PairOps<float> {
  static constexpr bool kPacked = true;
  static __device__ __forceinline__ void fma(float& a0, float& a1, float x0, float x1, float y0, float y1) {
    u64 a = pk2(a0, a1);
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a) : "l"(pk2(x0, x1)), "l"(pk2(y0, y1)));
    unpk2(a, a0, a1);
  }
  static __device__ __forceinline__ void add(float& r0, float& r1, float a0, float a1, float b0, float b1) {
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a0, a1)), "l"(pk2(b0, b1)));
    unpk2(r, r0, r1);
  }
  static __device__ __forceinline__ void sub(float& r0, float& r1, float a0, float a1, float b0, float b1) {
    u64 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a0, a1)), "l"(pk2(b0, b1)));
    unpk2(r, r0, r1);
  }
};
#endif

template <typename T> __device__ __forceinline__ void zero_on_lane0(T& v, int lane) { if (lane == 0) v = T(0); }   // (ptxas turns a predicated 64-bit move into the same two selects)

template <int N, typename F, int I = 0>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (I < N) { f(std::integral_constant<int, I>{}); static_for<N, F, I + 1>(static_cast<F&&>(f)); }
}

template <typename T, int OP> struct SkewFinish {
  template <int BIG>
  static __device__ __forceinline__ T apply(T acc, T xc, T yc, int base, T p0, T p1, double tab) {
    if (OP == OP_GENERIC) return base_eval_call<T>(base, p0, p1, acc, xc, yc);
    return BaseOp<T, OP>::finish(acc, xc, yc, base, p0, p1);
  }
};
template <> struct SkewFinish<double, OP_RBF> {
  template <int BIG>
  static __device__ __forceinline__ double apply(double acc, double, double, int, double, double, double tab) {
#if GPSIG_SKEW_EXP2 && GPSIG_SKEW_EXP2_T2048
    if constexpr (BIG != 0) return exp2_tab16k(acc, (uint32_t)__double2loint(tab));
    return exp2_tab2048(acc, (uint32_t)__double2loint(tab));      // `tab` carries the table's shared-memory address
#elif GPSIG_SKEW_EXP2
    return exp2_tab32(acc, tab);
#else
    return exp_tab32(acc, tab);
#endif
  }
};

// rows per step (R) and warps per CTA (NW); tunables, overridable at build time for experiments.
// Measured at c2 (tools/time_trunc.py): float32 R=2: 8 warps 17.5 M pairs/s, 12 warps 20.2 M (168 registers), 14: 14.4 M,
// 16: 16.9 M; float64 (tile in tensor memory) R=2 x 8 warps 9.03 M, R=1 x 12 warps 8.87 M.
#ifndef GPSIG_SKEW_R_F64
#define GPSIG_SKEW_R_F64 2
#endif
#ifndef GPSIG_SKEW_NW_F64
#define GPSIG_SKEW_NW_F64 8
#endif
#ifndef GPSIG_SKEW_R_F32
#define GPSIG_SKEW_R_F32 2
#endif
#ifndef GPSIG_SKEW_NW_F32
#define GPSIG_SKEW_NW_F32 12
#endif
template <typename T> struct SkewCfg;
template <> struct SkewCfg<double> { static constexpr int R = GPSIG_SKEW_R_F64; static constexpr int NW = GPSIG_SKEW_NW_F64; };
template <> struct SkewCfg<float> { static constexpr int R = GPSIG_SKEW_R_F32; static constexpr int NW = GPSIG_SKEW_NW_F32; };
// warps per CTA of one instantiation: float32 shapes whose y tile sits in shared memory keep 8 (12 tiles do not fit)
template <typename T, int DP, int COLS> struct SkewNW {
  static constexpr int value = (sizeof(T) == 4 && !YPolicy<T, DP, COLS>::kInRegs) ? 8 : SkewCfg<T>::NW;
};
// y tile in tensor memory: float64 shapes whose tile does not fit in registers (and whose warps' tiles fit the 512 columns)
template <typename T, int DP, int COLS> struct SkewYTmem {
  static constexpr bool value = GPSIG_SKEW_TMEM && (sizeof(T) == 8 ? !YPolicy<T, DP, COLS>::kInRegs : (GPSIG_SKEW_TMEM_F32 != 0)) &&
                                DP * sizeof(T) >= 32 && SkewNW<T, DP, COLS>::value % 4 == 0 &&
                                (SkewNW<T, DP, COLS>::value / 4) * COLS * DP * (sizeof(T) / 4) <= 512;
};

// which instantiations read the duplicated layout: float32 with the y tile in registers (and not in tensor memory)
// ... and whose state leaves room for the pair registers at 168 registers per thread (no spills: up to 7 columns per
// lane and 5 levels; the 8-level / 8-column instantiations spill 16-484 bytes with it and keep the plain layout)
constexpr int kSkewDupMaxCols = 7, kSkewDupMaxLev = 5;
template <typename T, int DP, int COLS, int MLEV, int OP> struct SkewDup {
  static constexpr bool value = GPSIG_SKEW_DUP_F32 && sizeof(T) == 4 && YPolicy<T, DP, COLS>::kInRegs && !SkewYTmem<T, DP, COLS>::value &&
                                COLS <= kSkewDupMaxCols && MLEV <= kSkewDupMaxLev && OP != OP_GENERIC;
};

// float32 RBF: the packed coordinates are pre-scaled by sqrt(log2 e) (abi.cu), so the exponent
// arrives in base-2 units and exp() is a single MUFU.EX2
template <> struct SkewFinish<float, OP_RBF> {
  template <int BIG>
  static __device__ __forceinline__ float apply(float acc, float, float, int, float, float, double) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(acc));
    return r;
  }
};

template <typename T, int DP, int COLS>
constexpr size_t skew_ywarp_bytes() { return SkewYTmem<T, DP, COLS>::value ? 0 : trunc_ywarp_bytes<T, DP, COLS>(); }

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <typename T, int DP, int COLS, int MLEV, int OP, int BIG = 0, int L0D = 0>
__global__ void __launch_bounds__(SkewNW<T, DP, COLS>::value * 32, 1) trunc_skew_kernel(const TruncParams P, const int RS) {
  using V16 = typename Vec16<T>::type;
  constexpr int V = Vec16<T>::V;
  constexpr int KV = DP / V;
  constexpr int R = SkewCfg<T>::R;
  constexpr int NW = SkewNW<T, DP, COLS>::value;
  constexpr bool YREG = YPolicy<T, DP, COLS>::kInRegs;
  constexpr bool YTM = SkewYTmem<T, DP, COLS>::value;      // y tile in tensor memory (float64)
  constexpr bool DUP = SkewDup<T, DP, COLS, MLEV, OP>::value;        // x / y rows stored with duplicated coordinates (float32, y in registers)
  constexpr int CE = 32 / sizeof(T);                        // elements per 8-word tensor-memory load (4 doubles / 8 floats)
  constexpr int KC = DP / CE > 0 ? DP / CE : 1;             // chunks per column
  constexpr int TMW = COLS * DP * (sizeof(T) / 4);          // TMEM words (columns) per warp
  constexpr int TMCOLS = tmem_cols_pow2((NW / 4) * TMW);
  static_assert(!YTM || (NW % 4 == 0 && (NW / 4) * TMW <= 512), "tensor memory plan does not fit");
  constexpr int NB = kSkewBufs;
  constexpr bool TM16 = YTM && GPSIG_SKEW_TM16 != 0 && sizeof(T) == 8 && KC == 2;   // column-major tile, x16 loads
  using Op = BaseOp<T, OP>;

  extern __shared__ __align__(128) unsigned char skew_smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(skew_smem);       // [NB]
  uint64_t* empty_bar = full_bar + NB;                               // [NB]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // the systolic kernel is built for the differenced tile only (difference = False, a non-default option, runs
  // trunc_gram_kernel): as a compile-time constant the per-row branch disappears and the rows of a step are one basic block
  constexpr bool diff = true;
  const int L = P.L, L2 = P.L2;
  const size_t xbytes = P.xblk * sizeof(T);
  const size_t xslot = align_up_dev(xbytes);
  unsigned char* xring = skew_smem + 128;
  T* ysm = reinterpret_cast<T*>(xring + NB * xslot) + (size_t)warp * (COLS * 32 * DP);   // only if !YREG && !YTM

  const int i_lo = P.row_begin + blockIdx.y * P.ichunk;
  const int i_hi = min(P.row_end, i_lo + P.ichunk);
  const int npairs = i_hi - i_lo;
  const T* xg = static_cast<const T*>(P.xp);
  // K(X, X): every pair of this CTA has i > j -> its mirror image (j, i) is computed elsewhere (whole CTA exits)
  if (P.tri && i_lo >= (int)(blockIdx.x + 1) * NW) return;

  if (threadIdx.x == 0) {
    for (int b = 0; b < NB; ++b) { mbar_init(&full_bar[b], 1); mbar_init(&empty_bar[b], NW); }
    fence_mbar_init();
  }
  uint32_t tbase = 0;
  if (YTM) tbase = tmem_alloc(reinterpret_cast<uint32_t*>(skew_smem + 96), TMCOLS);   // includes the block barrier
  else __syncthreads();
  // this warp's tile: its lane quarter (warp % 4), column block (warp / 4)
  const uint32_t ta = tbase + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * TMW);
  if (threadIdx.x == 0) {
    for (int it = 0; it < NB && it < npairs; ++it) {
      mbar_expect_tx(&full_bar[it], (uint32_t)xbytes);
      tma_load_1d(xring + it * xslot, xg + (size_t)(i_lo + it) * P.xblk, (uint32_t)xbytes, &full_bar[it]);
    }
  }

  const int j = blockIdx.x * NW + warp;
  const bool active = j < P.N2;
  const T* yblk = static_cast<const T*>(P.yp) + (size_t)(active ? j : P.N2 - 1) * P.yblk;
  const T* ysqn = yblk + (size_t)L2 * P.yrs;

  // ---- this lane's columns of y (fixed for the whole kernel) ---------------------
  // Slots past the last column repeat it: their first differences equal their left
  // neighbour's, so their 2-D increments are exactly zero without any masking.
  // L0D: lane 0 is a DUMMY lane to the left of column 0 and the columns start at lane 1.  All its slots repeat column 0
  // (as the slots past the last column repeat that one), so its 2-D increments and column sums are exactly 0, and what
  // it hands to lane 1 -- the start of every chain -- is what it received from ITSELF (shfl.up leaves lane 0 its own
  // value): the zeros its send registers start with.  The chain values need no per-step select for lane 0 any more;
  // only the first difference to the left of its first slot still does (`left` below).
  const int q0 = (L0D != 0 ? lane - 1 : lane) * COLS;
  T yc[COLS];
  V16 yreg[YREG ? COLS : 1][YREG ? KV : 1];
#pragma unroll
  for (int c = 0; c < COLS; ++c) {
    const int q = max(min(q0 + c, L2 - 1), 0);
    yc[c] = Op::col_const(ysqn[q]);
    if constexpr (YTM) {
#pragma unroll
      for (int kc = 0; kc < KC; ++kc) {
        union { V16 v[2]; uint32_t w[8]; } u;
        u.v[0] = *reinterpret_cast<const V16*>(yblk + (size_t)q * P.yrs + kc * CE);
        u.v[1] = *reinterpret_cast<const V16*>(yblk + (size_t)q * P.yrs + kc * CE + V);
        tmem_st8w(ta + (uint32_t)((TM16 ? c * KC + kc : kc * COLS + c) * 8), u.w);
      }
    } else {
#pragma unroll
      for (int kv = 0; kv < KV; ++kv) {
        V16 v;
        if constexpr (DUP) {                       // duplicated layout: coordinate k sits at 2k
          const float4 a = *reinterpret_cast<const float4*>(yblk + (size_t)q * P.yrs + 2 * kv * V);
          const float4 b = *reinterpret_cast<const float4*>(yblk + (size_t)q * P.yrs + 2 * kv * V + 4);
          float4 t = make_float4(a.x, a.z, b.x, b.z);
          v = *reinterpret_cast<V16*>(&t);
        } else v = *reinterpret_cast<const V16*>(yblk + (size_t)q * P.yrs + kv * V);
        if (YREG) yreg[YREG ? c : 0][YREG ? kv : 0] = v;
        else *reinterpret_cast<V16*>(ysm + ((size_t)(c * KV + kv) * 32 + lane) * V) = v;
      }
    }
  }
  if (YTM) tmem_st_wait();
  else if (!YREG) __syncwarp();
  // float32, duplicated x rows: the y tile as loop-invariant pairs of ADJACENT COLUMNS (y_c[k], y_{c+1}[k])
  [[maybe_unused]] u64 ypk[DUP ? (COLS / 2 > 0 ? COLS / 2 : 1) : 1][DUP ? DP : 1];
  if constexpr (DUP) {
#pragma unroll
    for (int c2 = 0; c2 < COLS / 2; ++c2)
#pragma unroll
      for (int k = 0; k < DP; ++k)
        ypk[c2][k] = pk2((float)vget(yreg[YREG ? 2 * c2 : 0][YREG ? k / V : 0], k % V), (float)vget(yreg[YREG ? 2 * c2 + 1 : 0][YREG ? k / V : 0], k % V));
  }
  double tab = 0.0;
  if (OP == OP_RBF && sizeof(T) == 8) tab = exp2((double)lane * (1.0 / 32.0));
  if constexpr (SkewExpTable<T, OP>::value) {
    // 2^(j/2048), j = 0..2047, behind the x ring and the y tiles; the address travels in the low word of `tab`
    double* etab = reinterpret_cast<double*>(xring + NB * xslot + (size_t)NW * skew_ywarp_bytes<T, DP, COLS>());
    constexpr int NE = BIG != 0 ? 16384 : 2048;
    for (int e = threadIdx.x; e < NE; e += NW * 32) etab[e] = exp2((double)e * (1.0 / NE));
    __syncthreads();
    tab = __hiloint2double(0, (int)smem_u32(etab));
  }

  // ---- state ------------------------------------------------------------------------
  // column sums C_l[q] of levels 0..MLEV-2; of the top level only the total is ever used: partial sums
  T C[MLEV - 1][COLS];
#pragma unroll
  for (int l = 0; l < MLEV - 1; ++l)
#pragma unroll
    for (int c = 0; c < COLS; ++c) C[l][c] = T(0);
  constexpr int NT = sizeof(T) == 8 ? GPSIG_SKEW_NT_F64 : COLS;            // float32: one per column (a 2-deep FFMA chain would stall on latency)
  T Ct[NT];
#pragma unroll
  for (int k = 0; k < NT; ++k) Ct[k] = T(0);
  T Mp[2][COLS];                                   // last base-kernel row, ping-pong over the step parity
#pragma unroll
  for (int c = 0; c < COLS; ++c) Mp[0][c] = Mp[1][c] = T(0);
  T send[R][MLEV], sendU[R];
#pragma unroll
  for (int rr = 0; rr < R; ++rr) {
    sendU[rr] = T(0);
#pragma unroll
    for (int l = 0; l < MLEV; ++l) send[rr][l] = T(0);
  }

  constexpr bool SADDR = sizeof(T) == 8 ? (GPSIG_SKEW_SADDR_F64 != 0) : (GPSIG_SKEW_SADDR_F32 != 0);
  uint32_t xring_s = smem_u32(xring);
  if constexpr (SADDR) asm volatile("" : "+r"(xring_s));      // opaque: kept in a register, not rebuilt
  const int SPS = (L + R - 1) / R;                 // steps per pair (>= 32, see skew_eligible)
  const int nsteps = npairs * SPS + 32;            // + finalisation step + pipeline drain
  int kk = 0, sseg = 0, ubuf = 0, ubuf_prev = NB - 1;   // warp-uniform: step within pair, pair and ring slot of lane 0
  const T p0 = T(P.p0), p1 = T(P.p1);
  const size_t rows_out = (size_t)(P.row_end - P.row_begin);

  auto step = [&](auto PAR, const int s) {
    constexpr int par = decltype(PAR)::value;
    // ---- ring: wait for the pair lane 0 enters now; hand back the slot lane 31 just left ----
#if GPSIG_SKEW_RING1
    // ONE warp-uniform test per step: at the first step of a pair the warp waits for that pair's rows and hands back
    // the slot of the pair before the previous one (lane 31 left it 69 steps ago); thread 0 refills it with the NEXT
    // pair, which then has a whole pair's time to arrive.
    if (kk == 0) {
      if (sseg < npairs) mbar_wait(&full_bar[ubuf], (sseg / NB) & 1);
      if (sseg >= 2) {
        const int dseg = sseg - 2;
        const int fslot = ubuf + 1 == NB ? 0 : ubuf + 1;        // slot of pair sseg - 2 == slot of pair sseg + 1 (NB = 3)
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[fslot]);
        const int nxt = dseg + NB;
        if (threadIdx.x == 0 && nxt < npairs) {
          mbar_wait(&empty_bar[fslot], (dseg / NB) & 1);
          mbar_expect_tx(&full_bar[fslot], (uint32_t)xbytes);
          tma_load_1d(xring + fslot * xslot, xg + (size_t)(i_lo + nxt) * P.xblk, (uint32_t)xbytes, &full_bar[fslot]);
        }
      }
    }
#else
    if (kk == 0 && sseg < npairs) mbar_wait(&full_bar[ubuf], (sseg / NB) & 1);
    if (kk == 31 && sseg >= 1) {
      const int dseg = sseg - 1;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[ubuf_prev]);
      const int nxt = dseg + NB;                   // the pair that reuses this slot
      if (threadIdx.x == 0 && nxt < npairs) {
        mbar_wait(&empty_bar[ubuf_prev], (dseg / NB) & 1);
        mbar_expect_tx(&full_bar[ubuf_prev], (uint32_t)xbytes);
        tma_load_1d(xring + ubuf_prev * xslot, xg + (size_t)(i_lo + nxt) * P.xblk, (uint32_t)xbytes, &full_bar[ubuf_prev]);
      }
    }
#endif
    // ---- chain values from the lane to the left (computed one step ago) -----------
    // (tensor-memory variant: the shuffles are issued inside the contraction below)
    T recv[R][MLEV], recvU[R];
    if constexpr (!YTM) {
#pragma unroll
      for (int rr = 0; rr < R; ++rr) {
        recvU[rr] = __shfl_up_sync(0xffffffffu, sendU[rr], 1);
#pragma unroll
        for (int l = 0; l < MLEV; ++l) {
          if (l < MLEV - 1 || rr == 0) recv[rr][l] = __shfl_up_sync(0xffffffffu, send[rr][l], 1);
          else recv[rr][l] = T(0);
          if (L0D == 0 && lane == 0) recv[rr][l] = T(0);
        }
      }
    }
    // ---- this lane's place in the stream: one step behind the lane to its left ----
    const bool behind = kk < lane;                            // still in the previous pair
    const int seg = sseg - (behind ? 1 : 0);
    const int rstep = kk - lane + (behind ? SPS : 0);
    const bool started = s >= lane;
    const bool live = started && seg < npairs;                // rows of a real pair (else: fill / finalisation / drain)
    const T* xb = reinterpret_cast<const T*>(xring + (live ? (behind ? ubuf_prev : ubuf) : 0) * xslot);
    const T* xsq = xb + (size_t)L * RS;
    const uint32_t xb_s = xring_s + (uint32_t)((live ? (behind ? ubuf_prev : ubuf) : 0) * xslot);
    auto ldx = [&](int row, int off) -> V16 {          // 16 bytes of x row `row` from element `off`
      if constexpr (SADDR) return lds16(xb_s + (uint32_t)((row * RS + off) * (int)sizeof(T)), T(0));
      else return *reinterpret_cast<const V16*>(xb + (size_t)row * RS + off);
    };
    const int r0 = live ? rstep * R : 0;

    // ---- base kernel on R rows x COLS columns -------------------------------------
    T Mc[R][COLS];
    // tensor-memory variant: the tile arrives in groups of GS columns x 4 dimensions; group g+1 is
    // requested before the FMAs of group g, its wait comes after them (wait::ld covers all loads in flight)
    constexpr int GS = sizeof(T) == 8 ? GPSIG_SKEW_GS_F64 : GPSIG_SKEW_GS_F32;
    constexpr int NG = (COLS + GS - 1) / GS;          // groups per 4-dimension chunk
    constexpr int NGT = NG * KC;
    constexpr int NCH = R + R * (MLEV - 1) + 1;       // chain values per step
    constexpr int CPG = (NCH + NGT - 1) / NGT;        // ... shuffled per group
    [[maybe_unused]] TmemQuad yq[2][GS];
    [[maybe_unused]] auto tm_issue = [&](auto G) {
      constexpr int g = decltype(G)::value;
      constexpr int kc = g / NG, c0 = (g % NG) * GS;
#pragma unroll
      for (int c = c0; c < c0 + GS && c < COLS; ++c) tmem_ld4(yq[g & 1][c - c0], ta + (uint32_t)((kc * COLS + c) * 8));
    };
    [[maybe_unused]] auto chain_slot = [&](auto K) {
      constexpr int k = decltype(K)::value;
      if constexpr (k < R) recvU[k] = shfl_up1_pinned(sendU[k]);
      else if constexpr (k < R + R * (MLEV - 1)) {
        constexpr int rr = (k - R) / (MLEV - 1), l = (k - R) % (MLEV - 1);
        recv[rr][l] = shfl_up1_pinned(send[rr][l]);
      } else if constexpr (k == NCH - 1) recv[0][MLEV - 1] = shfl_up1_pinned(send[0][MLEV - 1]);
    };
    [[maybe_unused]] TmemHex yh[2];
    if constexpr (TM16) tmem_ld8d(yh[0], ta);
    else if constexpr (YTM) tm_issue(std::integral_constant<int, 0>{});
    {
      T xc[R];
      T acc[R][COLS];
#pragma unroll
      for (int rr = 0; rr < R; ++rr) {
        const int row = min(r0 + rr, L - 1);
        if constexpr (SADDR) xc[rr] = Op::row_const(lds1(xb_s + (uint32_t)((L * RS + row) * (int)sizeof(T)), T(0)));
        else xc[rr] = Op::row_const(xsq[row]);
#pragma unroll
        for (int c = 0; c < COLS; ++c) acc[rr][c] = Op::init(xc[rr], yc[c]);
      }
      if constexpr (TM16) {
        // one 16-word load per column; the rows' coordinates stay in registers for the whole step
        V16 xr[R][KV];
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
          for (int kv = 0; kv < KV; ++kv)
            xr[rr][kv] = ldx(min(r0 + rr, L - 1), kv * V);
        constexpr int CPC = (NCH + COLS - 1) / COLS;       // chain values shuffled per column
        static_for<COLS>([&](auto CI) {
          constexpr int c = decltype(CI)::value;
          tmem_ld_wait();
          if (GPSIG_TMEM_TIE) tmem_tie(yh[c & 1]);
          if constexpr (c + 1 < COLS) tmem_ld8d(yh[(c + 1) & 1], ta + (uint32_t)((c + 1) * 16));
          static_for<CPC>([&](auto J) {
            constexpr int k = c * CPC + decltype(J)::value;
            if constexpr (k < NCH) chain_slot(std::integral_constant<int, k>{});
          });
          const TmemHex& cur = yh[c & 1];
#pragma unroll
          for (int rr = 0; rr < R; ++rr) {
            T a = acc[rr][c];
#pragma unroll
            for (int e = 0; e < DP; ++e)
              a = gfma<T>(vget(xr[rr][e / V], e % V), T(__hiloint2double((int)cur.w[2 * e + 1], (int)cur.w[2 * e])), a);
            acc[rr][c] = a;
          }
        });
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
          for (int l = 0; l < MLEV; ++l) {
            if (!(l < MLEV - 1 || rr == 0)) recv[rr][l] = T(0);
            else if constexpr (L0D == 0) zero_on_lane0(recv[rr][l], lane);
          }
      } else if constexpr (YTM) {
        V16 xa[R], xq[R];
        static_for<NGT>([&](auto G) {
          constexpr int g = decltype(G)::value;
          constexpr int kc = g / NG, c0 = (g % NG) * GS;
          if constexpr (g % NG == 0) {
#pragma unroll
            for (int rr = 0; rr < R; ++rr) {
              xa[rr] = ldx(min(r0 + rr, L - 1), kc * CE);
              xq[rr] = ldx(min(r0 + rr, L - 1), kc * CE + V);
            }
          }
          tmem_ld_wait();
#pragma unroll
          for (int c = c0; c < c0 + GS && c < COLS; ++c) if (GPSIG_TMEM_TIE) tmem_tie(yq[g & 1][c - c0]);
          if constexpr (g + 1 < NGT) tm_issue(std::integral_constant<int, g + 1>{});
          static_for<CPG>([&](auto J) {
            constexpr int k = g * CPG + decltype(J)::value;
            if constexpr (k < NCH) chain_slot(std::integral_constant<int, k>{});
          });
#pragma unroll
          for (int c = c0; c < c0 + GS && c < COLS; ++c) {
            const TmemQuad& cur = yq[g & 1][c - c0];
#pragma unroll
            for (int rr = 0; rr < R; ++rr) {
              T a = acc[rr][c];
#pragma unroll
              for (int e = 0; e < V; ++e) a = gfma<T>(vget(xa[rr], e), tmem_get<T>(cur, e), a);
#pragma unroll
              for (int e = 0; e < V; ++e) a = gfma<T>(vget(xq[rr], e), tmem_get<T>(cur, V + e), a);
              acc[rr][c] = a;
            }
          }
        });
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
          for (int l = 0; l < MLEV; ++l) {
            if (!(l < MLEV - 1 || rr == 0)) recv[rr][l] = T(0);
            else if constexpr (L0D == 0) zero_on_lane0(recv[rr][l], lane);
          }
      } else if constexpr (DUP) {
        // packed pairs: one FFMA2 advances two adjacent columns; x arrives as ready-made (x_k, x_k) pairs
        u64 accp[R][COLS / 2 > 0 ? COLS / 2 : 1];
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
          for (int c2 = 0; c2 < COLS / 2; ++c2) accp[rr][c2] = pk2((float)acc[rr][2 * c2], (float)acc[rr][2 * c2 + 1]);
#pragma unroll
        for (int j = 0; j < DP / 2; ++j) {
          ulonglong2 xd[R];
#pragma unroll
          for (int rr = 0; rr < R; ++rr)
            xd[rr] = *reinterpret_cast<const ulonglong2*>(xb + (size_t)min(r0 + rr, L - 1) * RS + 4 * j);
#pragma unroll
          for (int rr = 0; rr < R; ++rr) {
#pragma unroll
            for (int c2 = 0; c2 < COLS / 2; ++c2) {
              asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(accp[rr][c2]) : "l"(xd[rr].x), "l"(ypk[c2][2 * j]));
              asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(accp[rr][c2]) : "l"(xd[rr].y), "l"(ypk[c2][2 * j + 1]));
            }
            if constexpr ((COLS & 1) != 0) {
              float x0, x1, dmy;
              unpk2(xd[rr].x, x0, dmy);
              unpk2(xd[rr].y, x1, dmy);
              const V16 yv = yreg[YREG ? COLS - 1 : 0][YREG ? (2 * j) / V : 0];
              acc[rr][COLS - 1] = gfma<T>(T(x0), vget(yv, (2 * j) % V), acc[rr][COLS - 1]);
              acc[rr][COLS - 1] = gfma<T>(T(x1), vget(yv, (2 * j + 1) % V), acc[rr][COLS - 1]);
            }
          }
        }
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
#pragma unroll
          for (int c2 = 0; c2 < COLS / 2; ++c2) {
            float lo, hi;
            unpk2(accp[rr][c2], lo, hi);
            acc[rr][2 * c2] = T(lo); acc[rr][2 * c2 + 1] = T(hi);
          }
      } else {
#pragma unroll
      for (int kv = 0; kv < KV; ++kv) {
        V16 xv[R];
#pragma unroll
        for (int rr = 0; rr < R; ++rr)
          xv[rr] = ldx(min(r0 + rr, L - 1), kv * V);
        if constexpr (PairOps<T>::kPacked && YREG) {
          // adjacent columns as packed pairs (the y pairs are loop invariant: packed once by the compiler)
#pragma unroll
          for (int c = 0; c + 1 < COLS; c += 2) {
            const V16 ya = yreg[YREG ? c : 0][YREG ? kv : 0], yb = yreg[YREG ? c + 1 : 0][YREG ? kv : 0];
#pragma unroll
            for (int rr = 0; rr < R; ++rr)
#pragma unroll
              for (int e = 0; e < V; ++e)
                PairOps<T>::fma(acc[rr][c], acc[rr][c + 1], vget(xv[rr], e), vget(xv[rr], e), vget(ya, e), vget(yb, e));
          }
          if (COLS & 1) {
            const V16 yv = yreg[YREG ? COLS - 1 : 0][YREG ? kv : 0];
#pragma unroll
            for (int rr = 0; rr < R; ++rr)
#pragma unroll
              for (int e = 0; e < V; ++e) acc[rr][COLS - 1] = gfma<T>(vget(xv[rr], e), vget(yv, e), acc[rr][COLS - 1]);
          }
        } else {
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          V16 yv;
          if (YREG) yv = yreg[YREG ? c : 0][YREG ? kv : 0];
          else yv = *reinterpret_cast<const V16*>(ysm + ((size_t)(c * KV + kv) * 32 + lane) * V);
#pragma unroll
          for (int rr = 0; rr < R; ++rr)
#pragma unroll
            for (int e = 0; e < V; ++e) acc[rr][c] = gfma<T>(vget(xv[rr], e), vget(yv, e), acc[rr][c]);
        }
        }
      }
      }
#pragma unroll
      for (int rr = 0; rr < R; ++rr)
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          const T v = SkewFinish<T, OP>::template apply<BIG>(acc[rr][c], xc[rr], yc[c], P.base, p0, p1, tab);
          if (rr == R - 1) Mp[par ^ 1][c] = v; else Mc[rr][c] = v;
        }
    }

    // ---- recursion, row by row -----------------------------------------------------
#pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      T D[COLS];
      if (diff) {
        // 2-D increment (signature_algs.py:26).  No masks: row 0 of a pair is consumed by the
        // first-row branch below, padded rows/columns repeat their neighbour (increment 0).
        T U[COLS];
        auto curv = [&](int c) -> T { return (rr == R - 1) ? Mp[par ^ 1][c] : Mc[rr][c]; };
        auto prvv = [&](int c) -> T { return (rr == 0) ? Mp[par][c] : Mc[rr > 0 ? rr - 1 : 0][c]; };
#pragma unroll
        for (int c = 0; c + 1 < COLS; c += 2) PairOps<T>::sub(U[c], U[c + 1], curv(c), curv(c + 1), prvv(c), prvv(c + 1));
        if (COLS & 1) U[COLS - 1] = curv(COLS - 1) - prvv(COLS - 1);
        const T left = lane == 0 ? U[0] : recvU[rr];          // column -1 does not exist: D = 0 there
#pragma unroll
        for (int c = 0; c < COLS; ++c) D[c] = U[c] - (c > 0 ? U[c - 1] : left);
        sendU[rr] = U[COLS - 1];
      } else {
        const bool rowok = live && (r0 + rr < L);
#pragma unroll
        for (int c = 0; c < COLS; ++c) {
          const T cur = (rr == R - 1) ? Mp[par ^ 1][c] : Mc[rr][c];
          D[c] = (rowok && q0 + c < L2 && q0 + c >= 0) ? cur : T(0);
        }
      }

      constexpr bool FRF = sizeof(T) == 8 ? (GPSIG_SKEW_FRF_F64 != 0) : (GPSIG_SKEW_FRF_F32 != 0);
      const bool first = rr == 0 && started && kk == lane;          // first row of a pair (one lane per step)
      if constexpr (!FRF) {
      if (first) {
        // first row of a pair: the chain carries the level totals of the pair just finished
        T tot[MLEV];
#pragma unroll
        for (int l = 0; l < MLEV; ++l) {
          T e = recv[0][l];
          if (l < MLEV - 1) {
#pragma unroll
            for (int c = 0; c < COLS; ++c) e += C[l < MLEV - 1 ? l : 0][c];
          } else {
#pragma unroll
            for (int k = 0; k < NT; ++k) e += Ct[k];
          }
          tot[l] = e;
          send[0][l] = e;
        }
        if (lane == 31 && seg > 0 && active) {
          // ---- epilogue: normalise, scale, store (kernels.py:431-477) ----
          const int i = i_lo + seg - 1;
          const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
          T sum = T(0);
#pragma unroll
          for (int lvl = 0; lvl <= MLEV; ++lvl) {
            if (lvl <= P.m) {
              T v = lvl == 0 ? T(1) : tot[lvl > 0 ? lvl - 1 : 0];
              if (P.symmetric && i == j) v += T(P.sym_jitter);
              if (P.norm_x) {
                const T a = gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + i]);
                const T b = gsqrt<T>(static_cast<const T*>(P.norm_x2)[(size_t)lvl * P.N2 + j]);
                v = v / (a * b);
              }
              if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
              if (P.return_levels) static_cast<T*>(P.out)[(size_t)lvl * rows_out * P.N2 + rc] = v;
              else sum += v;
            }
          }
          if (!P.return_levels) static_cast<T*>(P.out)[rc] = sum;
        }
#pragma unroll
        for (int l = 1; l < MLEV - 1; ++l)
#pragma unroll
          for (int c = 0; c < COLS; ++c) C[l][c] = T(0);
#pragma unroll
        for (int k = 0; k < NT; ++k) Ct[k] = T(0);
#pragma unroll
        for (int c = 0; c < COLS; ++c) C[0][c] = diff ? T(0) : D[c];
      } else {
        // C_{l+1}[q] += D[p, q] * sum_{q' < q} C_l[q'], highest level first (C_l is read before row p updates it)
#pragma unroll
        for (int l = MLEV - 2; l >= 0; --l) {
          T e = recv[rr][l];
          if constexpr (PairOps<T>::kPacked && NT == COLS) {
            T t[COLS];
#pragma unroll
            for (int c = 0; c < COLS; ++c) { t[c] = e; e += C[l][c]; }
            const int ln = l + 1 < MLEV - 1 ? l + 1 : 0;
#pragma unroll
            for (int c = 0; c + 1 < COLS; c += 2) {
              if (l == MLEV - 2) PairOps<T>::fma(Ct[c], Ct[c + 1 < NT ? c + 1 : 0], D[c], D[c + 1], t[c], t[c + 1]);
              else PairOps<T>::fma(C[ln][c], C[ln][c + 1], D[c], D[c + 1], t[c], t[c + 1]);
            }
            if (COLS & 1) {
              if (l == MLEV - 2) Ct[(COLS - 1) % NT] = gfma<T>(D[COLS - 1], t[COLS - 1], Ct[(COLS - 1) % NT]);
              else C[ln][COLS - 1] = gfma<T>(D[COLS - 1], t[COLS - 1], C[ln][COLS - 1]);
            }
          } else {
#pragma unroll
          for (int c = 0; c < COLS; ++c) {
            const T t = e;
            e += C[l][c];
            if (l == MLEV - 2) Ct[c % NT] = gfma<T>(D[c], t, Ct[c % NT]);
            else C[l + 1 < MLEV - 1 ? l + 1 : 0][c] = gfma<T>(D[c], t, C[l + 1 < MLEV - 1 ? l + 1 : 0][c]);
          }
          }
          send[rr][l] = e;
        }
#pragma unroll
        for (int c = 0; c + 1 < COLS; c += 2) PairOps<T>::add(C[0][c], C[0][c + 1], C[0][c], C[0][c + 1], D[c], D[c + 1]);
        if (COLS & 1) C[0][COLS - 1] += D[COLS - 1];
      }
      } else {
        // The first row of a pair takes the common path as well: every level reads C_l (and the chain value from the
        // left, which at this row is the left lanes' TOTAL of the pair just finished) before it writes C_{l+1}, so
        // send[0][l] comes out as the level total, whatever D holds; what the row's D (new pair's row 0 minus the old
        // pair's last row) adds to C and Ct afterwards is wiped by the reset below.  The top level's total needs Ct
        // from before the row: summed here, once per step.
        T ctsum = T(0);
        if (rr == 0) {
#pragma unroll
          for (int k = 0; k < NT; ++k) ctsum += Ct[k];
        }
        // C_{l+1}[q] += D[p, q] * sum_{q' < q} C_l[q'], highest level first (C_l is read before row p updates it)
#pragma unroll
        for (int l = MLEV - 2; l >= 0; --l) {
          T e = recv[rr][l];
          if constexpr (PairOps<T>::kPacked && NT == COLS) {
            T t[COLS];
#pragma unroll
            for (int c = 0; c < COLS; ++c) { t[c] = e; e += C[l][c]; }
            const int ln = l + 1 < MLEV - 1 ? l + 1 : 0;
#pragma unroll
            for (int c = 0; c + 1 < COLS; c += 2) {
              if (l == MLEV - 2) PairOps<T>::fma(Ct[c], Ct[c + 1 < NT ? c + 1 : 0], D[c], D[c + 1], t[c], t[c + 1]);
              else PairOps<T>::fma(C[ln][c], C[ln][c + 1], D[c], D[c + 1], t[c], t[c + 1]);
            }
            if (COLS & 1) {
              if (l == MLEV - 2) Ct[(COLS - 1) % NT] = gfma<T>(D[COLS - 1], t[COLS - 1], Ct[(COLS - 1) % NT]);
              else C[ln][COLS - 1] = gfma<T>(D[COLS - 1], t[COLS - 1], C[ln][COLS - 1]);
            }
          } else {
#pragma unroll
          for (int c = 0; c < COLS; ++c) {
            const T t = e;
            e += C[l][c];
            if (l == MLEV - 2) Ct[c % NT] = gfma<T>(D[c], t, Ct[c % NT]);
            else C[l + 1 < MLEV - 1 ? l + 1 : 0][c] = gfma<T>(D[c], t, C[l + 1 < MLEV - 1 ? l + 1 : 0][c]);
          }
          }
          send[rr][l] = e;
        }
#pragma unroll
        for (int c = 0; c + 1 < COLS; c += 2) PairOps<T>::add(C[0][c], C[0][c + 1], C[0][c], C[0][c + 1], D[c], D[c + 1]);
        if (COLS & 1) C[0][COLS - 1] += D[COLS - 1];
        if (first) {
          const T toptot = recv[0][MLEV - 1] + ctsum;
          send[0][MLEV - 1] = toptot;
          if (lane == 31 && seg > 0 && active) {
            // ---- epilogue: normalise, scale, store (kernels.py:431-477) ----
            const int i = i_lo + seg - 1;
            const size_t rc = (size_t)(i - P.row_begin) * P.N2 + j;
            T sum = T(0);
#pragma unroll
            for (int lvl = 0; lvl <= MLEV; ++lvl) {
              if (lvl <= P.m) {
                T v = lvl == 0 ? T(1) : (lvl == MLEV ? toptot : send[0][lvl > 0 && lvl < MLEV ? lvl - 1 : 0]);
                if (P.symmetric && i == j) v += T(P.sym_jitter);
                if (P.norm_x) {
                  const T a = gsqrt<T>(static_cast<const T*>(P.norm_x)[(size_t)lvl * P.N + i]);
                  const T b = gsqrt<T>(static_cast<const T*>(P.norm_x2)[(size_t)lvl * P.N2 + j]);
                  v = v / (a * b);
                }
                if (P.level_scale) v *= static_cast<const T*>(P.level_scale)[lvl];
                if (P.return_levels) static_cast<T*>(P.out)[(size_t)lvl * rows_out * P.N2 + rc] = v;
                else sum += v;
              }
            }
            if (!P.return_levels) static_cast<T*>(P.out)[rc] = sum;
          }
#pragma unroll
          for (int l = 0; l < MLEV - 1; ++l)
#pragma unroll
            for (int c = 0; c < COLS; ++c) C[l][c] = T(0);
#pragma unroll
          for (int k = 0; k < NT; ++k) Ct[k] = T(0);
        }
      }
    }
    if (++kk == SPS) { kk = 0; ++sseg; ubuf_prev = ubuf; if (++ubuf == NB) ubuf = 0; }
  };

  for (int s = 0; s < nsteps; s += 2) {
    step(std::integral_constant<int, 0>{}, s);
    step(std::integral_constant<int, 1>{}, s + 1);
  }
  if (YTM) tmem_free(tbase, TMCOLS);
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
// Shapes this kernel takes (the rest runs trunc_gram_kernel): one column pass, rows long
// enough (>= 32 steps per pair) that a lane-skewed warp touches at most two consecutive pairs.

inline bool skew_eligible(int L, int L2, int elem) { return L >= 32 * (elem == 8 ? GPSIG_SKEW_R_F64 : GPSIG_SKEW_R_F32) + 1 && L2 <= 256; }

template <typename T, int DP, int COLS, int MLEV, int OP, int BIG = 0, int L0D = 0>
int launch_skew_inst(TruncParams P, cudaStream_t stream) {
  if constexpr (L0D == 0 && (sizeof(T) == 8 ? GPSIG_SKEW_L0D_F64 != 0 : GPSIG_SKEW_L0D_F32 != 0)) {
    // the columns fit on lanes 1..31: lane 0 becomes the dummy lane
    if (P.L2 <= 31 * COLS) return launch_skew_inst<T, DP, COLS, MLEV, OP, BIG, 1>(P, stream);
  }
  const int RS = P.xrs;
  const size_t xslot = align_up_dev(P.xblk * sizeof(T));
  const size_t smem0 = 128 + kSkewBufs * xslot + (size_t)SkewNW<T, DP, COLS>::value * skew_ywarp_bytes<T, DP, COLS>();
  if constexpr (BIG == 0 && GPSIG_SKEW_EXP2_T16K != 0 && SkewExpTable<T, OP>::value) {
    // the large exp2 table whenever it fits next to the x ring
    if (smem0 + kSkewExpTableBytesBig <= kMaxSmem) return launch_skew_inst<T, DP, COLS, MLEV, OP, 1, L0D>(P, stream);
  }
  const size_t smem = smem0 + (SkewExpTable<T, OP>::value ? (BIG != 0 ? kSkewExpTableBytesBig : kSkewExpTableBytes) : 0);
  if (P.dry_run) return smem > kMaxSmem ? GPSIG_ERR_UNSUPPORTED : GPSIG_OK;       // the caller then takes trunc_gram_kernel
  if (smem > kMaxSmem)
    GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "sequence too long for the on-chip plan: needs %zu bytes of shared memory (L=%d)", smem, P.L);
  auto kern = trunc_skew_kernel<T, DP, COLS, MLEV, OP, BIG, L0D>;
  GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid;
  constexpr int NW = SkewNW<T, DP, COLS>::value;
  grid.x = (P.N2 + NW - 1) / NW;
  const int rows = P.row_end - P.row_begin;
  // >= ~24 waves of CTAs over the 148 SMs, but >= 16 pairs per warp so the 31-step pipeline fill stays < 1 %
  int want_y = (24 * 148 + (int)grid.x - 1) / (int)grid.x;
  if (P.tri) want_y *= 2;                         // half of the CTAs exit at once; smaller chunks also waste less along the diagonal
  int ichunk = (rows + want_y - 1) / want_y;
  if (ichunk < 16) ichunk = 16;
  if (ichunk > rows) ichunk = rows;
  if (!P.tri) {
    // All CTAs take the same time (one CTA per SM, ichunk pairs per warp), so the launch runs in whole waves: pick the
    // row split y whose last wave is full, i.e. minimise waves(y) x (pairs per chunk + pipeline fill) around the
    // value above (c2, 250 x 15 CTAs = 25.3 waves of 134 pairs -> 250 x 13 = 21.96 waves of 154: 2.8 % less).
    const int sms = device_sm_count();
    const int ymax = rows / 16 > 0 ? rows / 16 : 1;
    double best = 1e300;
    int best_y = 0;
    for (int y = 1; y <= ymax && y <= 4 * want_y + 8; ++y) {
      const int ch = (rows + y - 1) / y;
      const int ny = (rows + ch - 1) / ch;
      const long ctas = (long)grid.x * ny;
      const double waves = (double)((ctas + sms - 1) / sms);
      const double cost = waves * ((double)ch + 0.32) * (ctas < sms ? (double)sms / ctas : 1.0);   // fewer CTAs than SMs: idle SMs
      if (cost < best * 0.99) { best = cost; best_y = y; }      // a finer split only for a real gain: every row chunk re-streams its x rows past L2
    }
    if (best_y > 0) ichunk = (rows + best_y - 1) / best_y;
  }
  P.ichunk = ichunk;
  grid.y = (rows + ichunk - 1) / ichunk;
  prof_begin(stream);
  kern<<<grid, NW * 32, smem, stream>>>(P, RS);
  prof_end(stream);
  GPSIG_LAUNCH_CHECK("trunc_skew_kernel");
  return GPSIG_OK;
}

#ifdef GPSIG_FAST_BUILD
#define GPSIG_SKEW_COLS(X) X(2) X(7)
#define GPSIG_SKEW_MLEV(X) X(4) X(5)
#else
#define GPSIG_SKEW_COLS(X) X(2) X(4) X(6) X(7) X(8)
#define GPSIG_SKEW_MLEV(X) X(4) X(5) X(8)
#endif

template <typename T, int OP, int DP, int COLS>
int launch_skew_mlev(const TruncParams& P, cudaStream_t stream) {
#define GPSIG_CASE(M_) if (P.m <= M_) return launch_skew_inst<T, DP, COLS, M_, OP>(P, stream);
  GPSIG_SKEW_MLEV(GPSIG_CASE)
#undef GPSIG_CASE
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported by this build", P.m);
}

// host mirror of the dispatch below: does the instantiation that takes (d, L2, elem) read the duplicated layout?
inline bool skew_uses_dup(int d, int L2, int elem, int m, int base) {
  if (!GPSIG_SKEW_DUP_F32 || elem != 4 || (base != GPSIG_BASE_LINEAR && base != GPSIG_BASE_RBF)) return false;
  const int need = (L2 + 31) / 32, dp = padded_dim(d);
  int cols = 0, mlev = 0;
#define GPSIG_CASE(C_) if (cols == 0 && need <= C_) cols = C_;
  GPSIG_SKEW_COLS(GPSIG_CASE)
#undef GPSIG_CASE
#define GPSIG_CASE(M_) if (mlev == 0 && m <= M_) mlev = M_;
  GPSIG_SKEW_MLEV(GPSIG_CASE)
#undef GPSIG_CASE
  return cols > 0 && cols * dp * 4 <= 256 && !(GPSIG_SKEW_TMEM && GPSIG_SKEW_TMEM_F32) && cols <= kSkewDupMaxCols && mlev > 0 &&
         mlev <= kSkewDupMaxLev;
}

template <typename T, int OP, int DP>
int launch_trunc_skew(const TruncParams& P, cudaStream_t stream) {
  const int need = (P.L2 + 31) / 32;
#define GPSIG_CASE(C_) if (need <= C_) return launch_skew_mlev<T, OP, DP, C_>(P, stream);
  GPSIG_SKEW_COLS(GPSIG_CASE)
#undef GPSIG_CASE
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "no systolic kernel shape for L2=%d", P.L2);
}

}  // namespace gpsig
