// float32 LIFT on the 5th-generation tensor cores (north-star subsystem 1): the base-kernel tile
// M[p, q] = kappa(x_i[p], y_j[q]) (gpsig/kernels.py:225-231, 946-957) as tcgen05.mma kind::tf32 -- see lifted_kernels.cuh.
//
// One accumulator tile = one sequence pair (128 rows of x_i) x (32*COLS columns of y_j), K = the state dimension.
// float32 accuracy from TF32 tensor cores by the 3-term split  x = hi + lo  (hi = x rounded to TF32, lo = x - hi rounded
// to TF32):  x.y ~= hi.hi + hi.lo + lo.hi  -- three MMAs per K step on two stored operand halves (22 mantissa bits,
// error on par with an fp32 FMA chain; tests/test_gpu_large_d.py holds it to the float32 tolerance 1e-5).
//
// Operands are pre-split and pre-tiled by lift_pack_mma_kernel into the canonical K-major, no-swizzle UMMA layout
// (core matrix = 8 rows x 16 bytes, contiguous), K stages outermost, so one K stage of an operand is ONE contiguous
// block = one 1-D TMA bulk copy (cp.async.bulk, SASS UBLKCP) into a 4-deep shared-memory ring (full/empty mbarriers).
// Warp roles (64 + 128 kMmaEpq threads, one CTA per SM, persistent over streams of pairs):
//   warp 0   TMA producer          warp 1   MMA issuer (one elected lane; tcgen05.mma, tcgen05.commit -> mbarriers)
//   the others: epilogue: tcgen05.ld their 32 accumulator rows (tensor memory, double buffered: the MMAs of tile n+1
//            overlap the epilogue of tile n; two warps per lane quarter split the columns), base-kernel epilogue, store
//            in the skewed order of the recursion kernel.
#pragma once
#include "lifted_kernels.cuh"
#include "trunc_skew_kernel.cuh"      // tensor-memory / mbarrier helpers

namespace gpsig {

constexpr int kMmaKC = 16;            // K elements per pipeline stage (2 MMA K-steps of 8)
constexpr int kMmaStages = 4;
#ifndef GPSIG_MMA_EPQ
#define GPSIG_MMA_EPQ 2     // measured: 4 per quarter is slower (51.5 against 43.8 ms at N=1000 L=100 d=126)
#endif
constexpr int kMmaEpq = GPSIG_MMA_EPQ;                 // epilogue warps per lane quarter (they split the column groups)
constexpr int kMmaThreads = 64 + 128 * kMmaEpq;       // TMA warp, MMA warp, 4 * kMmaEpq epilogue warps

struct LiftMmaOperands {
  const void* xop; const void* yop;      // pre-split operands (floats), or NULL: the CUDA-core lift runs instead
  size_t xop_blk, yop_blk;               // elements per sequence: MT * 2 * KP * 128  /  2 * KP * 32 * COLS
  int KP;                                // padded K (multiple of kMmaKC)
};

__host__ __device__ inline int lift_mma_kp(int d) { return (d + kMmaKC - 1) / kMmaKC * kMmaKC; }
__host__ __device__ inline size_t lift_mma_a_elems(int L, int d) { return (size_t)((L + 127) / 128) * 2 * lift_mma_kp(d) * 128; }
__host__ __device__ inline size_t lift_mma_b_elems(int L2, int d) { return (size_t)2 * lift_mma_kp(d) * 32 * lift_cols(L2); }

__device__ __forceinline__ float tf32_rna(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return __uint_as_float(r);
}

// exp(x) for the RBF epilogue: MUFU.EX2 on x * log2(e) with the product carried in two terms, so that the error does
// not grow with |x| the way __expf's does (|x| ~ |x_p|^2 reaches 10-30 at large d): ~2 ulp, 6 instructions
__device__ __forceinline__ float exp_2term(float x) {
  const float t = x * 1.4426950216293335f;                       // log2(e) rounded to float
  float r = fmaf(x, 1.4426950216293335f, -t);                    // exact residual of the product
  r = fmaf(x, 1.9259629911e-8f, r);                              // + x * (log2(e) - float(log2(e)))
  const float e = exp2_approx(t);
  return fmaf(e, r * 0.6931471805599453f, e);
}

// One thread per (sequence, tile row, 4 consecutive k).  Block of one sequence:
//   [row tile (A only)][K stage][half: hi, lo][KC/4 chunks][R rows][4 floats],   R = 128 (A) or 32*COLS (B)
// A rows past L are zero, B rows past L2 repeat the last point (the padded columns of the tile repeat the last one).
static __global__ void lift_pack_mma_kernel(const float* __restrict__ packed, size_t blk, int DP, int N, int L, int d,
                                     int R, int ntiles, int KP, int repeat_last, float* __restrict__ dst, size_t dst_blk) {
  const int KQ = KP / 4;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t per = (size_t)ntiles * R * KQ;
  if (idx >= (size_t)N * per) return;
  const int n = (int)(idx / per);
  size_t rem = idx - (size_t)n * per;
  const int kq = (int)(rem % KQ); rem /= KQ;
  const int r = (int)(rem % R);
  const int tile = (int)(rem / R);
  int row = tile * R + r;
  // A (repeat_last == 0): the last row tile keeps only its rows up to L, rounded up to a core matrix (lift_mma_rows)
  int Rt = R;
  if (!repeat_last) { const int left = L - tile * R; Rt = left >= R ? R : (left + 7) / 8 * 8; if (r >= Rt) return; }
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row >= L && repeat_last) row = L - 1;
  if (row < L) {
    const float* src = packed + (size_t)n * blk + (size_t)row * DP + 4 * kq;
    float t[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) t[e] = (4 * kq + e < d) ? src[e] : 0.f;
    v = make_float4(t[0], t[1], t[2], t[3]);
  }
  const float4 hi = make_float4(tf32_rna(v.x), tf32_rna(v.y), tf32_rna(v.z), tf32_rna(v.w));
  const float4 lo = make_float4(tf32_rna(v.x - hi.x), tf32_rna(v.y - hi.y), tf32_rna(v.z - hi.z), tf32_rna(v.w - hi.w));
  constexpr int CQ = kMmaKC / 4;                   // 16-byte chunks per stage and half
  const int stage = kq / CQ, ch = kq % CQ;
  float* base = dst + (size_t)n * dst_blk + (size_t)tile * (2 * (size_t)KP * R);
  const size_t off_hi = (((size_t)stage * 2 + 0) * CQ + ch) * Rt + r;
  const size_t off_lo = (((size_t)stage * 2 + 1) * CQ + ch) * Rt + r;
  reinterpret_cast<float4*>(base)[off_hi] = hi;
  reinterpret_cast<float4*>(base)[off_lo] = lo;
}

// shared-memory matrix descriptor: K-major, no swizzle; LBO = bytes between the two 16-byte K chunks of one MMA
// (= 16 * R in the layout above), SBO = bytes between 8-row groups (= 128); version 1 (Blackwell)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t (&r)[16], uint32_t addr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(addr));
}

// rows of A transferred for row tile mt: the last tile stops at L (rounded up to a core matrix); the MMA still reads 128
// rows (whatever follows in shared memory) but the accumulator rows past L are never read back
__host__ __device__ inline int lift_mma_rows(int L, int mt) { const int r = L - 128 * mt; return r >= 128 ? 128 : (r + 7) / 8 * 8; }

// BRES: the B operand (y_j, all K stages, both halves) stays resident in shared memory for the pairs of a stream (they
// share j) and only A streams through the ring: operand traffic per tile 2 -> ~1.1 operands.  A CTA takes whole streams.
template <int COLS, bool BRES>
__global__ void __launch_bounds__(kMmaThreads, 1) lift_mma_kernel(const LiftParams P, const LiftMmaOperands Op) {
  constexpr int LP2 = 32 * COLS;                       // N of the MMA
  constexpr int KC = kMmaKC, NS = kMmaStages;
  constexpr uint32_t A_STAGE = 2u * KC * 128 * 4;      // bytes of a full stage: hi + lo halves
  constexpr uint32_t B_STAGE = 2u * KC * LP2 * 4;
  constexpr uint32_t B_HALF = B_STAGE / 2;
  constexpr uint32_t SLOT = A_STAGE + (BRES ? 0u : B_STAGE);
  constexpr int TMCOLS = tmem_cols_pow2(2 * LP2);
  // instruction descriptor: D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9, 10-12 = 2), K-major both, N >> 3 at 17, M >> 4 at 24
  constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(LP2 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

  extern __shared__ __align__(128) unsigned char mma_smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(mma_smem);        // [NS]
  uint64_t* empty_bar = full_bar + NS;                               // [NS]
  uint64_t* tfull_bar = empty_bar + NS;                              // [2]
  uint64_t* tempty_bar = tfull_bar + 2;                              // [2]
  uint64_t* bfull_bar = tempty_bar + 2;                              // resident B: loaded / free again
  uint64_t* bfree_bar = bfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bfree_bar + 1);
  float* ysm = reinterpret_cast<float*>(mma_smem + 128);             // [2][LP2] column constants of the tile in flight
  unsigned char* stages = mma_smem + 128 + 2 * LP2 * 4;              // 128-byte aligned
  unsigned char* bres = stages + (size_t)NS * SLOT;                  // BRES: [KP / KC][B_STAGE]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int L = P.L, L2 = P.L2, DP = P.DP;
  const int MT = (L + 127) / 128;
  const int nk = Op.KP / KC;                                         // K stages per tile

  if (threadIdx.x == 0) {
    for (int b = 0; b < NS; ++b) { mbar_init(&full_bar[b], 1); mbar_init(&empty_bar[b], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4 * kMmaEpq); }
    mbar_init(bfull_bar, 1); mbar_init(bfree_bar, 1);
    fence_mbar_init();
  }
  const uint32_t tbase = tmem_alloc(tmem_slot, TMCOLS);              // warp 0 allocates; includes the block barrier

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int st = 0; uint32_t ph = 0, bph = 0;
      for (int sg = blockIdx.x; sg < P.nstreams; sg += gridDim.x) {
        const int np = lift_stream_pairs(P, sg);
        if (np <= 0) continue;
        int i0, j;
        lift_pair(P, sg, 0, i0, j);
        const float* b = static_cast<const float*>(Op.yop) + (size_t)j * Op.yop_blk;
        if (BRES) {
          mbar_wait(bfree_bar, bph ^ 1);                              // the MMAs of the previous stream have read B
          mbar_expect_tx(bfull_bar, (uint32_t)nk * B_STAGE);
          for (int k = 0; k < nk; ++k) tma_load_1d(bres + (size_t)k * B_STAGE, b + (size_t)k * (B_STAGE / 4), B_STAGE, bfull_bar);
          bph ^= 1;
        }
        for (int n = 0; n < np; ++n)
          for (int mt = 0; mt < MT; ++mt) {
            const uint32_t a_stage = 2u * KC * (uint32_t)lift_mma_rows(L, mt) * 4;
            const float* a = static_cast<const float*>(Op.xop) + (size_t)(i0 + n) * Op.xop_blk + (size_t)mt * (2 * (size_t)Op.KP * 128);
            for (int k = 0; k < nk; ++k) {
              mbar_wait(&empty_bar[st], ph ^ 1);
              unsigned char* sa = stages + (size_t)st * SLOT;
              mbar_expect_tx(&full_bar[st], a_stage + (BRES ? 0u : B_STAGE));
              tma_load_1d(sa, a + (size_t)k * (a_stage / 4), a_stage, &full_bar[st]);
              if (!BRES) tma_load_1d(sa + A_STAGE, b + (size_t)k * (B_STAGE / 4), B_STAGE, &full_bar[st]);
              if (++st == NS) { st = 0; ph ^= 1; }
            }
          }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      int st = 0; uint32_t ph = 0, bph = 0; int buf = 0; uint32_t tph = 0;
      for (int sg = blockIdx.x; sg < P.nstreams; sg += gridDim.x) {
        const int np = lift_stream_pairs(P, sg);
        if (np <= 0) continue;
        if (BRES) { mbar_wait(bfull_bar, bph); bph ^= 1; }
        for (int n = 0; n < np; ++n)
          for (int mt = 0; mt < MT; ++mt) {
            const uint32_t RA = (uint32_t)lift_mma_rows(L, mt);
            mbar_wait(&tempty_bar[buf], tph ^ 1);                     // the epilogue has drained this accumulator
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t dcol = tbase + (uint32_t)(buf * LP2);
            for (int k = 0; k < nk; ++k) {
              mbar_wait(&full_bar[st], ph);
              asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
              const uint32_t sa = smem_u32(stages + (size_t)st * SLOT);
              const uint32_t sb = BRES ? smem_u32(bres + (size_t)k * B_STAGE) : sa + A_STAGE;
#pragma unroll
              for (int ks = 0; ks < KC / 8; ++ks) {
                // one K step = 8 elements = two 16-byte chunks; chunk c of a half starts at c * R * 16 bytes
                const uint64_t a_hi = umma_desc(sa + (uint32_t)(2 * ks) * RA * 16, RA * 16, 128);
                const uint64_t a_lo = umma_desc(sa + KC * RA * 4 + (uint32_t)(2 * ks) * RA * 16, RA * 16, 128);
                const uint64_t b_hi = umma_desc(sb + (uint32_t)(2 * ks) * LP2 * 16, LP2 * 16, 128);
                const uint64_t b_lo = umma_desc(sb + B_HALF + (uint32_t)(2 * ks) * LP2 * 16, LP2 * 16, 128);
                umma_tf32(dcol, a_hi, b_hi, IDESC, (k | ks) != 0);
                umma_tf32(dcol, a_hi, b_lo, IDESC, 1);
                umma_tf32(dcol, a_lo, b_hi, IDESC, 1);
              }
              umma_commit(&empty_bar[st]);                            // the stage is free once these MMAs have read it
              if (++st == NS) { st = 0; ph ^= 1; }
            }
            umma_commit(&tfull_bar[buf]);                             // accumulator complete
            if (++buf == 2) { buf = 0; tph ^= 1; }
          }
        if (BRES) umma_commit(bfree_bar);                             // every MMA that reads this stream's B has completed
      }
    }
  } else {
    // ===== epilogue: 8 warps; warps w and w + 4 share accumulator rows 32 * (w % 4) .. + 31 and split the column groups =====
    const int wq = warp & 3, half = (warp - 2) >> 2;                  // half: which share of the column groups
    const int et = threadIdx.x - 64;                                  // index among the epilogue threads
    int buf = 0; uint32_t tph = 0;
    const float p0 = (float)P.p0, p1 = (float)P.p1;
    // MODE 0: linear; 1: RBF on coordinates pre-scaled by sqrt(log2 e) (abi.cu): 2^(x.y - |x|^2/2 - |y|^2/2) is two
    // adds and one MUFU.EX2 per element; 2: any other base kernel
    auto run = [&](auto MODE) {
      constexpr int mode = decltype(MODE)::value;
      for (int sg = blockIdx.x; sg < P.nstreams; sg += gridDim.x) {
        const int np = lift_stream_pairs(P, sg);
        if (np <= 0) continue;
        int i0, j;
        lift_pair(P, sg, 0, i0, j);
        const float* ysq = static_cast<const float*>(P.yp) + (size_t)j * P.yblk + (size_t)L2 * DP;
        float* S = static_cast<float*>(P.scratch) + (size_t)sg * P.stream_elems;
        for (int n = 0; n < np; ++n) {
          const float* xsq = static_cast<const float*>(P.xp) + (size_t)(i0 + n) * P.xblk + (size_t)L * DP;
          for (int mt = 0; mt < MT; ++mt) {
            float* ys = ysm + buf * LP2;
            for (int q = et; q < LP2; q += 128 * kMmaEpq) { const float y = ysq[min(q, L2 - 1)]; ys[q] = mode == 1 ? -0.5f * y : y; }
            const int p = mt * 128 + wq * 32 + lane;
            float xs = p < L ? xsq[p] : 0.f;
            if (mode == 1) xs *= -0.5f;
            asm volatile("bar.sync 1, %0;" ::"n"(128 * kMmaEpq) : "memory");           // the tile's column constants are staged
            mbar_wait(&tfull_bar[buf], tph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t ta = tbase + ((uint32_t)(wq * 32) << 16) + (uint32_t)(buf * LP2);
            constexpr int NG = LP2 / 16;
            constexpr int NGW = NG / kMmaEpq;                         // column groups of 16 per warp
            static_assert(NG % kMmaEpq == 0, "column groups must split evenly over the epilogue warps");
            uint32_t r[2][16];
            tmem_ld16(r[0], ta + (uint32_t)(half * 16));
#pragma unroll
            for (int gi = 0; gi < NGW; ++gi) {
              const int g = half + kMmaEpq * gi;
              tmem_ld_wait();
              if (gi + 1 < NGW) tmem_ld16(r[(gi + 1) & 1], ta + (uint32_t)((g + kMmaEpq) * 16));  // next group in flight during the math
              float v[16];
              const float4* yv = reinterpret_cast<const float4*>(ys + g * 16);
#pragma unroll
              for (int e4 = 0; e4 < 4; ++e4) {
                const float4 y4 = yv[e4];
                const float yy[4] = {y4.x, y4.y, y4.z, y4.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const int e = 4 * e4 + k;
                  const float dot = __uint_as_float(r[gi & 1][e]);
                  if (mode == 0) v[e] = dot;
                  else if (mode == 1) v[e] = exp2_approx((dot + xs) + yy[k]);
                  else v[e] = base_eval_call<float>(P.base, p0, p1, dot, xs, yy[k]);
                }
              }
              if (p < L) {
                // skewed order of the recursion kernel: (p, q), lane t = q / COLS -> row n L + p + t, slot t
#pragma unroll
                for (int e0 = 0; e0 < 16; e0 += (COLS < 4 ? COLS : 4)) {
                  const int q = g * 16 + e0;
                  const int t = q / COLS, c = q - t * COLS;
                  float* dst = S + ((size_t)(n * L + p + t) * 32 + t) * COLS + c;
                  if constexpr (COLS >= 4) *reinterpret_cast<float4*>(dst) = make_float4(v[e0], v[e0 + 1], v[e0 + 2], v[e0 + 3]);
                  else *reinterpret_cast<float2*>(dst) = make_float2(v[e0], v[e0 + 1]);
                }
              }
            }
            // this warp has read its share of the accumulator: hand the buffer back to the MMA issuer
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[buf]);
            if (++buf == 2) { buf = 0; tph ^= 1; }
          }
        }
      }
    };
    if (P.base == GPSIG_BASE_LINEAR) run(std::integral_constant<int, 0>{});
    else if (P.base == GPSIG_BASE_RBF) run(std::integral_constant<int, 1>{});
    else run(std::integral_constant<int, 2>{});
  }
  tmem_free(tbase, TMCOLS);
}

inline size_t lift_mma_smem(int COLS, int KP, bool bres) {
  const size_t LP2 = 32 * COLS;
  const size_t a = 2 * kMmaKC * 128 * 4, b = 2 * kMmaKC * LP2 * 4;
  return 128 + 2 * LP2 * 4 + (size_t)kMmaStages * (a + (bres ? 0 : b)) + (bres ? (size_t)(KP / kMmaKC) * b : 0);
}

template <int COLS>
int launch_lift_mma(const LiftParams& P, const LiftMmaOperands& op, cudaStream_t s) {
  // B resident when it fits next to the A ring (debug switch GPSIG_MMA_NO_BRES: always stream both operands)
  const bool bres = lift_mma_smem(COLS, op.KP, true) <= kMaxSmem && getenv("GPSIG_MMA_NO_BRES") == nullptr;
  const size_t smem = lift_mma_smem(COLS, op.KP, bres);
  int sms = 148;
  { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
  const unsigned grid = (unsigned)(P.nstreams < sms ? P.nstreams : sms);
  if (bres) {
    auto kern = lift_mma_kernel<COLS, true>;
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kMmaThreads, smem, s>>>(P, op);
  } else {
    auto kern = lift_mma_kernel<COLS, false>;
    GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kMmaThreads, smem, s>>>(P, op);
  }
  GPSIG_LAUNCH_CHECK("lift_mma_kernel");
  return GPSIG_OK;
}

}  // namespace gpsig
