// Microbenchmarks, second set (B200, sm_100a): can tensor memory serve as per-lane private storage
// for the column tile of the systolic Gram kernel (tcgen05.st once, tcgen05.ld every step), and what
// do the packed FP32 instructions (FFMA2 / FADD2) buy?
//   LDTM / STTM throughput and round-trip correctness, LDTM interleaved with DFMA (interference with
//   the FP64 pipe), LDS.128 for comparison, FFMA vs FFMA2 issue rate.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench2 tools/microbench2.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ITERS 2048

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t tmem_alloc_all(uint32_t* slot) {
  if ((threadIdx.x >> 5) == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot;
}
__device__ __forceinline__ void tmem_free_all(uint32_t base) {
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(base) : "memory");
}

__device__ __forceinline__ void ldtm16(uint32_t (&r)[16], uint32_t addr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(addr) : "memory");
}
__device__ __forceinline__ void sttm16(uint32_t addr, const uint32_t (&r)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                 "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// each warp owns columns [ (warp/4)*CPW, +CPW ) of its lane quarter
__device__ __forceinline__ uint32_t warp_taddr(uint32_t base, int cpw) {
  const int warp = threadIdx.x >> 5;
  return base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * cpw);
}

// round trip: every lane stores 112 words of its own pattern, reads them back
__global__ void k_roundtrip(int* bad) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc_all(&slot);
  const int cpw = 512 / ((blockDim.x / 32 + 3) / 4);
  const uint32_t ta = warp_taddr(base, cpw);
  uint32_t r[16];
  for (int k = 0; k < 7; ++k) {
#pragma unroll
    for (int e = 0; e < 16; ++e) r[e] = threadIdx.x * 1000003u + (k * 16 + e) * 7919u + blockIdx.x;
    sttm16(ta + k * 16, r);
  }
  wait_st();
  int nb = 0;
  for (int k = 6; k >= 0; --k) {
    ldtm16(r, ta + k * 16);
    wait_ld();
#pragma unroll
    for (int e = 0; e < 16; ++e) nb += r[e] != threadIdx.x * 1000003u + (k * 16 + e) * 7919u + blockIdx.x;
  }
  if (nb) atomicAdd(bad, nb);
  tmem_free_all(base);
}

// LDTM throughput: 7 x16 loads (= one column tile of the f64 kernel: 7 columns x 8 doubles) per iteration
template <int NDFMA>
__global__ void k_ldtm(double* out, double a, double b) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc_all(&slot);
  const int cpw = 512 / ((blockDim.x / 32 + 3) / 4);
  const uint32_t ta = warp_taddr(base, cpw);
  uint32_t acc = 0;
  double v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  uint32_t r[2][16];
  ldtm16(r[0], ta);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      wait_ld();
      ldtm16(r[(k + 1) & 1], ta + ((k + 1) % 7) * 16);
#pragma unroll
      for (int e = 0; e < 16; ++e) acc ^= r[k & 1][e];
#pragma unroll
      for (int f = 0; f < NDFMA; ++f) v[f % 8] = fma(v[f % 8], a, b);
    }
  }
  wait_ld();
  double s = acc;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
  tmem_free_all(base);
}

__global__ void k_sttm(double* out) {
  __shared__ uint32_t slot;
  const uint32_t base = tmem_alloc_all(&slot);
  const int cpw = 512 / ((blockDim.x / 32 + 3) / 4);
  const uint32_t ta = warp_taddr(base, cpw);
  uint32_t r[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) r[e] = threadIdx.x + e;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 7; ++k) { sttm16(ta + k * 16, r); r[k] += 1; }
  }
  wait_st();
  if (r[0] == 12345u) out[0] = r[0];
  tmem_free_all(base);
}

// LDS.128 with the lane-interleaved layout of the current kernel (conflict free): 28 per iteration
template <int NDFMA>
__global__ void k_lds(double* out, double a, double b) {
  extern __shared__ __align__(16) unsigned char sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint4* tile = reinterpret_cast<uint4*>(sm) + (size_t)warp * 28 * 32;
  for (int k = 0; k < 28; ++k) tile[k * 32 + lane] = make_uint4(k, lane, warp, 1);
  __syncwarp();
  uint32_t acc = 0;
  double v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 28; ++k) {
      uint4 q;
      asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(smem_u32(&tile[k * 32 + lane])));
      acc ^= q.x ^ q.y ^ q.z ^ q.w;
      if ((k & 3) == 3) {
#pragma unroll
        for (int f = 0; f < NDFMA; ++f) v[f % 8] = fma(v[f % 8], a, b);
      }
    }
  }
  double s = acc;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
}

__global__ void k_ffma(float* out, float a, float b) {
  float v[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = fmaf(v[k], a, b);
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += v[k];
  if (s == 123.456f) out[0] = s;
}

__device__ __forceinline__ unsigned long long ffma2(unsigned long long x, unsigned long long y, unsigned long long z) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(y), "l"(z));
  return r;
}
__device__ __forceinline__ unsigned long long fadd2(unsigned long long x, unsigned long long y) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
  return r;
}
// MODE 0: FFMA2 only; 1: FADD2 only; 2: one FFMA2 + one scalar integer op (does the packed op free issue slots?)
template <int MODE>
__global__ void k_ffma2(float* out, unsigned long long a, unsigned long long b) {
  unsigned long long v[8];
  unsigned w[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { v[k] = threadIdx.x + k; w[k] = threadIdx.x * k; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (MODE == 1) v[k] = fadd2(v[k], a);
      else v[k] = ffma2(v[k], a, b);
      if (MODE == 2) w[k] = (w[k] ^ (unsigned)it) + 0x9e3779b9u;
    }
  }
  unsigned long long s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k] + w[k];
  if (s == 123456ull) out[0] = (float)s;
}

template <typename F>
float time_ms(F launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); cudaDeviceSynchronize();
  cudaEventRecord(e0);
  launch();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(err));
  return ms;
}

int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount;
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  printf("device %s, %d SMs, %.3f GHz nominal\n", prop.name, sms, ghz);
  double* dout; cudaMalloc(&dout, 64);
  int* bad; cudaMalloc(&bad, 4); cudaMemset(bad, 0, 4);
  for (int threads : {128, 256, 512}) {
    k_roundtrip<<<sms, threads>>>(bad);
    int hb = -1; cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
    printf("[%2d warps/SM] TMEM st/ld round trip: %d mismatches (%s)\n", threads / 32, hb, cudaGetErrorString(cudaGetLastError()));
  }
  for (int threads : {128, 256, 384, 512}) {
    const int w = threads / 32;
    const double tile_bytes = 7.0 * 16 * 4 * 32;            // per warp per iteration
    float ms = time_ms([&] { k_ldtm<0><<<sms, threads>>>(dout, 1.0, 0.5); });
    printf("[%2d warps/SM] LDTM.x16 alone: %.1f B/clk/SM\n", w, tile_bytes * w * ITERS / (ms * 1e-3) / (ghz * 1e9));
    ms = time_ms([&] { k_ldtm<16><<<sms, threads>>>(dout, 1.0, 0.5); });
    printf("[%2d warps/SM] LDTM.x16 + 16 DFMA each: %.1f B/clk/SM, DFMA %.2f TFLOP/s\n", w,
           tile_bytes * w * ITERS / (ms * 1e-3) / (ghz * 1e9), 2.0 * 16 * 7 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_sttm<<<sms, threads>>>(dout); });
    printf("[%2d warps/SM] STTM.x16 alone: %.1f B/clk/SM\n", w, tile_bytes * w * ITERS / (ms * 1e-3) / (ghz * 1e9));
    const size_t smem = (size_t)w * 28 * 32 * 16;
    cudaFuncSetAttribute(k_lds<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_lds<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    ms = time_ms([&] { k_lds<0><<<sms, threads, smem>>>(dout, 1.0, 0.5); });
    printf("[%2d warps/SM] LDS.128 alone: %.1f B/clk/SM\n", w, tile_bytes * w * ITERS / (ms * 1e-3) / (ghz * 1e9));
    ms = time_ms([&] { k_lds<16><<<sms, threads, smem>>>(dout, 1.0, 0.5); });
    printf("[%2d warps/SM] LDS.128 + 16 DFMA per 4: %.1f B/clk/SM, DFMA %.2f TFLOP/s\n", w,
           tile_bytes * w * ITERS / (ms * 1e-3) / (ghz * 1e9), 2.0 * 16 * 7 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12);
  }
  float* fout; cudaMalloc(&fout, 64);
  for (int threads : {256, 512, 1024}) {
    const int w = threads / 32;
    float ms = time_ms([&] { k_ffma<<<sms, threads>>>(fout, 1.0f, 0.5f); });
    printf("[%2d warps/SM] FFMA   %.2f TFLOP/s\n", w, 2.0 * 16 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_ffma2<0><<<sms, threads>>>(fout, 0x3f8000003f800000ull, 0x3f0000003f000000ull); });
    printf("[%2d warps/SM] FFMA2  %.2f TFLOP/s  (%.2f warp-instr/clk/SM)\n", w, 2.0 * 2 * 8 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12,
           8.0 * ITERS * w / (ms * 1e-3) / (ghz * 1e9));
    ms = time_ms([&] { k_ffma2<1><<<sms, threads>>>(fout, 0x3f8000003f800000ull, 0x3f0000003f000000ull); });
    printf("[%2d warps/SM] FADD2  %.2f T adds/s  (%.2f warp-instr/clk/SM)\n", w, 2.0 * 8 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12,
           8.0 * ITERS * w / (ms * 1e-3) / (ghz * 1e9));
    ms = time_ms([&] { k_ffma2<2><<<sms, threads>>>(fout, 0x3f8000003f800000ull, 0x3f0000003f000000ull); });
    printf("[%2d warps/SM] FFMA2 + 2 ALU ops each: FFMA2 part %.2f TFLOP/s (%.2f FFMA2/clk/SM)\n", w,
           2.0 * 2 * 8 * ITERS * threads * (double)sms / (ms * 1e-3) / 1e12, 8.0 * ITERS * w / (ms * 1e-3) / (ghz * 1e9));
  }
  return 0;
}
