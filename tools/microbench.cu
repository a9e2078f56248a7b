// Microbenchmarks that decide the design of the fused Gram kernel on B200 (sm_100a):
//   FP64 FMA peak, FP64 tensor (mma.sync m8n8k4.f64) peak, the two interleaved in one warp
//   (do they share a pipe?), SHFL throughput, cost of exp(double) / expf.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <math.h>

#define ITERS 4096

__global__ void k_dfma(double* out, double a, double b) {
  double v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fma(v[k], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
}

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void k_dmma(double* out, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int k = 0; k < 8; ++k) { c[k][0] = threadIdx.x; c[k][1] = k; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) dmma(c[k][0], c[k][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += c[k][0] + c[k][1];
  if (s == 123.456) out[0] = s;
}

// NF fma per NM mma, interleaved
template <int NF, int NM>
__global__ void k_mix(double* out, double a, double b) {
  double c[4][2], v[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) { c[k][0] = threadIdx.x; c[k][1] = k; }
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < NM; ++k) dmma(c[k % 4][0], c[k % 4][1], a, b);
#pragma unroll
    for (int k = 0; k < NF; ++k) v[k % 8] = fma(v[k % 8], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) s += c[k][0] + c[k][1];
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
}

__global__ void k_shfl(double* out) {
  unsigned v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __shfl_up_sync(0xffffffffu, v[k], 1) + 1;
  }
  unsigned s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 12345u) out[0] = s;
}

__global__ void k_exp(double* out, double a) {
  double v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = -1e-3 * (threadIdx.x + k);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = exp(v[k]) * a;     // a < 0 keeps the argument negative and O(1)
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
}
__global__ void k_expf(double* out, float a) {
  float v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = -1e-3f * (threadIdx.x + k);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = expf(v[k]) * a;
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) s += v[k];
  if (s == 123.456f) out[0] = s;
}
__global__ void k_ffma(double* out, float a, float b) {
  float v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], a, b);
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k];
  if (s == 123.456f) out[0] = s;
}

template <typename F>
float timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  double* out; cudaMalloc(&out, 64);
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount;
  printf("device %s, %d SMs\n", p.name, nsm);
  for (int wps = 4; wps <= 32; wps *= 2) {       // warps per SM
    const int blocks = nsm * (wps / 4 > 0 ? wps / 4 : 1), threads = 128;
    const double nthreads = (double)blocks * threads;
    float t;
    t = timeit([&] { k_dfma<<<blocks, threads>>>(out, 0.999, 1e-7); });
    printf("[%2d warps/SM] DFMA      %.2f TFLOP/s\n", wps, 2.0 * nthreads * 8 * ITERS / t / 1e9);
    t = timeit([&] { k_dmma<<<blocks, threads>>>(out, 0.999, 1e-7); });
    printf("[%2d warps/SM] DMMA      %.2f TFLOP/s\n", wps, 2.0 * (nthreads / 32) * 256 * 8 * ITERS / t / 1e9);
    t = timeit([&] { k_mix<8, 1><<<blocks, threads>>>(out, 0.999, 1e-7); });
    printf("[%2d warps/SM] mix 8 DFMA + 1 DMMA: DFMA part %.2f TFLOP/s, DMMA part %.2f TFLOP/s\n", wps,
           2.0 * nthreads * 8 * ITERS / t / 1e9, 2.0 * (nthreads / 32) * 256 * 1 * ITERS / t / 1e9);
    t = timeit([&] { k_mix<8, 2><<<blocks, threads>>>(out, 0.999, 1e-7); });
    printf("[%2d warps/SM] mix 8 DFMA + 2 DMMA: DFMA part %.2f TFLOP/s, DMMA part %.2f TFLOP/s\n", wps,
           2.0 * nthreads * 8 * ITERS / t / 1e9, 2.0 * (nthreads / 32) * 256 * 2 * ITERS / t / 1e9);
    t = timeit([&] { k_mix<8, 4><<<blocks, threads>>>(out, 0.999, 1e-7); });
    printf("[%2d warps/SM] mix 8 DFMA + 4 DMMA: DFMA part %.2f TFLOP/s, DMMA part %.2f TFLOP/s\n", wps,
           2.0 * nthreads * 8 * ITERS / t / 1e9, 2.0 * (nthreads / 32) * 256 * 4 * ITERS / t / 1e9);
    t = timeit([&] { k_ffma<<<blocks, threads>>>(out, 0.999f, 1e-7f); });
    printf("[%2d warps/SM] FFMA      %.2f TFLOP/s\n", wps, 2.0 * nthreads * 8 * ITERS / t / 1e9);
    t = timeit([&] { k_shfl<<<blocks, threads>>>(out); });
    printf("[%2d warps/SM] SHFL.32   %.1f G warp-instr/s  (%.2f per clk per SM at 1.965 GHz)\n", wps,
           (nthreads / 32) * 8 * ITERS / t / 1e6, (nthreads / 32) * 8 * ITERS / t / 1e6 / nsm / 1.965);
    t = timeit([&] { k_exp<<<blocks, threads>>>(out, -0.7); });
    printf("[%2d warps/SM] exp(f64)  %.1f G/s  -> %.1f DFMA-equivalents each at 18.6 T DFMA/s\n", wps,
           nthreads * 4 * ITERS / t / 1e6, 18.6e3 / (nthreads * 4 * ITERS / t / 1e6));
    t = timeit([&] { k_expf<<<blocks, threads>>>(out, -0.7f); });
    printf("[%2d warps/SM] expf      %.1f G/s  -> %.1f FFMA-equivalents each at 37.2 T FFMA/s\n", wps,
           nthreads * 4 * ITERS / t / 1e6, 37.2e3 / (nthreads * 4 * ITERS / t / 1e6));
  }
  return 0;
}
