// Microbenchmark 3 (B200, sm_100a): does an FP64 instruction occupy the warp scheduler's issue port for
// its two pipe cycles, or can integer / FP32 instructions issue in the shadow of a DFMA?
// Per loop iteration: 8 independent DFMA + NA ALU ops (IADD3/LOP3 chains) [+ NF FFMA].
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench3 tools/microbench3.cu
#include <cuda_runtime.h>
#include <stdio.h>
#define ITERS 4096

template <int NA, int NF>
__global__ void k_mix(double* out, double a, double b, unsigned m) {
  double v[8];
  unsigned w[8];
  float f[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { v[k] = threadIdx.x + k; w[k] = threadIdx.x * (k + 1); f[k] = threadIdx.x + 0.5f * k; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = fma(v[k], a, b);
      if (k < NA) w[k] = (w[k] ^ m) + (unsigned)it;
      if (k + 8 < NA) w[k] = (w[k] & m) + 0x9e3779b9u;
      if (k < NF) f[k] = fmaf(f[k], (float)a, (float)b);
    }
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k] + w[k] + f[k];
  if (s == 123.456) out[0] = s;
}

// dependent-issue latency of DFMA: NCH independent chains per warp, all instructions FP64
template <int NCH>
__global__ void k_lat(double* out, double a, double b) {
  double v[NCH];
#pragma unroll
  for (int k = 0; k < NCH; ++k) v[k] = threadIdx.x + k;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int k = 0; k < NCH; ++k) v[k] = fma(v[k], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < NCH; ++k) s += v[k];
  if (s == 123.456) out[0] = s;
}

template <typename F> float time_ms(F launch) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); cudaDeviceSynchronize();
  cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

template <int NA, int NF> void run(int sms, int threads, double* d) {
  float ms = time_ms([&] { k_mix<NA, NF><<<sms, threads>>>(d, 1.0, 0.5, 0x5a5a5a5au); });
  const double dfma = 8.0 * ITERS * (threads / 32) * sms / (ms * 1e-3);   // warp-instr/s
  printf("[%2d warps/SM] 8 DFMA + %2d ALU(x2 ops) + %d FFMA per iter: DFMA %.2f TFLOP/s, total %.2f warp-instr/clk/SM\n", threads / 32, NA, NF,
         dfma * 64 / 1e12, (8.0 + 2.0 * NA + NF) / 8.0 * dfma / sms / 1.965e9);
}

int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount;
  double* d; cudaMalloc(&d, 64);
  for (int threads : {32, 128, 256, 768}) {
    float ms = time_ms([&] { k_lat<1><<<sms, threads>>>(d, 1.0, 0.5); });
    printf("[%2d warps/SM] 1 dependent DFMA chain per warp: %.1f cycles per DFMA per warp\n", threads / 32, ms * 1e-3 * 1.965e9 / (8.0 * ITERS));
    ms = time_ms([&] { k_lat<2><<<sms, threads>>>(d, 1.0, 0.5); });
    printf("[%2d warps/SM] 2 chains per warp: %.1f cycles per DFMA pair per warp\n", threads / 32, ms * 1e-3 * 1.965e9 / (8.0 * ITERS));
    ms = time_ms([&] { k_lat<4><<<sms, threads>>>(d, 1.0, 0.5); });
    printf("[%2d warps/SM] 4 chains per warp: %.1f cycles per 4 DFMA per warp\n", threads / 32, ms * 1e-3 * 1.965e9 / (8.0 * ITERS));
  }
  for (int threads : {256, 512, 1024}) {
    run<0, 0>(sms, threads, d);
    run<2, 0>(sms, threads, d);
    run<4, 0>(sms, threads, d);
    run<8, 0>(sms, threads, d);
    run<16, 0>(sms, threads, d);
    run<0, 8>(sms, threads, d);
    run<4, 8>(sms, threads, d);
  }
  return 0;
}
