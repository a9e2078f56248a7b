// Microbenchmark 4 (B200, sm_100a): what ONE extra non-FP64 instruction costs next to a stream of DFMA at the occupancy
// of the systolic Gram kernel (2 warps per scheduler).  Per loop iteration: 8 independent DFMA + 8 instructions of one
// kind (inline PTX, so that the SASS opcode is the one named).  Output: scheduler cycles per iteration (8 DFMA alone =
// 16 pipe cycles) and the marginal cost per extra instruction.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/microbench4 tools/microbench4.cu
#include <cuda_runtime.h>
#include <stdio.h>
#define ITERS 4096

enum { K_NONE, K_IMAD, K_LOP3, K_SHF, K_FSEL, K_VIMNMX, K_SHFL, K_LDS64, K_FFMA, K_IADD3, K_MOV, K_LDS128B, K_CS2R };
static const char* kname[] = {"none", "IMAD", "LOP3", "SHF", "FSEL", "VIMNMX", "SHFL", "LDS.64 (random)", "FFMA", "IADD3", "MOV", "LDS.128 (broadcast)", "CS2R"};

template <int KIND, int NX>
__global__ void k_mix(double* out, double a, double b, unsigned m) {
  __shared__ double tab[2048];
  for (int e = threadIdx.x; e < 2048; e += blockDim.x) tab[e] = e;
  __syncthreads();
  const unsigned tb = (unsigned)__cvta_generic_to_shared(tab);
  double v[8];
  unsigned w[8];
  float f[8];
  double acc = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) { v[k] = threadIdx.x + k; w[k] = threadIdx.x * (k + 1) * 2654435761u; f[k] = threadIdx.x + 0.5f * k; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = fma(v[k], a, b);
      if (k < NX) {
        if (KIND == K_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(w[k]) : "r"(m), "r"(it));
        if (KIND == K_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(w[k]) : "r"(m), "r"(it));
        if (KIND == K_SHF) asm volatile("shf.r.wrap.b32 %0, %0, %1, 11;" : "+r"(w[k]) : "r"(m));
        if (KIND == K_FSEL) asm volatile("{.reg .pred p; setp.ne.u32 p, %2, 0; selp.f32 %0, %0, %1, p;}" : "+f"(f[k]) : "f"(0.f), "r"(m));
        if (KIND == K_VIMNMX) asm volatile("min.u32 %0, %0, %1;" : "+r"(w[k]) : "r"(m));
        if (KIND == K_SHFL) asm volatile("shfl.sync.up.b32 %0, %0, 1, 0, 0xffffffff;" : "+r"(w[k]));
        if (KIND == K_LDS64) { double t; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t) : "r"(tb + ((w[k] + it * 2654435761u) >> 21 << 3))); acc += 0; w[k] ^= __double2loint(t); }
        if (KIND == K_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[k]) : "f"((float)a), "f"((float)b));
        if (KIND == K_IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(w[k]) : "r"(m));
        if (KIND == K_MOV) asm volatile("mov.b32 %0, %1;" : "=r"(w[k]) : "r"(w[(k + 1) & 7]));
        if (KIND == K_LDS128B) { double2 t; asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(t.x), "=d"(t.y) : "r"(tb + ((it + k) & 127) * 16)); w[k] ^= __double2loint(t.x); }
        if (KIND == K_CS2R) asm volatile("mov.b64 %0, 0;" : "=d"(acc));
      }
    }
  }
  double s = acc;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += v[k] + w[k] + f[k];
  if (s == 123.456) out[0] = s;
}

template <typename F> float time_ms(F launch) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); cudaDeviceSynchronize();
  cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

static double base_cyc[33];
template <int KIND, int NX> void run(int sms, int threads, double* d) {
  float ms = time_ms([&] { k_mix<KIND, NX><<<sms, threads>>>(d, 1.0, 0.5, 0x5a5a5a5au); });
  const int wps = threads / 128;                        // warps per scheduler
  const double cyc = ms * 1e-3 * 1.965e9 / ITERS / wps;   // scheduler cycles per warp-iteration
  if (KIND == K_NONE) base_cyc[wps] = cyc;
  printf("[%d warps/scheduler] 8 DFMA + %d %-20s: %6.2f cycles per iteration, %+.2f per extra instruction\n", wps, NX, kname[KIND], cyc,
         NX ? (cyc - base_cyc[wps]) / NX : 0.0);
}

template <int KIND> void run_kind(int sms, double* d) {
  run<KIND, 4>(sms, 256, d);
  run<KIND, 8>(sms, 256, d);
  run<KIND, 8>(sms, 512, d);
}

int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const int sms = prop.multiProcessorCount;
  double* d; cudaMalloc(&d, 64);
  run<K_NONE, 0>(sms, 256, d);
  run<K_NONE, 0>(sms, 512, d);
  run_kind<K_IMAD>(sms, d); run_kind<K_LOP3>(sms, d); run_kind<K_SHF>(sms, d); run_kind<K_FSEL>(sms, d);
  run_kind<K_VIMNMX>(sms, d); run_kind<K_IADD3>(sms, d); run_kind<K_MOV>(sms, d); run_kind<K_FFMA>(sms, d);
  run_kind<K_SHFL>(sms, d); run_kind<K_LDS64>(sms, d); run_kind<K_LDS128B>(sms, d); run_kind<K_CS2R>(sms, d);
  return 0;
}
