// ffma2_bench.cu — DEVELOPER TOOL: issue rate and latency of sm_100's packed FFMA2 against scalar FFMA (profiles/r2_experiments.txt E18).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o exp/ffma2_bench tools/micro/ffma2_bench.cu
#include <cuda_runtime.h>
#include <cstdio>

__device__ __forceinline__ void fma2(float& x, float& y, float s, float c) {
  unsigned long long a, b, m, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(x), "f"(y));
  asm("mov.b64 %0, {%1, %1};" : "=l"(b) : "f"(s));
  asm("mov.b64 %0, {%1, %1};" : "=l"(m) : "f"(c));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(m));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(d));
}

// kMode 0: 16 independent FFMA chains; 1: 8 independent FFMA2 chains (same 16 values); 2: 16 FFMA + 32 FMNMX; 3: 8 FFMA2 + 32 FMNMX
template <int kMode>
__global__ void __launch_bounds__(256) k_rate(float* out, int iters, float b, float c) {
  float f[16], g[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) f[k] = (float)(threadIdx.x + k) * 1e-3f, g[k] = f[k] + 1.f;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (kMode == 0 || kMode == 2) {
#pragma unroll
        for (int k = 0; k < 16; ++k) f[k] = fmaf(f[k], b, c);
      } else {
#pragma unroll
        for (int k = 0; k < 16; k += 2) fma2(f[k], f[k + 1], b, c);
      }
      if (kMode >= 2) {
#pragma unroll
        for (int k = 0; k < 16; ++k) g[k] = fmaxf(fminf(g[k], f[k]), f[(k + 1) & 15]);
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += f[k] + g[k];
  if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// one dependent chain per thread, one warp per SM-quarter: cycles per instruction = latency
template <int kMode>
__global__ void k_lat(float* out, long long* cycles, int iters, float b, float c) {
  float x = threadIdx.x * 1e-3f, y = x + 1.f;
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      if (kMode == 0) x = fmaf(x, b, c);
      else fma2(x, y, b, c);
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cycles[0] = t1 - t0;
  if (x + y == 12345.678f) out[0] = x;
}

template <int kMode>
static double rate(int sms, float* d_out) {
  const int blocks = sms * 8, threads = 256, iters = 1 << 14;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k_rate<kMode><<<blocks, threads>>>(d_out, iters, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep) best = best < ms ? best : ms;
  }
  return (double)blocks * threads * iters / (best * 1e-3);  // loop iterations per second over all threads
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float* d_out;
  long long* d_cyc;
  cudaMalloc(&d_out, (size_t)sms * 8 * 256 * 4);
  cudaMalloc(&d_cyc, 8);
  const double r0 = rate<0>(sms, d_out), r1 = rate<1>(sms, d_out), r2 = rate<2>(sms, d_out), r3 = rate<3>(sms, d_out);
  std::printf("per thread-iteration: 32 scalar FMAs (or 16 FFMA2), optionally + 64 FMNMX\n");
  std::printf("FFMA  only : %.2f T FMA/s   %.2f T warp-lane-instr/s\n", r0 * 32 / 1e12, r0 * 32 / 1e12);
  std::printf("FFMA2 only : %.2f T FMA/s   %.2f T warp-lane-instr/s (an FFMA2 counted once)\n", r1 * 32 / 1e12, r1 * 16 / 1e12);
  std::printf("FFMA  + 2 FMNMX per FMA : %.2f T FMA/s   %.2f T lane-instr/s\n", r2 * 32 / 1e12, r2 * 96 / 1e12);
  std::printf("FFMA2 + 2 FMNMX per FMA : %.2f T FMA/s   %.2f T lane-instr/s (an FFMA2 counted once)\n", r3 * 32 / 1e12, r3 * 80 / 1e12);
  long long cyc;
  k_lat<0><<<1, 32>>>(d_out, d_cyc, 4096, 0.999f, 1e-3f);
  cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  std::printf("dependent FFMA  latency: %.2f cycles\n", (double)cyc / (4096.0 * 16));
  k_lat<1><<<1, 32>>>(d_out, d_cyc, 4096, 0.999f, 1e-3f);
  cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  std::printf("dependent FFMA2 latency: %.2f cycles\n", (double)cyc / (4096.0 * 16));
  std::printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
