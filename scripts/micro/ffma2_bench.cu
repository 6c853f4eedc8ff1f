// Microbenchmark: issue rate of packed fma.rn.f32x2 (FFMA2) against scalar FFMA on sm_100a,
// alone and mixed with ALU-pipe work.  nvcc -gencode arch=compute_100a,code=sm_100a -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ float fma1(float a, float b, float c) {
  float r;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
template <int MODE> __global__ void k(float *out, int iters, float s) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 1e-3f + i;
  unsigned ia = threadIdx.x, ib = 0x9e3779b9u;
  const float m = s, a = 1e-3f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) { // 16 scalar FFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fma1(acc[i], m, a);
    } else if (MODE == 1) { // 8 FFMA2 = same 16 lane-FMAs
      u64 mm, aa;
      asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
      asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        u64 v;
        asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(acc[2 * i]), "f"(acc[2 * i + 1]));
        v = fma2(v, mm, aa);
        asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[2 * i]), "=f"(acc[2 * i + 1]) : "l"(v));
      }
    } else if (MODE == 2) { // 8 FFMA2 + 8 ALU (LOP3/IADD) per iteration
      u64 mm, aa;
      asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
      asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        u64 v;
        asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(acc[2 * i]), "f"(acc[2 * i + 1]));
        v = fma2(v, mm, aa);
        asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[2 * i]), "=f"(acc[2 * i + 1]) : "l"(v));
        ia = (ia ^ ib) + 0x1234567u * (i + 1);
      }
    } else { // 16 FFMA + 8 ALU
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fma1(acc[i], m, a);
#pragma unroll
      for (int i = 0; i < 8; ++i) ia = (ia ^ ib) + 0x1234567u * (i + 1);
    }
  }
  float t = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) t += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = t + (float)ia;
}
template <int MODE> void run(const char *name, float *out, int sms) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000, blocks = sms * 4, threads = 512;
  k<MODE><<<blocks, threads>>>(out, 100, 0.999f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters, 0.999f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double lane_fma = (double)blocks * threads * iters * 16;
  printf("%-28s %.3f ms  %.1f TFMA-lanes/s = %.2f lane-FMA/clk/SM @1.965GHz\n", name, ms,
         lane_fma / ms * 1e-9, lane_fma / (ms * 1e-3) / (sms * 1.965e9));
}
int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  float *out;
  cudaMalloc(&out, p.multiProcessorCount * 4 * 512 * 4);
  run<0>("16 FFMA", out, p.multiProcessorCount);
  run<1>("8 FFMA2", out, p.multiProcessorCount);
  run<3>("16 FFMA + 8 ALU", out, p.multiProcessorCount);
  run<2>("8 FFMA2 + 8 ALU", out, p.multiProcessorCount);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
