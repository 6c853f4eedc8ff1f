// pipe_model.cu — cycles per warp instruction per SM sub-partition (scheduler) on sm_100a for the
// instruction classes of the cost kernels, alone and mixed, at 8 warps per scheduler with 8
// independent chains per thread.  Feeds the throughput model in DESIGN.md section 5.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_model pipe_model.cu && ./pipe_model
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define FMA2(v, m, a) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(m), "l"(a))
#define FMA1(v, m, a) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(v) : "f"(m), "f"(a))
#define LOP(v, b) asm volatile("xor.b32 %0, %0, %1;" : "+r"(v) : "r"(b))
#define IMAD(v, m, a) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v) : "r"(m), "r"(a))
#define IMIN(v, b) asm volatile("min.s32 %0, %0, %1;" : "+r"(v) : "r"(b))
#define FMIN(v, b) asm volatile("min.f32 %0, %0, %1;" : "+f"(v) : "f"(b))
#define FMIN3(v, b, c) asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(v) : "f"(b), "f"(c))
#define F2I(r, v) asm volatile("cvt.rzi.u32.f32 %0, %1;" : "=r"(r) : "f"(v))
#define ISETP_SEL(r, a, b) asm volatile("{ .reg .pred p; setp.lt.u32 p, %1, %2; selp.u32 %0, %1, %2, p; }" : "=r"(r) : "r"(a), "r"(b))

template <int MODE> __global__ void __launch_bounds__(1024) k(float *out, int iters, float s, long long *cycles) {
  u64 v2[8];
  float v1[8];
  unsigned vi[8];
  const float m = s, a = 1e-3f;
  u64 mm, aa;
  asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
  asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v1[i] = threadIdx.x * 1e-3f + i;
    vi[i] = threadIdx.x * 7 + i;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v2[i]) : "f"(v1[i]), "f"(v1[i] + 0.5f));
  }
  unsigned ib = 0x9e3779b9u ^ blockIdx.x, sink = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { FMA2(v2[i], mm, aa); FMA2(v2[i], mm, aa); }
      if (MODE == 1) { FMA1(v1[i], m, a); FMA1(v1[i], m, a); }
      if (MODE == 2) { LOP(vi[i], ib); LOP(vi[i], ib); }
      if (MODE == 3) { IMAD(vi[i], ib, ib); IMAD(vi[i], ib, ib); }
      if (MODE == 4) { unsigned r; F2I(r, v1[i]); sink ^= r; F2I(r, v1[(i + 1) & 7]); sink += r; }
      if (MODE == 5) { FMIN(v1[i], m); FMIN(v1[i], a); }
      if (MODE == 6) { IMIN(vi[i], ib); IMIN(vi[i], ib); }
      if (MODE == 7) { FMA2(v2[i], mm, aa); LOP(vi[i], ib); }
      if (MODE == 8) { FMA2(v2[i], mm, aa); FMA1(v1[i], m, a); }
      if (MODE == 9) { FMA2(v2[i], mm, aa); IMAD(vi[i], ib, ib); }
      if (MODE == 10) { FMA1(v1[i], m, a); LOP(vi[i], ib); }
      if (MODE == 11) { FMIN3(v1[i], m, a); FMIN3(v1[i], a, m); }
      if (MODE == 12) { FMA1(v1[i], m, a); FMIN(v1[(i + 4) & 7], m); }
      if (MODE == 13) { FMA2(v2[i], mm, aa); LOP(vi[i], ib); LOP(vi[i], ib); }      // 1 : 2
      if (MODE == 14) { FMA2(v2[i], mm, aa); FMA2(v2[i], mm, aa); LOP(vi[i], ib); } // 2 : 1
      if (MODE == 15) { unsigned r; ISETP_SEL(r, vi[i], ib); vi[i] = r; ISETP_SEL(r, vi[i], sink); vi[i] = r; }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  float t = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v2[i]));
    t += v1[i] + lo + hi + (float)vi[i];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = t + (float)sink;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char *name, int per_iter, float *out, long long *cyc, int sms) {
  const int iters = 4000, threads = 1024; // 32 warps per SM = 8 per scheduler, one CTA per SM
  k<MODE><<<sms, threads>>>(out, 50, 0.999f, cyc);
  cudaDeviceSynchronize();
  k<MODE><<<sms, threads>>>(out, iters, 0.999f, cyc);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("%-34s FAILED\n", name); return; }
  long long h[256], mx = 0;
  cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
  for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
  const double instr_per_sched = (double)iters * per_iter * 8; // 8 warps per scheduler
  printf("%-34s %7.3f cycles per warp-instruction per scheduler\n", name, mx / instr_per_sched);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  float *out;
  long long *cyc;
  cudaMalloc(&out, sizeof(float) * sms * 1024);
  cudaMalloc(&cyc, sizeof(long long) * 256);
  printf("%s, %d SMs\n", p.name, sms);
  run<0>("FFMA2", 16, out, cyc, sms);
  run<1>("FFMA", 16, out, cyc, sms);
  run<2>("LOP3 (ALU)", 16, out, cyc, sms);
  run<3>("IMAD", 16, out, cyc, sms);
  run<4>("F2I.TRUNC (XU)", 16, out, cyc, sms);
  run<5>("FMNMX", 16, out, cyc, sms);
  run<6>("IMNMX", 16, out, cyc, sms);
  run<15>("ISETP+SEL pair", 32, out, cyc, sms);
  run<11>("FMNMX3 (3-input min)", 16, out, cyc, sms);
  run<7>("FFMA2 + LOP3 (1:1)", 16, out, cyc, sms);
  run<13>("FFMA2 + 2 LOP3 (1:2)", 24, out, cyc, sms);
  run<14>("2 FFMA2 + LOP3 (2:1)", 24, out, cyc, sms);
  run<8>("FFMA2 + FFMA (1:1)", 16, out, cyc, sms);
  run<9>("FFMA2 + IMAD (1:1)", 16, out, cyc, sms);
  run<10>("FFMA + LOP3 (1:1)", 16, out, cyc, sms);
  run<12>("FFMA + FMNMX (1:1)", 16, out, cyc, sms);
  return 0;
}
