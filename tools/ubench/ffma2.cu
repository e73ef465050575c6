// microbenchmark: FFMA vs FFMA2 issue/pipe throughput on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float m = 1.0001f, c = 0.5f;
  unsigned long long p0, p1, p2, p3, pm, pc;
  asm("mov.b64 %0, {%1,%2};" : "=l"(p0) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p1) : "f"(a2), "f"(a3));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p2) : "f"(a4), "f"(a5));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p3) : "f"(a6), "f"(a7));
  asm("mov.b64 %0, {%1,%2};" : "=l"(pm) : "f"(m), "f"(m));
  asm("mov.b64 %0, {%1,%2};" : "=l"(pc) : "f"(c), "f"(c));
  int x = threadIdx.x;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (MODE == 0) {  // 8 independent FFMA
        a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
        a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
      } else if (MODE == 1) {  // 4 FFMA2 (same flops as mode 0)
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p0) : "l"(pm), "l"(pc));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p1) : "l"(pm), "l"(pc));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2) : "l"(pm), "l"(pc));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p3) : "l"(pm), "l"(pc));
      } else if (MODE == 2) {  // 4 FFMA2 + 4 integer ALU ops (can they co-issue in the freed slots?)
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p0) : "l"(pm), "l"(pc));
        x = (x ^ (x << 1)) + 3;
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p1) : "l"(pm), "l"(pc));
        x = (x ^ (x >> 1)) + 5;
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2) : "l"(pm), "l"(pc));
        x = (x ^ (x << 2)) + 7;
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p3) : "l"(pm), "l"(pc));
        x = (x ^ (x >> 2)) + 9;
      } else {  // 8 FFMA + the same integer ops
        a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c);
        x = (x ^ (x << 1)) + 3;
        a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
        x = (x ^ (x >> 1)) + 5;
        a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c);
        x = (x ^ (x << 2)) + 7;
        a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        x = (x ^ (x >> 2)) + 9;
      }
    }
  }
  float r = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + (float)x;
  float lo, hi;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p0)); r += lo + hi;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p1)); r += lo + hi;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p2)); r += lo + hi;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p3)); r += lo + hi;
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d) {
  const int iters = 4096;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<148 * 8, 256>>>(d, 16);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  k<MODE><<<148 * 8, 256>>>(d, iters);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double fma = 148.0 * 8 * 256 * (double)iters * 16 * 8;
  printf("%-28s %.3f ms  %.1f TFLOP/s (FMA = 2 flop)\n", name, ms, 2 * fma / ms / 1e9);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("8 FFMA", d); run<1>("4 FFMA2", d); run<2>("4 FFMA2 + 4 int", d); run<3>("8 FFMA + 4 int", d);
  return 0;
}
