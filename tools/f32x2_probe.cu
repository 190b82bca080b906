// f32x2_probe.cu -- does the B200 issue packed FP32 (FFMA2 / FADD2, sm_100) at the rate of scalar FFMA?
// Per scheduler and clock: warp-instructions issued for (a) scalar FFMA, (b) FFMA2, (c) each mixed 1:1 with IADD.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/f32x2_probe tools/f32x2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, int iters, float a, float b) {
  float2 x[8];
  int n[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = make_float2(a + i + threadIdx.x, b + i); n[i] = i + threadIdx.x; }
  const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.999f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0 || MODE == 2) { x[i].x = fmaf(x[i].x, a, b); x[i].y = fmaf(x[i].y, a2.y, b2.y); }
      else x[i] = __ffma2_rn(x[i], a2, b2);
      if (MODE >= 2) { n[i] = n[i] * 3 + it; }
    }
  }
  float s = 0; int m = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { s += x[i].x + x[i].y; m += n[i]; }
  if (s == 12345.678f || m == 123457) out[0] = s + m;
}

template <int MODE> void run(const char* name, int flops_per_iter_thread, int inst_per_iter) {
  float* out; cudaMalloc(&out, 4);
  const int iters = 20000, blocks = 148 * 2, threads = 512;
  k<MODE><<<blocks, threads>>>(out, 100, 1.0001f, 0.5f);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double cycles = ms * 1e-3 * clk * 1e3;
  const double warp_inst = (double)blocks * threads / 32 * iters * inst_per_iter;
  printf("%-28s %.3f ms  %.2f warp-inst/clk/SM  %.1f FMA lanes/clk/SM\n", name, ms, warp_inst / cycles / 148,
         (double)blocks * threads * iters * flops_per_iter_thread / cycles / 148);
  cudaFree(out);
}

int main() {
  run<0>("scalar FFMA x16", 16, 16);
  run<1>("FFMA2 x8", 16, 8);
  run<2>("scalar FFMA x16 + IMAD x8", 16, 24);
  run<3>("FFMA2 x8 + IMAD x8", 16, 16);
  return 0;
}
