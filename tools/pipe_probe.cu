// pipe_probe.cu -- measures what the B200 SM actually sustains for the instruction mixes the fbank
// kernel is made of (FP64 issue rate vs ILP and warps per scheduler, F2F conversions, 128-bit
// shared-memory traffic next to FP64).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipe_probe tools/pipe_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP, int OP>
__global__ void fp64_chain(double* out, int iters, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (OP == 0) x[i] = fma(x[i], a, b);
      else if (OP == 1) x[i] = x[i] + b;
      else x[i] = x[i] * a;
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  if (s == 12345.678) out[0] = s;
}

template <int ILP>
__global__ void f2f_chain(float* out, int iters, float a) {
  float x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      double d;
      asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(x[i]));
      asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(x[i]) : "d"(d));
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  if (s == 12345.678f) out[0] = s;
}

// FP64 FMAs interleaved with 128-bit shared loads/stores (the exchange pattern): per iteration
// NF DFMAs + 1 LDS.128 + 1 STS.128 per thread
template <int NF>
__global__ void fp64_lds(double* out, int iters, double a, double b) {
  extern __shared__ double2 sm[];
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = a + i + threadIdx.x;
  double2* p = sm + threadIdx.x;
  for (int it = 0; it < iters; ++it) {
    double2 v = p[(it & 7) * blockDim.x];
#pragma unroll
    for (int k = 0; k < NF; ++k) x[k & 7] = fma(x[k & 7], a, b);
    x[0] += v.x; x[1] += v.y;
    p[((it + 3) & 7) * blockDim.x] = make_double2(x[2], x[3]);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 12345.678) out[0] = s;
}

template <class F> float time_ms(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const double hz = clk_khz * 1e3;
  const int sms = pr.multiProcessorCount;
  double* out; cudaMalloc(&out, 1024);
  const int iters = 8192;
  printf("%s, %d SMs, %.0f MHz (cycles below assume this clock)\n", pr.name, sms, hz / 1e6);
  printf("# FP64: warp-instructions per cycle per SM sub-partition (scheduler)\n");
#define RUN(ILP, OP, THREADS) { float ms = time_ms([&] { fp64_chain<ILP, OP><<<sms, THREADS>>>(out, iters, 1.0000001, 1e-9); }); \
    double wi = (double)iters * ILP * (THREADS / 32); double cyc = ms * 1e-3 * hz; \
    printf("op %s ilp %d warps/sched %d: %.3f winst/clk/sched, %.1f cycles per dependent step\n", OP == 0 ? "DFMA" : OP == 1 ? "DADD" : "DMUL", ILP, THREADS / 128, wi / 4 / cyc, cyc / iters); }
  RUN(1, 0, 128) RUN(2, 0, 128) RUN(4, 0, 128) RUN(8, 0, 128)
  RUN(1, 0, 512) RUN(2, 0, 512) RUN(4, 0, 512) RUN(8, 0, 512) RUN(16, 0, 512)
  RUN(4, 1, 512) RUN(8, 1, 512) RUN(4, 2, 512) RUN(8, 2, 512)
  RUN(8, 0, 1024)
  printf("# F2F f32->f64->f32 round trips (2 conversions each)\n");
#define RUNC(ILP, THREADS) { float* o2 = (float*)out; float ms = time_ms([&] { f2f_chain<ILP><<<sms, THREADS>>>(o2, iters, 1.5f); }); \
    double wi = 2.0 * iters * ILP * (THREADS / 32); double cyc = ms * 1e-3 * hz; \
    printf("F2F ilp %d warps/sched %d: %.3f winst/clk/sched\n", ILP, THREADS / 128, wi / 4 / cyc); }
  RUNC(4, 512) RUNC(8, 512)
  printf("# FP64 + 128-bit shared load/store per iteration, 512 threads\n");
#define RUNL(NF) { cudaFuncSetAttribute(fp64_lds<NF>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 512 * 16); \
    float ms = time_ms([&] { fp64_lds<NF><<<sms, 512, 8 * 512 * 16>>>(out, iters, 1.0000001, 1e-9); }); \
    double cyc = ms * 1e-3 * hz; \
    printf("NF %d: %.1f cycles/iter/SM; DFMA %.3f winst/clk/sched; smem %.1f B/clk/SM\n", NF, cyc / iters, (double)(NF + 2) * 16 / 4 / (cyc / iters), 2.0 * 512 * 16 / (cyc / iters)); }
  RUNL(4) RUNL(8) RUNL(16) RUNL(32)
  return 0;
}
