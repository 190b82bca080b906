// afe_platform.h -- build-mode shim for the audio front end kernels.
//
// The kernels are written as "phased block programs": a kernel object exposes
//     static constexpr int kThreads, kPhases;
//     long  num_items() const;                                   // tiles of work
//     void  phase(int ph, long item, unsigned char* smem, int tid) const;
// with the rule that NO per-thread state crosses a phase boundary and, inside a phase, a
// shared-memory word written by one thread is not touched by any other thread.
//
//   * nvcc build (product, sm_100a): afe_launch<> runs the phases as one persistent CUDA
//     kernel, one __syncthreads() between phases.
//   * g++ -DAFE_EMULATE build (TEST ONLY, tests/_emu/libafe_emu.so): afe_launch<> runs
//     "for phase: for tid:" on the host, so the very same index arithmetic is checked
//     against the oracle in the GPU-less container.  The product package never loads it.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cmath>
#include <cstring>

#if defined(__CUDACC__) && !defined(AFE_EMULATE)
#define AFE_CUDA 1
#include <cuda_runtime.h>
#define AFE_HD __host__ __device__ __forceinline__
#define AFE_RESTRICT __restrict__
#else
#define AFE_CUDA 0
#define AFE_HD inline
#define AFE_RESTRICT __restrict__
typedef void* cudaStream_t;
#endif

namespace afe {

constexpr int kNumSMsB200 = 148;

// ---------------------------------------------------------------- small math helpers
template <typename T> struct cplx { T re, im; };

AFE_HD float fast_log(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __logf(x);              // lg2.approx * ln2: abs err ~2e-7 on our range, << 1e-4 budget
#else
  return std::log(x);
#endif
}
AFE_HD float fast_log10(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __log10f(x);
#else
  return std::log10(x);
#endif
}
AFE_HD float fast_sqrt(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __fsqrt_rn(x);
#else
  return std::sqrt(x);
#endif
}
AFE_HD float fma_f(float a, float b, float c) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return std::fma(a, b, c);
#endif
}
// read-only global load (tables): L1-cached on the device
template <typename T> AFE_HD T ldg(const T* p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

// overloads __ldg lacks: 64-bit ints and pointers
AFE_HD int64_t ldg(const int64_t* p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return (int64_t)__ldg(reinterpret_cast<const long long*>(p));
#else
  return *p;
#endif
}
AFE_HD const float* ldg(const float* const* p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return reinterpret_cast<const float*>(__ldg(reinterpret_cast<const unsigned long long*>(p)));
#else
  return *p;
#endif
}

template <typename T> AFE_HD T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> AFE_HD T tmax(T a, T b) { return a > b ? a : b; }

// ---------------------------------------------------------------- memory + launch plumbing
int  dev_set(int device);                         // cudaSetDevice (no-op in emulation)
void* dev_alloc(size_t bytes);                    // nullptr on failure
void  dev_free(void* p);
int   dev_upload(void* dst, const void* src, size_t bytes);               // blocking H2D
int   dev_upload_async(void* dst, const void* src, size_t bytes, cudaStream_t s);  // src pinned/pooled
int   dev_memset_async(void* dst, int v, size_t bytes, cudaStream_t s);
void* host_pinned_alloc(size_t bytes);
void  host_pinned_free(void* p);
const char* dev_last_error_string();
int   dev_check_last();                           // 0 = ok
void  note_error(const char* what, const char* detail);   // records the thread's last error

}  // namespace afe

#if AFE_CUDA
// ------------------------------------------------------------------ CUDA driver
namespace afe {
template <class K>
__global__ void __launch_bounds__(K::kThreads, K::kMinBlocks) phased_kernel(const K k) {
  extern __shared__ __align__(16) unsigned char afe_smem[];
  const long n = k.num_items();
  for (long item = blockIdx.x; item < n; item += gridDim.x) {
#pragma unroll
    for (int ph = 0; ph < K::kPhases; ++ph) {
      k.phase(ph, item, afe_smem, (int)threadIdx.x);
      __syncthreads();
    }
  }
}
template <class K>
int afe_launch(const K& k, size_t smem_bytes, int max_blocks, cudaStream_t stream) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  const void* fn = (const void*)phased_kernel<K>;
  if (smem_bytes > 48 * 1024) {
    const cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) { note_error("cudaFuncSetAttribute(max dynamic smem)", cudaGetErrorString(e)); return -1; }
  }
  long grid = n < (long)max_blocks ? n : (long)max_blocks;
  phased_kernel<K><<<(unsigned)grid, K::kThreads, smem_bytes, stream>>>(k);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { note_error("kernel launch", cudaGetErrorString(e)); return -1; }
  return 0;
}
}  // namespace afe
#else
// ------------------------------------------------------------------ host emulation driver
#include <vector>
namespace afe {
template <class K>
int afe_launch(const K& k, size_t smem_bytes, int max_blocks, cudaStream_t) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  std::vector<double> backing((smem_bytes + 64) / sizeof(double) + 2);   // 8-byte aligned, zeroed
  unsigned char* smem = reinterpret_cast<unsigned char*>(backing.data());
  long grid = n < (long)max_blocks ? n : (long)max_blocks;
  for (long block = 0; block < grid; ++block)
    for (long item = block; item < n; item += grid)
      for (int ph = 0; ph < K::kPhases; ++ph)
        for (int tid = 0; tid < K::kThreads; ++tid) k.phase(ph, item, smem, tid);
  return 0;
}
}  // namespace afe
#endif
