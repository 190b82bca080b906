// afe_platform.h -- build-mode shim for the audio front end kernels.
//
// The kernels are written as "phased block programs": a kernel object exposes
//     static constexpr int kThreads, kPhases;
//     long  num_items() const;                                   // tiles of work
//     Item  prepare(long item) const;                            // per-item context, computed once
//     void  phase(int ph, const Item&, unsigned char* smem, int tid) const;
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
#include <cstdio>
#include <cstdlib>

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
struct alignas(8) float2 { float x, y; };
struct alignas(16) double2 { double x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(8) short4 { short x, y, z, w; };
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

// L2-coherent 128-bit load (data this block wrote earlier in the same launch)
AFE_HD float4 ldcg4(const float4* p) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __ldcg(p);
#else
  return *p;
#endif
}

// two adjacent table entries (16-byte aligned for double, 8-byte for float) in ONE load
AFE_HD void ldg2(const double* p, double& a, double& b) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  const double2 v = __ldg(reinterpret_cast<const double2*>(p));
#else
  const double2 v = *reinterpret_cast<const double2*>(p);
#endif
  a = v.x; b = v.y;
}
AFE_HD void ldg2(const float* p, float& a, float& b) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  const float2 v = __ldg(reinterpret_cast<const float2*>(p));
#else
  const float2 v = *reinterpret_cast<const float2*>(p);
#endif
  a = v.x; b = v.y;
}

// kernels without per-item context
struct PlainItem { long item; int buf; int bulk; };
#define AFE_PLAIN_ITEM                                   \
  using Item = ::afe::PlainItem;                         \
  static constexpr bool kAsyncStage = false;             \
  AFE_HD Item prepare(long item) const { return Item{item, 0, 0}; }

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

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier helpers --------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "AFE_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra AFE_DONE;\n\t"
      "bra AFE_WAIT;\n\t"
      "AFE_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// Kernels with kAsyncStage provide
//   bool stage_src(const Item&, const float*& src, int& nfloats)   true -> the item's staging
//        (phase 0) is one contiguous 16-byte-aligned copy of nfloats (multiple of 4) floats
//   Item next(const Item& cur, long item)   the context of item `item` = cur.item + 1, derived
//        incrementally (no search: the next tile of the same waveform, or the next waveform)
//   kStageOff / kStageBytes: two staging buffers back to back at kStageOff; Item::buf picks one.
// A block owns a CONTIGUOUS range of items.  Thread 0 derives item i+1 from item i, publishes
// it through shared memory and issues its TMA copy while everybody computes item i; the wait
// on the copy's mbarrier is the only synchronisation before phase 1.  Items whose staging is
// not a plain copy (edges, unaligned pointers) run phase 0 cooperatively.
template <class K>
__global__ void __launch_bounds__(K::kThreads, K::kMinBlocks) phased_kernel(const K k) {
  extern __shared__ __align__(16) unsigned char afe_smem[];
  const long n = k.num_items();
  const int tid = (int)threadIdx.x;
  if constexpr (K::kAsyncStage) {
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ typename K::Item items[2];
    const long per = (n + gridDim.x - 1) / gridDim.x;
    long item = (long)blockIdx.x * per;
    const long end = (item + per < n) ? item + per : n;
    if (item >= end) return;
    if (tid == 0) {
      mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); fence_mbar_init();
      typename K::Item c0 = k.prepare(item);
      c0.buf = 0;
      const float* src; int nf;
      c0.bulk = k.stage_src(c0, src, nf) ? 1 : 0;
      items[0] = c0;
      if (c0.bulk) {
        mbar_expect_tx(&bars[0], (uint32_t)nf * 4u);
        bulk_g2s(afe_smem + K::kStageOff, src, (uint32_t)nf * 4u, &bars[0]);
      }
    }
    __syncthreads();
    uint32_t par[2] = {0u, 0u};
    for (int it = 0; item < end; ++item, ++it) {
      const int buf = it & 1;
      const typename K::Item cur = items[buf];
      if (tid == 0 && item + 1 < end) {
        typename K::Item nx = k.next(cur, item + 1);
        nx.buf = buf ^ 1;
        const float* src; int nf;
        nx.bulk = k.stage_src(nx, src, nf) ? 1 : 0;
        items[buf ^ 1] = nx;
        if (nx.bulk) {
          fence_proxy_async();            // earlier generic-proxy reads of that buffer are done (barriers)
          mbar_expect_tx(&bars[buf ^ 1], (uint32_t)nf * 4u);
          bulk_g2s(afe_smem + K::kStageOff + (size_t)(buf ^ 1) * K::kStageBytes, src, (uint32_t)nf * 4u, &bars[buf ^ 1]);
        }
      }
      if (cur.bulk) { mbar_wait(&bars[buf], par[buf]); par[buf] ^= 1u; }
      else { k.phase(0, cur, afe_smem, tid); __syncthreads(); }
#pragma unroll
      for (int ph = 1; ph < K::kPhases; ++ph) {
        k.phase(ph, cur, afe_smem, tid);
        __syncthreads();
      }
      if constexpr (K::kTailPhases > 0) {     // block-uniform extra phases (e.g. the last tile of a waveform)
        if (k.tail(cur)) {
#pragma unroll
          for (int ph = K::kPhases; ph < K::kPhases + K::kTailPhases; ++ph) {
            k.phase(ph, cur, afe_smem, tid);
            __syncthreads();
          }
        }
      }
    }
  } else {
    for (long item = blockIdx.x; item < n; item += gridDim.x) {
      const typename K::Item ic = k.prepare(item);
#pragma unroll
      for (int ph = 0; ph < K::kPhases; ++ph) {
        k.phase(ph, ic, afe_smem, tid);
        __syncthreads();
      }
    }
  }
}
// blocks of an async (contiguous-range) launch over n items; block i owns items [i*per, (i+1)*per),
// per = ceil(n / grid).  Host code that needs to know the ranges (fused normalisation) calls this too.
inline long afe_async_grid(long n, int max_blocks) { return n < (long)max_blocks ? n : (long)max_blocks; }
template <class K>
int afe_launch(const K& k, size_t smem_bytes, int max_blocks, cudaStream_t stream) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  const void* fn = (const void*)phased_kernel<K>;
  if (smem_bytes > 48 * 1024) {
    const cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) { note_error("cudaFuncSetAttribute(max dynamic smem)", cudaGetErrorString(e)); return -1; }
    // all of the SM's configurable L1/shared array as shared memory: kernels sized for several resident
    // blocks per SM (K::kMinBlocks) only get them with the maximum carve-out
    cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    static const bool debug = std::getenv("AFE_DEBUG_OCCUPANCY") != nullptr;
    if (debug) {
      int nb = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, K::kThreads, smem_bytes);
      std::fprintf(stderr, "[afe] %s: %d threads, %zu B dynamic smem, %d resident blocks/SM\n", __PRETTY_FUNCTION__, K::kThreads, smem_bytes, nb);
    }
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
// emulation: at most 3 "blocks", so ranges are long and next() crosses waveforms
inline long afe_async_grid(long n, int max_blocks) {
  long g = n < (long)max_blocks ? n : (long)max_blocks;
  return g > 3 ? 3 : g;
}
template <class K>
int afe_launch(const K& k, size_t smem_bytes, int max_blocks, cudaStream_t) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  std::vector<double> backing((smem_bytes + 64) / sizeof(double) + 2);   // 8-byte aligned, zeroed
  unsigned char* smem = reinterpret_cast<unsigned char*>(backing.data());
  long grid = n < (long)max_blocks ? n : (long)max_blocks;
  if constexpr (K::kAsyncStage) {
    grid = afe_async_grid(n, max_blocks);
    const long per = (n + grid - 1) / grid;              // contiguous ranges, items derived by next()
    for (long block = 0; block < grid; ++block) {
      const long first = block * per, end = first + per < n ? first + per : n;
      typename K::Item ic{};
      for (long item = first; item < end; ++item) {
        ic = (item == first) ? k.prepare(item) : k.next(ic, item);
        ic.buf = (int)((item - first) & 1);              // exercise both staging buffers
        ic.bulk = 0;
        for (int ph = 0; ph < K::kPhases; ++ph)
          for (int tid = 0; tid < K::kThreads; ++tid) k.phase(ph, ic, smem, tid);
        if constexpr (K::kTailPhases > 0) {
          if (k.tail(ic))
            for (int ph = K::kPhases; ph < K::kPhases + K::kTailPhases; ++ph)
              for (int tid = 0; tid < K::kThreads; ++tid) k.phase(ph, ic, smem, tid);
        }
      }
    }
  } else {
    for (long block = 0; block < grid; ++block)
      for (long item = block; item < n; item += grid) {
        typename K::Item ic = k.prepare(item);
        for (int ph = 0; ph < K::kPhases; ++ph)
          for (int tid = 0; tid < K::kThreads; ++tid) k.phase(ph, ic, smem, tid);
      }
  }
  return 0;
}
}  // namespace afe
#endif
