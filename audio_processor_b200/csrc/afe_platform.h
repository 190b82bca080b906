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
#include <utility>
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
struct alignas(16) int4 { int x, y, z, w; };
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
// a / b to ~2 ulp (one reciprocal + one multiply): per-frame scalars whose contract is 1e-4 relative
AFE_HD float fast_div(float a, float b) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __fdividef(a, b);
#else
  return a / b;
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

// exact float -> double of a value that is >= 0 (magnitudes, powers): three integer instructions instead of
// the quarter-rate F2F conversion (8.5+ cycles of the XU pipe per warp instruction on the B200).
// Zero, infinity and NaN map like the cast; denormals do not occur (producers flush them).
AFE_HD double nonneg_to_double(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  const unsigned b = __float_as_uint(x);
  unsigned hi = (b >> 3) + 0x38000000u;                  // exponent rebias 127 -> 1023, top 20 mantissa bits
  hi = (b >= 0x7f800000u) ? (hi | 0x7ff00000u) : hi;     // inf / NaN keep an all-ones exponent
  hi = b ? hi : 0u;
  return __hiloint2double((int)hi, (int)(b << 29));
#else
  return (double)x;
#endif
}

// The same in two integer instructions for accumulations that tolerate +0 -> 2^-127 (5.9e-39): no zero, infinity
// or NaN cases.  The caller tracks the largest bit pattern it converted and poisons its sums if that was not finite.
AFE_HD double nonneg_to_double_biased(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  const unsigned b = __float_as_uint(x);
  return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
#else
  return x == 0.f ? 0x1p-127 : (double)x;
#endif
}
AFE_HD float bits_float(unsigned b) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __uint_as_float(b);
#else
  float x; std::memcpy(&x, &b, 4); return x;
#endif
}
AFE_HD unsigned float_bits(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __float_as_uint(x);
#else
  unsigned b; std::memcpy(&b, &x, 4); return b;
#endif
}

// sum / min / max over the 32 lanes of a warp as the fixed binary tree ((v0+v1)+(v2+v3))+...; every lane
// gets the result.  tree32_* are the host-emulation twins over an array of the 32 lane values.
#if AFE_CUDA && defined(__CUDA_ARCH__)
__device__ __forceinline__ double warp_tree_sum(double v) {
#pragma unroll
  for (int w = 1; w < 32; w <<= 1) v += __shfl_xor_sync(0xffffffffu, v, w);
  return v;
}
__device__ __forceinline__ double warp_tree_min(double v) {
#pragma unroll
  for (int w = 1; w < 32; w <<= 1) { const double o = __shfl_xor_sync(0xffffffffu, v, w); v = o < v ? o : v; }
  return v;
}
__device__ __forceinline__ double warp_tree_max(double v) {
#pragma unroll
  for (int w = 1; w < 32; w <<= 1) { const double o = __shfl_xor_sync(0xffffffffu, v, w); v = o > v ? o : v; }
  return v;
}
#endif
// 16-lane segments of a warp (two per warp): inclusive Hillis-Steele scan and xor-butterfly reductions.
// The array forms below run the same additions in the same order (host emulation twins).
#if AFE_CUDA
__device__ __forceinline__ double seg16_scan(double v, int r) {
#pragma unroll
  for (int d = 1; d < 16; d <<= 1) { const double t = __shfl_up_sync(0xffffffffu, v, d, 16); if (r >= d) v += t; }
  return v;
}
__device__ __forceinline__ double seg16_sum(double v) {
#pragma unroll
  for (int w = 1; w < 16; w <<= 1) v += __shfl_xor_sync(0xffffffffu, v, w, 16);
  return v;
}
__device__ __forceinline__ float seg16_sum(float v) {
#pragma unroll
  for (int w = 1; w < 16; w <<= 1) v += __shfl_xor_sync(0xffffffffu, v, w, 16);
  return v;
}
__device__ __forceinline__ float seg16_max(float v) {
#pragma unroll
  for (int w = 1; w < 16; w <<= 1) { const float o = __shfl_xor_sync(0xffffffffu, v, w, 16); v = o > v ? o : v; }
  return v;
}
#endif
template <typename T> AFE_HD void scan16(T (&a)[16]) {
  for (int d = 1; d < 16; d <<= 1)
    for (int j = 15; j >= d; --j) a[j] = a[j] + a[j - d];
}
template <typename T> AFE_HD T tree16_sum(T (&a)[16]) {
  for (int w = 1; w < 16; w <<= 1)
    for (int i = 0; i < 16; i += 2 * w) a[i] = a[i] + a[i + w];
  return a[0];
}

AFE_HD double tree32_sum(double (&a)[32]) {
  for (int w = 1; w < 32; w <<= 1)
    for (int i = 0; i < 32; i += 2 * w) a[i] = a[i] + a[i + w];
  return a[0];
}

// a value every lane of the warp holds (e.g. the warp's role): broadcasting it from lane 0 tells the
// compiler so, and loads indexed by it from the kernel parameters become uniform constant loads
AFE_HD int warp_uniform(int v) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  return __shfl_sync(0xffffffffu, v, 0);
#else
  return v;
#endif
}

// optional per-block prologue  void block_init(unsigned char* smem, int tid) const  (tables -> shared memory)
template <class K, class = void> struct has_block_init { static constexpr bool value = false; };
template <class K>
struct has_block_init<K, decltype(std::declval<const K&>().block_init((unsigned char*)nullptr, 0))> {
  static constexpr bool value = true;
};

// optional per-block epilogue  void block_flush(unsigned char* smem, int tid) const  -- work a kernel deferred from its last
// item into "the next item's first phase" (AcousticTileKernel: the tile record), run once after the block's last item
template <class K, class = void> struct has_block_flush { static constexpr bool value = false; };
template <class K>
struct has_block_flush<K, decltype(std::declval<const K&>().block_flush((unsigned char*)nullptr, 0))> {
  static constexpr bool value = true;
};

// the thread that derives the next item and issues its TMA copy while the block works on the current
// one: K::kPrefetchThread if the kernel names a thread whose warp idles in phase 1, else thread 0
// optional two-segment staging  bool stage_src2(const Item&, const float* (&src)[2], int (&nfloats)[2], int (&dst_float_off)[2]) const
// (e.g. the two rows of a stereo tile): both copies complete on the item's one mbarrier
template <class K, class = void> struct has_stage2 { static constexpr bool value = false; };
template <class K> struct has_stage2<K, decltype((void)K::kStageSegments)> { static constexpr bool value = true; };
template <class K, class = void> struct prefetch_thread { static constexpr int value = 0; };
template <class K> struct prefetch_thread<K, decltype((void)K::kPrefetchThread)> { static constexpr int value = K::kPrefetchThread; };

// kernels without per-item context
struct PlainItem { long item; int buf; int bulk; };
#define AFE_PLAIN_ITEM                                   \
  using Item = ::afe::PlainItem;                         \
  static constexpr bool kAsyncStage = false;             \
  AFE_HD Item prepare(long item) const { return Item{item, 0, 0}; }

template <typename T> AFE_HD T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> AFE_HD T tmax(T a, T b) { return a > b ? a : b; }

// ---------------------------------------------------------------- NVTX range of a C-ABI entry point
// (the reference has loguru lines only; a timeline tool shows these ranges around the launches)
#if AFE_CUDA
}  // namespace afe
#include <nvtx3/nvToolsExt.h>
namespace afe {
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
#else
struct NvtxRange { explicit NvtxRange(const char*) {} };
#endif
#define AFE_NVTX(name) ::afe::NvtxRange afe_nvtx_range_(name)

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
// kTrace (debug, AFE_PHASE_TRACE=1): thread 0 adds the cycles between consecutive barriers to
// trace[block * 16 + phase] (slot 0 = wait for the staged tile, slot 15 = items processed)
// sets it.bulk and, when the item's staging is a plain copy, arms the mbarrier and issues the TMA bulk copy / copies
template <class K>
__device__ __forceinline__ void afe_issue_stage(const K& k, typename K::Item& it, unsigned char* buf, uint64_t* bar) {
  if constexpr (has_stage2<K>::value) {
    const float* src[2]; int nf[2], off[2];
    it.bulk = k.stage_src2(it, src, nf, off) ? 1 : 0;
    if (it.bulk) {
      mbar_expect_tx(bar, (uint32_t)(nf[0] + nf[1]) * 4u);
      bulk_g2s(buf + (size_t)off[0] * 4u, src[0], (uint32_t)nf[0] * 4u, bar);
      bulk_g2s(buf + (size_t)off[1] * 4u, src[1], (uint32_t)nf[1] * 4u, bar);
    }
  } else {
    const float* src; int nf;
    it.bulk = k.stage_src(it, src, nf) ? 1 : 0;
    if (it.bulk) {
      mbar_expect_tx(bar, (uint32_t)nf * 4u);
      bulk_g2s(buf, src, (uint32_t)nf * 4u, bar);
    }
  }
}

template <class K, bool kTrace = false>
__global__ void __launch_bounds__(K::kThreads, K::kMinBlocks) phased_kernel(const K k, unsigned long long* trace = nullptr) {
  extern __shared__ __align__(16) unsigned char afe_smem[];
  const long n = k.num_items();
  const int tid = (int)threadIdx.x;
  if constexpr (has_block_init<K>::value) { k.block_init(afe_smem, tid); __syncthreads(); }
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
      afe_issue_stage(k, c0, afe_smem + K::kStageOff, &bars[0]);
      items[0] = c0;
    }
    __syncthreads();
    uint32_t parbits = 0u;                 // bit b = phase parity of staging buffer b's mbarrier
    long long tacc[24] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // kTrace: [0..7] own work, [8..15] work + barrier wait, [16] items
    for (int it = 0; item < end; ++item, ++it) {
      const int buf = it & 1;
      long long t0 = 0, tp = 0;
      if constexpr (kTrace) tp = clock64();
      const typename K::Item cur = items[buf];
      if (tid == prefetch_thread<K>::value && item + 1 < end) {
        typename K::Item nx = k.next(cur, item + 1);
        nx.buf = buf ^ 1;
        // the buffer's last generic-proxy accesses (phase 1 reads, or the phase 0 stores of an unaligned tile) were
        // ordered before this point by the block barriers; the proxy fence orders them before the async-proxy write
        fence_proxy_async();
        afe_issue_stage(k, nx, afe_smem + K::kStageOff + (size_t)(buf ^ 1) * K::kStageBytes, &bars[buf ^ 1]);
        items[buf ^ 1] = nx;
      }
      if (cur.bulk) { mbar_wait(&bars[buf], (parbits >> buf) & 1u); parbits ^= 1u << buf; }
      else { k.phase(0, cur, afe_smem, tid); __syncthreads(); }
      if constexpr (kTrace) { t0 = clock64(); tacc[0] += t0 - tp; tacc[16] += 1; }
#pragma unroll
      for (int ph = 1; ph < K::kPhases; ++ph) {
        k.phase(ph, cur, afe_smem, tid);
        if constexpr (kTrace) { tacc[ph] += clock64() - t0; }            // this warp's own work in the phase
        __syncthreads();
        if constexpr (kTrace) { const long long t1 = clock64(); tacc[8 + (ph < 7 ? ph : 7)] += t1 - t0; t0 = t1; }   // work + barrier wait
      }
      if constexpr (K::kTailPhases > 0) {     // block-uniform extra phases (e.g. the last tile of a waveform)
        if (k.tail(cur)) {
#pragma unroll
          for (int ph = K::kPhases; ph < K::kPhases + K::kTailPhases; ++ph) {
            k.phase(ph, cur, afe_smem, tid);
            __syncthreads();
            if constexpr (kTrace) { const long long t1 = clock64(); tacc[7] += t1 - t0; tacc[15] += t1 - t0; t0 = t1; }
          }
        }
      }
    }
    if constexpr (has_block_flush<K>::value) k.block_flush(afe_smem, tid);     // the last item's barrier has been passed
    if constexpr (kTrace) {     // lane 0 of every warp: trace[(block * 32 + warp) * 16 + slot]
      if ((tid & 31) == 0) for (int i = 0; i < 24; ++i) trace[((size_t)blockIdx.x * 32 + (tid >> 5)) * 24 + i] = (unsigned long long)tacc[i];
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
  if constexpr (K::kAsyncStage) {
    static const bool tracing = std::getenv("AFE_PHASE_TRACE") != nullptr;
    if (tracing) {      // debug only: synchronous, prints the mean cycles per item each phase took (per block, thread 0)
      const void* tfn = (const void*)phased_kernel<K, true>;
      cudaFuncSetAttribute(tfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
      cudaFuncSetAttribute(tfn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
      unsigned long long* tr = nullptr;
      const size_t nslots = (size_t)grid * 32 * 24;
      cudaMallocManaged(&tr, sizeof(unsigned long long) * nslots);
      std::memset(tr, 0, sizeof(unsigned long long) * nslots);
      phased_kernel<K, true><<<(unsigned)grid, K::kThreads, smem_bytes, stream>>>(k, tr);
      cudaStreamSynchronize(stream);
      // per phase: this-warp work (min / mean / max over the block's warps) and wall = work + barrier wait (warp 0)
      const int nw = K::kThreads / 32;
      double wmin[8] = {0}, wmean[8] = {0}, wmax[8] = {0}, wall[8] = {0}, items = 0;
      for (long b = 0; b < grid; ++b) {
        const double it_b = (double)tr[(size_t)b * 32 * 24 + 16];
        items += it_b;
        for (int p = 0; p < 8; ++p) {
          double mn = 1e30, mx = 0, sm = 0;
          for (int w = 0; w < nw; ++w) { const double v = (double)tr[((size_t)b * 32 + w) * 24 + p]; mn = v < mn ? v : mn; mx = v > mx ? v : mx; sm += v; }
          wmin[p] += mn; wmax[p] += mx; wmean[p] += sm / nw;
          wall[p] += (double)tr[((size_t)b * 32) * 24 + 8 + p];
        }
      }
      std::fprintf(stderr, "[afe trace] %s: %ld blocks, %.0f items; cycles per item, phase: warp work min/mean/max -> wall\n", __PRETTY_FUNCTION__, grid, items);
      double tot = wmean[0] / items;
      std::fprintf(stderr, "[afe trace]   prologue+wait %.0f/%.0f/%.0f\n", wmin[0] / items, wmean[0] / items, wmax[0] / items);
      for (int p = 1; p < 8; ++p) {
        if (wall[p] == 0) continue;
        std::fprintf(stderr, "[afe trace]   phase %d%s: %.0f/%.0f/%.0f -> %.0f\n", p, (p == 7 && K::kTailPhases > 0) ? " (tail)" : "", wmin[p] / items, wmean[p] / items, wmax[p] / items, wall[p] / items);
        tot += wall[p] / items;
      }
      std::fprintf(stderr, "[afe trace]   total %.0f\n", tot);
      cudaFree(tr);
      return 0;
    }
  }
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
      if constexpr (has_block_init<K>::value)
        for (int tid = 0; tid < K::kThreads; ++tid) k.block_init(smem, tid);
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
      if constexpr (has_block_flush<K>::value)
        for (int tid = 0; tid < K::kThreads; ++tid) k.block_flush(smem, tid);
    }
  } else {
    for (long block = 0; block < grid; ++block) {
      if constexpr (has_block_init<K>::value)
        for (int tid = 0; tid < K::kThreads; ++tid) k.block_init(smem, tid);
      for (long item = block; item < n; item += grid) {
        typename K::Item ic = k.prepare(item);
        for (int ph = 0; ph < K::kPhases; ++ph)
          for (int tid = 0; tid < K::kThreads; ++tid) k.phase(ph, ic, smem, tid);
      }
    }
  }
  return 0;
}
}  // namespace afe
#endif
