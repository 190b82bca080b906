// afe_runtime.cu -- memory / error plumbing and the afe_ctx lifecycle.
// Compiled by nvcc (product) and by g++ -x c++ -DAFE_EMULATE (test-only emulation).
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "afe_ctx.h"

namespace afe {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) { set_error(msg); return code; }
void note_error(const char* what, const char* detail) { set_error(std::string(what) + ": " + detail); }

#if AFE_CUDA
static int cuda_fail(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error(std::string(what) + ": " + cudaGetErrorString(e));
  return -1;
}
int dev_set(int device) { return cuda_fail(cudaSetDevice(device), "cudaSetDevice"); }
void* dev_alloc(size_t bytes) {
  void* p = nullptr;
  if (cuda_fail(cudaMalloc(&p, bytes), "cudaMalloc") != 0) return nullptr;
  return p;
}
void dev_free(void* p) { if (p) cudaFree(p); }
int dev_upload(void* dst, const void* src, size_t bytes) {
  return cuda_fail(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice), "cudaMemcpy H2D");
}
// Descriptor blocks travel in KERNEL PARAMETERS (through the launch queue), not through the
// copy engine: a small cudaMemcpyAsync on the compute stream queues behind -- and, while it
// waits for its own stream, blocks -- the bulk waveform uploads that run on another stream,
// which serialises upload and compute (measured: no overlap at all).  16 KiB per launch.
struct ParamBlob { uint4 q[1024]; };
__global__ void __launch_bounds__(256) blob_store_kernel(const __grid_constant__ ParamBlob blob, unsigned char* dst, int nbytes) {
  const int nq = nbytes >> 4;
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  for (int i = (int)threadIdx.x; i < nq; i += (int)blockDim.x) d4[i] = blob.q[i];
  const unsigned char* tail = reinterpret_cast<const unsigned char*>(blob.q);
  for (int i = (nq << 4) + (int)threadIdx.x; i < nbytes; i += (int)blockDim.x) dst[i] = tail[i];
}
int dev_upload_async(void* dst, const void* src, size_t bytes, cudaStream_t s) {
  if (bytes > (4u << 20) || (reinterpret_cast<uintptr_t>(dst) & 15u) != 0)
    return cuda_fail(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s), "cudaMemcpyAsync H2D");
  ParamBlob blob;
  const unsigned char* p = static_cast<const unsigned char*>(src);
  unsigned char* d = static_cast<unsigned char*>(dst);
  for (size_t off = 0; off < bytes; off += sizeof(ParamBlob)) {
    const size_t n = bytes - off < sizeof(ParamBlob) ? bytes - off : sizeof(ParamBlob);
    std::memcpy(&blob, p + off, n);
    blob_store_kernel<<<1, 256, 0, s>>>(blob, d + off, (int)n);
  }
  return cuda_fail(cudaGetLastError(), "descriptor upload launch");
}
int dev_memset_async(void* dst, int v, size_t bytes, cudaStream_t s) {
  return cuda_fail(cudaMemsetAsync(dst, v, bytes, s), "cudaMemsetAsync");
}
void* host_pinned_alloc(size_t bytes) {
  void* p = nullptr;
  if (cuda_fail(cudaMallocHost(&p, bytes), "cudaMallocHost") != 0) return nullptr;
  return p;
}
void host_pinned_free(void* p) { if (p) cudaFreeHost(p); }
int dev_check_last() { return cuda_fail(cudaGetLastError(), "kernel launch"); }
#else
int dev_set(int) { return 0; }
void* dev_alloc(size_t bytes) { return std::calloc(1, bytes ? bytes : 1); }
void dev_free(void* p) { std::free(p); }
int dev_upload(void* dst, const void* src, size_t bytes) { std::memcpy(dst, src, bytes); return 0; }
int dev_upload_async(void* dst, const void* src, size_t bytes, cudaStream_t) { std::memcpy(dst, src, bytes); return 0; }
int dev_memset_async(void* dst, int v, size_t bytes, cudaStream_t) { std::memset(dst, v, bytes); return 0; }
void* host_pinned_alloc(size_t bytes) { return std::malloc(bytes ? bytes : 1); }
void host_pinned_free(void* p) { std::free(p); }
int dev_check_last() { return 0; }
#endif
const char* dev_last_error_string() { return g_last_error.c_str(); }

}  // namespace afe

using namespace afe;

afe::StagingBlock* afe_ctx::staging_acquire(size_t bytes) {
  std::lock_guard<std::mutex> lock(mu);
  for (auto& b : staging) {
    if (b.cap < bytes || b.state == StagingBlock::kReserved) continue;
    if (b.state == StagingBlock::kSubmitted) {
#if AFE_CUDA
      if (cudaEventQuery((cudaEvent_t)b.event) != cudaSuccess) continue;
#endif
    }
    b.state = StagingBlock::kReserved;            // owned by this call until staging_submit
    return &b;
  }
  if (staging.capacity() < 256) staging.reserve(256);     // keep pointers stable
  if (staging.size() >= 256) { set_error("staging pool exhausted"); return nullptr; }
  StagingBlock nb;
  nb.cap = bytes < 65536 ? 65536 : bytes;
  nb.host = host_pinned_alloc(nb.cap);
  if (!nb.host) return nullptr;
#if AFE_CUDA
  cudaEvent_t ev;
  if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { host_pinned_free(nb.host); return nullptr; }
  nb.event = ev;
#endif
  nb.state = StagingBlock::kReserved;
  staging.push_back(nb);
  return &staging.back();
}

void afe_ctx::staging_submit(afe::StagingBlock* blk, cudaStream_t stream) {
  std::lock_guard<std::mutex> lock(mu);
#if AFE_CUDA
  cudaEventRecord((cudaEvent_t)blk->event, stream);
  blk->state = StagingBlock::kSubmitted;
#else
  (void)stream;
  blk->state = StagingBlock::kFree;
#endif
}

const afe::DeviceTaps* afe_ctx::get_taps(int orig_hz, int new_hz) {
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair(orig_hz, new_hz);
  auto it = taps.find(key);
  if (it != taps.end()) return &it->second;
  DeviceTaps dt;
  dt.host = make_resample_taps(orig_hz, new_hz);
  dt.d_tapsT = upload(dt.host.tapsT);
  if (!dt.d_tapsT) return nullptr;
  return &taps.emplace(key, std::move(dt)).first->second;
}

void* afe_ctx::prof_event() {
#if AFE_CUDA
  if (!prof_free.empty()) { void* e = prof_free.back(); prof_free.pop_back(); return e; }
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
  return e;
#else
  return nullptr;
#endif
}

namespace afe {
ProfScope::ProfScope(afe_ctx* ctx, int kernel_id, cudaStream_t stream) : c(ctx), id(kernel_id), s(stream) {
  std::lock_guard<std::mutex> lock(c->mu);
  c->prof_launches[id]++;
#if AFE_CUDA
  if (c->profiling) {
    void* e0 = c->prof_event();
    e1 = c->prof_event();
    if (e0 && e1) {
      cudaEventRecord((cudaEvent_t)e0, s);
      c->prof_recs.push_back({id, e0, e1});
    } else {
      e1 = nullptr;
    }
  }
#endif
}
ProfScope::~ProfScope() {
#if AFE_CUDA
  if (e1) cudaEventRecord((cudaEvent_t)e1, s);
#endif
}
}  // namespace afe

extern "C" {

static const char* kKernelNames[AFE_K_COUNT] = {"resample", "mix_stereo", "fbank", "fbank_stats", "fbank_norm",
                                                "acoustic_tile", "acoustic_max", "acoustic_mfcc", "acoustic_final"};
const char* afe_kernel_name(int id) { return (id >= 0 && id < AFE_K_COUNT) ? kKernelNames[id] : "?"; }

int afe_profile_enable(afe_ctx* ctx, int enabled) {
  if (!ctx) return fail(AFE_ERR_INVALID, "afe_profile_enable: ctx is NULL");
  std::lock_guard<std::mutex> lock(ctx->mu);
  ctx->profiling = enabled != 0;
  return AFE_OK;
}

int afe_profile_reset(afe_ctx* ctx) {
  if (!ctx) return fail(AFE_ERR_INVALID, "afe_profile_reset: ctx is NULL");
  std::lock_guard<std::mutex> lock(ctx->mu);
  for (auto& r : ctx->prof_recs) { ctx->prof_free.push_back(r.e0); ctx->prof_free.push_back(r.e1); }
  ctx->prof_recs.clear();
  for (int i = 0; i < AFE_K_COUNT; ++i) ctx->prof_launches[i] = 0;
  return AFE_OK;
}

int afe_profile_read(afe_ctx* ctx, double* h_ms, int64_t* h_launches) {
  if (!ctx || !h_ms || !h_launches) return fail(AFE_ERR_INVALID, "afe_profile_read: NULL argument");
  std::lock_guard<std::mutex> lock(ctx->mu);
  for (int i = 0; i < AFE_K_COUNT; ++i) { h_ms[i] = 0.0; h_launches[i] = ctx->prof_launches[i]; }
#if AFE_CUDA
  for (auto& r : ctx->prof_recs) {
    if (cudaEventSynchronize((cudaEvent_t)r.e1) != cudaSuccess) return fail(AFE_ERR_CUDA, "afe_profile_read: event sync failed");
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, (cudaEvent_t)r.e0, (cudaEvent_t)r.e1) != cudaSuccess)
      return fail(AFE_ERR_CUDA, "afe_profile_read: elapsed time failed");
    h_ms[r.id] += (double)ms;
  }
#endif
  return AFE_OK;
}

const char* afe_version(void) { return "afe-b200 0.1.0"; }
int afe_backend(void) { return AFE_CUDA ? 1 : 2; }
const char* afe_last_error(void) { return dev_last_error_string(); }
int afe_num_sms(const afe_ctx* ctx) { return ctx ? ctx->num_sms : 0; }

int afe_create(int device, afe_ctx** out) {
  if (!out) return fail(AFE_ERR_INVALID, "afe_create: out is NULL");
  *out = nullptr;
#if AFE_CUDA
  int prev = 0;
  cudaGetDevice(&prev);
  if (dev_set(device) != 0) return AFE_ERR_CUDA;
#endif
  afe_ctx* c = new afe_ctx();
  c->device = device;
#if AFE_CUDA
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && sms > 0) c->num_sms = sms;
#endif
  bool ok = fbank_static_setup(c) == 0;
  {
    // window tables are role-major for the FFT's stage 1: t[role][n1] = w[16 n1 + role] (afe_rfft512.h)
    std::vector<double> w = make_povey_window();
    std::vector<double> w64(16 * 26, 0.0); std::vector<float> w32(16 * 26, 0.f);
    for (size_t i = 0; i < w.size(); ++i) {
      const size_t j = (i % 16) * 26 + i / 16;
      w64[j] = w[i] * 16384.0; w32[j] = (float)(w[i] * 16384.0);                                          // 32768 / 2
    }
    ok = ok && (c->d_povey64 = c->upload(w64)) && (c->d_povey32 = c->upload(w32));
    std::vector<double> tw = make_twiddle512();
    std::vector<float> tw32(tw.begin(), tw.end());
    ok = ok && (c->d_tw64 = c->upload(tw)) && (c->d_tw32 = c->upload(tw32));
    SparseBank mel = make_kaldi_mel_bank();
    ok = ok && (c->d_mel_start = c->upload(mel.start)) && (c->d_mel_count = c->upload(mel.count)) &&
         (c->d_mel_offset = c->upload(mel.offset)) && (c->d_mel_w = c->upload(mel.w));
    { const MelSlots ms = make_mel_slots(mel, kKaldiSlotLen, 5);
      if (!ms.ok) note_error("afe_create", "Kaldi mel bank does not fit the kernels' slot lengths");
      ok = ok && ms.ok && (c->kaldi_slots.w = c->upload(ms.w)) && (c->kaldi_slots.start = c->upload(ms.start)); }
    std::vector<double> hn = make_hann_periodic();
    std::vector<float> hn32(16 * 26, 0.f);
    for (size_t i = 0; i < hn.size(); ++i) hn32[(i % 16) * 26 + i / 16] = (float)(hn[i] * 0.5);
    ok = ok && (c->d_hann32 = c->upload(hn32));
    SparseBank htk = make_htk_mel_bank();
    ok = ok && (c->d_htk_start = c->upload(htk.start)) && (c->d_htk_count = c->upload(htk.count)) &&
         (c->d_htk_offset = c->upload(htk.offset)) && (c->d_htk_w = c->upload(htk.w));
    { const MelSlots ms = make_mel_slots(htk, kHtkSlotLen, 8);
      if (!ms.ok) note_error("afe_create", "HTK mel bank does not fit the kernels' slot lengths");
      ok = ok && ms.ok && (c->htk_slots.w = c->upload(ms.w)) && (c->htk_slots.start = c->upload(ms.start)); }
    c->h_dct = make_dct_matrix();
    ok = ok && (c->d_dct = c->upload(c->h_dct));
    c->log_floor = (float)std::log((double)kMelFloor);
  }
#if AFE_CUDA
  cudaSetDevice(prev);
#endif
  if (!ok) { afe_destroy(c); return AFE_ERR_CUDA; }
  *out = c;
  return AFE_OK;
}

void afe_destroy(afe_ctx* c) {
  if (!c) return;
  for (void* p : c->owned) dev_free(p);
#if AFE_CUDA
  for (auto& r : c->prof_recs) { cudaEventDestroy((cudaEvent_t)r.e0); cudaEventDestroy((cudaEvent_t)r.e1); }
  for (void* e : c->prof_free) cudaEventDestroy((cudaEvent_t)e);
#endif
  for (auto& b : c->staging) {
#if AFE_CUDA
    if (b.event) cudaEventDestroy((cudaEvent_t)b.event);
#endif
    host_pinned_free(b.host);
  }
  delete c;
}

}  // extern "C"
