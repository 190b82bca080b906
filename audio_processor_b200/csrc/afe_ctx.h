// afe_ctx.h -- internal: the context object behind the C ABI (include/afe.h).
#pragma once
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/afe.h"
#include "afe_platform.h"
#include "afe_tables.h"

namespace afe {

void set_error(const std::string& msg);          // thread-local message for afe_last_error
int fail(int code, const std::string& msg);      // set_error + return code

struct DeviceTaps {                              // one resampler bank resident on the device
  ResampleTaps host;
  float* d_tapsT = nullptr;                      // [K][N]
};

// Pinned host staging for per-call descriptor uploads.  A block is reusable once the event
// recorded after its H2D copy has completed; the pool grows on demand (re-entrant).
struct StagingBlock {
  void* host = nullptr;
  size_t cap = 0;
  void* event = nullptr;                         // cudaEvent_t
  // kFree -> kReserved (a caller is filling it; its event is still the previous, completed one and must not be
  // consulted) -> kSubmitted (event recorded after the copy: reusable once the event has completed) -> ...
  enum State { kFree = 0, kReserved = 1, kSubmitted = 2 };
  int state = kFree;
};

}  // namespace afe

namespace afe { struct ProfScope; }

struct afe_ctx {
  int device = 0;
  int num_sms = afe::kNumSMsB200;
  // fbank tables
  double* d_povey64 = nullptr;  float* d_povey32 = nullptr;     // [400] povey * 32768 / 2
  double* d_tw64 = nullptr;     float* d_tw32 = nullptr;        // [16][16][2]
  int32_t *d_mel_start = nullptr, *d_mel_count = nullptr, *d_mel_offset = nullptr;
  float* d_mel_w = nullptr;
  // acoustic tables
  float* d_hann32 = nullptr;                                    // [400] hann_periodic / 2
  int32_t *d_htk_start = nullptr, *d_htk_count = nullptr, *d_htk_offset = nullptr;
  float* d_htk_w = nullptr;
  float* d_dct = nullptr;                                       // [13][128]
  std::vector<float> h_dct;                                     // host copy: rides in the MFCC kernel's parameters
  float log_floor = 0.f;
  struct MelSlotPtr { const float* w; const int32_t* start; };   // slot-form banks (afe_mel_slots.h): [16][40], [16][8]
  MelSlotPtr kaldi_slots{}, htk_slots{};
  std::mutex mu;
  std::map<std::pair<int, int>, afe::DeviceTaps> taps;          // (orig, new) -> bank
  std::vector<afe::StagingBlock> staging;
  std::vector<void*> owned;                                     // device allocations to free
  void* fbank_static_scratch = nullptr;                         // device address of afe_fbank.cu's static scratch array
  // optional per-kernel timing (afe_profile_*): (kernel id, start event, stop event)
  bool profiling = false;
  struct ProfRec { int id; void* e0; void* e1; };
  std::vector<ProfRec> prof_recs;
  std::vector<void*> prof_free;                                 // recycled events
  int64_t prof_launches[AFE_K_COUNT] = {0};
  void* prof_event();

  // returns a pinned block of >= bytes (nullptr on failure); call staging_submit after the copy
  afe::StagingBlock* staging_acquire(size_t bytes);
  void staging_submit(afe::StagingBlock* blk, cudaStream_t stream);
  const afe::DeviceTaps* get_taps(int orig_hz, int new_hz);
  template <typename T> T* upload(const std::vector<T>& v) {
    T* p = static_cast<T*>(afe::dev_alloc(sizeof(T) * (v.empty() ? 1 : v.size())));
    if (!p) return nullptr;
    owned.push_back(p);
    if (!v.empty() && afe::dev_upload(p, v.data(), sizeof(T) * v.size()) != 0) return nullptr;
    return p;
  }
};

namespace afe {
int fbank_static_setup(afe_ctx* ctx);            // afe_fbank.cu: cudaGetSymbolAddress of the static scratch (device current)
// brackets one kernel launch with events when profiling is on; always counts the launch
struct ProfScope {
  afe_ctx* c; int id; cudaStream_t s; void* e1 = nullptr;
  ProfScope(afe_ctx* ctx, int kernel_id, cudaStream_t stream);
  ~ProfScope();
};
}  // namespace afe
