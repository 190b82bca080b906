// afe_fbank_kernels.h -- Kaldi/SeamlessM4T fbank as phased block programs (afe_platform.h).
//
// Replaces the per-frame numpy loop of transformers audio_utils.spectrogram (:788-804) as
// called through SeamlessM4TFeatureExtractor (feature_extraction_seamless_m4t.py:112-139)
// from services/audio_processor.py:253-261 and services/speech_segment.py:226-231.
//
// One block = 16 roles (warps) x 32 frames (lanes); one work item = a tile of 32 consecutive
// frames of one utterance (5360 samples).  Linear path (pre-emphasis, DC removal, window,
// FFT) in T; power / mel / log in float.
#pragma once
#include "afe_mel_gen.h"
#include "afe_rfft512.h"
#include "afe_tables.h"

namespace afe {

constexpr int kTileSamples = (kTileFrames - 1) * kHop + kFrame;     // 5360
constexpr int kGroups16 = kTileSamples / 16;                        // 335 sums of 16 samples
constexpr int kOutStride = kMel + 1;                                // 81: conflict-free rows

AFE_HD int phys_idx(int i) { return i + (i >> 5); }                 // 1 pad word per 32: odd lane stride

struct BatchDesc {            // device-resident description of a batch of mono waveforms
  const float* const* wave;   // [B] device pointers
  const int64_t* len;         // [B] samples
  const int32_t* frames;      // [B] fbank frames T_b
  const int64_t* tile_prefix; // [B+1] exclusive prefix of ceil(T_b / 32)
  const int64_t* out_row;     // [B] first frame row of utterance b in the feature buffer
  const int32_t* tp;          // [B] padded frame count of utterance b's group (even)
  int B;
  long ntiles;
};

// locate (utterance, tile-in-utterance) of a global tile index: binary search over the prefix
AFE_HD void locate_tile(const BatchDesc& d, long item, int& b, int& t) {
  int lo = 0, hi = d.B;       // invariant: prefix[lo] <= item < prefix[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (ldg(d.tile_prefix + mid) <= item) lo = mid; else hi = mid;
  }
  b = lo;
  t = (int)(item - ldg(d.tile_prefix + lo));
}

// interval-form triangular bank on the device (afe_tables.h SparseBank::up/dn/iv_*)
struct MelTables {
  const float* up; const float* dn;               // [257]
  const int32_t* iv_start; const int32_t* iv_count;   // [F+1]
  const int32_t* role_iv;                         // [17]
};

constexpr int kMelIv = kMel + 1;                  // 81 intervals

template <typename T> struct FbankSmem {
  static constexpr size_t kRawBytes = sizeof(float) * kTileSamples;                 // 21440 per staging buffer
  static constexpr size_t kRawOff = 0;                                              // float[2][5360]
  static constexpr size_t kSumOff = kRawOff + 2 * kRawBytes;                        // double[336]
  static constexpr size_t kDOff = kSumOff + sizeof(double) * 336;                   // T[5528] | sP + sOut
  static constexpr size_t kDBytesT = sizeof(T) * (kTileSamples + kTileSamples / 32 + 8);
  static constexpr size_t kPBytes = sizeof(float) * (kBins * kTileFrames);          // 32896
  static constexpr size_t kOutBytes = sizeof(float) * (kTileFrames * kOutStride);   // 10368
  static constexpr size_t kDBytes = (kDBytesT > kPBytes + kOutBytes ? kDBytesT : kPBytes + kOutBytes);
  static constexpr size_t kExOff = (kDOff + kDBytes + 15) / 16 * 16;                // cplx<T>[16*16*32] | sU,sD
  static constexpr size_t kExBytesT = sizeof(cplx<T>) * 16 * 16 * kTileFrames;
  static constexpr size_t kUDBytes = sizeof(float) * 2 * kMelIv * kTileFrames;      // 20736
  static constexpr size_t kBytes = kExOff + (kExBytesT > kUDBytes ? kExBytesT : kUDBytes);
};

template <typename T> struct FbankKernel {
  static constexpr int kThreads = kRoles * kTileFrames;   // 512
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 7;
  static constexpr bool kAsyncStage = true;
  static constexpr size_t kStageOff = FbankSmem<T>::kRawOff;
  static constexpr size_t kStageBytes = FbankSmem<T>::kRawBytes;

  struct Item { long item; int buf; int bulk; int b; int f0; int nvalid; int count; const float* src; int64_t out_row; };

  BatchDesc d;
  float* out;              // [rows][80] raw log-mel lands at its final position
  double* partials;        // [ntiles][80][2]  (sum, sum of squares) per mel bin
  const T* window;         // [400]  povey * 32768 / 2
  const T* tw512;          // [16][16][2]
  MelTables mel;
  float log_floor;         // float32(log(float64(mel_floor)))

  AFE_HD long num_items() const { return d.ntiles; }

  AFE_HD Item prepare(long item) const {
    Item it;
    int t;
    locate_tile(d, item, it.b, t);
    it.item = item; it.buf = 0; it.bulk = 0;
    it.f0 = t * kTileFrames;
    it.nvalid = tmin(kTileFrames, ldg(d.frames + it.b) - it.f0);
    it.count = (it.nvalid - 1) * kHop + kFrame;
    it.src = ldg(d.wave + it.b) + (int64_t)it.f0 * kHop;
    it.out_row = ldg(d.out_row + it.b);
    return it;
  }
  // the tile is one contiguous run of `count` samples: a single TMA bulk copy when 16-byte aligned
  AFE_HD bool stage_src(const Item& it, const float*& src, int& nfloats) const {
    src = it.src; nfloats = it.count;
    return (reinterpret_cast<uintptr_t>(it.src) & 15u) == 0;
  }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    using S = FbankSmem<T>;
    float* raw = reinterpret_cast<float*>(smem + S::kRawOff + (size_t)it.buf * S::kRawBytes);
    double* sum16 = reinterpret_cast<double*>(smem + S::kSumOff);
    T* dbuf = reinterpret_cast<T*>(smem + S::kDOff);
    float* sP = reinterpret_cast<float*>(smem + S::kDOff);
    float* sOut = reinterpret_cast<float*>(smem + S::kDOff + S::kPBytes);
    cplx<T>* ex = reinterpret_cast<cplx<T>*>(smem + S::kExOff);
    float* sU = reinterpret_cast<float*>(smem + S::kExOff);
    float* sD = sU + kMelIv * kTileFrames;
    const int role = tid >> 5, f = tid & 31;

    switch (ph) {
      case 0: {   // cooperative staging (the CUDA driver replaces it by a TMA bulk copy when it can)
        for (int i = tid; i < it.count; i += kThreads) raw[i] = it.src[i];
      } break;
      case 1: {   // pre-emphasis in T into the padded layout + 16-sample partial sums
        if (it.count == kTileSamples) {          // full tile: no bounds checks
          for (int i = tid; i < kTileSamples; i += kThreads) {
            const T x = (T)raw[i];
            const T xm = (i > 0) ? (T)raw[i - 1] : T(0);
            dbuf[phys_idx(i)] = x - T(0.97) * xm;
          }
        } else {
          for (int i = tid; i < kTileSamples; i += kThreads) {
            const T x = (i < it.count) ? (T)raw[i] : T(0);
            const T xm = (i > 0 && i - 1 < it.count) ? (T)raw[i - 1] : T(0);
            dbuf[phys_idx(i)] = x - T(0.97) * xm;
          }
        }
        if (tid < kGroups16) {
          double s = 0.0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int i = tid * 16 + ((j + tid) & 15);          // rotated start: 2-way instead of 16-way conflicts
            s += (i < it.count) ? (double)raw[i] : 0.0;
          }
          sum16[tid] = s;
        }
      } break;
      case 2: {   // DC removal + window + stage 1 of the real FFT
        double s = 0.0;
#pragma unroll
        for (int g = 0; g < 25; ++g) s += sum16[10 * f + g];
        const T c = (T(1) - T(0.97)) * (T)(s / (double)kFrame);
        T r[25];
#pragma unroll
        for (int n1 = 0; n1 < 25; ++n1) {
          const int n = 16 * n1 + role;
          r[n1] = ldg(window + n) * (dbuf[phys_idx(kHop * f + n)] - c);
        }
        rfft512_stage1<T>(r, role, f, tw512, ex);
      } break;
      case 3:
        rfft512_stage2<T>(role, f, ex, sP);
        break;
      case 4:     // triangular mel, interval form, generated straight-line code (afe_mel_gen.h)
        mel_intervals_kaldi(role, f, sP, sU, sD);
        break;
      case 5: {   // filter m = U[m] + D[m+1]; floor; log
#pragma unroll
        for (int m = role; m < kMel; m += kRoles) {
          const float acc = sU[m * kTileFrames + f] + sD[(m + 1) * kTileFrames + f];
          sOut[f * kOutStride + m] = (acc <= kMelFloor) ? log_floor : fast_log(acc);
        }
      } break;
      case 6: {   // per-bin partial statistics (float64) + coalesced store of the raw log-mel
        if (tid < kMel) {
          double s = 0.0, ss = 0.0;
          for (int j = 0; j < it.nvalid; ++j) {
            const double v = (double)sOut[j * kOutStride + tid];
            s += v; ss += v * v;
          }
          partials[(it.item * kMel + tid) * 2] = s;
          partials[(it.item * kMel + tid) * 2 + 1] = ss;
        }
        float* dst = out + (it.out_row + it.f0) * kMel;
        for (int i = tid; i < it.nvalid * kMel; i += kThreads) dst[i] = sOut[(i / kMel) * kOutStride + (i % kMel)];
      } break;
    }
  }
};

// mean / 1/sqrt(var+1e-7) per (utterance, mel bin): deterministic in-order reduction of the
// tile partials.  One item = one utterance, 128 threads (80 active).
struct FbankStatsKernel {
  static constexpr int kThreads = 128;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 1;
  BatchDesc d;
  const double* partials;
  float* stats;            // [B][80][2] = (mean, denominator sqrt(var + 1e-7))
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    if (tid >= kMel) return;
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    const int T_b = ldg(d.frames + item);
    double s = 0.0, ss = 0.0;
    for (long t = t0; t < t1; ++t) { s += partials[(t * kMel + tid) * 2]; ss += partials[(t * kMel + tid) * 2 + 1]; }
    const double mean = s / (double)T_b;
    const double var = (ss - s * mean) / (double)(T_b - 1);      // ddof = 1; T_b == 1 -> 0/0 = NaN
    const float varf = (float)var;
    stats[(item * kMel + tid) * 2] = (float)mean;
    stats[(item * kMel + tid) * 2 + 1] = std::sqrt(varf + 1e-7f);
  }
};

// in-place normalisation, right padding and the stride-2 attention mask.
// One item = 2048 consecutive floats of one utterance's (Tp x 80) plane.
struct FbankNormKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 4;
  static constexpr int kPhases = 1;
  static constexpr int kItemFloats = 2048;
  BatchDesc d;
  float* feat;             // [rows][80]
  int32_t* mask;           // [rows/2]
  const float* stats;
  float padding_value;
  long items_per_utt;      // ceil(max Tp * 80 / 2048)
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return items_per_utt * d.B; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    const int b = (int)(item / items_per_utt);
    const long chunk = item % items_per_utt;
    const int T_b = ldg(d.frames + b);
    const int64_t plane = (int64_t)ldg(d.tp + b) * kMel;
    const int64_t row0 = ldg(d.out_row + b);
    float* base = feat + row0 * kMel;
    for (int v = tid; v < kItemFloats; v += kThreads) {
      const int64_t idx = chunk * kItemFloats + v;
      if (idx >= plane) break;
      const int64_t t = idx / kMel;
      const int m = (int)(idx % kMel);
      float y = padding_value;
      if (t < T_b) y = (base[idx] - ldg(stats + ((int64_t)b * kMel + m) * 2)) / ldg(stats + ((int64_t)b * kMel + m) * 2 + 1);
      base[idx] = y;
      // the stacked mask takes the ODD frame's bit: mask[b][j] = (2j+1 < T_b)
      if (m == 0 && (t & 1)) mask[(row0 >> 1) + (t >> 1)] = (t < T_b) ? 1 : 0;
    }
  }
};

}  // namespace afe
