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

template <typename T> struct FbankSmem {
  static constexpr size_t kRawOff = 0;                                              // float[5360]
  static constexpr size_t kSumOff = kRawOff + sizeof(float) * kTileSamples;         // float[336]
  static constexpr size_t kDOff = kSumOff + sizeof(float) * 336;                    // T[5528] | sP + sOut
  static constexpr size_t kDBytesT = sizeof(T) * (kTileSamples + kTileSamples / 32 + 8);
  static constexpr size_t kPBytes = sizeof(float) * (kBins * kTileFrames);          // 32896
  static constexpr size_t kOutBytes = sizeof(float) * (kTileFrames * kOutStride);   // 10368
  static constexpr size_t kDBytes = (kDBytesT > kPBytes + kOutBytes ? kDBytesT : kPBytes + kOutBytes);
  static constexpr size_t kExOff = (kDOff + kDBytes + 15) / 16 * 16;                // cplx<T>[16*16*32]
  static constexpr size_t kBytes = kExOff + sizeof(cplx<T>) * 16 * 16 * kTileFrames;
};

template <typename T> struct FbankKernel {
  static constexpr int kThreads = kRoles * kTileFrames;   // 512
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 6;

  BatchDesc d;
  float* out;              // [rows][80] raw log-mel lands at its final position
  double* partials;        // [ntiles][80][2]  (sum, sum of squares) per mel bin
  const T* window;         // [400]  povey * 32768 / 2
  const T* tw512;          // [16][16][2]
  const int32_t* mel_start; const int32_t* mel_count; const int32_t* mel_offset; const float* mel_w;
  float log_floor;         // float32(log(float64(mel_floor)))

  AFE_HD long num_items() const { return d.ntiles; }

  AFE_HD void phase(int ph, long item, unsigned char* smem, int tid) const {
    using S = FbankSmem<T>;
    float* raw = reinterpret_cast<float*>(smem + S::kRawOff);
    float* sum16 = reinterpret_cast<float*>(smem + S::kSumOff);
    T* dbuf = reinterpret_cast<T*>(smem + S::kDOff);
    float* sP = reinterpret_cast<float*>(smem + S::kDOff);
    float* sOut = reinterpret_cast<float*>(smem + S::kDOff + S::kPBytes);
    cplx<T>* ex = reinterpret_cast<cplx<T>*>(smem + S::kExOff);
    int b, t;
    locate_tile(d, item, b, t);
    const int T_b = ldg(d.frames + b);
    const int f0 = t * kTileFrames;
    const int nvalid = tmin(kTileFrames, T_b - f0);
    const int role = tid >> 5, f = tid & 31;

    switch (ph) {
      case 0: {   // stage the tile's samples (zeros beyond the last valid frame)
        const float* src = ldg(d.wave + b) + (int64_t)f0 * kHop;
        const int count = (nvalid - 1) * kHop + kFrame;
        for (int i = tid; i < kTileSamples; i += kThreads) raw[i] = (i < count) ? src[i] : 0.f;
      } break;
      case 1: {   // pre-emphasis in T into the padded layout + 16-sample partial sums
        for (int i = tid; i < kTileSamples; i += kThreads) {
          const T x = (T)raw[i];
          const T xm = (i > 0) ? (T)raw[i - 1] : T(0);
          dbuf[phys_idx(i)] = x - T(0.97) * xm;
        }
        if (tid < kGroups16) {
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) s += raw[tid * 16 + j];
          sum16[tid] = s;
        }
      } break;
      case 2: {   // DC removal + window + stage 1 of the real FFT
        float s = 0.f;
#pragma unroll
        for (int g = 0; g < 25; ++g) s += sum16[10 * f + g];
        const T c = (T(1) - T(0.97)) * ((T)s / T(kFrame));
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
      case 4: {   // sparse triangular mel + floor + log; role w owns filters w, w+16, ..
#pragma unroll 1
        for (int m = role; m < kMel; m += kRoles) {
          const int k0 = ldg(mel_start + m), cnt = ldg(mel_count + m), off = ldg(mel_offset + m);
          float acc = 0.f;
          for (int i = 0; i < cnt; ++i) acc = fma_f(ldg(mel_w + off + i), sP[(k0 + i) * kTileFrames + f], acc);
          sOut[f * kOutStride + m] = (acc <= kMelFloor) ? log_floor : fast_log(acc);
        }
      } break;
      case 5: {   // per-bin partial statistics (float64) + coalesced store of the raw log-mel
        if (tid < kMel) {
          double s = 0.0, ss = 0.0;
          for (int j = 0; j < nvalid; ++j) {
            const double v = (double)sOut[j * kOutStride + tid];
            s += v; ss += v * v;
          }
          partials[(item * kMel + tid) * 2] = s;
          partials[(item * kMel + tid) * 2 + 1] = ss;
        }
        float* dst = out + (ldg(d.out_row + b) + f0) * kMel;
        for (int i = tid; i < nvalid * kMel; i += kThreads) dst[i] = sOut[(i / kMel) * kOutStride + (i % kMel)];
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
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, long item, unsigned char*, int tid) const {
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
  AFE_HD long num_items() const { return items_per_utt * d.B; }
  AFE_HD void phase(int, long item, unsigned char*, int tid) const {
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
