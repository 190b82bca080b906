// afe_acoustic_kernels.h -- the 38 acoustic features of
// AcousticSentimentAnalyzer._extract_acoustic_features (services/sentiment_analysis.py:807-894)
// as phased block programs.  float32 throughout like the reference (torch.stft / MFCC).
//
//   AcousticTileKernel   one item = 32 STFT frames of one waveform: unfold-frame RMS / ZCR,
//                        centred reflect-padded Hann STFT -> |X|, centroid, bandwidth,
//                        rolloff, in-tile spectral flux, 128 HTK-mel dB rows (to workspace),
//                        per-tile float64 partial statistics and boundary spectra.
//   AcousticMaxKernel    per waveform: global max dB (top_db = 80 reference point).
//   AcousticMfccKernel   one item = the same tile: clamp at max-80, DCT-II -> 13 MFCC,
//                        per-tile float64 partials.
//   AcousticFinalKernel  per waveform: ordered reduction of partials -> 38 floats.
#pragma once
#include "afe_fbank_kernels.h"

namespace afe {

constexpr int kAcRaw = kTileSamples + 200 + 8;    // 5568 staged samples: STFT span + unfold tail + 1
constexpr int kDbStride = kAcMel + 1;             // 129
constexpr int kTileRec = 16;                      // doubles per tile record
constexpr int kBoundary = 260;                    // floats per boundary spectrum (257 used)

struct AcDesc {
  const float* const* wave;   // [B]
  const int64_t* len;         // [B]
  const int32_t* frames1;     // [B] unfold frames  F1 = (L-400)/160 + 1   (0 if L < 400)
  const int32_t* frames2;     // [B] STFT frames    F2 = 1 + L/160         (0 if L < 400)
  const int64_t* tile_prefix; // [B+1] prefix of ceil(F2/32)
  const int64_t* row_prefix;  // [B+1] prefix of F2 (rows of the mel-dB workspace)
  int B;
  long ntiles;
};

AFE_HD void ac_locate(const AcDesc& d, long item, int& b, int& t) {
  int lo = 0, hi = d.B;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (ldg(d.tile_prefix + mid) <= item) lo = mid; else hi = mid;
  }
  b = lo;
  t = (int)(item - ldg(d.tile_prefix + lo));
}

struct AcSmem {
  // region A: raw + padded copy, later the magnitudes sM[257][32]
  static constexpr size_t kRawOff = 0;                                           // float[5568]
  static constexpr size_t kDOff = kRawOff + sizeof(float) * kAcRaw;              // float[5536]
  static constexpr size_t kABytes = sizeof(float) * (kAcRaw + 5536);             // 44416 >= 32896
  static constexpr size_t kSmallOff = kABytes;
  //   small arrays (floats): sq16[336] zc16[336] lastx[32] parts_e[512] parts_c[512] red[512] pf[8][32]
  static constexpr int kSq = 0, kZc = 336, kLast = 672, kPe = 704, kPc = 1216, kRed = 1728, kPf = 2240;
  static constexpr size_t kSmallBytes = sizeof(float) * (2240 + 8 * 32);
  static constexpr size_t kPOff = (kSmallOff + kSmallBytes + 15) / 16 * 16;      // float sP[257*32]
  static constexpr size_t kExOff = (kPOff + sizeof(float) * kBins * kTileFrames + 15) / 16 * 16;
  static constexpr size_t kBytes = kExOff + sizeof(cplx<float>) * 16 * 16 * kTileFrames;   // ex | sDb[32][129]
};

struct AcousticTileKernel {
  static constexpr int kThreads = 512;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 7;

  AcDesc d;
  const float* hann;        // [400] periodic Hann / 2
  const float* tw512;       // [16][16][2]
  const int32_t* mel_start; const int32_t* mel_count; const int32_t* mel_offset; const float* mel_w;
  float* meldb;             // [rows][128]
  double* tile_rec;         // [ntiles][16]
  float* boundary;          // [ntiles][2][260]  first / last valid frame magnitudes

  AFE_HD long num_items() const { return d.ntiles; }

  AFE_HD void phase(int ph, long item, unsigned char* smem, int tid) const {
    float* raw = reinterpret_cast<float*>(smem + AcSmem::kRawOff);
    float* dbuf = reinterpret_cast<float*>(smem + AcSmem::kDOff);
    float* sM = reinterpret_cast<float*>(smem + AcSmem::kRawOff);
    float* small = reinterpret_cast<float*>(smem + AcSmem::kSmallOff);
    float* sP = reinterpret_cast<float*>(smem + AcSmem::kPOff);
    cplx<float>* ex = reinterpret_cast<cplx<float>*>(smem + AcSmem::kExOff);
    float* sDb = reinterpret_cast<float*>(smem + AcSmem::kExOff);
    float* sq16 = small + AcSmem::kSq;  float* zc16 = small + AcSmem::kZc;  float* lastx = small + AcSmem::kLast;
    float* pe = small + AcSmem::kPe;    float* pc = small + AcSmem::kPc;    float* red = small + AcSmem::kRed;
    float* pf = small + AcSmem::kPf;    // [0]=centroid [1]=bandwidth [2]=rolloff [3]=flux [4]=rms [5]=zcr [6]=dbmax
    int b, t;
    ac_locate(d, item, b, t);
    const int64_t L = ldg(d.len + b);
    const int F1 = ldg(d.frames1 + b), F2 = ldg(d.frames2 + b);
    const int f0 = t * kTileFrames;
    const int nvalid = tmin(kTileFrames, F2 - f0);
    const int role = tid >> 5, f = tid & 31;

    switch (ph) {
      case 0: {   // stage y[reflect(160 f0 - 200 + i)]
        const float* src = ldg(d.wave + b);
        const int64_t g0 = (int64_t)f0 * kHop - 200;
        for (int i = tid; i < kAcRaw; i += kThreads) {
          int64_t g = g0 + i;
          if (g < 0) g = -g;
          if (g >= L) g = 2 * (L - 1) - g;
          raw[i] = (g >= 0 && g < L) ? src[g] : 0.f;
        }
      } break;
      case 1: {   // padded copy for the STFT + 16-sample partials of the unfold frames
        for (int i = tid; i < kTileSamples; i += kThreads) dbuf[phys_idx(i)] = raw[i];
        if (tid < kGroups16) {
          float s = 0.f, z = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float a = raw[200 + tid * 16 + j], c = raw[200 + tid * 16 + j + 1];
            s = fma_f(a, a, s);
            z += ((a > 0.f && c < 0.f) || (a < 0.f && c > 0.f)) ? 1.f : 0.f;
          }
          sq16[tid] = s; zc16[tid] = z;
        }
        if (tid < kTileFrames) {   // the pair (399, 400) of unfold frame tid is NOT part of the frame
          const float a = raw[200 + kHop * tid + 399], c = raw[200 + kHop * tid + 400];
          lastx[tid] = ((a > 0.f && c < 0.f) || (a < 0.f && c > 0.f)) ? 1.f : 0.f;
        }
      } break;
      case 2: {   // Hann window + stage 1 of the real FFT
        float r[25];
#pragma unroll
        for (int n1 = 0; n1 < 25; ++n1) {
          const int n = 16 * n1 + role;
          r[n1] = ldg(hann + n) * dbuf[phys_idx(kHop * f + n)];
        }
        rfft512_stage1<float>(r, role, f, tw512, ex);
      } break;
      case 3:
        rfft512_stage2<float>(role, f, ex, sP);
        break;
      case 4: {   // magnitudes, energy / first-moment partials, 128 HTK mel -> dB
        float e = 0.f, c = 0.f;
        for (int k = role; k < kBins; k += kRoles) {
          const float m = fast_sqrt(sP[k * kTileFrames + f]);
          sM[k * kTileFrames + f] = m;
          e += m; c = fma_f(m, (float)k, c);
        }
        pe[role * 32 + f] = e; pc[role * 32 + f] = c;
        float mx = -3.0e38f;
#pragma unroll 1
        for (int m = role; m < kAcMel; m += kRoles) {
          const int k0 = ldg(mel_start + m), cnt = ldg(mel_count + m), off = ldg(mel_offset + m);
          float acc = 0.f;
          for (int i = 0; i < cnt; ++i) acc = fma_f(ldg(mel_w + off + i), sP[(k0 + i) * kTileFrames + f], acc);
          const float db = 10.f * fast_log10(tmax(acc, 1e-10f));
          sDb[f * kDbStride + m] = db;
          mx = tmax(mx, db);
        }
        red[tid] = mx;
      } break;
      case 5: {   // per-frame scalars, one role each
        if (role == 0 || role == 2) {
          float e = 0.f, c = 0.f;
#pragma unroll
          for (int r = 0; r < kRoles; ++r) { e += pe[r * 32 + f]; c += pc[r * 32 + f]; }
          const float energy = e + 1e-8f;
          const float cen = 31.25f * c / energy;
          if (role == 0) {
            pf[0 * 32 + f] = cen;
            double acc = 0.0;
            for (int k = 0; k < kBins; ++k) acc += (double)sM[k * kTileFrames + f];
            const float thr = 0.85f * ((float)acc + 1e-8f);
            acc = 0.0;
            int below = 0;
            for (int k = 0; k < kBins; ++k) { acc += (double)sM[k * kTileFrames + f]; below += ((float)acc < thr) ? 1 : 0; }
            pf[2 * 32 + f] = 31.25f * (float)(below >= kBins ? 0 : below);
          } else {
            float s = 0.f;
            for (int k = 0; k < kBins; ++k) {
              const float df = 31.25f * (float)k - cen;
              s = fma_f(sM[k * kTileFrames + f], df * df, s);
            }
            pf[1 * 32 + f] = fast_sqrt(s / energy);
          }
        } else if (role == 1) {
          float s = 0.f;
          if (f + 1 < nvalid)
            for (int k = 0; k < kBins; ++k) s += tmax(sM[k * kTileFrames + f + 1] - sM[k * kTileFrames + f], 0.f);
          pf[3 * 32 + f] = s;
        } else if (role == 3) {
          float s = 0.f, z = 0.f;
#pragma unroll
          for (int g = 0; g < 25; ++g) { s += sq16[10 * f + g]; z += zc16[10 * f + g]; }
          pf[4 * 32 + f] = fast_sqrt(s / 400.f);
          pf[5 * 32 + f] = (z - lastx[f]) / 400.f;
        } else if (role == 4) {
          float mx = -3.0e38f;
#pragma unroll
          for (int r = 0; r < kRoles; ++r) mx = tmax(mx, red[r * 32 + f]);
          pf[6 * 32 + f] = mx;
        }
      } break;
      case 6: {   // tile record, boundary spectra, mel-dB rows
        double* rec = tile_rec + item * kTileRec;
        if (tid < 5) {          // 0 rms 1 zcr (unfold frames) 2 centroid 3 bandwidth 4 rolloff
          const int src = (tid == 0) ? 4 : (tid == 1) ? 5 : (tid - 2);
          const int n = (tid < 2) ? tmax(0, tmin(kTileFrames, F1 - f0)) : nvalid;
          double s = 0.0, ss = 0.0, mn = 3.0e38, mxv = -3.0e38;
          for (int j = 0; j < n; ++j) {
            const double v = (double)pf[src * 32 + j];
            s += v; ss += v * v; mn = v < mn ? v : mn; mxv = v > mxv ? v : mxv;
          }
          rec[2 * tid] = s; rec[2 * tid + 1] = ss;
          if (tid == 0) { rec[11] = mn; rec[12] = mxv; }
        } else if (tid == 5) {
          double s = 0.0;
          for (int j = 0; j + 1 < nvalid; ++j) { const double v = (double)pf[3 * 32 + j]; s += v * v; }
          rec[10] = s;
        } else if (tid == 6) {
          float mx = -3.0e38f;
          for (int j = 0; j < nvalid; ++j) mx = tmax(mx, pf[6 * 32 + j]);
          rec[13] = (double)mx;
        }
        if (tid < kBins) {
          float* bd = boundary + item * 2 * kBoundary;
          bd[tid] = sM[tid * kTileFrames + 0];
          bd[kBoundary + tid] = sM[tid * kTileFrames + (nvalid - 1)];
        }
        float* dst = meldb + (ldg(d.row_prefix + b) + f0) * kAcMel;
        for (int i = tid; i < nvalid * kAcMel; i += kThreads) dst[i] = sDb[(i >> 7) * kDbStride + (i & 127)];
      } break;
    }
  }
};

struct AcousticMaxKernel {
  static constexpr int kThreads = 32;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 1;
  AcDesc d; const double* tile_rec; float* gmax;
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, long item, unsigned char*, int tid) const {
    if (tid != 0) return;
    double mx = -3.0e38;
    for (long t = ldg(d.tile_prefix + item); t < ldg(d.tile_prefix + item + 1); ++t) {
      const double v = tile_rec[t * kTileRec + 13];
      mx = v > mx ? v : mx;
    }
    gmax[item] = (float)mx;
  }
};

struct AcousticMfccKernel {
  static constexpr int kThreads = 512;
  static constexpr int kMinBlocks = 2;
  static constexpr int kPhases = 3;
  static constexpr size_t kSmemBytes = sizeof(float) * (kTileFrames * kDbStride + 16 * 32);
  AcDesc d; const float* meldb; const float* gmax; const float* dct; double* mfcc_rec;   // [ntiles][13][2]
  AFE_HD long num_items() const { return d.ntiles; }
  AFE_HD void phase(int ph, long item, unsigned char* smem, int tid) const {
    float* sD = reinterpret_cast<float*>(smem);
    float* sC = sD + kTileFrames * kDbStride;
    int b, t;
    ac_locate(d, item, b, t);
    const int F2 = ldg(d.frames2 + b);
    const int f0 = t * kTileFrames;
    const int nvalid = tmin(kTileFrames, F2 - f0);
    if (ph == 0) {
      const float cut = ldg(gmax + b) - 80.f;
      const float* src = meldb + (ldg(d.row_prefix + b) + f0) * kAcMel;
      for (int i = tid; i < kTileFrames * kAcMel; i += kThreads)
        sD[(i >> 7) * kDbStride + (i & 127)] = (i < nvalid * kAcMel) ? tmax(src[i], cut) : 0.f;
    } else if (ph == 1) {
      const int c = tid >> 5, f = tid & 31;
      if (c < kMfcc) {
        float acc = 0.f;
        for (int n = 0; n < kAcMel; ++n) acc = fma_f(sD[f * kDbStride + n], ldg(dct + c * kAcMel + n), acc);
        sC[c * 32 + f] = acc;
      }
    } else if (tid < kMfcc) {
      double s = 0.0, ss = 0.0;
      for (int j = 0; j < nvalid; ++j) { const double v = (double)sC[tid * 32 + j]; s += v; ss += v * v; }
      mfcc_rec[(item * kMfcc + tid) * 2] = s;
      mfcc_rec[(item * kMfcc + tid) * 2 + 1] = ss;
    }
  }
};

struct AcousticFinalKernel {
  static constexpr int kThreads = 32;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 1;
  AcDesc d; const double* tile_rec; const double* mfcc_rec; const float* boundary;
  float* out;              // [B][38]
  int32_t* status;         // [B]
  AFE_HD static void mean_std(double s, double ss, int n, float& mean, float& sd) {
    const double m = s / (double)n;
    mean = (float)m;
    const double var = (ss - s * m) / (double)(n - 1);          // n == 1 -> 0/0 = NaN like torch.std
    sd = (float)std::sqrt(var < 0.0 ? 0.0 : var);
  }
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, long item, unsigned char*, int tid) const {
    float* o = out + item * kAcFeat;
    const int F1 = ldg(d.frames1 + item), F2 = ldg(d.frames2 + item);
    if (F2 <= 0) {                                   // 'Audio too short'
      if (tid == 0) status[item] = 1;
      for (int i = tid; i < kAcFeat; i += kThreads) o[i] = 0.f;
      return;
    }
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    if (tid == 0) status[item] = 0;
    if (tid < 5) {              // rms, zcr over F1 frames; centroid, bandwidth, rolloff over F2
      double s = 0.0, ss = 0.0, mn = 3.0e38, mx = -3.0e38;
      for (long t = t0; t < t1; ++t) {
        s += tile_rec[t * kTileRec + 2 * tid]; ss += tile_rec[t * kTileRec + 2 * tid + 1];
        if (tid == 0) {
          const double a = tile_rec[t * kTileRec + 11], c = tile_rec[t * kTileRec + 12];
          mn = a < mn ? a : mn; mx = c > mx ? c : mx;
        }
      }
      float mean, sd;
      mean_std(s, ss, tid < 2 ? F1 : F2, mean, sd);
      const int slot = (tid == 0) ? 0 : (tid == 1) ? 3 : (5 + 2 * (tid - 2));
      o[slot] = mean; o[slot + 1] = sd;
      if (tid == 0) o[2] = (float)(mx - mn);
    } else if (tid == 5) {      // spectral-flux energy: in-tile pairs + tile-boundary pairs
      double s = 0.0;
      for (long t = t0; t < t1; ++t) {
        s += tile_rec[t * kTileRec + 10];
        if (t + 1 < t1) {
          const float* last = boundary + t * 2 * kBoundary + kBoundary;
          const float* first = boundary + (t + 1) * 2 * kBoundary;
          float fl = 0.f;
          for (int k = 0; k < kBins; ++k) fl += tmax(first[k] - last[k], 0.f);
          s += (double)fl * (double)fl;
        }
      }
      o[11] = (float)s;
    } else if (tid >= 6 && tid < 6 + kMfcc) {
      const int c = tid - 6;
      double s = 0.0, ss = 0.0;
      for (long t = t0; t < t1; ++t) { s += mfcc_rec[(t * kMfcc + c) * 2]; ss += mfcc_rec[(t * kMfcc + c) * 2 + 1]; }
      float mean, sd;
      mean_std(s, ss, F2, mean, sd);
      o[12 + 2 * c] = mean; o[13 + 2 * c] = sd;
    }
  }
};

}  // namespace afe
