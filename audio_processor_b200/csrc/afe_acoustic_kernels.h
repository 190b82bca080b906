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
  const int32_t* out_index;   // [B] output row of waveform b (nullptr: b)
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

constexpr int kHtkIv = kAcMel + 1;                // 129 intervals

struct AcSmem {
  static constexpr size_t kRawBytes = sizeof(float) * kAcRaw;                    // 22272 per staging buffer
  static constexpr size_t kRawOff = 0;                                           // float[2][5568]
  static constexpr size_t kDOff = kRawOff + 2 * kRawBytes;                       // float[5536] padded copy
  static constexpr size_t kSmallOff = kDOff + sizeof(float) * 5536;
  // small arrays, in floats from kSmallOff (epart is double[512], 8-byte aligned)
  static constexpr int kEp = 0;            // double[16][32] contiguous-range magnitude sums
  static constexpr int kSq = 1024, kZc = 1360, kLast = 1696;        // [336] [336] [32]
  static constexpr int kPc = 1728, kRed = 2240, kCnt = 2752, kBw = 3264, kFx = 3776;   // [512] each
  static constexpr int kPf = 4288;         // [8][32] per-frame scalars
  static constexpr size_t kSmallBytes = sizeof(float) * (4288 + 8 * 32);
  static constexpr size_t kPOff = (kSmallOff + kSmallBytes + 15) / 16 * 16;      // float sP[257*32] -> magnitudes in place
  static constexpr size_t kExOff = (kPOff + sizeof(float) * kBins * kTileFrames + 15) / 16 * 16;
  //   ex region after stage 2: sU[129][32] sD[129][32] sDb[32][129]
  static constexpr size_t kDbOff = kExOff + sizeof(float) * 2 * kHtkIv * kTileFrames;
  static constexpr size_t kBytes = kExOff + sizeof(cplx<float>) * 16 * 16 * kTileFrames;
  static_assert(kDbOff + sizeof(float) * kTileFrames * kDbStride <= kBytes, "ex region too small");
};

struct AcousticTileKernel {
  static constexpr int kThreads = 512;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 9;
  static constexpr bool kAsyncStage = true;
  static constexpr size_t kStageOff = AcSmem::kRawOff;
  static constexpr size_t kStageBytes = AcSmem::kRawBytes;

  struct Item { long item; int buf; int bulk; int b; int f0; int nvalid; int F1; int F2; int64_t L; const float* wave; int64_t row0; };

  AcDesc d;
  const float* hann;        // [400] periodic Hann / 2
  const float* tw512;       // [16][16][2]
  float* meldb;             // [rows][128]
  double* tile_rec;         // [ntiles][16]
  float* boundary;          // [ntiles][2][260]  first / last valid frame magnitudes

  AFE_HD long num_items() const { return d.ntiles; }

  AFE_HD Item prepare(long item) const {
    Item it;
    int t;
    ac_locate(d, item, it.b, t);
    it.item = item; it.buf = 0; it.bulk = 0;
    fill(it, it.b, t * kTileFrames);
    return it;
  }
  AFE_HD void fill(Item& it, int b, int f0) const {
    it.b = b; it.f0 = f0;
    it.F2 = ldg(d.frames2 + b);
    it.nvalid = tmin(kTileFrames, it.F2 - f0);
    it.F1 = ldg(d.frames1 + b);
    it.L = ldg(d.len + b);
    it.wave = ldg(d.wave + b);
    it.row0 = ldg(d.row_prefix + b);
  }
  // item cur.item + 1 without a search: next tile of the same waveform, else the next waveform with frames
  AFE_HD Item next(const Item& cur, long item) const {
    Item it = cur;
    it.item = item;
    if (cur.f0 + kTileFrames < cur.F2) {
      it.f0 = cur.f0 + kTileFrames;
      it.nvalid = tmin(kTileFrames, cur.F2 - it.f0);
      return it;
    }
    int b = cur.b + 1;
    while (b + 1 < d.B && ldg(d.frames2 + b) <= 0) ++b;
    fill(it, b, 0);
    return it;
  }
  // interior tiles (no reflection) are one contiguous aligned copy
  AFE_HD bool stage_src(const Item& it, const float*& src, int& nfloats) const {
    const int64_t g0 = (int64_t)it.f0 * kHop - 200;
    src = it.wave + g0; nfloats = kAcRaw;
    return g0 >= 0 && g0 + kAcRaw <= it.L && (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
  }

  AFE_HD static bool crossing(float a, float c) { return (a > 0.f && c < 0.f) || (a < 0.f && c > 0.f); }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    float* raw = reinterpret_cast<float*>(smem + AcSmem::kRawOff + (size_t)it.buf * AcSmem::kRawBytes);
    float* dbuf = reinterpret_cast<float*>(smem + AcSmem::kDOff);
    float* small = reinterpret_cast<float*>(smem + AcSmem::kSmallOff);
    float* sP = reinterpret_cast<float*>(smem + AcSmem::kPOff);        // power, then magnitude in place
    cplx<float>* ex = reinterpret_cast<cplx<float>*>(smem + AcSmem::kExOff);
    float* sU = reinterpret_cast<float*>(smem + AcSmem::kExOff);
    float* sD = sU + kHtkIv * kTileFrames;
    float* sDb = reinterpret_cast<float*>(smem + AcSmem::kDbOff);
    double* epart = reinterpret_cast<double*>(small + AcSmem::kEp);
    float* sq16 = small + AcSmem::kSq;  float* zc16 = small + AcSmem::kZc;  float* lastx = small + AcSmem::kLast;
    float* pc = small + AcSmem::kPc;    float* red = small + AcSmem::kRed;
    float* cntp = small + AcSmem::kCnt; float* bwp = small + AcSmem::kBw;   float* fxp = small + AcSmem::kFx;
    float* pf = small + AcSmem::kPf;    // [0]=centroid [1]=bandwidth [2]=rolloff [3]=flux [4]=rms [5]=zcr [6]=dbmax [7]=energy
    const int role = tid >> 5, f = tid & 31;
    const int kb = role * 16, ke = (role == 15) ? kBins : kb + 16;       // contiguous bin range of this role

    switch (ph) {
      case 0: {   // stage y[reflect(160 f0 - 200 + i)] (edge tiles; interior tiles arrive by TMA)
        const int64_t g0 = (int64_t)it.f0 * kHop - 200;
        for (int i = tid; i < kAcRaw; i += kThreads) {
          int64_t g = g0 + i;
          if (g < 0) g = -g;
          if (g >= it.L) g = 2 * (it.L - 1) - g;
          raw[i] = (g >= 0 && g < it.L) ? it.wave[g] : 0.f;
        }
      } break;
      case 1: {   // padded copy for the STFT + 16-sample partials of the unfold frames
        for (int i = tid; i < kTileSamples; i += kThreads) dbuf[phys_idx(i)] = raw[i];
        if (tid < kGroups16) {
          float s = 0.f, z = 0.f;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int i = 200 + tid * 16 + ((j + tid) & 15);           // rotated: 2-way bank conflicts
            const float a = raw[i], c = raw[i + 1];
            s = fma_f(a, a, s);
            z += crossing(a, c) ? 1.f : 0.f;
          }
          sq16[tid] = s; zc16[tid] = z;
        }
        if (tid < kTileFrames)     // the pair (399, 400) of unfold frame tid is NOT part of the frame
          lastx[tid] = crossing(raw[200 + kHop * tid + 399], raw[200 + kHop * tid + 400]) ? 1.f : 0.f;
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
      case 4:     // 128 HTK mel filters, interval form (reads the POWER spectrum)
        mel_intervals_htk(role, f, sP, sU, sD);
        break;
      case 5: {   // dB rows; magnitudes in place + contiguous-range sums
        float mx = -3.0e38f;
#pragma unroll
        for (int m = role; m < kAcMel; m += kRoles) {
          const float acc = sU[m * kTileFrames + f] + sD[(m + 1) * kTileFrames + f];
          const float db = 10.f * fast_log10(tmax(acc, 1e-10f));
          sDb[f * kDbStride + m] = db;
          mx = tmax(mx, db);
        }
        red[tid] = mx;
        double e = 0.0;
        float c = 0.f;
        for (int k = kb; k < ke; ++k) {
          const float m = fast_sqrt(sP[k * kTileFrames + f]);
          sP[k * kTileFrames + f] = m;
          e += (double)m; c = fma_f(m, (float)k, c);
        }
        epart[role * 32 + f] = e; pc[role * 32 + f] = c;
      } break;
      case 6: {   // per-frame totals, then this role's share of rolloff / bandwidth / flux
        double tot = 0.0, pre = 0.0;
        float c = 0.f;
#pragma unroll
        for (int r = 0; r < kRoles; ++r) {
          if (r == role) pre = tot;
          tot += epart[r * 32 + f]; c += pc[r * 32 + f];
        }
        const float total = (float)tot;                  // cum_energy[-1] (float64 accumulate like ATen's cumsum)
        const float energy = total + 1e-8f;
        const float cen = 31.25f * c / energy;
        const float thr = 0.85f * (total + 1e-8f);
        double acc = pre;
        int below = 0;
        float bw = 0.f, fx = 0.f;
        const bool has_next = (f + 1 < it.nvalid);
        for (int k = kb; k < ke; ++k) {
          const float m = sP[k * kTileFrames + f];
          acc += (double)m;
          below += ((float)acc < thr) ? 1 : 0;
          const float df = 31.25f * (float)k - cen;
          bw = fma_f(m, df * df, bw);
          if (has_next) fx += tmax(sP[k * kTileFrames + f + 1] - m, 0.f);
        }
        cntp[role * 32 + f] = (float)below; bwp[role * 32 + f] = bw; fxp[role * 32 + f] = fx;
        if (role == 0) { pf[0 * 32 + f] = cen; pf[7 * 32 + f] = energy; }
        if (role == 1) {
          float s = 0.f, z = 0.f;
#pragma unroll
          for (int g = 0; g < 25; ++g) { s += sq16[10 * f + g]; z += zc16[10 * f + g]; }
          pf[4 * 32 + f] = fast_sqrt(s / 400.f);
          pf[5 * 32 + f] = (z - lastx[f]) / 400.f;
        }
        if (role == 2) {
          float mx = -3.0e38f;
#pragma unroll
          for (int r = 0; r < kRoles; ++r) mx = tmax(mx, red[r * 32 + f]);
          pf[6 * 32 + f] = mx;
        }
      } break;
      case 7: {   // combine the 16 partials per frame
        if (role < 3) {
          float s = 0.f;
          const float* src = (role == 0) ? cntp : (role == 1) ? bwp : fxp;
#pragma unroll
          for (int r = 0; r < kRoles; ++r) s += src[r * 32 + f];
          if (role == 0) pf[2 * 32 + f] = 31.25f * ((s >= (float)kBins) ? 0.f : s);
          else if (role == 1) pf[1 * 32 + f] = fast_sqrt(s / pf[7 * 32 + f]);
          else pf[3 * 32 + f] = s;
        }
      } break;
      case 8: {   // tile record, boundary spectra, mel-dB rows
        double* rec = tile_rec + it.item * kTileRec;
        if (tid < 5) {          // 0 rms 1 zcr (unfold frames) 2 centroid 3 bandwidth 4 rolloff
          const int src = (tid == 0) ? 4 : (tid == 1) ? 5 : (tid - 2);
          const int n = (tid < 2) ? tmax(0, tmin(kTileFrames, it.F1 - it.f0)) : it.nvalid;
          double s = 0.0, ss = 0.0, mn = 3.0e38, mxv = -3.0e38;
          for (int j = 0; j < n; ++j) {
            const double v = (double)pf[src * 32 + j];
            s += v; ss += v * v; mn = v < mn ? v : mn; mxv = v > mxv ? v : mxv;
          }
          rec[2 * tid] = s; rec[2 * tid + 1] = ss;
          if (tid == 0) { rec[11] = mn; rec[12] = mxv; }
        } else if (tid == 5) {
          double s = 0.0;
          for (int j = 0; j + 1 < it.nvalid; ++j) { const double v = (double)pf[3 * 32 + j]; s += v * v; }
          rec[10] = s;
        } else if (tid == 6) {
          float mx = -3.0e38f;
          for (int j = 0; j < it.nvalid; ++j) mx = tmax(mx, pf[6 * 32 + j]);
          rec[13] = (double)mx;
        }
        if (tid >= 64 && tid < 64 + kBins) {
          const int k = tid - 64;
          float* bd = boundary + it.item * 2 * kBoundary;
          bd[k] = sP[k * kTileFrames + 0];
          bd[kBoundary + k] = sP[k * kTileFrames + (it.nvalid - 1)];
        }
        float* dst = meldb + (it.row0 + it.f0) * kAcMel;
        for (int i = tid; i < it.nvalid * kAcMel; i += kThreads) dst[i] = sDb[(i >> 7) * kDbStride + (i & 127)];
      } break;
    }
  }
};

struct AcousticMaxKernel {
  static constexpr int kThreads = 32;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 1;
  AcDesc d; const double* tile_rec; float* gmax;
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    if (tid != 0) return;
    double mx = -3.0e38;
    for (long t = ldg(d.tile_prefix + item); t < ldg(d.tile_prefix + item + 1); ++t) {
      const double v = tile_rec[t * kTileRec + 13];
      mx = v > mx ? v : mx;
    }
    gmax[item] = (float)mx;
  }
};

// One item = one 32-frame tile; 4 warps, lane = frame, warp w owns mel bins [32w, 32w+32).
// The DCT coefficients are immediates of generated straight-line FFMAs (afe_mel_gen.h; sm_100
// has no constant-bank ALU operands, a table in the parameters costs one LDCU per FFMA): per
// frame 128 shared loads (as 128-bit words) and 13 x 128 FFMAs, no table loads.
struct AcousticMfccKernel {
  static constexpr int kThreads = 128;
  static constexpr int kMinBlocks = 8;
  static constexpr int kPhases = 4;
  static constexpr int kRow = kAcMel + 4;           // 132: rows stay 16-byte aligned, LDS.128 by lane = frame is conflict-free
  static constexpr int kWarps = kThreads / 32;
  static constexpr size_t kSmemBytes =
      sizeof(float) * (kTileFrames * kRow + kWarps * kMfcc * 32 + kMfcc * 32 + kThreads);
  static constexpr bool kAsyncStage = false;
  AcDesc d; const float* meldb; const float* gmax; const float* boundary;
  double* mfcc_rec;        // [ntiles][14][2]: 13 x (sum, sum sq) + (boundary flux^2, 0)

  struct Item { long item; int buf; int bulk; int b; int f0; int nvalid; int has_next; float cut; int64_t row0; };
  AFE_HD long num_items() const { return d.ntiles; }
  AFE_HD Item prepare(long item) const {
    Item it;
    int t;
    ac_locate(d, item, it.b, t);
    it.item = item; it.buf = 0; it.bulk = 0;
    it.f0 = t * kTileFrames;
    it.nvalid = tmin(kTileFrames, ldg(d.frames2 + it.b) - it.f0);
    it.has_next = (item + 1 < ldg(d.tile_prefix + it.b + 1)) ? 1 : 0;   // another tile of the same waveform follows
    it.cut = ldg(gmax + it.b) - 80.f;
    it.row0 = ldg(d.row_prefix + it.b);
    return it;
  }

  template <int W> AFE_HD void partial_dct(const float* AFE_RESTRICT sD, float* AFE_RESTRICT sC, int f) const {
    float acc[kMfcc];
#pragma unroll
    for (int c = 0; c < kMfcc; ++c) acc[c] = 0.f;
    const float4* row = reinterpret_cast<const float4*>(sD + f * kRow + 32 * W);
    if (W == 0) dct13_partial_0(row, acc);
    else if (W == 1) dct13_partial_1(row, acc);
    else if (W == 2) dct13_partial_2(row, acc);
    else dct13_partial_3(row, acc);
#pragma unroll
    for (int c = 0; c < kMfcc; ++c) sC[(W * kMfcc + c) * 32 + f] = acc[c];
  }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    float* sD = reinterpret_cast<float*>(smem);
    float* sC = sD + kTileFrames * kRow;         // [4][13][32] partial dot products
    float* sV = sC + kWarps * kMfcc * 32;        // [13][32] coefficients
    float* sX = sV + kMfcc * 32;                 // [128] boundary flux partials
    if (ph == 0) {        // clamp at max - 80 dB while staging the tile (coalesced 128-bit loads)
      const float4* src = reinterpret_cast<const float4*>(meldb + (it.row0 + it.f0) * kAcMel);
      const int nq = it.nvalid * (kAcMel / 4);
      for (int i = tid; i < kTileFrames * (kAcMel / 4); i += kThreads) {
        float4 v = float4{0.f, 0.f, 0.f, 0.f};
        if (i < nq) {
          v = src[i];
          v.x = tmax(v.x, it.cut); v.y = tmax(v.y, it.cut); v.z = tmax(v.z, it.cut); v.w = tmax(v.w, it.cut);
        }
        *reinterpret_cast<float4*>(sD + (i >> 5) * kRow + (i & 31) * 4) = v;
      }
      // spectral flux across the tile boundary: relu(first frame of tile+1 - last frame of this tile)
      float fl = 0.f;
      if (it.has_next) {
        const float* last = boundary + it.item * 2 * kBoundary + kBoundary;
        const float* first = boundary + (it.item + 1) * 2 * kBoundary;
        for (int q = tid; q < kBins; q += kThreads) fl += tmax(first[q] - last[q], 0.f);
      }
      sX[tid] = fl;
    } else if (ph == 1) {
      const int f = tid & 31;
      switch (tid >> 5) {
        case 0: partial_dct<0>(sD, sC, f); break;
        case 1: partial_dct<1>(sD, sC, f); break;
        case 2: partial_dct<2>(sD, sC, f); break;
        default: partial_dct<3>(sD, sC, f); break;
      }
    } else if (ph == 2) {
      for (int i = tid; i < kMfcc * 32; i += kThreads) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) v += sC[w * kMfcc * 32 + i];
        sV[i] = v;
      }
    } else if (tid < kMfcc) {
      double s = 0.0, ss = 0.0;
      for (int j = 0; j < it.nvalid; ++j) { const double v = (double)sV[tid * 32 + j]; s += v; ss += v * v; }
      mfcc_rec[(it.item * 14 + tid) * 2] = s;
      mfcc_rec[(it.item * 14 + tid) * 2 + 1] = ss;
    } else if (tid == 32) {
      float fl = 0.f;
      for (int j = 0; j < kThreads; ++j) fl += sX[j];
      mfcc_rec[(it.item * 14 + 13) * 2] = (double)fl * (double)fl;
      mfcc_rec[(it.item * 14 + 13) * 2 + 1] = 0.0;
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
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    const long orow = d.out_index ? (long)ldg(d.out_index + item) : item;
    float* o = out + orow * kAcFeat;
    const int F1 = ldg(d.frames1 + item), F2 = ldg(d.frames2 + item);
    if (F2 <= 0) {                                   // 'Audio too short'
      if (tid == 0) status[orow] = 1;
      for (int i = tid; i < kAcFeat; i += kThreads) o[i] = 0.f;
      return;
    }
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    if (tid == 0) status[orow] = 0;
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
      for (long t = t0; t < t1; ++t) s += tile_rec[t * kTileRec + 10] + mfcc_rec[(t * 14 + 13) * 2];
      o[11] = (float)s;
    } else if (tid >= 6 && tid < 6 + kMfcc) {
      const int c = tid - 6;
      double s = 0.0, ss = 0.0;
      for (long t = t0; t < t1; ++t) { s += mfcc_rec[(t * 14 + c) * 2]; ss += mfcc_rec[(t * 14 + c) * 2 + 1]; }
      float mean, sd;
      mean_std(s, ss, F2, mean, sd);
      o[12 + 2 * c] = mean; o[13 + 2 * c] = sd;
    }
  }
};

}  // namespace afe
