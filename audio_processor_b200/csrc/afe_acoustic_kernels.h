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

struct AcSmem {
  static constexpr size_t kRawBytes = sizeof(float) * kAcRaw;                    // 22272 per staging buffer
  static constexpr size_t kRawOff = 0;                                           // float[2][5568]
  static constexpr size_t kDOff = kRawOff + 2 * kRawBytes;                       // float[5536] padded copy
  static constexpr size_t kKeepOff = kDOff + sizeof(float) * 5536;               // sq16[336] zc16[336] lastx[32]
  static constexpr int kSq = 0, kZc = 336, kLast = 672;
  static constexpr size_t kPOff = (kKeepOff + sizeof(float) * 704 + 15) / 16 * 16;      // float sP[257*32] power
  static constexpr size_t kMOff = kPOff + sizeof(float) * kBins * kTileFrames;           // float sM[257*32] magnitude
  static constexpr size_t kExOff = (kMOff + sizeof(float) * kBins * kTileFrames + 15) / 16 * 16;
  // the exchange region after stage 2: sDb[32][129] | double part[3][16][33] | float fx,red [16][33] | pf[8][32]
  // (role stride 33: odd, so lane = frame accesses of a role row are conflict-free whatever the row)
  static constexpr int kRS = 33, kPS = 16 * kRS;
  static constexpr size_t kDbOff = kExOff;
  static constexpr size_t kPartOff = (kDbOff + sizeof(float) * kTileFrames * kDbStride + 15) / 16 * 16;
  static constexpr size_t kFxOff = kPartOff + sizeof(double) * 3 * kPS;
  static constexpr size_t kRedOff = kFxOff + sizeof(float) * kPS;
  static constexpr size_t kPfOff = kRedOff + sizeof(float) * kPS;
  static constexpr size_t kThrOff = kPfOff + sizeof(float) * 8 * 32;
  static constexpr size_t kDsumOff = kThrOff + sizeof(float) * 64;                   // float thr[32] | int rc[32] at kThrOff, double lo[32] here: the rolloff hand-over
  // int64 stash[2]: the tile whose record is still owed {item index + 1 (0: none), n1 | nvalid << 16} -- written in the last
  // phase of a tile, consumed by warps 11..15 in phase 1 of the block's next tile (or by block_flush after its last)
  static constexpr size_t kStashOff = kDsumOff + sizeof(double) * 32;
  static constexpr size_t kHannOff = kExOff + sizeof(cplx<float>) * 16 * 16 * kTileFrames;   // float hann[16][26] (block_init)
  static constexpr size_t kTwOff = kHannOff + sizeof(float) * kRoles * kWinRow;               // float tw512[16][16][2]
  static constexpr size_t kMelWOff = (kTwOff + sizeof(float) * kRoles * 32 + 15) / 16 * 16;      // float melw[16][40] slot-form HTK bank (block_init)
  static constexpr size_t kMelSOff = kMelWOff + sizeof(float) * kRoles * kMelSlotRow;            // int32 mels[16][8]
  static constexpr size_t kBytes = kMelSOff + sizeof(int32_t) * kRoles * kMelSlotStarts;
  static_assert(kDsumOff % 8 == 0 && kStashOff + 16 <= kHannOff, "ex region too small");
};

// sqrt for magnitudes: one MUFU (max relative error ~1 ulp); the reference's float32 |X| carries
// the rounding noise of its own FFT, far above that.
AFE_HD float mag_sqrt(float x) {
#if AFE_CUDA && defined(__CUDA_ARCH__)
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return std::sqrt(x);
#endif
}

struct AcousticTileKernel {
  static constexpr int kThreads = 512;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 6;
  static constexpr int kTailPhases = 0;
  static constexpr int kRecordWarp0 = 11;                 // warps 11..15 have no 16-sample groups in phase 1
  static constexpr bool kAsyncStage = true;
  static constexpr int kRS = AcSmem::kRS, kPS = AcSmem::kPS;
  static constexpr size_t kStageOff = AcSmem::kRawOff;
  static constexpr size_t kStageBytes = AcSmem::kRawBytes;
  static constexpr int kPrefetchThread = kThreads - 32;   // lane 0 of the last warp: the lightest one in phase 1

  struct Item { long item; int buf; int bulk; int b; int f0; int nvalid; int F1; int F2; int64_t L; const float* wave; int64_t row0; };

  AcDesc d;
  const float* hann;        // [16][26] role-major: periodic Hann[16 n1 + role] / 2
  const float* tw512;       // [16][16][2]
  const float* mel_w;       // [16][40] slot-form HTK bank (afe_mel_slots.h)
  const int32_t* mel_start; // [16][8]
  float* meldb;             // [rows][128]
  double* tile_rec;         // [ntiles][16]
  float* boundary;          // [ntiles][2][260]  first / last valid frame magnitudes

  AFE_HD long num_items() const { return d.ntiles; }
  // once per block: the FFT tables move to shared memory (see FbankKernel::block_init)
  AFE_HD void block_init(unsigned char* smem, int tid) const {
    float* w = reinterpret_cast<float*>(smem + AcSmem::kHannOff);
    float* t = reinterpret_cast<float*>(smem + AcSmem::kTwOff);
    for (int i = tid; i < kRoles * kWinRow; i += kThreads) w[i] = ldg(hann + i);
    for (int i = tid; i < kRoles * 32; i += kThreads) t[i] = ldg(tw512 + i);
    float* mw = reinterpret_cast<float*>(smem + AcSmem::kMelWOff);
    int32_t* ms = reinterpret_cast<int32_t*>(smem + AcSmem::kMelSOff);
    for (int i = tid; i < kRoles * kMelSlotRow; i += kThreads) mw[i] = ldg(mel_w + i);
    for (int i = tid; i < kRoles * kMelSlotStarts; i += kThreads) ms[i] = ldg(mel_start + i);
    if (tid == 0) reinterpret_cast<int64_t*>(smem + AcSmem::kStashOff)[0] = 0;      // no record owed yet
  }
  // the record of the block's last tile (every other tile's is written under phase 1 of the tile after it)
  AFE_HD void block_flush(unsigned char* smem, int tid) const { owed_record(smem, tid); }

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

  // (sum, sum of squares) of n <= 32 per-frame scalars, 4 independent chains when the tile is full
  AFE_HD static void stat32(const float* v, int n, double& s, double& ss) {
    double a[4] = {0.0, 0.0, 0.0, 0.0}, q[4] = {0.0, 0.0, 0.0, 0.0};
    if (n == kTileFrames) {
#pragma unroll
      for (int j = 0; j < kTileFrames; ++j) { const double x = (double)v[j]; a[j & 3] += x; q[j & 3] += x * x; }
    } else {
      for (int j = 0; j < n; ++j) { const double x = (double)v[j]; a[0] += x; q[0] += x * x; }
    }
    s = (a[0] + a[1]) + (a[2] + a[3]); ss = (q[0] + q[1]) + (q[2] + q[3]);
  }

  // second half of the rolloff of one frame: the bin inside role rc where the cumulative energy (lo_c before the
  // role's bins) reaches thr; 31.25 Hz per bin.  rc < 0: no bin reaches the threshold (argmax of all-False is 0).
  AFE_HD static float rolloff_bin(const float* sM, int rc, double lo_c, float thr, int f) {
    if (rc < 0) return 0.f;
    const float* mp = sM + (16 * rc) * kTileFrames + f;
    double acc = lo_c;
    int below = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      acc += nonneg_to_double_biased(mp[j * kTileFrames]);      // +0 -> 2^-127: far below any threshold step
      below += ((float)acc < thr) ? 1 : 0;
    }
    if (rc == 15) below += ((float)(acc + nonneg_to_double(mp[16 * kTileFrames])) < thr) ? 1 : 0;   // bin 256
    return 31.25f * (float)(16 * rc + below);
  }

  // ---- tile record: one warp per statistic (lane = frame, fixed-tree reductions); statistic 4 (rolloff) first finds the
  // crossing bin inside the role phase 5 located.  0 rms 1 zcr (unfold frames) 2 centroid 3 bandwidth 4 rolloff: (sum, sum sq);
  // 5 flux energy; 6 rms range; 7 max dB.  Reads pf / the rolloff hand-over / sM of the tile it belongs to, which stay
  // untouched until phase 3 / phase 5 of the next tile.
  AFE_HD void record_stat(int stat, int f, long rec_item, int n1, int nvalid, unsigned char* smem) const {
    const float* sM = reinterpret_cast<const float*>(smem + AcSmem::kMOff);
    const float* roll_thr = reinterpret_cast<const float*>(smem + AcSmem::kThrOff);
    const int* roll_rc = reinterpret_cast<const int*>(smem + AcSmem::kThrOff) + 32;
    const double* roll_lo = reinterpret_cast<const double*>(smem + AcSmem::kDsumOff);
    const float* pf = reinterpret_cast<const float*>(smem + AcSmem::kPfOff);
    double* rec = tile_rec + rec_item * kTileRec;
    const int src = (stat == 0) ? 4 : (stat == 1) ? 5 : (stat < 5) ? (stat - 2) : (stat == 5) ? 3 : (stat == 6) ? 4 : 6;
    const int n = (stat < 2 || stat == 6) ? n1 : (stat == 5) ? nvalid - 1 : nvalid;
#if AFE_CUDA && defined(__CUDA_ARCH__)
    const double v = (double)((stat == 4) ? rolloff_bin(sM, roll_rc[f], roll_lo[f], roll_thr[f], f) : pf[src * 32 + f]);
    if (stat < 6) {
      const double x = (f < n) ? v : 0.0;
      const double s = warp_tree_sum(x), ss = warp_tree_sum(x * x);
      if (f == 0) {
        if (stat < 5) { rec[2 * stat] = s; rec[2 * stat + 1] = ss; } else rec[10] = ss;
      }
    } else if (stat == 6) {
      const double mn = warp_tree_min((f < n) ? v : 3.0e38), mxv = warp_tree_max((f < n) ? v : -3.0e38);
      if (f == 0) { rec[11] = mn; rec[12] = mxv; }
    } else {
      const double mxv = warp_tree_max((f < n) ? v : -3.0e38);
      if (f == 0) rec[13] = (double)(float)mxv;
    }
#else
    if (f == 0) {   // host emulation: lane 0 walks the same trees
      double a[32], q[32], mn = 3.0e38, mxv = -3.0e38;
      for (int j = 0; j < 32; ++j) {
        const double v = (j >= n) ? 0.0 : (double)((stat == 4) ? rolloff_bin(sM, roll_rc[j], roll_lo[j], roll_thr[j], j) : pf[src * 32 + j]);
        a[j] = v; q[j] = v * v;
        if (j < n) { mn = v < mn ? v : mn; mxv = v > mxv ? v : mxv; }
      }
      if (stat < 5) { rec[2 * stat] = tree32_sum(a); rec[2 * stat + 1] = tree32_sum(q); }
      else if (stat == 5) rec[10] = tree32_sum(q);
      else if (stat == 6) { rec[11] = mn; rec[12] = mxv; }
      else rec[13] = (double)(float)mxv;
    }
#endif
  }
  // the record the stash says is owed, by warps 11..15: 11 rolloff (the long dependent chain), 12: rms + zcr, 13: centroid +
  // bandwidth, 14: flux + rms range, 15: max dB.  Warp-uniform branches only (the trees are full-warp shuffles).
  AFE_HD void owed_record(unsigned char* smem, int tid) const {
    const int w = tid >> 5, f = tid & 31;
    if (w < kRecordWarp0) return;
    const int64_t* stash = reinterpret_cast<const int64_t*>(smem + AcSmem::kStashOff);
    const int64_t owed = stash[0];
    if (owed == 0) return;
    const int n1 = (int)(stash[1] & 0xffff), nvalid = (int)(stash[1] >> 16);
    const long rec_item = (long)(owed - 1);
    switch (w - kRecordWarp0) {
      case 0: record_stat(4, f, rec_item, n1, nvalid, smem); break;
      case 1: record_stat(0, f, rec_item, n1, nvalid, smem); record_stat(1, f, rec_item, n1, nvalid, smem); break;
      case 2: record_stat(2, f, rec_item, n1, nvalid, smem); record_stat(3, f, rec_item, n1, nvalid, smem); break;
      case 3: record_stat(5, f, rec_item, n1, nvalid, smem); record_stat(6, f, rec_item, n1, nvalid, smem); break;
      default: record_stat(7, f, rec_item, n1, nvalid, smem); break;
    }
  }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    float* raw = reinterpret_cast<float*>(smem + AcSmem::kRawOff + (size_t)it.buf * AcSmem::kRawBytes);
    float* dbuf = reinterpret_cast<float*>(smem + AcSmem::kDOff);
    float* keep = reinterpret_cast<float*>(smem + AcSmem::kKeepOff);
    float* sP = reinterpret_cast<float*>(smem + AcSmem::kPOff);        // power
    float* sM = reinterpret_cast<float*>(smem + AcSmem::kMOff);        // magnitude
    cplx<float>* ex = reinterpret_cast<cplx<float>*>(smem + AcSmem::kExOff);
    float* sDb = reinterpret_cast<float*>(smem + AcSmem::kDbOff);
    double* part = reinterpret_cast<double*>(smem + AcSmem::kPartOff);  // [0] e  [1] sum m k  [2] sum m k^2, each [role * 33 + f]
    float* fxp = reinterpret_cast<float*>(smem + AcSmem::kFxOff);
    float* red = reinterpret_cast<float*>(smem + AcSmem::kRedOff);
    float* roll_thr = reinterpret_cast<float*>(smem + AcSmem::kThrOff);                    // [32] rolloff threshold of frame f
    int* roll_rc = reinterpret_cast<int*>(smem + AcSmem::kThrOff) + 32;                    // [32] role holding the crossing (-1: none)
    double* roll_lo = reinterpret_cast<double*>(smem + AcSmem::kDsumOff);                  // [32] cumulative energy before that role
    float* pf = reinterpret_cast<float*>(smem + AcSmem::kPfOff);       // [0]=centroid [1]=bandwidth [2]=rolloff [3]=flux [4]=rms [5]=zcr [6]=dbmax
    float* sq16 = keep + AcSmem::kSq;  float* zc16 = keep + AcSmem::kZc;  float* lastx = keep + AcSmem::kLast;
    const int role = tid >> 5, f = tid & 31;
    const int kb = role * 16, ke = (role == 15) ? kBins : kb + 16;       // contiguous bin range of this role

    switch (ph) {
      case 0: {   // stage y[reflect(160 f0 - 200 + i)] (edge tiles; interior tiles arrive by TMA)
        const int64_t g0 = (int64_t)it.f0 * kHop - 200;
        constexpr int kRounds = (kAcRaw + kThreads - 1) / kThreads;     // 11 loads in flight, then the stores
        float v[kRounds];
#pragma unroll
        for (int j = 0; j < kRounds; ++j) {
          int64_t g = g0 + tid + kThreads * j;
          if (g < 0) g = -g;
          if (g >= it.L) g = 2 * (it.L - 1) - g;
          v[j] = (tid + kThreads * j < kAcRaw && g >= 0 && g < it.L) ? ldg(it.wave + g) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < kRounds; ++j)
          if (tid + kThreads * j < kAcRaw) raw[tid + kThreads * j] = v[j];
      } break;
      case 1: {   // padded copy for the STFT + 16-sample partials of the unfold frames
        {         // 128-bit loads; the four words of chunk v land at 4 v + (v >> 3) + e (conflict-free both ways)
          const float4* r4 = reinterpret_cast<const float4*>(raw);
#pragma unroll
          for (int j = 0; j < (kTileSamples / 4 + kThreads - 1) / kThreads; ++j) {
            const int v = tid + kThreads * j;
            if (v < kTileSamples / 4) {
              const float4 x = r4[v];
              float* dp = dbuf + 4 * v + (v >> 3);
              dp[0] = x.x; dp[1] = x.y; dp[2] = x.z; dp[3] = x.w;
            }
          }
        }
        if ((tid >> 5) <= (kGroups16 - 1) >> 5) {     // warps 0..10 (whole warps: the neighbour exchange below is a shuffle)
          // group t = unfold samples [16 t, 16 t + 16) = raw[200 + 16 t ...]: sum of squares and the number of sign
          // changes over the pairs (i - 1, i) ENDING in the group.  Four 128-bit loads in an order rotated by
          // rot(t) (conflict-free, see FbankKernel::group16); the sample before the group is the neighbour lane's
          // last one.  Signs as exact small floats: s = [x > 0] - [x < 0], n = s^2 = [x != 0]; a pair crosses iff
          // s_a s_c = -1, so crossings = (sum n_a n_c - sum s_a s_c) / 2 -- two FMAs per pair, no branches, zeros
          // never count (sentiment_analysis.py:830-832).
          const int t = tmin(tid, kGroups16 - 1);
          const int rot = ((t >> 1) + 2 * (t & 1)) & 3;
          const float4* g4 = reinterpret_cast<const float4*>(raw + 200 + 16 * t);
          float x[4][4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 v = g4[(q + rot) & 3];
            x[q][0] = v.x; x[q][1] = v.y; x[q][2] = v.z; x[q][3] = v.w;
          }
          float xprev;
#if AFE_CUDA && defined(__CUDA_ARCH__)
          {
            const float last = (rot == 0) ? x[3][3] : (rot == 1) ? x[2][3] : (rot == 2) ? x[1][3] : x[0][3];   // chunk 3
            xprev = __shfl_up_sync(0xffffffffu, last, 1);
            if ((tid & 31) == 0) xprev = raw[200 + 16 * t - 1];
          }
#else
          xprev = raw[200 + 16 * t - 1];
#endif
          float sg[4][4], nz[4][4];
          float sq0 = 0.f, sq1 = 0.f, sq2 = 0.f, sq3 = 0.f;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              sg[q][e] = ((x[q][e] > 0.f) ? 1.f : 0.f) - ((x[q][e] < 0.f) ? 1.f : 0.f);
              nz[q][e] = sg[q][e] * sg[q][e];
            }
            sq0 = fma_f(x[q][0], x[q][0], sq0); sq1 = fma_f(x[q][1], x[q][1], sq1);
            sq2 = fma_f(x[q][2], x[q][2], sq2); sq3 = fma_f(x[q][3], x[q][3], sq3);
          }
          const float sprev = ((xprev > 0.f) ? 1.f : 0.f) - ((xprev < 0.f) ? 1.f : 0.f), nprev = sprev * sprev;
          float zn0 = 0.f, zs0 = 0.f, zn1 = 0.f, zs1 = 0.f;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int cq = (q + rot) & 3;                         // chunk held by slot q; slot q - 1 holds chunk cq - 1
            const float sb = (cq == 0) ? sprev : sg[(q + 3) & 3][3];
            const float nb = (cq == 0) ? nprev : nz[(q + 3) & 3][3];
            zn0 = fma_f(nb, nz[q][0], zn0);       zs0 = fma_f(sb, sg[q][0], zs0);
            zn1 = fma_f(nz[q][0], nz[q][1], zn1); zs1 = fma_f(sg[q][0], sg[q][1], zs1);
            zn0 = fma_f(nz[q][1], nz[q][2], zn0); zs0 = fma_f(sg[q][1], sg[q][2], zs0);
            zn1 = fma_f(nz[q][2], nz[q][3], zn1); zs1 = fma_f(sg[q][2], sg[q][3], zs1);
          }
          if (tid < kGroups16) {
            sq16[t] = (sq0 + sq1) + (sq2 + sq3);
            zc16[t] = 0.5f * ((zn0 + zn1) - (zs0 + zs1));
          }
        }
        if (tid < kTileFrames)     // the pair (-1, 0) of unfold frame tid is counted by its first group but is NOT part of the frame
          lastx[tid] = crossing(raw[200 + kHop * tid - 1], raw[200 + kHop * tid]) ? 1.f : 0.f;
        // the record of the block's PREVIOUS tile, by the warps that have no groups above: a one-warp dependent chain
        // (rolloff) that used to be a phase of its own with 15 warps waiting at its barrier
        owed_record(smem, tid);
      } break;
      case 2: {   // Hann window + stage 1 of the real FFT
        float r[25];
        const float* col = dbuf + (kHop + kHop / 32) * f + role;  // phys_idx(160 f + 16 n1 + role) = 165 f + role + 16 n1 + (n1 >> 1)
        const float* wrow = reinterpret_cast<const float*>(smem + AcSmem::kHannOff) + role * kWinRow;
        const float* twrow = reinterpret_cast<const float*>(smem + AcSmem::kTwOff) + role * 32;
        windowed_column_fn<float>([wrow](int n1) { return wrow[n1]; }, r, [col](int n1) { return col[16 * n1 + (n1 >> 1)]; });
        rfft512_stage1_tw<float, kTileFrames>(r, role, f, [twrow](int k1, float& cc, float& ss) { cc = twrow[2 * k1]; ss = twrow[2 * k1 + 1]; }, ex);
      } break;
      case 3:     // stage 2: power (mel bank) and magnitude (spectral features)
        rfft512_stage2<float, true>(role, f, ex, sP, sM);
        break;
      case 4: {   // 128 HTK mel filters (slot form) -> dB rows; moments / flux of this role's contiguous bins
        float mx = -3.0e38f;
        float* row = sDb + f * kDbStride;
        mel_slots_htk<kTileFrames>(role, f, sP, reinterpret_cast<const float*>(smem + AcSmem::kMelWOff),
                                   reinterpret_cast<const int32_t*>(smem + AcSmem::kMelSOff), [row, &mx](int m, float acc) {
          const float db = 10.f * fast_log10(tmax(acc, 1e-10f));
          row[m] = db;
          mx = tmax(mx, db);
        });
        red[role * kRS + f] = mx;
        // moments over k = kb + j with the LOCAL index j, as running sums taken from the top bin down:
        //   a += m_j; b += a; q += b   =>   a = sum m, b = sum (j+1) m_j, q = sum (j+1)(j+2)/2 m_j
        // three float64 additions per bin, no index arithmetic (float32 magnitudes: the sums are exact to ~1e-16)
        double a = 0.0, b = 0.0, q = 0.0;
        float fx = 0.f;
        unsigned top = 0u;                                   // largest bit pattern seen (>= 0 floats order like integers)
        const bool has_next = (f + 1 < it.nvalid);
        // a ROLLED loop: the phase runs once per tile, so its cost on the B200 is the instruction fetch of
        // cold code -- 17 unrolled copies of this body were slower than the loop-carried FP64 chains
        {
          const float* mp = sM + kb * kTileFrames + f;
          const int nb = ke - kb;
#pragma unroll 4
          for (int j = nb - 1; j >= 0; --j) {
            const float m = mp[j * kTileFrames];
            top = tmax(top, float_bits(m));
            a += nonneg_to_double_biased(m); b += a; q += b;
            if (has_next) fx += tmax(mp[j * kTileFrames + 1] - m, 0.f);
          }
        }
        if (top > 0x7f7fffffu) a = b = q = (double)bits_float(0x7fc00000u);   // an infinite / NaN magnitude poisons the frame
        double e = a, c = b - a, c2 = (q + q) - 3.0 * b + a;   // sum m, sum j m, sum j^2 m
        {   // sum m k = kb e + c ; sum m k^2 = kb^2 e + 2 kb c + c2   (all exact products)
          const double kd = (double)kb;
          c2 = c2 + kd * (2.0 * c + kd * e);
          c = c + kd * e;
        }
        {
          const int idx = role * kRS + f;
          part[idx] = e; part[kPS + idx] = c; part[2 * kPS + idx] = c2; fxp[idx] = fx;
        }
      } break;
      case 5: {   // per-frame scalars: warp = quantity, lane = frame.  Every warp reads the 16 role partials of its frame
                  // into registers and reduces them there as fixed trees (tree16_sum / scan16: depth 4, the additions
                  // of the earlier 16-lane-segment shuffle version, so the same bits); the warps that need the total
                  // energy re-derive it with the same additions, so no value crosses warps and the phase has no
                  // shuffles (the shuffle version spent 1 900 cycles per tile on ~900 of them).
        const int fa = f;
        if (role == 0 || role == 1) {            // unfold frame fa = 25 groups of 16 samples: sum of squares / sign changes
          const float* g = ((role == 0) ? sq16 : zc16) + 10 * fa;
          float v[16];
#pragma unroll
          for (int r = 0; r < 16; ++r) v[r] = (r < 9) ? g[r] + g[16 + r] : g[r];
          const float t = tree16_sum(v);
          if (role == 0) pf[4 * 32 + fa] = mag_sqrt(t * (1.f / 400.f));
          else pf[5 * 32 + fa] = (t - lastx[fa]) * (1.f / 400.f);
        } else if (role == 2 || role == 3) {     // centroid / bandwidth from the three moments
          double e[16], c[16];
#pragma unroll
          for (int r = 0; r < 16; ++r) { e[r] = part[r * kRS + fa]; c[r] = part[kPS + r * kRS + fa]; }
          const double tot = tree16_sum(e), C = tree16_sum(c);
          const float energy = (float)tot + 1e-8f;
          const float cen = fast_div(31.25f * (float)C, energy);
          if (role == 2) pf[0 * 32 + fa] = cen;
          else {   // sum m (31.25 k - cen)^2 from the three moments, in float64
#pragma unroll
            for (int r = 0; r < 16; ++r) c[r] = part[2 * kPS + r * kRS + fa];
            const double C2 = tree16_sum(c);
            const double cd = (double)cen;
            const double num = (31.25 * 31.25) * C2 - 2.0 * 31.25 * cd * C + cd * cd * tot;
            pf[1 * 32 + fa] = mag_sqrt(fast_div((float)(num > 0.0 ? num : 0.0), energy));
          }
        } else if (role == 4) {                  // rolloff, first half: the role whose bins hold the 85 % crossing of the
                                                 // cumulative energy (float compare of in-order float64 sums, like ATen's
                                                 // cumsum against the float32 threshold); the last phase finds the bin
          double tot = 0.0;
#pragma unroll
          for (int r = 0; r < 16; ++r) tot += part[r * kRS + fa];
          const float thr = 0.85f * ((float)tot + 1e-8f);
          int rc = -1;
          double run = 0.0, lo_c = 0.0;
#pragma unroll
          for (int r = 0; r < 16; ++r) {         // the same additions again: run == the prefix of tot
            const double nxt = run + part[r * kRS + fa];
            if (rc < 0 && !((float)nxt < thr)) { rc = r; lo_c = run; }
            run = nxt;
          }
          roll_thr[fa] = thr; roll_rc[fa] = rc; roll_lo[fa] = lo_c;
        } else if (role == 5) {                  // in-tile spectral flux
          float v[16];
#pragma unroll
          for (int r = 0; r < 16; ++r) v[r] = fxp[r * kRS + fa];
          pf[3 * 32 + fa] = tree16_sum(v);
        } else if (role == 6) {                  // largest dB of the frame
          float a = red[fa];
#pragma unroll
          for (int r = 1; r < 16; ++r) a = tmax(a, red[r * kRS + fa]);
          pf[6 * 32 + fa] = a;
        } else {                                 // warps 7..15 (otherwise idle here): the tile's mel-dB rows and boundary
                                                 // spectra go to global memory while the scalar warps work
          constexpr int kFirst = 7 * 32, kN = kThreads - kFirst;
          // (scalar on purpose: 128-bit stores fed by four bank-conflicted loads measured slower); five loads in
          // flight per round
          float* dst = meldb + (it.row0 + it.f0) * kAcMel;
          const int total = it.nvalid * kAcMel;
          for (int i0 = tid - kFirst; i0 < total; i0 += 5 * kN) {
            float v[5];
#pragma unroll
            for (int u = 0; u < 5; ++u) { const int i = i0 + u * kN; v[u] = (i < total) ? sDb[(i >> 7) * kDbStride + (i & 127)] : 0.f; }
#pragma unroll
            for (int u = 0; u < 5; ++u) { const int i = i0 + u * kN; if (i < total) dst[i] = v[u]; }
          }
          // magnitudes of the first and the last valid frame (the same frame when the tile holds only one), straight
          // from the stage-2 output (514 bank-conflicted words per tile, off the critical path)
          for (int i = tid - kFirst; i < 2 * kBins; i += kN) {
            const int second = i >= kBins, k = i - second * kBins;
            boundary[(it.item * 2 + second) * kBoundary + k] = sM[k * kTileFrames + (second ? it.nvalid - 1 : 0)];
          }
        }
        if (tid == kThreads - 1) {              // this tile's record is owed: stash what it needs (consumed after the barrier)
          int64_t* stash = reinterpret_cast<int64_t*>(smem + AcSmem::kStashOff);
          const int n1 = tmax(0, tmin(kTileFrames, it.F1 - it.f0));
          stash[0] = (int64_t)it.item + 1;
          stash[1] = (int64_t)n1 | ((int64_t)it.nvalid << 16);
        }
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
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    double mx = -3.0e38;
#if AFE_CUDA && defined(__CUDA_ARCH__)
    for (long t = t0 + tid; t < t1; t += kThreads) {       // lane-strided: the loads of a waveform's tiles are in flight together
      const double v = tile_rec[t * kTileRec + 13];
      mx = v > mx ? v : mx;
    }
    mx = warp_tree_max(mx);
#else
    if (tid != 0) return;
    for (long t = t0; t < t1; ++t) {
      const double v = tile_rec[t * kTileRec + 13];
      mx = v > mx ? v : mx;
    }
#endif
    if (tid == 0) gmax[item] = (float)mx;
  }
};

// One item = one 32-frame tile; 4 warps, lane = frame, warp w owns mel bins [32w, 32w+32).
// The DCT coefficients are immediates of generated straight-line FFMAs (afe_dct_gen.h; sm_100
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
    if (ph == 0) {        // clamp at max - 80 dB while staging the tile (coalesced 128-bit loads, all issued first:
                          // one load per loop round left the phase bound by the L2 / DRAM round trip)
      const float4* src = reinterpret_cast<const float4*>(meldb + (it.row0 + it.f0) * kAcMel);
      const int nq = it.nvalid * (kAcMel / 4);
      constexpr int kRounds = kTileFrames * (kAcMel / 4) / kThreads;     // 8
      float4 v[kRounds];
#pragma unroll
      for (int j = 0; j < kRounds; ++j) {
        const int i = tid + kThreads * j;
        v[j] = (i < nq) ? ldcg4(src + i) : float4{0.f, 0.f, 0.f, 0.f};
      }
      // spectral flux across the tile boundary: relu(first frame of tile+1 - last frame of this tile)
      float b0[3], b1[3];
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int q = tid + kThreads * j;
        const bool on = it.has_next && q < kBins;
        b0[j] = on ? ldg(boundary + (it.item + 1) * 2 * kBoundary + q) : 0.f;
        b1[j] = on ? ldg(boundary + it.item * 2 * kBoundary + kBoundary + q) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < kRounds; ++j) {
        const int i = tid + kThreads * j;
        float4 c = v[j];
        if (i < nq) { c.x = tmax(c.x, it.cut); c.y = tmax(c.y, it.cut); c.z = tmax(c.z, it.cut); c.w = tmax(c.w, it.cut); }
        *reinterpret_cast<float4*>(sD + (i >> 5) * kRow + (i & 31) * 4) = c;
      }
      float fl = 0.f;
#pragma unroll
      for (int j = 0; j < 3; ++j) fl += tmax(b0[j] - b1[j], 0.f);
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
  static constexpr int kPhases = 2;
  AcDesc d; const double* tile_rec; const double* mfcc_rec; const float* boundary;
  float* out;              // [B][38]
  int32_t* status;         // [B]
  // optional StandardScaler epilogue (analyze_batch_sentiment, sentiment_analysis.py:736-752): float64 rows
  // (x - mean) / scale of the feature LIST the reference hands to scaler.transform -- tempo finished,
  // zero features for a failed waveform, padded / truncated to n_features
  const double* sc_mean; const double* sc_scale; double* scaled; int n_features;
  // the reference's tempo from the flux energy (:869-875, see acoustic.py:_tempo), in float64 like Python
  AFE_HD static double tempo(double flux_energy, int64_t num_samples) {
    const int64_t hop = 160, f2 = 1 + num_samples / hop;
    const int64_t max_lag = flux_energy > 0.0 ? f2 - 2 : 0;
    const double period = max_lag > 0 ? (double)(max_lag * hop) / 16000.0 : 0.0;
    return period > 0.0 ? 60.0 / period : 0.0;
  }
  // s += p[t * stride], ss += p[t * stride + 1] for t in [t0, t1), in tile order, eight tiles' loads in flight
  AFE_HD static void sum_pairs(const double* p, int stride, long t0, long t1, double& s, double& ss) {
    for (long t = t0; t < t1; t += 8) {
      double a[8], b[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) { const bool on = t + u < t1; a[u] = on ? p[(t + u) * stride] : 0.0; b[u] = on ? p[(t + u) * stride + 1] : 0.0; }
#pragma unroll
      for (int u = 0; u < 8; ++u) if (t + u < t1) { s += a[u]; ss += b[u]; }
    }
  }
  AFE_HD static void mean_std(double s, double ss, int n, float& mean, float& sd) {
    const double m = s / (double)n;
    mean = (float)m;
    const double var = (ss - s * m) / (double)(n - 1);          // n == 1 -> 0/0 = NaN like torch.std
    sd = (float)std::sqrt(var < 0.0 ? 0.0 : var);
  }
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return d.B; }
  AFE_HD void phase(int ph, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    const long orow = d.out_index ? (long)ldg(d.out_index + item) : item;
    float* o = out + orow * kAcFeat;
    const int F1 = ldg(d.frames1 + item), F2 = ldg(d.frames2 + item);
    if (ph == 1) {
      if (scaled == nullptr) return;
      const int64_t L = ldg(d.len + item);
      for (int i = tid; i < n_features; i += kThreads) {
        double x = 0.0;
        if (F2 > 0 && i < kAcFeat) {
          x = (double)o[i];
          if (i == 11) x = tempo(x, L);
        }
        scaled[orow * n_features + i] = (x - ldg(sc_mean + i)) / ldg(sc_scale + i);
      }
      return;
    }
    if (F2 <= 0) {                                   // 'Audio too short'
      if (tid == 0) status[orow] = 1;
      for (int i = tid; i < kAcFeat; i += kThreads) o[i] = 0.f;
      return;
    }
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    if (tid == 0) status[orow] = 0;
    if (tid < 5) {              // rms, zcr over F1 frames; centroid, bandwidth, rolloff over F2
      double s = 0.0, ss = 0.0, mn = 3.0e38, mx = -3.0e38;
      sum_pairs(tile_rec + 2 * tid, kTileRec, t0, t1, s, ss);
      if (tid == 0) {
        // min / max of the per-tile rms extremes, batched the same way
        for (long t = t0; t < t1; t += 8) {
          double a[8], c[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) { const bool on = t + u < t1; a[u] = on ? tile_rec[(t + u) * kTileRec + 11] : 3.0e38; c[u] = on ? tile_rec[(t + u) * kTileRec + 12] : -3.0e38; }
#pragma unroll
          for (int u = 0; u < 8; ++u) { mn = a[u] < mn ? a[u] : mn; mx = c[u] > mx ? c[u] : mx; }
        }
      }
      float mean, sd;
      mean_std(s, ss, tid < 2 ? F1 : F2, mean, sd);
      const int slot = (tid == 0) ? 0 : (tid == 1) ? 3 : (5 + 2 * (tid - 2));
      o[slot] = mean; o[slot + 1] = sd;
      if (tid == 0) o[2] = (float)(mx - mn);
    } else if (tid == 5) {      // spectral-flux energy: in-tile pairs + tile-boundary pairs
      double s = 0.0;
      for (long t = t0; t < t1; t += 8) {      // eight tiles' loads in flight, added in tile order
        double a[8], c[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { const bool on = t + u < t1; a[u] = on ? tile_rec[(t + u) * kTileRec + 10] : 0.0; c[u] = on ? mfcc_rec[((t + u) * 14 + 13) * 2] : 0.0; }
#pragma unroll
        for (int u = 0; u < 8; ++u) if (t + u < t1) s += a[u] + c[u];
      }
      o[11] = (float)s;
    } else if (tid >= 6 && tid < 6 + kMfcc) {
      const int c = tid - 6;
      double s = 0.0, ss = 0.0;
      sum_pairs(mfcc_rec + 2 * c, 28, t0, t1, s, ss);
      float mean, sd;
      mean_std(s, ss, F2, mean, sd);
      o[12 + 2 * c] = mean; o[13 + 2 * c] = sd;
    }
  }
};

}  // namespace afe
