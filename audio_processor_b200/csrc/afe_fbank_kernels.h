// afe_fbank_kernels.h -- Kaldi/SeamlessM4T fbank (afe_platform.h block programs + the two-group tile pipeline).
//
// Replaces the per-frame numpy loop of transformers audio_utils.spectrogram (:788-804) as
// called through SeamlessM4TFeatureExtractor (feature_extraction_seamless_m4t.py:112-139)
// from services/audio_processor.py:253-261 and services/speech_segment.py:226-231.
//
// One work item = a tile of kFbTile = 16 consecutive frames of one utterance; 16 roles x 16 frames = 256 threads
// (two roles per warp, lanes 0-15 / 16-31) work on a tile stage.  Linear path (pre-emphasis, DC removal, window,
// FFT) in T; power / mel / log in float.  See FbankPipeKernel below for the pipeline.
#pragma once
#include "afe_dct_gen.h"
#include "afe_mel_slots.h"
#include "afe_rfft512.h"
#include "afe_tables.h"

namespace afe {

constexpr int kFbTile = 16;                                         // frames per fbank tile
constexpr int kTileSamples = (kTileFrames - 1) * kHop + kFrame;     // 5360 (acoustic tile, 32 frames)
constexpr int kGroups16 = kTileSamples / 16;                        // 335 sums of 16 samples
constexpr int kOutStride = kMel + 1;                                // 81: conflict-free rows

AFE_HD int phys_idx(int i) { return i + (i >> 5); }                 // 1 pad word per 32: odd lane stride

struct BatchDesc {            // device-resident description of a batch of mono waveforms
  const float* const* wave;   // [B] device pointers
  const int64_t* len;         // [B] samples
  const int32_t* frames;      // [B] fbank frames T_b
  const int64_t* tile_prefix; // [B+1] exclusive prefix of ceil(T_b / kFbTile)
  const int64_t* out_row;     // [B] first frame row of utterance b in the feature buffer
  const int32_t* tp;          // [B] padded frame count of utterance b's group (even)
  const int32_t* whole;       // [B] 1: all tiles of b belong to one block of the fbank launch, which
                              //        normalises b itself; 0: FbankStatsKernel / FbankNormKernel do
  const int32_t* sel;         // [nsel] utterances left to the stats / norm kernels
  int nsel;
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

// (mean, 1 / denominator) of one mel bin from its float64 sum and sum of squares over n frames --
// the reciprocal is formed once per (utterance, bin) so that the normalisation of the utterance's T x 80 values
// is a subtract and a multiply (within 1.5 ulp of the reference's divide; the contract is 1e-3 absolute) --
// numpy's mean / var(ddof=1) / sqrt(var + 1e-7) of feature_extraction_seamless_m4t.py:283-288
AFE_HD void fbank_mean_den(double s, double ss, int n, float& mean_out, float& den_out) {
  const double mean = s / (double)n;
  const double var = (ss - s * mean) / (double)(n - 1);      // ddof = 1; n == 1 -> 0/0 = NaN
  const float varf = (float)var;
  mean_out = (float)mean;
  den_out = 1.f / std::sqrt(varf + 1e-7f);
}

// in-place (x - mean) * (1 / den) over 128-bit words [q0, q1) of one utterance's (tp x 80) plane, right
// padding with padding_value, stride-2 mask (the ODD frame's bit: mask[j] = (2j+1 < n)).
// stat = [80][2] (mean, 1 / den).  Shared by the fused tail phase and FbankNormKernel.
template <int kU>                 // words in flight per thread
AFE_HD void fbank_normalise_plane(float* feat, int32_t* mask, int64_t row0, int tp, int n, const float* stat,
                                  float padding_value, int tid, int nthreads, int64_t q0, int64_t q1) {
  constexpr int kQ = kMel / 4;     // 20 words per frame
                                   // kU: the loads of a round are issued before its stores
                                   // (one load per round left the loop bound by the L2 round trip)
  float4* base = reinterpret_cast<float4*>(feat + row0 * kMel);
  for (int64_t r4 = q0 + tid; r4 < q1; r4 += (int64_t)kU * nthreads) {
    float4 x[kU];
    int64_t t[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i4 = r4 + (int64_t)u * nthreads;
      t[u] = i4 / kQ;
      x[u] = (i4 < q1 && t[u] < n) ? ldcg4(base + i4) : float4{0.f, 0.f, 0.f, 0.f};
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t i4 = r4 + (int64_t)u * nthreads;
      if (i4 >= q1) break;
      const int m4 = (int)(i4 - t[u] * kQ);
      float4 y = float4{padding_value, padding_value, padding_value, padding_value};
      if (t[u] < n) {
        const float* st = stat + 8 * m4;
        y.x = (x[u].x - st[0]) * st[1]; y.y = (x[u].y - st[2]) * st[3];
        y.z = (x[u].z - st[4]) * st[5]; y.w = (x[u].w - st[6]) * st[7];
      }
      base[i4] = y;
      if (m4 == 0 && (t[u] & 1)) mask[(row0 >> 1) + (t[u] >> 1)] = (t[u] < n) ? 1 : 0;
    }
  }
  (void)tp;
}

// ------------------------------------------------------------------------------------------------------------
// FbankPipeKernel: the fbank tile pipeline as TWO warp groups working on neighbouring tiles.
//
// One block per SM, 512 threads = group S1 (warps 0-7) + group S2 (warps 8-15); a tile is 16 consecutive frames
// of one utterance, a group is 16 roles x 16 frames (two roles per warp).  In step s
//     S1:  FFT stage 1 of tile s      (dbuf[s & 1], dc[s & 1] -> ex[s & 1]),  then the pre-emphasis pass of tile
//          s + 1 (TMA-staged raw samples -> dbuf[(s + 1) & 1], frame DC terms -> dc[(s + 1) & 1])
//     S2:  FFT stage 2 of tile s - 1  (ex[(s - 1) & 1] -> power sP), mel + log (-> sOut), tile statistics + store
// and ONE block-wide barrier ends the step (the old kernel ran the six phases of a tile one after the other, nine
// barriers per tile, every phase a single-pipe burst: loads, then FP64, then stores).  The FP64-heavy stage 1 of
// one tile now runs under the shared-memory-heavy stage 2 / mel / statistics phases of the tile before it; inside
// S2 the three phases are ordered by a named barrier (bar.sync 1, 256), inside S1 the frame sums by bar.sync 2.
// The exchange buffer, the pre-emphasised samples and the DC terms are double-buffered, the raw staging buffer is
// single: thread 0 issues the TMA bulk copy of tile s + 2 right after the barrier of step s, when the pre-emphasis
// pass of tile s + 1 has read it, and the copy lands under stage 1 of tile s + 1.
// The host emulation (tests only) runs the same member functions group after group, sub-phase after sub-phase.
#ifndef AFE_FB_SWAP
#define AFE_FB_SWAP 0          // 1: S1 = warps 8-15 (the SM's issue arbiter prefers the higher warp ids)
#endif
#ifndef AFE_FB_DEFER
#define AFE_FB_DEFER 0         // 1: S2 runs the statistics + store of tile s - 2 between stage 2 and mel of tile s - 1
#endif
template <typename T> struct FbankPipeSmem {
  static constexpr int TF = kFbTile;
  static constexpr int kSamples = (TF - 1) * kHop + kFrame;                         // 2800
  static constexpr int kG16 = kSamples / 16;                                        // 175 sums of 16 samples
  static constexpr int kG80 = kG16 / 5;                                             // 35 sums of 80 samples = 2 TF + 3
  static constexpr int kDStride = (kSamples + kSamples / 32 + 8 + 1) / 2 * 2;        // padded tile, elements of T
  static constexpr size_t kRawOff = 0;                                              // float raw[kSamples] (TMA destination)
  static constexpr size_t kSumOff = kRawOff + sizeof(float) * kSamples;             // double psum80[kG80 + 1]
  static constexpr size_t kDcOff = kSumOff + sizeof(double) * (kG80 + 1);           // T dc[2][TF]
  static constexpr size_t kAccOff = kDcOff + sizeof(double) * 2 * TF;               // double acc[80][2]: running (sum, sum sq)
  static constexpr size_t kStatOff = kAccOff + sizeof(double) * 160;                // float stat[80][2]: (mean, 1 / denominator)
  static constexpr size_t kDOff = kStatOff + sizeof(float) * 160;                   // T dbuf[2][kDStride]
  static constexpr size_t kPOff = kDOff + 2 * sizeof(T) * kDStride;                 // float sP[257][TF]
  static constexpr size_t kOutOff = kPOff + sizeof(float) * kBins * TF;             // float sOut[TF][81]
  static constexpr int kOutBufs = AFE_FB_DEFER == 1 ? 2 : 1;
  static constexpr size_t kOutBytes = sizeof(float) * TF * kOutStride;
  static constexpr size_t kExOff = (kOutOff + kOutBufs * kOutBytes + 15) / 16 * 16;   // cplx<T> ex[2][16][16][TF]
  static constexpr size_t kExBytes = sizeof(cplx<T>) * 16 * 16 * TF;
  static constexpr size_t kWinOff = kExOff + 2 * kExBytes;                          // T window[16][26]
  static constexpr size_t kTwOff = kWinOff + sizeof(T) * kRoles * kWinRow;          // T tw512[16][16][2]
  static constexpr size_t kMelWOff = kTwOff + sizeof(T) * kRoles * 32;              // float melw[16][40]
  static constexpr size_t kMelSOff = kMelWOff + sizeof(float) * kRoles * kMelSlotRow;   // int32 mels[16][8]
  static constexpr size_t kBytes = kMelSOff + sizeof(int32_t) * kRoles * kMelSlotStarts;
  static_assert(kMelWOff % 16 == 0 && kExOff % 16 == 0 && kDOff % 16 == 0, "128-bit rows");
  static_assert(kSamples % 80 == 0 && kG80 == 2 * TF + 3, "tile samples must split into groups of 16 and 80");
  static_assert(kBytes <= 227 * 1024, "one block per SM");
};

template <typename T> struct FbankPipeKernel {
  using S = FbankPipeSmem<T>;
  static constexpr int TF = kFbTile;
  static constexpr int kGroup = kRoles * TF;              // 256 threads per warp group
  static constexpr int kThreads = 2 * kGroup;
  static constexpr int kSamples = S::kSamples;
  static constexpr int kG16 = S::kG16;
  static constexpr int kG80 = S::kG80;
  static constexpr int kSumLanes = 2 * kMel;              // S2 threads [0, 160): two per mel bin (8 frames each)
  static_assert((kG16 + 29) / 30 * 32 <= kGroup && kSumLanes < kGroup, "phase thread budgets");

  struct Item { long item; int b; int f0; int nfr; int nvalid; int count; int whole; int tp; int bulk;
                const float* src; int64_t out_row; };

  BatchDesc d;
  float* out;              // [rows][80] raw log-mel lands at its final position
  double* partials;        // [ntiles][80][2]  (sum, sum of squares) per mel bin
  const T* window;         // [16][26] role-major: povey[16 n1 + role] * 32768 / 2   (copied to shared memory
  const T* tw512;          // [16][16][2]                                            once per block: block_init)
  const float* mel_w;      // [16][40] slot-form Kaldi bank (afe_mel_slots.h), copied to shared memory by block_init
  const int32_t* mel_start;   // [16][8]
  float log_floor;         // float32(log(float64(mel_floor)))
  int32_t* mask;           // [rows/2]   (fused normalisation only)
  float padding_value;

  AFE_HD long num_items() const { return d.ntiles; }
  // the last tile of a waveform whose tiles all ran in this block: the block normalises it
  AFE_HD bool tail(const Item& it) const { return it.whole != 0 && it.f0 + TF >= it.nfr; }

  // once per block: the FFT and mel tables move to shared memory (a table read is then a broadcast load)
  AFE_HD void block_init(unsigned char* smem, int tid) const {
    T* w = reinterpret_cast<T*>(smem + S::kWinOff);
    T* t = reinterpret_cast<T*>(smem + S::kTwOff);
    for (int i = tid; i < kRoles * kWinRow; i += kThreads) w[i] = ldg(window + i);
    for (int i = tid; i < kRoles * 32; i += kThreads) t[i] = ldg(tw512 + i);
    float* mw = reinterpret_cast<float*>(smem + S::kMelWOff);
    int32_t* ms = reinterpret_cast<int32_t*>(smem + S::kMelSOff);
    for (int i = tid; i < kRoles * kMelSlotRow; i += kThreads) mw[i] = ldg(mel_w + i);
    for (int i = tid; i < kRoles * kMelSlotStarts; i += kThreads) ms[i] = ldg(mel_start + i);
  }

  AFE_HD void fill(Item& it, int b, int f0) const {
    it.b = b; it.f0 = f0;
    it.whole = d.whole ? ldg(d.whole + b) : 0;
    it.tp = ldg(d.tp + b);
    it.nfr = ldg(d.frames + b);
    it.nvalid = tmin(TF, it.nfr - f0);
    it.count = (it.nvalid - 1) * kHop + kFrame;
    it.src = ldg(d.wave + b) + (int64_t)f0 * kHop;
    it.out_row = ldg(d.out_row + b);
    // the tile is one contiguous run of `count` samples: a single TMA bulk copy when 16-byte aligned
    it.bulk = (reinterpret_cast<uintptr_t>(it.src) & 15u) == 0 ? 1 : 0;
  }
  AFE_HD Item prepare(long item) const {
    Item it;
    int b, t;
    locate_tile(d, item, b, t);
    it.item = item;
    fill(it, b, t * TF);
    return it;
  }
  // item cur.item + 1: the next tile of the same utterance, else the first tile of the next
  // utterance that has frames (no search; utterance data is reloaded only on a crossing)
  AFE_HD Item next(const Item& cur, long item) const {
    Item it = cur;
    it.item = item;
    if (cur.f0 + TF < cur.nfr) {
      it.f0 = cur.f0 + TF;
      it.nvalid = tmin(TF, cur.nfr - it.f0);
      it.count = (it.nvalid - 1) * kHop + kFrame;
      it.src = cur.src + (int64_t)TF * kHop;
      it.bulk = (reinterpret_cast<uintptr_t>(it.src) & 15u) == 0 ? 1 : 0;
      return it;
    }
    int b = cur.b + 1;
    while (b + 1 < d.B && ldg(d.frames + b) <= 0) ++b;
    fill(it, b, 0);
    return it;
  }

  // ---- S1, pre-emphasis pass -------------------------------------------------------------------------------
  // cooperative staging of a tile the TMA cannot move (unaligned source)
  AFE_HD void pre_copy(const Item& it, unsigned char* smem, int g) const {
    float* raw = reinterpret_cast<float*>(smem + S::kRawOff);
    for (int i = g; i < it.count; i += kGroup) raw[i] = it.src[i];
  }
  // Samples [16 t, 16 t + 16) of the tile: pre-emphasis x[n] - 0.97 x[n-1] in T into the padded layout and the
  // float64 sum of the 16 raw samples.  The four 128-bit chunks are visited in an order rotated by rot(t), which
  // makes both the 128-bit loads and the 64-bit stores (element 16 t + j at 16 t + (t >> 1) + j) bank-conflict
  // free; every sample is converted to T once.  slot q of v holds chunk (q + rot) & 3.
  AFE_HD static int group16_rot(int t) { return ((t >> 1) + 2 * (t & 1)) & 3; }
  AFE_HD static void group16_load(const float* raw, int t, float4 (&v)[4]) {
    const int rot = group16_rot(t);
    const float4* g4 = reinterpret_cast<const float4*>(raw + 16 * t);
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = g4[(q + rot) & 3];
  }
  AFE_HD static float group16_last(const float4 (&v)[4], int t) {   // sample 16 t + 15: chunk 3 sits in slot (3 - rot) & 3
    const int rot = group16_rot(t);
    return rot == 0 ? v[3].w : rot == 1 ? v[2].w : rot == 2 ? v[1].w : v[0].w;
  }
  // `prev` = sample 16 t - 1 (the neighbour lane's last sample; 0 for the first group: window[0] is 0)
  AFE_HD static double group16_full(const float4 (&v)[4], T* dp, int t, float prev) {   // dp = dbuf + 16 t + (t >> 1)
    const int rot = group16_rot(t);
    T x[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { x[q][0] = (T)v[q].x; x[q][1] = (T)v[q].y; x[q][2] = (T)v[q].z; x[q][3] = (T)v[q].w; }
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { s0 += (double)x[q][0]; s1 += (double)x[q][1]; s2 += (double)x[q][2]; s3 += (double)x[q][3]; }
    const T xp = (T)prev;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int cq = (q + rot) & 3;                            // chunk held by slot q; slot q-1 holds chunk cq-1
      const T xm = (cq == 0) ? xp : x[(q + 3) & 3][3];
      T* o = dp + 4 * cq;
      o[0] = x[q][0] - T(0.97) * xm;
      o[1] = x[q][1] - T(0.97) * x[q][0];
      o[2] = x[q][2] - T(0.97) * x[q][1];
      o[3] = x[q][3] - T(0.97) * x[q][2];
    }
    return (s0 + s1) + (s2 + s3);
  }
  // a group that reaches past the tile's valid samples (zeros beyond `count`)
  AFE_HD static double group16_edge(const float* raw, T* dbuf, int t, int count, bool store) {
    double s = 0.0;
    for (int j = 0; j < 16; ++j) {
      const int i = 16 * t + j;
      const float xv = (i < count) ? raw[i] : 0.f;
      if (store) {
        const float xmv = (i > 0 && i - 1 < count) ? raw[i - 1] : 0.f;
        dbuf[phys_idx(i)] = (T)xv - T(0.97) * (T)xmv;
      }
      s += (double)xv;
    }
    return s;
  }
  // lane l < 30 of warp w owns group t = 30 w + l of 16 samples; five neighbouring lanes make one sum of 80
  AFE_HD void pre_a(const Item& it, int par, unsigned char* smem, int g) const {
    const float* raw = reinterpret_cast<const float*>(smem + S::kRawOff);
    double* psum80 = reinterpret_cast<double*>(smem + S::kSumOff);
    T* dbuf = reinterpret_cast<T*>(smem + S::kDOff) + (size_t)par * S::kDStride;
    const int l = g & 31, t = 30 * (g >> 5) + l;
    const bool active = l < 30 && t < kG16;
    const bool full = active && 16 * t + 16 <= it.count;
    float4 v[4] = {float4{0.f, 0.f, 0.f, 0.f}, float4{0.f, 0.f, 0.f, 0.f}, float4{0.f, 0.f, 0.f, 0.f}, float4{0.f, 0.f, 0.f, 0.f}};
#if AFE_CUDA && defined(__CUDA_ARCH__)
    if ((g >> 5) * 30 >= kG16) return;                 // whole warps without groups
    if (full) group16_load(raw, t, v);
    // the sample before a full group is the neighbour lane's last sample: one shuffle instead of a 15-way
    // bank-conflicted scalar load (lane 0 of a warp, one lane, reads it)
    float prev = __shfl_up_sync(0xffffffffu, group16_last(v, t), 1);
    if (l == 0) prev = (full && t > 0) ? raw[16 * t - 1] : 0.f;
    double s16 = 0.0;
    if (full) s16 = group16_full(v, dbuf + 16 * t + (t >> 1), t, prev);
    else if (active) s16 = group16_edge(raw, dbuf, t, it.count, true);
    const double a = s16 + __shfl_down_sync(0xffffffffu, s16, 1);      // lanes 5 g, 5 g + 2: pairs
    const double b = a + __shfl_down_sync(0xffffffffu, a, 2);          // lane 5 g: (g0 + g1) + (g2 + g3)
    const double c5 = b + __shfl_down_sync(0xffffffffu, s16, 4);       //           ... + g4
    if (active && l % 5 == 0) psum80[t / 5] = c5;
#else
    if (!active) return;
    auto sum16 = [&](int tt, bool store) {
      if (16 * tt + 16 > it.count) return group16_edge(raw, dbuf, tt, it.count, store);
      float4 u[4];
      group16_load(raw, tt, u);
      if (store) return group16_full(u, dbuf + 16 * tt + (tt >> 1), tt, tt > 0 ? raw[16 * tt - 1] : 0.f);
      T scratch[16];                                 // same sums, stores into a scratch row
      return group16_full(u, scratch, tt, 0.f);
    };
    const double s16 = sum16(t, true);
    (void)full; (void)v;
    if (l % 5 == 0) {   // host emulation runs the lanes one after the other: same tree, recomputed
      double q[5] = {s16, 0.0, 0.0, 0.0, 0.0};
      for (int m = 1; m < 5; ++m) q[m] = sum16(t + m, false);
      psum80[t / 5] = ((q[0] + q[1]) + (q[2] + q[3])) + q[4];
    }
#endif
  }
  // frame f = samples [160 f, 160 f + 400) = five sums of 80: dc[f] = (1 - 0.97) * mean(frame)
  AFE_HD void pre_b(int par, unsigned char* smem, int g) const {
    if (g >= TF) return;
    const double* p5 = reinterpret_cast<const double*>(smem + S::kSumOff) + 2 * g;
    const double tot = ((p5[0] + p5[1]) + (p5[2] + p5[3])) + p5[4];
    reinterpret_cast<T*>(smem + S::kDcOff)[par * TF + g] = (T)((1.0 - 0.97) * (tot * (1.0 / (double)kFrame)));
  }

  // ---- S1, FFT stage 1: DC removal + window + the 16 real 32-point column transforms --------------------------
  AFE_HD void stage1(int par, unsigned char* smem, int g) const {
    const int role = g / TF, f = g % TF;
    const T* dbuf = reinterpret_cast<const T*>(smem + S::kDOff) + (size_t)par * S::kDStride;
    cplx<T>* ex = reinterpret_cast<cplx<T>*>(smem + S::kExOff + (size_t)par * S::kExBytes);
    const T c = reinterpret_cast<const T*>(smem + S::kDcOff)[par * TF + f];
    T r[25];
    const T* col = dbuf + (kHop + kHop / 32) * f + role;      // phys_idx(160 f + 16 n1 + role) = 165 f + role + 16 n1 + (n1 >> 1)
    const T* wrow = reinterpret_cast<const T*>(smem + S::kWinOff) + role * kWinRow;
    const T* twrow = reinterpret_cast<const T*>(smem + S::kTwOff) + role * 32;
    windowed_column_fn<T>([wrow](int n1) { return wrow[n1]; }, r,
                          [col, c](int n1) { return col[16 * n1 + (n1 >> 1)] - c; });
    rfft512_stage1_tw<T, TF>(r, role, f, [twrow](int k1, T& cc, T& ss) { cc = twrow[2 * k1]; ss = twrow[2 * k1 + 1]; }, ex);
  }

  // ---- S2 ----------------------------------------------------------------------------------------------------
  AFE_HD void stage2(int par, unsigned char* smem, int g) const {
    const cplx<T>* ex = reinterpret_cast<const cplx<T>*>(smem + S::kExOff + (size_t)par * S::kExBytes);
    rfft512_stage2<T, false, TF>(g / TF, g % TF, ex, reinterpret_cast<float*>(smem + S::kPOff));
  }
  // triangular mel bank (slot form: one code path for all roles) + floor + log
  AFE_HD void mel(int opar, unsigned char* smem, int g) const {
    const int role = g / TF, f = g % TF;
    float* row = reinterpret_cast<float*>(smem + S::kOutOff + (size_t)opar * S::kOutBytes) + f * kOutStride;
    const float lf = log_floor;
    mel_slots_kaldi<TF>(role, f, reinterpret_cast<const float*>(smem + S::kPOff),
                        reinterpret_cast<const float*>(smem + S::kMelWOff),
                        reinterpret_cast<const int32_t*>(smem + S::kMelSOff), [row, lf](int m, float acc) {
      row[m] = (acc <= kMelFloor) ? lf : fast_log(acc);
    });
  }
  // (sum, sum of squares) in float64 of mel bin m over half h (8 frames) of the tile's valid frames
  AFE_HD static void half_sums(const float* sOut, int m, int h, int nvalid, double& s, double& ss) {
    constexpr int kQ = TF / 2;
    double s0 = 0.0, s1 = 0.0, q0 = 0.0, q1 = 0.0;
    const float* col = sOut + (h * kQ) * kOutStride + m;
    if ((h + 1) * kQ <= nvalid) {
#pragma unroll
      for (int i = 0; i < kQ; i += 2) {
        const double a = (double)col[i * kOutStride], b = (double)col[(i + 1) * kOutStride];
        s0 += a; q0 += a * a; s1 += b; q1 += b * b;
      }
    } else {
      for (int i = 0; i < kQ && h * kQ + i < nvalid; ++i) {
        const double a = (double)col[i * kOutStride];
        s0 += a; q0 += a * a;
      }
    }
    s = s0 + s1; ss = q0 + q1;
  }
  // per-bin partial statistics (float64) + coalesced store of the raw log-mel
  AFE_HD void stats_store(const Item& it, int opar, unsigned char* smem, int g) const {
    const float* sOut = reinterpret_cast<const float*>(smem + S::kOutOff + (size_t)opar * S::kOutBytes);
    if (g < kSumLanes) {
      const int m = g >> 1, h = g & 1;
      double ts, tq;
      half_sums(sOut, m, h, it.nvalid, ts, tq);
#if AFE_CUDA && defined(__CUDA_ARCH__)
      ts += __shfl_xor_sync(0xffffffffu, ts, 1); tq += __shfl_xor_sync(0xffffffffu, tq, 1);
#else
      if (h == 0) { double s1, q1; half_sums(sOut, m, 1, it.nvalid, s1, q1); ts = ts + s1; tq = tq + q1; }
#endif
      if (h != 0) return;
      if (!it.whole) {
        partials[(it.item * kMel + m) * 2] = ts;
        partials[(it.item * kMel + m) * 2 + 1] = tq;
      } else {
        // same in-order float64 reduction over the tiles as FbankStatsKernel, kept in shared memory
        double* acc = reinterpret_cast<double*>(smem + S::kAccOff);
        const double as = (it.f0 == 0) ? 0.0 + ts : acc[2 * m] + ts;
        const double aq = (it.f0 == 0) ? 0.0 + tq : acc[2 * m + 1] + tq;
        acc[2 * m] = as; acc[2 * m + 1] = aq;
        if (it.f0 + TF >= it.nfr) {
          float* stat = reinterpret_cast<float*>(smem + S::kStatOff);
          fbank_mean_den(as, aq, it.nfr, stat[2 * m], stat[2 * m + 1]);
        }
      }
    } else {
      // rows are 320 bytes: 20 128-bit words per frame, destination 16-byte aligned
      float4* dst = reinterpret_cast<float4*>(out + (it.out_row + it.f0) * kMel);
      for (int i = g - kSumLanes; i < it.nvalid * (kMel / 4); i += kGroup - kSumLanes) {
        const int row = i / (kMel / 4), c4 = i - row * (kMel / 4);
        const float* sp = sOut + row * kOutStride + 4 * c4;
        dst[i] = float4{sp[0], sp[1], sp[2], sp[3]};
      }
    }
  }
  // tail: normalise the whole utterance in place (its raw rows are this block's own, L2-hot); all 512 threads
  AFE_HD void tail_norm(const Item& it, unsigned char* smem, int tid) const {
    const float* stat = reinterpret_cast<const float*>(smem + S::kStatOff);
    fbank_normalise_plane<8>(out, mask, it.out_row, it.tp, it.nfr, stat, padding_value, tid, kThreads, 0,
                             (int64_t)it.tp * (kMel / 4));
  }
};

#if AFE_CUDA
__device__ __forceinline__ void group_bar(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }

// kTrace (debug, AFE_PHASE_TRACE=1): lane 0 of every warp adds up the cycles from each step's start to its events
template <class K, bool kTrace>
__global__ void __launch_bounds__(K::kThreads, 1) fbank_pipe_kernel(const K k, unsigned long long* trace) {
  extern __shared__ __align__(16) unsigned char afe_smem[];
  using Item = typename K::Item;
  constexpr int kRing = 8;                 // item ring: tile t is created in step t - 2 and last read in step t + 2
  constexpr int kLag = AFE_FB_DEFER ? 2 : 1;       // S2's statistics + store run kLag steps behind stage 1
  __shared__ __align__(8) uint64_t bar_raw;
  __shared__ Item items[kRing];
  const long n = k.num_items();
  const int tid = (int)threadIdx.x;
  const long per = (n + gridDim.x - 1) / gridDim.x;
  const long first = (long)blockIdx.x * per;
  const long end = (first + per < n) ? first + per : n;
  if (first >= end) return;
  unsigned char* raw = afe_smem + K::S::kRawOff;
  const bool s1 = AFE_FB_SWAP ? tid >= K::kGroup : tid < K::kGroup;
  const int g = tid % K::kGroup;
  const bool lead = s1 && g == 0;          // S1's first thread keeps the item ring and issues the TMA copies
  k.block_init(afe_smem, tid);
  if (lead) {
    mbar_init(&bar_raw, 1); fence_mbar_init();
    Item c0 = k.prepare(first);
    if (c0.bulk) { mbar_expect_tx(&bar_raw, (uint32_t)c0.count * 4u); bulk_g2s(raw, c0.src, (uint32_t)c0.count * 4u, &bar_raw); }
    items[first % kRing] = c0;
    if (first + 1 < end) items[(first + 1) % kRing] = k.next(c0, first + 1);
  }
  __syncthreads();
  uint32_t rawpar = 0u;
  long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long t0 = 0;
  (void)tacc; (void)t0;
  // the pre-emphasis pass of a tile (S1 group): wait for / make the staged samples, sums, DC terms
  auto pre = [&](const Item& it, int par) {
    if (it.bulk) { mbar_wait(&bar_raw, rawpar); rawpar ^= 1u; }
    else { k.pre_copy(it, afe_smem, g); group_bar(2); }
    if constexpr (kTrace) tacc[2] += clock64() - t0;
    k.pre_a(it, par, afe_smem, g);
    if constexpr (kTrace) tacc[3] += clock64() - t0;
    group_bar(2);
    if constexpr (kTrace) tacc[4] += clock64() - t0;
    k.pre_b(par, afe_smem, g);
  };
  auto issue = [&](const Item& nx) {
    if (nx.bulk) { fence_proxy_async(); mbar_expect_tx(&bar_raw, (uint32_t)nx.count * 4u); bulk_g2s(raw, nx.src, (uint32_t)nx.count * 4u, &bar_raw); }
  };
  if (s1) pre(items[first % kRing], (int)(first & 1));
  __syncthreads();
  if (lead && first + 1 < end) issue(items[(first + 1) % kRing]);
  for (long s = first; s < end + kLag; ++s) {
    if constexpr (kTrace) t0 = clock64();
    // kTrace: tacc[i] += cycles from the step's start to event i of this warp (S1: 0 stage 1 done, 1 at the step barrier;
    // S2: 0 stage 2 done, 1 second work item done / past the first group barrier, 2..4 the following events);
    // 5: past the step barrier, 6: past the tail, 7: steps
    if (s1) {
      if (lead && s + 2 < end) items[(s + 2) % kRing] = k.next(items[(s + 1) % kRing], s + 2);
      if (s < end) k.stage1((int)(s & 1), afe_smem, g);
      if constexpr (kTrace) tacc[0] += clock64() - t0;
      if (s + 1 < end) pre(items[(s + 1) % kRing], (int)((s + 1) & 1));
      if constexpr (kTrace) tacc[1] += clock64() - t0;
    } else {
      const long t = s - 1;                   // tile in FFT stage 2 / mel
      const bool on = t >= first && t < end;
#if AFE_FB_DEFER == 2
      // mel + statistics of tile s - 2 first, stage 2 of tile s - 1 last: S2's FP64 work falls under S1's pre-emphasis pass
      const long u = s - 2;
      if (u >= first) k.mel(0, afe_smem, g);
      if constexpr (kTrace) tacc[0] += clock64() - t0;
      group_bar(1);
      if constexpr (kTrace) tacc[1] += clock64() - t0;
      if (u >= first) k.stats_store(items[u % kRing], 0, afe_smem, g);
      if constexpr (kTrace) tacc[2] += clock64() - t0;
      if (on) k.stage2((int)(t & 1), afe_smem, g);
      if constexpr (kTrace) { tacc[3] += clock64() - t0; tacc[4] += clock64() - t0; }
#elif AFE_FB_DEFER
      const long u = s - 2;                   // tile in statistics + store
      if (on) k.stage2((int)(t & 1), afe_smem, g);
      if constexpr (kTrace) tacc[0] += clock64() - t0;
      if (u >= first) k.stats_store(items[u % kRing], (int)(u & 1), afe_smem, g);
      if constexpr (kTrace) tacc[1] += clock64() - t0;
      group_bar(1);
      if constexpr (kTrace) tacc[2] += clock64() - t0;
      if (on) k.mel((int)(t & 1), afe_smem, g);
      if constexpr (kTrace) { tacc[3] += clock64() - t0; tacc[4] += clock64() - t0; }
#else
      if (on) {
        k.stage2((int)(t & 1), afe_smem, g);
        if constexpr (kTrace) tacc[0] += clock64() - t0;
        group_bar(1);
        if constexpr (kTrace) tacc[1] += clock64() - t0;
        k.mel(0, afe_smem, g);
        if constexpr (kTrace) tacc[2] += clock64() - t0;
        group_bar(1);
        if constexpr (kTrace) tacc[3] += clock64() - t0;
        k.stats_store(items[t % kRing], 0, afe_smem, g);
        if constexpr (kTrace) tacc[4] += clock64() - t0;
      }
#endif
    }
    __syncthreads();
    if constexpr (kTrace) { tacc[5] += clock64() - t0; tacc[7] += 1; }
    const long done = s - kLag;               // its statistics are complete, its rows stored
    if (done >= first && k.tail(items[done % kRing])) {
      k.tail_norm(items[done % kRing], afe_smem, tid);
      __syncthreads();
    }
    if constexpr (kTrace) tacc[6] += clock64() - t0;
    if (lead && s + 2 < end) issue(items[(s + 2) % kRing]);    // the raw buffer is free: the pre-emphasis pass of tile s + 1 has read it
  }
  if constexpr (kTrace) {
    if ((tid & 31) == 0) for (int i = 0; i < 8; ++i) trace[((size_t)blockIdx.x * 16 + (tid >> 5)) * 8 + i] = (unsigned long long)tacc[i];
  }
}

template <class K>
int afe_launch_fbank_pipe(const K& k, int max_blocks, cudaStream_t stream) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  const size_t smem_bytes = K::S::kBytes;
  const long grid = afe_async_grid(n, max_blocks);
  static const bool tracing = std::getenv("AFE_PHASE_TRACE") != nullptr;
  if (tracing) {      // debug only: synchronous; mean cycles per tile of each sub-phase, per warp group
    const void* tfn = (const void*)fbank_pipe_kernel<K, true>;
    cudaFuncSetAttribute(tfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    unsigned long long* tr = nullptr;
    const size_t nslots = (size_t)grid * 16 * 8;
    cudaMallocManaged(&tr, sizeof(unsigned long long) * nslots);
    std::memset(tr, 0, sizeof(unsigned long long) * nslots);
    fbank_pipe_kernel<K, true><<<(unsigned)grid, K::kThreads, smem_bytes, stream>>>(k, tr);
    cudaStreamSynchronize(stream);
    double acc[16][8] = {{0}}, steps = 0;
    for (long b = 0; b < grid; ++b) {
      steps += (double)tr[(size_t)b * 16 * 8 + 7];
      for (int w = 0; w < 16; ++w)
        for (int i = 0; i < 7; ++i) acc[w][i] += (double)tr[((size_t)b * 16 + w) * 8 + i];
    }
    std::fprintf(stderr, "[afe trace] %s: %ld blocks, %.0f steps; mean cycles from the step's start, per warp\n", __PRETTY_FUNCTION__, grid, steps);
    for (int w = 0; w < 16; ++w) {
      if ((w < 8) != (AFE_FB_SWAP != 0)) std::fprintf(stderr, "[afe trace]   S1 warp %2d: stage1 done %5.0f  raw ready %5.0f  pre_a done %5.0f  past its barrier %5.0f  at step barrier %5.0f  past it %5.0f  past tail %5.0f\n", w,
                              acc[w][0] / steps, acc[w][2] / steps, acc[w][3] / steps, acc[w][4] / steps, acc[w][1] / steps, acc[w][5] / steps, acc[w][6] / steps);
      else std::fprintf(stderr, "[afe trace]   S2 warp %2d: events %5.0f %5.0f %5.0f %5.0f %5.0f  past step barrier %5.0f  past tail %5.0f\n", w,
                        acc[w][0] / steps, acc[w][1] / steps, acc[w][2] / steps, acc[w][3] / steps, acc[w][4] / steps, acc[w][5] / steps, acc[w][6] / steps);
    }
    cudaFree(tr);
    return 0;
  }
  const void* fn = (const void*)fbank_pipe_kernel<K, false>;
  const cudaError_t e0 = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e0 != cudaSuccess) { note_error("cudaFuncSetAttribute(max dynamic smem)", cudaGetErrorString(e0)); return -1; }
  fbank_pipe_kernel<K, false><<<(unsigned)grid, K::kThreads, smem_bytes, stream>>>(k, nullptr);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { note_error("kernel launch", cudaGetErrorString(e)); return -1; }
  return 0;
}
#else
// host emulation (tests only): the same member functions, group after group, sub-phase after sub-phase
template <class K>
int afe_launch_fbank_pipe(const K& k, int max_blocks, cudaStream_t) {
  using Item = typename K::Item;
  const long n = k.num_items();
  if (n <= 0) return 0;
  std::vector<double> backing((K::S::kBytes + 64) / sizeof(double) + 2);
  unsigned char* smem = reinterpret_cast<unsigned char*>(backing.data());
  const long grid = afe_async_grid(n, max_blocks);
  const long per = (n + grid - 1) / grid;
  for (long block = 0; block < grid; ++block) {
    const long first = block * per, end = first + per < n ? first + per : n;
    if (first >= end) continue;
    for (int tid = 0; tid < K::kThreads; ++tid) k.block_init(smem, tid);
    constexpr int kRing = 8;
    constexpr int kLag = AFE_FB_DEFER ? 2 : 1;
    Item items[kRing];
    items[first % kRing] = k.prepare(first);
    auto pre = [&](const Item& it, int par) {
      for (int g = 0; g < K::kGroup; ++g) k.pre_copy(it, smem, g);
      for (int g = 0; g < K::kGroup; ++g) k.pre_a(it, par, smem, g);
      for (int g = 0; g < K::kGroup; ++g) k.pre_b(par, smem, g);
    };
    pre(items[first % kRing], (int)(first & 1));
    for (long s = first; s < end + kLag; ++s) {
      if (s + 1 < end) items[(s + 1) % kRing] = k.next(items[s % kRing], s + 1);
      if (s < end) for (int g = 0; g < K::kGroup; ++g) k.stage1((int)(s & 1), smem, g);
      if (s + 1 < end) pre(items[(s + 1) % kRing], (int)((s + 1) & 1));
      const long t = s - 1, u = s - kLag;
      const bool on = t >= first && t < end;
      const int opar = AFE_FB_DEFER == 1 ? (int)(t & 1) : 0;
      if (AFE_FB_DEFER == 2 && u >= first) {
        for (int g = 0; g < K::kGroup; ++g) k.mel(0, smem, g);
        for (int g = 0; g < K::kGroup; ++g) k.stats_store(items[u % kRing], 0, smem, g);
      }
      if (on) for (int g = 0; g < K::kGroup; ++g) k.stage2((int)(t & 1), smem, g);
      if (AFE_FB_DEFER == 1 && u >= first) for (int g = 0; g < K::kGroup; ++g) k.stats_store(items[u % kRing], (int)(u & 1), smem, g);
      if (AFE_FB_DEFER != 2 && on) for (int g = 0; g < K::kGroup; ++g) k.mel(opar, smem, g);
      if (!AFE_FB_DEFER && on) for (int g = 0; g < K::kGroup; ++g) k.stats_store(items[t % kRing], 0, smem, g);
      if (u >= first && k.tail(items[u % kRing])) for (int tid = 0; tid < K::kThreads; ++tid) k.tail_norm(items[u % kRing], smem, tid);
    }
  }
  return 0;
}
#endif

// mean / 1/sqrt(var+1e-7) per (utterance, mel bin): deterministic in-order reduction of the
// tile partials.  One item = one utterance, 128 threads (80 active).
struct FbankStatsKernel {
  static constexpr int kThreads = 128;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 1;
  BatchDesc d;
  const double* partials;
  float* stats;            // [B][80][2] = (mean, 1 / sqrt(var + 1e-7))
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return d.nsel; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = ldg(d.sel + it.item);
    if (tid >= kMel) return;
    const long t0 = ldg(d.tile_prefix + item), t1 = ldg(d.tile_prefix + item + 1);
    const int T_b = ldg(d.frames + item);
    double s = 0.0, ss = 0.0;
    for (long t = t0; t < t1; t += 8) {        // eight tiles' partials in flight, added in tile order
      double a[8], b[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const bool on = t + u < t1;
        a[u] = on ? partials[((t + u) * kMel + tid) * 2] : 0.0;
        b[u] = on ? partials[((t + u) * kMel + tid) * 2 + 1] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) if (t + u < t1) { s += a[u]; ss += b[u]; }
    }
    fbank_mean_den(s, ss, T_b, stats[(item * kMel + tid) * 2], stats[(item * kMel + tid) * 2 + 1]);
  }
};

// in-place normalisation, right padding and the stride-2 attention mask for the utterances the
// fbank kernel did not finish itself.  One item = 8192 consecutive floats of one utterance's
// (Tp x 80) plane, moved as 128-bit words (planes start 16-byte aligned: rows are 320 bytes).
struct FbankNormKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 4;
  static constexpr int kPhases = 1;
  static constexpr int kItemFloats = 8192;      // 8 words per thread in flight
  BatchDesc d;
  float* feat;             // [rows][80]
  int32_t* mask;           // [rows/2]
  const float* stats;      // [B][80][2] = (mean, 1 / denominator)
  float padding_value;
  long items_per_utt;      // ceil(max Tp * 80 / kItemFloats)
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return items_per_utt * d.nsel; }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const int b = ldg(d.sel + (int)(it.item / items_per_utt));
    const long chunk = it.item % items_per_utt;
    const int tp = ldg(d.tp + b);
    const int64_t plane4 = (int64_t)tp * (kMel / 4);
    const int64_t q0 = chunk * (kItemFloats / 4);
    if (q0 >= plane4) return;
    fbank_normalise_plane<kItemFloats / 4 / kThreads>(feat, mask, ldg(d.out_row + b), tp, ldg(d.frames + b), stats + (int64_t)b * kMel * 2,
                          padding_value, tid, kThreads, q0, tmin<int64_t>(plane4, q0 + kItemFloats / 4));
  }
};

}  // namespace afe
