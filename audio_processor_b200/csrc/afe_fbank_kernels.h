// afe_fbank_kernels.h -- Kaldi/SeamlessM4T fbank as phased block programs (afe_platform.h).
//
// Replaces the per-frame numpy loop of transformers audio_utils.spectrogram (:788-804) as
// called through SeamlessM4TFeatureExtractor (feature_extraction_seamless_m4t.py:112-139)
// from services/audio_processor.py:253-261 and services/speech_segment.py:226-231.
//
// One block = 16 roles x TF frames; one work item = a tile of TF consecutive frames of one
// utterance.  TF = kFbTile.  TF = 32: one role per warp, 128 KB float64 exchange buffer, one block
// per SM.  TF = 16: two roles share a warp (lanes 0-15 / 16-31), 64 KB exchange, two blocks per
// SM -- measured on B200 (profiles/README.md, v12/v13): the second block hides ~11 % of the FFT
// phases' stalls but the half-warp roles double the role-specialised mel code, net zero, so
// TF = 32 stays.  Linear path (pre-emphasis, DC removal, window, FFT) in T; power / mel / log
// in float.
#pragma once
#include "afe_dct_gen.h"
#include "afe_mel_slots.h"
#include "afe_rfft512.h"
#include "afe_tables.h"

namespace afe {

constexpr int kFbTile = 32;                                         // frames per fbank tile
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

// float32 sum of n copies of v exactly as numpy's pairwise add.reduce forms it along contiguous memory
// (numpy/_core/src/umath/loops_utils.h.src: < 8 sequential; <= 128 eight accumulators, a fixed tree, the
// remainder sequentially; above that halves rounded down to a multiple of 8).  With equal addends the eight
// accumulators are equal and their tree is an exact multiplication by 8.  Iterative post-order walk of the
// halving tree (depth <= 40); only ever called for the constant bins of a silent utterance.
#if AFE_CUDA
#define AFE_NOINLINE __noinline__
#else
#define AFE_NOINLINE __attribute__((noinline))
#endif
AFE_NOINLINE
#if AFE_CUDA
__host__ __device__
#endif
inline float numpy_pairwise_sum_const(float v, int64_t n) {
  int64_t size[48];
  float left[48];
  signed char state[48];          // 0: descend left, 1: left done -> descend right, 2: both done
  int sp = 0;
  size[0] = n; state[0] = 0;
  float ret = 0.f;
  while (sp >= 0) {
    const int64_t m = size[sp];
    if (m <= 128) {
      float res;
      if (m < 8) {
        res = 0.f;
        for (int64_t i = 0; i < m; ++i) res += v;
      } else {
        float r = v;
        int64_t i = 8;
        for (; i < m - (m % 8); i += 8) r += v;
        res = 8.f * r;
        for (; i < m; ++i) res += v;
      }
      ret = res; --sp;
      continue;
    }
    int64_t n2 = m / 2; n2 -= n2 % 8;
    if (state[sp] == 0) { state[sp] = 1; size[sp + 1] = n2; state[sp + 1] = 0; ++sp; }
    else if (state[sp] == 1) { left[sp] = ret; state[sp] = 2; size[sp + 1] = m - n2; state[sp + 1] = 0; ++sp; }
    else { ret = left[sp] + ret; --sp; }
  }
  return ret;
}

// (mean, 1 / denominator) of one mel bin from its float64 sum and sum of squares over n frames --
// the reciprocal is formed once per (utterance, bin) so that the normalisation of the utterance's T x 80 values
// is a subtract and a multiply (within 1.5 ulp of the reference's divide; the contract is 1e-3 absolute) --
// numpy's mean / var(ddof=1) / sqrt(var + 1e-7) of feature_extraction_seamless_m4t.py:257-262.
// A bin that is constant over the utterance (digital silence: every frame at the log floor) has variance 0 and the
// reference's output is its OWN float32 rounding noise over sqrt(1e-7): mean32 = pairwise_sum(v, n) / n differs from
// v by a few ulp, and (v - mean32) / 3.16e-4 reaches 6e-3.  For such bins (float64 variance < 1e-10: the values
// agree to ~10 ulp) the float32 mean and variance are formed the way numpy forms them, bit for bit.
AFE_HD void fbank_mean_den(double s, double ss, int n, float& mean_out, float& den_out) {
  const double mean = s / (double)n;
  const double var = (ss - s * mean) / (double)(n - 1);      // ddof = 1; n == 1 -> 0/0 = NaN
  if (var < 1e-10 && n > 1) {
    const float v = (float)mean;
    const float mean32 = (float)((double)numpy_pairwise_sum_const(v, n) / (double)n);
    const float d = v - mean32;
    const float var32 = (float)((double)numpy_pairwise_sum_const(d * d, n) / (double)(n - 1));
    mean_out = mean32;
    den_out = 1.f / std::sqrt(var32 + 1e-7f);
    return;
  }
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

// shared-memory carve of one fbank block (TF frames per tile)
template <typename T, int TF> struct FbankSmem {
  static constexpr int kSamples = (TF - 1) * kHop + kFrame;                         // 2800 (TF 16) / 5360 (TF 32)
  static constexpr int kG16 = kSamples / 16;                                        // sums of 16 samples
  static constexpr size_t kRawBytes = sizeof(float) * kSamples;                     // per staging buffer
  static constexpr size_t kRawOff = 0;                                              // float[2][kSamples]
  static constexpr int kG80 = kG16 / 5;                                             // sums of 80 samples: 2 TF + 3
  static constexpr size_t kSumOff = kRawOff + 2 * kRawBytes;                        // double psum80[kG80] (+ spare)
  static constexpr size_t kAccOff = kSumOff + sizeof(double) * ((kG80 + 2) / 2 * 2);   // double acc[80][2]: running (sum, sum sq)
  static constexpr size_t kStatOff = kAccOff + sizeof(double) * 160;                // float stat[80][2]: (mean, 1 / denominator)
  static constexpr size_t kDOff = kStatOff + sizeof(float) * 160;                   // T[padded tile] | sP + sOut
  static constexpr size_t kDBytesT = sizeof(T) * (kSamples + kSamples / 32 + 8);
  static constexpr size_t kPBytes = sizeof(float) * (kBins * TF);
  static constexpr size_t kOutBytes = sizeof(float) * (TF * kOutStride);
  static constexpr size_t kDBytes = (kDBytesT > kPBytes + kOutBytes ? kDBytesT : kPBytes + kOutBytes);
  static constexpr size_t kExOff = (kDOff + kDBytes + 15) / 16 * 16;                // cplx<T>[16*16*TF]
  static constexpr size_t kExBytesT = sizeof(cplx<T>) * 16 * 16 * TF;
  static constexpr size_t kWinOff = kExOff + kExBytesT;                             // T window[16][26] (block_init)
  static constexpr size_t kTwOff = kWinOff + sizeof(T) * kRoles * kWinRow;          // T tw512[16][16][2]
  static constexpr size_t kMelWOff = kTwOff + sizeof(T) * kRoles * 32;              // float melw[16][40]  slot-form mel bank (block_init)
  static constexpr size_t kMelSOff = kMelWOff + sizeof(float) * kRoles * kMelSlotRow;   // int32 mels[16][8]
  static constexpr size_t kBytes = kMelSOff + sizeof(int32_t) * kRoles * kMelSlotStarts;
  static_assert(kMelWOff % 16 == 0, "mel weight rows are read as 128-bit words");
  static_assert(kSamples % 80 == 0 && kG80 == 2 * TF + 3, "tile samples must split into groups of 16 and 80");
};

template <typename T, int TF = kFbTile> struct FbankKernel {
  using S = FbankSmem<T, TF>;
  static constexpr int kTile = TF;
  static constexpr int kThreads = kRoles * TF;            // 256 (TF 16)
  static constexpr int kMinBlocks = (TF <= 16) ? 2 : 1;   // resident blocks per SM the launch is sized for
  static constexpr int kPhases = 6;
  static constexpr int kTailPhases = 1;         // normalisation of a waveform this block owns entirely
  static constexpr bool kAsyncStage = true;
  static constexpr size_t kStageOff = S::kRawOff;
  static constexpr size_t kStageBytes = S::kRawBytes;
  static constexpr int kSamples = S::kSamples;
  static constexpr int kG16 = S::kG16;
  static constexpr int kStoreFirst = 4 * kMel;  // threads [320, kThreads) move the tile's rows to global memory
  static constexpr int kG80 = S::kG80;
  static constexpr int kPrefetchThread = kThreads - 32;   // lane 0 of the last warp: it has no group in phase 1
  static_assert((kG16 + 29) / 30 * 32 <= kThreads - 32 && kStoreFirst < kThreads, "phase thread budgets");

  struct Item { long item; int buf; int bulk; int b; int f0; int nfr; int nvalid; int count; int whole; int tp;
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
  // the last tile of a waveform whose tiles all ran in this block: normalise it here
  AFE_HD bool tail(const Item& it) const { return it.whole != 0 && it.f0 + TF >= it.nfr; }
  // once per block, before its first item: the FFT tables move to shared memory (a table read is then a
  // 23-cycle broadcast load instead of an L1 round trip queued behind the exchange stores)
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
  }
  AFE_HD Item prepare(long item) const {
    Item it;
    int b, t;
    locate_tile(d, item, b, t);
    it.item = item; it.buf = 0; it.bulk = 0;
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
      return it;
    }
    int b = cur.b + 1;
    while (b + 1 < d.B && ldg(d.frames + b) <= 0) ++b;
    fill(it, b, 0);
    return it;
  }
  // the tile is one contiguous run of `count` samples: a single TMA bulk copy when 16-byte aligned
  AFE_HD bool stage_src(const Item& it, const float*& src, int& nfloats) const {
    src = it.src; nfloats = it.count;
    return (reinterpret_cast<uintptr_t>(it.src) & 15u) == 0;
  }

  // (sum, sum of squares) in float64 of mel bin m over quarter q4 of the tile's valid frames
  AFE_HD static void quarter_sums(const float* sOut, int m, int q4, int nvalid, double& s, double& ss) {
    constexpr int kQ = TF / 4;
    double s0 = 0.0, s1 = 0.0, q0 = 0.0, q1 = 0.0;
    const float* col = sOut + (q4 * kQ) * kOutStride + m;
    if ((q4 + 1) * kQ <= nvalid) {
#pragma unroll
      for (int i = 0; i < kQ; i += 2) {
        const double a = (double)col[i * kOutStride], b = (double)col[(i + 1) * kOutStride];
        s0 += a; q0 += a * a; s1 += b; q1 += b * b;
      }
    } else {
      for (int i = 0; i < kQ && q4 * kQ + i < nvalid; ++i) {
        const double a = (double)col[i * kOutStride];
        s0 += a; q0 += a * a;
      }
    }
    s = s0 + s1; ss = q0 + q1;
  }

  // Samples [16 t, 16 t + 16) of the tile: pre-emphasis x[n] - 0.97 x[n-1] in T into the padded layout
  // (kStore) and the float64 sum of the 16 raw samples.  The four 128-bit chunks are visited in an
  // order rotated by rot(t), which makes both the 128-bit loads (8 lanes per wavefront) and the 64-bit
  // stores (16 lanes per wavefront, element 16 t + j at 16 t + (t >> 1) + j) bank-conflict free, and
  // every sample is converted to T exactly once.
  template <bool kStore>
  AFE_HD static double group16(const float* raw, T* dbuf, int t, int count) {
    if (16 * t + 16 <= count) {
      const int rot = ((t >> 1) + 2 * (t & 1)) & 3;
      const float4* g4 = reinterpret_cast<const float4*>(raw + 16 * t);
      T x[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 v = g4[(q + rot) & 3];
        x[q][0] = (T)v.x; x[q][1] = (T)v.y; x[q][2] = (T)v.z; x[q][3] = (T)v.w;
      }
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) { s0 += (double)x[q][0]; s1 += (double)x[q][1]; s2 += (double)x[q][2]; s3 += (double)x[q][3]; }
      if (kStore) {
        const T xp = (t > 0) ? (T)raw[16 * t - 1] : T(0);      // the sample before the group (frame sample 0 has window 0)
        T* dp = dbuf + 16 * t + (t >> 1);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int cq = (q + rot) & 3;                         // chunk held by slot q; slot q-1 holds chunk cq-1
          const T xm = (cq == 0) ? xp : x[(q + 3) & 3][3];
          T* o = dp + 4 * cq;
          o[0] = x[q][0] - T(0.97) * xm;
          o[1] = x[q][1] - T(0.97) * x[q][0];
          o[2] = x[q][2] - T(0.97) * x[q][1];
          o[3] = x[q][3] - T(0.97) * x[q][2];
        }
      }
      return (s0 + s1) + (s2 + s3);
    }
    double s = 0.0;
    for (int j = 0; j < 16; ++j) {
      const int i = 16 * t + j;
      const float xv = (i < count) ? raw[i] : 0.f;
      if (kStore) {
        const float xmv = (i > 0 && i - 1 < count) ? raw[i - 1] : 0.f;
        dbuf[phys_idx(i)] = (T)xv - T(0.97) * (T)xmv;
      }
      s += (double)xv;
    }
    return s;
  }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    float* raw = reinterpret_cast<float*>(smem + S::kRawOff + (size_t)it.buf * S::kRawBytes);
    double* psum80 = reinterpret_cast<double*>(smem + S::kSumOff);   // [kG80] sums of 80 samples
    T* dbuf = reinterpret_cast<T*>(smem + S::kDOff);
    float* sP = reinterpret_cast<float*>(smem + S::kDOff);
    float* sOut = reinterpret_cast<float*>(smem + S::kDOff + S::kPBytes);
    cplx<T>* ex = reinterpret_cast<cplx<T>*>(smem + S::kExOff);
    const int role = tid / TF, f = tid % TF;

    switch (ph) {
      case 0: {   // cooperative staging (the CUDA driver replaces it by a TMA bulk copy when it can)
        for (int i = tid; i < it.count; i += kThreads) raw[i] = it.src[i];
      } break;
      case 1: {   // pre-emphasis in T into the padded layout + the sums of 80 samples the frame means need
        // lane l < 30 of warp w owns group t = 30 w + l of 16 samples; five neighbouring lanes make one sum of 80
        const int l = tid & 31, t = 30 * (tid >> 5) + l;
        const bool active = l < 30 && t < kG16;
        const double s16 = active ? group16<true>(raw, dbuf, t, it.count) : 0.0;
#if AFE_CUDA && defined(__CUDA_ARCH__)
        const double a = s16 + __shfl_down_sync(0xffffffffu, s16, 1);      // lanes 5 g, 5 g + 2: pairs
        const double b = a + __shfl_down_sync(0xffffffffu, a, 2);          // lane 5 g: (g0 + g1) + (g2 + g3)
        const double c5 = b + __shfl_down_sync(0xffffffffu, s16, 4);       //           ... + g4
        if (active && l % 5 == 0) psum80[t / 5] = c5;
#else
        if (active && l % 5 == 0) {   // host emulation runs the lanes one after the other: same tree, recomputed
          double g[5] = {s16, 0.0, 0.0, 0.0, 0.0};
          for (int m = 1; m < 5; ++m) g[m] = group16<false>(raw, dbuf, t + m, it.count);
          psum80[t / 5] = ((g[0] + g[1]) + (g[2] + g[3])) + g[4];
        }
#endif
      } break;
      case 2: {   // DC removal + window + stage 1 of the real FFT
        const double* p5 = psum80 + 2 * f;      // frame f = samples [160 f, 160 f + 400) = five sums of 80
        const double tot = ((p5[0] + p5[1]) + (p5[2] + p5[3])) + p5[4];
        const T c = (T)((1.0 - 0.97) * (tot * (1.0 / (double)kFrame)));      // (1 - 0.97) * mean(frame)
        T r[25];
        const T* col = dbuf + (kHop + kHop / 32) * f + role;      // phys_idx(160 f + 16 n1 + role) = 165 f + role + 16 n1 + (n1 >> 1)
        const T* wrow = reinterpret_cast<const T*>(smem + S::kWinOff) + role * kWinRow;
        const T* twrow = reinterpret_cast<const T*>(smem + S::kTwOff) + role * 32;
        windowed_column_fn<T>([wrow](int n1) { return wrow[n1]; }, r,
                              [col, c](int n1) { return col[16 * n1 + (n1 >> 1)] - c; });
        rfft512_stage1_tw<T, TF>(r, role, f, [twrow](int k1, T& cc, T& ss) { cc = twrow[2 * k1]; ss = twrow[2 * k1 + 1]; }, ex);
      } break;
      case 3:
        rfft512_stage2<T, false, TF>(role, f, ex, sP);
        break;
      case 4: {   // triangular mel bank (slot form: one code path for all roles) + floor + log
        float* row = sOut + f * kOutStride;
        const float lf = log_floor;
        mel_slots_kaldi<TF>(role, f, sP, reinterpret_cast<const float*>(smem + S::kMelWOff),
                            reinterpret_cast<const int32_t*>(smem + S::kMelSOff), [row, lf](int m, float acc) {
          row[m] = (acc <= kMelFloor) ? lf : fast_log(acc);
        });
      } break;
      case 5: {   // per-bin partial statistics (float64) + coalesced store of the raw log-mel
        if (tid < 4 * kMel) {
          // four lanes per mel bin, TF / 4 frames each, combined as (p0 + p1) + (p2 + p3)
          const int m = tid >> 2, q4 = tid & 3;
          double ts, tq;
          quarter_sums(sOut, m, q4, it.nvalid, ts, tq);
#if AFE_CUDA && defined(__CUDA_ARCH__)
          ts += __shfl_xor_sync(0xffffffffu, ts, 1); tq += __shfl_xor_sync(0xffffffffu, tq, 1);
          ts += __shfl_xor_sync(0xffffffffu, ts, 2); tq += __shfl_xor_sync(0xffffffffu, tq, 2);
#else
          if (q4 == 0) {     // host emulation: the same tree, the other three quarters recomputed here
            double s1, q1, s2, q2, s3, q3;
            quarter_sums(sOut, m, 1, it.nvalid, s1, q1);
            quarter_sums(sOut, m, 2, it.nvalid, s2, q2);
            quarter_sums(sOut, m, 3, it.nvalid, s3, q3);
            ts = (ts + s1) + (s2 + s3); tq = (tq + q1) + (q2 + q3);
          }
#endif
          if (q4 != 0) break;
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
        } else if (tid >= kStoreFirst) {
          // rows are 320 bytes: 20 128-bit words per frame, destination 16-byte aligned
          float4* dst = reinterpret_cast<float4*>(out + (it.out_row + it.f0) * kMel);
          for (int i = tid - kStoreFirst; i < it.nvalid * (kMel / 4); i += kThreads - kStoreFirst) {
            const int row = i / (kMel / 4), c4 = i - row * (kMel / 4);
            const float* sp = sOut + row * kOutStride + 4 * c4;
            dst[i] = float4{sp[0], sp[1], sp[2], sp[3]};
          }
        }
      } break;
      case 6: {   // tail: normalise the whole utterance in place (its raw rows are this block's own, L2-hot)
        const float* stat = reinterpret_cast<const float*>(smem + S::kStatOff);
        fbank_normalise_plane<8>(out, mask, it.out_row, it.tp, it.nfr, stat, padding_value, tid, kThreads, 0,
                              (int64_t)it.tp * (kMel / 4));
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
