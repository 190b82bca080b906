// afe_fbank_pipe.h -- the fbank tile pipeline as THREE free-running warp groups (experimental twin of FbankKernel;
// selected with AFE_FBANK_PIPE=1, see afe_fbank.cu).  Same arithmetic, same in-order reductions, bit-identical output.
//
// One block per SM, 768 threads = three groups of 8 warps; a tile is kPipeTile = 16 consecutive frames of one utterance
// and a group is 16 roles x 16 frames (two roles per warp).  Every group loops over the block's tiles:
//     AUX: pre-emphasis pass of tile j (TMA-staged raw samples -> dbuf, frame DC terms -> dc), then the statistics and
//          the store of the raw log-mel rows of tile j - 3 and, after the last tile of an utterance the block owns, the
//          utterance normalisation
//     S1:  FFT stage 1 of tile t   (dbuf, dc -> registers -> exchange buffer ex)
//     S2:  FFT stage 2 of tile t   (ex -> registers -> power sP), then mel + log (sP -> sOut[t & 1])
// No block-wide barrier in the loop.  dbuf and ex are SINGLE buffers guarded by full / empty named barriers
// (bar.arrive by the side that is done, bar.sync by the side that needs it, 512 threads each): the consumer pulls its
// inputs into registers, releases the buffer at once and computes from registers, so the producer refills the buffer
// under the consumer's arithmetic.  sP is private to S2; sOut is double-buffered towards AUX.  The raw staging buffer
// is double: the TMA bulk copy of tile j + 2 is issued as soon as the pre-emphasis pass of tile j has read its buffer.
#pragma once
#include "afe_fbank_kernels.h"

namespace afe {

constexpr int kPipeTile = 16;

template <typename T> struct FbankPipeSmem {
  static constexpr int TF = kPipeTile;
  static constexpr int kSamples = (TF - 1) * kHop + kFrame;                         // 2800
  static constexpr int kG16 = kSamples / 16;                                        // 175 sums of 16 samples
  static constexpr int kG80 = kG16 / 5;                                             // 35 sums of 80 samples = 2 TF + 3
  static constexpr int kDStride = (kSamples + kSamples / 32 + 8 + 1) / 2 * 2;        // padded tile, elements of T
  static constexpr size_t kRawBytes = sizeof(float) * kSamples;
  static constexpr size_t kRawOff = 0;                                              // float raw[2][kSamples] (TMA destinations)
  static constexpr size_t kSumOff = kRawOff + 2 * kRawBytes;                        // double psum80[kG80 + 1]
  static constexpr size_t kDcOff = kSumOff + sizeof(double) * (kG80 + 1);           // T dc[TF]
  static constexpr size_t kAccOff = kDcOff + sizeof(double) * TF;                   // double acc[80][2]: running (sum, sum sq)
  static constexpr size_t kStatOff = kAccOff + sizeof(double) * 160;                // float stat[80][2]: (mean, 1 / denominator)
  static constexpr size_t kDOff = kStatOff + sizeof(float) * 160;                   // T dbuf[kDStride]
  static constexpr size_t kPOff = kDOff + sizeof(T) * kDStride;                     // float sP[257][TF]
  static constexpr size_t kOutBytes = sizeof(float) * TF * kOutStride;
  static constexpr size_t kOutOff = kPOff + sizeof(float) * kBins * TF;             // float sOut[2][TF][81]
  static constexpr size_t kExOff = (kOutOff + 2 * kOutBytes + 15) / 16 * 16;        // cplx<T> ex[16][16][TF]
  static constexpr size_t kExBytes = sizeof(cplx<T>) * 16 * 16 * TF;
  static constexpr size_t kWinOff = kExOff + kExBytes;                              // T window[16][26]
  static constexpr size_t kTwOff = kWinOff + sizeof(T) * kRoles * kWinRow;          // T tw512[16][16][2]
  static constexpr size_t kMelWOff = kTwOff + sizeof(T) * kRoles * 32;              // float melw[16][40]
  static constexpr size_t kMelSOff = kMelWOff + sizeof(float) * kRoles * kMelSlotRow;   // int32 mels[16][8]
  static constexpr size_t kBytes = kMelSOff + sizeof(int32_t) * kRoles * kMelSlotStarts;
  static_assert(kMelWOff % 16 == 0 && kExOff % 16 == 0 && kDOff % 16 == 0 && kRawBytes % 16 == 0, "128-bit rows");
  static_assert(kSamples % 80 == 0 && kG80 == 2 * TF + 3, "tile samples must split into groups of 16 and 80");
  static_assert(kBytes <= 227 * 1024, "one block per SM");
};

template <typename T> struct FbankPipeKernel {
  using S = FbankPipeSmem<T>;
  using Real = T;
  static constexpr int TF = kPipeTile;
  static constexpr int kGroup = kRoles * TF;              // 256 threads per warp group
  static constexpr int kThreads = 3 * kGroup;             // AUX (warps 0-7), S2 (8-15), S1 (16-23: the issue arbiter prefers high warp ids)
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
  AFE_HD void pre_copy(const Item& it, int rb, unsigned char* smem, int g) const {
    float* raw = reinterpret_cast<float*>(smem + S::kRawOff + (size_t)rb * S::kRawBytes);
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
  AFE_HD void pre_a(const Item& it, int rb, unsigned char* smem, int g) const {
    const float* raw = reinterpret_cast<const float*>(smem + S::kRawOff + (size_t)rb * S::kRawBytes);
    double* psum80 = reinterpret_cast<double*>(smem + S::kSumOff);
    T* dbuf = reinterpret_cast<T*>(smem + S::kDOff);
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
  AFE_HD void pre_b(unsigned char* smem, int g) const {
    if (g >= TF) return;
    const double* p5 = reinterpret_cast<const double*>(smem + S::kSumOff) + 2 * g;
    const double tot = ((p5[0] + p5[1]) + (p5[2] + p5[3])) + p5[4];
    reinterpret_cast<T*>(smem + S::kDcOff)[g] = (T)((1.0 - 0.97) * (tot * (1.0 / (double)kFrame)));
  }

  // ---- S1, FFT stage 1: DC removal + window + the 16 real 32-point column transforms --------------------------
  // the 25 pre-emphasised samples of (role, frame) and the frame's DC term -> registers (dbuf is free afterwards)
  AFE_HD void stage1_load(unsigned char* smem, int g, T (&r)[25], T& c) const {
    const int role = g / TF, f = g % TF;
    const T* col = reinterpret_cast<const T*>(smem + S::kDOff) + (kHop + kHop / 32) * f + role;   // phys_idx(160 f + 16 n1 + role)
    c = reinterpret_cast<const T*>(smem + S::kDcOff)[f];
#pragma unroll
    for (int n1 = 0; n1 < 25; ++n1) r[n1] = col[16 * n1 + (n1 >> 1)];
  }
  AFE_HD void stage1_compute(unsigned char* smem, int g, T (&r)[25], T c) const {
    const int role = g / TF, f = g % TF;
    cplx<T>* ex = reinterpret_cast<cplx<T>*>(smem + S::kExOff);
    const T* wrow = reinterpret_cast<const T*>(smem + S::kWinOff) + role * kWinRow;
    const T* twrow = reinterpret_cast<const T*>(smem + S::kTwOff) + role * 32;
#pragma unroll
    for (int n1 = 0; n1 < 25; ++n1) r[n1] = wrow[n1] * (r[n1] - c);
    rfft512_stage1_tw<T, TF>(r, role, f, [twrow](int k1, T& cc, T& ss) { cc = twrow[2 * k1]; ss = twrow[2 * k1 + 1]; }, ex);
  }

  // ---- S2 ----------------------------------------------------------------------------------------------------
  // the 16 exchange values of (role, frame) -> registers (ex is free afterwards)
  AFE_HD void stage2_load(unsigned char* smem, int g, cplx<T> (&v)[16]) const {
    const int role = g / TF, f = g % TF;
    const cplx<T>* ex = reinterpret_cast<const cplx<T>*>(smem + S::kExOff);
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) v[n2] = ex[(role * 16 + n2) * TF + f];
  }
  AFE_HD void stage2_compute(unsigned char* smem, int g, const cplx<T> (&v)[16]) const {
    rfft512_stage2_regs<T, false, TF>(g / TF, g % TF, v, reinterpret_cast<float*>(smem + S::kPOff));
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
  // tail: normalise the whole utterance in place (its raw rows are this block's own, L2-hot); the AUX group
  AFE_HD void tail_norm(const Item& it, unsigned char* smem, int g) const {
    const float* stat = reinterpret_cast<const float*>(smem + S::kStatOff);
    fbank_normalise_plane<8>(out, mask, it.out_row, it.tp, it.nfr, stat, padding_value, g, kGroup, 0,
                             (int64_t)it.tp * (kMel / 4));
  }
};

#ifndef AFE_PIPE_LAG
#define AFE_PIPE_LAG 2         // AUX runs the statistics + store of tile j - AFE_PIPE_LAG after the pre-emphasis pass of tile j
#endif
#if AFE_CUDA
// named barriers (ids 1..15; id 0 is __syncthreads)
__device__ __forceinline__ void nbar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void nbar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
// (setmaxnreg per warp group was tried: ptxas 12.9 segfaults on this kernel with it, so all groups run at 80 registers)
enum : int { kBarAux = 1, kBarS2, kBarDFull, kBarDEmpty, kBarXFull, kBarXEmpty, kBarOFull /* +0, +1 */, kBarOFull1,
             kBarOEmpty /* +0, +1 */, kBarOEmpty1 };

// kTrace (debug, AFE_PHASE_TRACE=1): lane 0 of every warp adds up the cycles it waited at each hand-over
// (a volatile shared load after the barrier makes the clock read wait for the barrier's release)
template <class K, bool kTrace>
__global__ void __launch_bounds__(K::kThreads, 1) fbank_pipe_kernel(const K k, unsigned long long* trace) {
  extern __shared__ __align__(16) unsigned char afe_smem[];
  using Item = typename K::Item;
  using S = typename K::S;
  using R = typename K::Real;
  constexpr int kRing = 8;                 // item ring: tile j is created at j - 2 and last read at j + AFE_PIPE_LAG
  constexpr int kLag = AFE_PIPE_LAG;
  constexpr int kPair = 2 * K::kGroup;     // threads of a hand-over barrier (producer group + consumer group)
  constexpr int kG = K::kGroup;
  __shared__ __align__(8) uint64_t bar_raw[2];
  __shared__ Item items[kRing];
  const long n = k.num_items();
  const int tid = (int)threadIdx.x;
  const long per = (n + gridDim.x - 1) / gridDim.x;
  const long first = (long)blockIdx.x * per;
  const long end = (first + per < n) ? first + per : n;
  if (first >= end) return;
  const int grp = tid / kG;                // 0: AUX, 1: S2, 2: S1
  const int g = tid % kG;
  const bool lead = tid == 0;              // keeps the item ring and issues the TMA copies
  auto issue = [&](const Item& nx, int rb) {
    if (nx.bulk) {
      fence_proxy_async();
      mbar_expect_tx(&bar_raw[rb], (uint32_t)nx.count * 4u);
      bulk_g2s(afe_smem + S::kRawOff + (size_t)rb * S::kRawBytes, nx.src, (uint32_t)nx.count * 4u, &bar_raw[rb]);
    }
  };
  k.block_init(afe_smem, tid);
  if (lead) {
    mbar_init(&bar_raw[0], 1); mbar_init(&bar_raw[1], 1); fence_mbar_init();
    Item c0 = k.prepare(first);
    items[first % kRing] = c0;
    issue(c0, (int)(first & 1));
    if (first + 1 < end) {
      const Item c1 = k.next(c0, first + 1);
      items[(first + 1) % kRing] = c1;
      issue(c1, (int)((first + 1) & 1));
    }
  }
  __syncthreads();
  long long tw[6] = {0, 0, 0, 0, 0, 0};
  long long tc = 0;
  (void)tw; (void)tc;
  volatile unsigned char* probe = afe_smem + S::kStatOff;
  auto tick = [&]() { if constexpr (kTrace) tc = clock64(); };
  auto tock = [&](int i) { if constexpr (kTrace) { (void)*probe; tw[i] += clock64() - tc; } };
  if constexpr (kTrace) tw[5] = -clock64();
  if (grp == 2) {
    // ---------------- S1: FFT stage 1 ----------------
    for (long t = first; t < end; ++t) {
      R r[25], c;
      tick(); nbar_sync(kBarDFull, kPair); tock(0);
      k.stage1_load(afe_smem, g, r, c);
      if (t + 1 < end) nbar_arrive(kBarDEmpty, kPair);
      tick(); if (t > first) nbar_sync(kBarXEmpty, kPair); tock(1);
      k.stage1_compute(afe_smem, g, r, c);
      nbar_arrive(kBarXFull, kPair);
    }
  } else if (grp == 1) {
    // ---------------- S2: FFT stage 2, mel + log ----------------
    for (long t = first; t < end; ++t) {
      const int ob = (int)(t & 1);
      cplx<R> v[16];
      tick(); nbar_sync(kBarXFull, kPair); tock(0);
      k.stage2_load(afe_smem, g, v);
      if (t + 1 < end) nbar_arrive(kBarXEmpty, kPair);
      k.stage2_compute(afe_smem, g, v);
      nbar_sync(kBarS2, kG);                                 // sP complete
      tick(); if (t >= first + 2) nbar_sync(kBarOEmpty + ob, kPair); tock(1);
      k.mel(ob, afe_smem, g);
      nbar_arrive(kBarOFull + ob, kPair);
      nbar_sync(kBarS2, kG);                                 // sP read: the next tile's stage 2 may overwrite it
    }
  } else {
    // ---------------- AUX: pre-emphasis of tile j, statistics + store of tile j - kLag ----------------
    uint32_t rawpar = 0u;                  // bit b: phase parity of raw buffer b's mbarrier
    for (long j = first; j < end + kLag; ++j) {
      if (j < end) {
        const Item& it = items[j % kRing];
        const int rb = (int)(j & 1);
        tick();
        if (it.bulk) { mbar_wait(&bar_raw[rb], (rawpar >> rb) & 1u); rawpar ^= 1u << rb; }
        else { k.pre_copy(it, rb, afe_smem, g); nbar_sync(kBarAux, kG); }
        tock(0);
        tick(); if (j > first) nbar_sync(kBarDEmpty, kPair); tock(1);
        k.pre_a(it, rb, afe_smem, g);
        nbar_sync(kBarAux, kG);
        if (lead && j + 2 < end) {          // raw[rb] is free: stage tile j + 2
          const Item nx = k.next(items[(j + 1) % kRing], j + 2);
          items[(j + 2) % kRing] = nx;
          issue(nx, rb);
        }
        k.pre_b(afe_smem, g);
        nbar_arrive(kBarDFull, kPair);
      }
      const long u = j - kLag;
      if (u >= first) {
        const Item& iu = items[u % kRing];
        const int ob = (int)(u & 1);
        tick(); nbar_sync(kBarOFull + ob, kPair); tock(2);
        k.stats_store(iu, ob, afe_smem, g);
        if (u + 2 < end) nbar_arrive(kBarOEmpty + ob, kPair);
        if (k.tail(iu)) {
          tick();
          nbar_sync(kBarAux, kG);           // stat[] complete
          k.tail_norm(iu, afe_smem, g);
          tock(3);
        }
      }
      nbar_sync(kBarAux, kG);               // psum80, acc / stat and the item ring turn over
    }
  }
  if constexpr (kTrace) {
    tw[5] += clock64();
    if ((tid & 31) == 0) for (int i = 0; i < 6; ++i) trace[((size_t)blockIdx.x * 24 + (tid >> 5)) * 6 + i] = (unsigned long long)tw[i];
  }
}

template <class K>
int afe_launch_fbank_pipe(const K& k, int max_blocks, cudaStream_t stream) {
  const long n = k.num_items();
  if (n <= 0) return 0;
  const size_t smem_bytes = K::S::kBytes;
  const long grid = afe_async_grid(n, max_blocks);
  static const bool tracing = std::getenv("AFE_PHASE_TRACE") != nullptr;
  if (tracing) {      // debug only: synchronous; mean cycles per tile each warp group waited at its hand-overs
    const void* tfn = (const void*)fbank_pipe_kernel<K, true>;
    cudaFuncSetAttribute(tfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    unsigned long long* tr = nullptr;
    const size_t nslots = (size_t)grid * 24 * 6;
    cudaMallocManaged(&tr, sizeof(unsigned long long) * nslots);
    std::memset(tr, 0, sizeof(unsigned long long) * nslots);
    fbank_pipe_kernel<K, true><<<(unsigned)grid, K::kThreads, smem_bytes, stream>>>(k, tr);
    cudaStreamSynchronize(stream);
    double acc[3][6] = {{0}};
    for (long b = 0; b < grid; ++b)
      for (int w = 0; w < 24; ++w)
        for (int i = 0; i < 6; ++i) acc[w >> 3][i] += (double)tr[((size_t)b * 24 + w) * 6 + i] / 8.0;
    const double tiles = (double)n;
    std::fprintf(stderr, "[afe trace] fbank pipe: %ld blocks, %.0f tiles; mean cycles per tile (16 frames) and warp\n", grid, tiles);
    std::fprintf(stderr, "[afe trace]   S1 : total %.0f  wait dbuf-full %.0f  wait ex-empty %.0f\n", acc[2][5] / tiles, acc[2][0] / tiles, acc[2][1] / tiles);
    std::fprintf(stderr, "[afe trace]   S2 : total %.0f  wait ex-full %.0f  wait out-empty %.0f\n", acc[1][5] / tiles, acc[1][0] / tiles, acc[1][1] / tiles);
    std::fprintf(stderr, "[afe trace]   AUX: total %.0f  wait raw %.0f  wait dbuf-empty %.0f  wait out-full %.0f  utterance normalisation %.0f\n",
                 acc[0][5] / tiles, acc[0][0] / tiles, acc[0][1] / tiles, acc[0][2] / tiles, acc[0][3] / tiles);
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
// host emulation (tests only): the same member functions, tile after tile
template <class K>
int afe_launch_fbank_pipe(const K& k, int max_blocks, cudaStream_t) {
  using Item = typename K::Item;
  using R = typename K::Real;
  const long n = k.num_items();
  if (n <= 0) return 0;
  std::vector<double> backing((K::S::kBytes + 64) / sizeof(double) + 2);
  unsigned char* smem = reinterpret_cast<unsigned char*>(backing.data());
  const long grid = afe_async_grid(n, max_blocks);
  const long per = (n + grid - 1) / grid;
  std::vector<R> regs((size_t)K::kGroup * 26);
  std::vector<cplx<R>> xr((size_t)K::kGroup * 16);
  for (long block = 0; block < grid; ++block) {
    const long first = block * per, end = first + per < n ? first + per : n;
    if (first >= end) continue;
    for (int tid = 0; tid < K::kThreads; ++tid) k.block_init(smem, tid);
    Item it = k.prepare(first);
    for (long t = first; t < end; ++t) {
      if (t > first) it = k.next(it, t);
      const int rb = (int)(t & 1), ob = (int)(t & 1);
      for (int g = 0; g < K::kGroup; ++g) k.pre_copy(it, rb, smem, g);
      for (int g = 0; g < K::kGroup; ++g) k.pre_a(it, rb, smem, g);
      for (int g = 0; g < K::kGroup; ++g) k.pre_b(smem, g);
      for (int g = 0; g < K::kGroup; ++g) {
        R r[25], c;
        k.stage1_load(smem, g, r, c);
        for (int i = 0; i < 25; ++i) regs[(size_t)g * 26 + i] = r[i];
        regs[(size_t)g * 26 + 25] = c;
      }
      for (int g = 0; g < K::kGroup; ++g) {
        R r[25];
        for (int i = 0; i < 25; ++i) r[i] = regs[(size_t)g * 26 + i];
        k.stage1_compute(smem, g, r, regs[(size_t)g * 26 + 25]);
      }
      for (int g = 0; g < K::kGroup; ++g) {
        cplx<R> v[16];
        k.stage2_load(smem, g, v);
        for (int i = 0; i < 16; ++i) xr[(size_t)g * 16 + i] = v[i];
      }
      for (int g = 0; g < K::kGroup; ++g) {
        cplx<R> v[16];
        for (int i = 0; i < 16; ++i) v[i] = xr[(size_t)g * 16 + i];
        k.stage2_compute(smem, g, v);
      }
      for (int g = 0; g < K::kGroup; ++g) k.mel(ob, smem, g);
      for (int g = 0; g < K::kGroup; ++g) k.stats_store(it, ob, smem, g);
      if (k.tail(it)) for (int g = 0; g < K::kGroup; ++g) k.tail_norm(it, smem, g);
    }
  }
  return 0;
}
#endif

}  // namespace afe
