// afe_resample_kernels.h -- polyphase Hann-windowed-sinc resampling + channel split.
//
// Replaces torchaudio.transforms.Resample as used by AudioProcessor.get_resampler /
// split_audio (services/audio_processor.py:99-131): y[n*N + j] = sum_k h[j][k] xp[n*O + k]
// with xp = zero-pad(x, width, width + O).  Every output is the in-order float32 FMA chain
// acc = fma(h[j][k], xp[.], acc), k = 0..K-1 from acc = 0 -- measured to be what torch's
// conv1d does, which makes this stage bit-exact against the reference.
#pragma once
#include "afe_platform.h"

namespace afe {

// ------------------------------------------------------------------ 1:2 upsampler (8 k -> 16 k)
// O = 1, N = 2, K = 15 (width 7).  One item = 1024 input positions of one call (STEREO: both
// rows + mixed = (a + c) * 0.5 in the same pass; mono: one row).  256 threads; thread t owns
// the position pairs q = t and t + 256; a pair yields 4 consecutive outputs, so a warp's
// store is one contiguous 512-byte run of 128-bit words (coalesced), and the tile is staged
// with 128-bit loads.  A launch covers a whole batch of calls.
struct UpCall {               // one call of a batched launch (device array, or inline for one call)
  const float* in;            // [rows][L]
  float* out;                 // [rows][2L]
  float* mixed;               // [2L] or nullptr
  int64_t L;
  int64_t item0;              // first item of this call in the launch
};

template <bool STEREO> struct Upsample2xKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 4;
  static constexpr int kPhases = 2;
  static constexpr int kPos = 1024;                 // input positions per item
  static constexpr int kTile = kPos + 16;           // staged samples per row: x[n0-8 .. n0+1032)
  static constexpr int kRows = STEREO ? 2 : 1;
  static constexpr size_t kStageBytes = sizeof(float) * kTile * kRows;      // one staging buffer: the rows' tiles back to back
  // stereo launches: a block owns a contiguous run of items and the rows of item i+1 arrive by two TMA bulk copies
  // while item i is computed (float input; PCM input is converted cooperatively in phase 0)
  static constexpr bool kAsyncStage = STEREO;
  static constexpr int kStageSegments = 2;
  static constexpr int kTailPhases = 0;
  static constexpr size_t kStageOff = 0;
  static constexpr size_t kSmemBytes = kStageBytes * (STEREO ? 2 : 1);

  const UpCall* calls;    // device array [ncalls] (nullptr: `one`)
  UpCall one;
  int ncalls;
  int rows;               // rows per call (STEREO: 2)
  int pcm16;              // 1: inputs are int16 PCM, planar [rows][L]; 2: interleaved frames [L][rows] as in a wav
                          // file (torchaudio.load's x / 32768 and the de-interleave happen while staging)
  long nitems;
  float h[2][15];

  struct Item { long item; int buf; int bulk; int ci; const float* src; float* dst; float* mix; int64_t L; int64_t n0; };

  AFE_HD long num_items() const { return nitems; }

  AFE_HD Item prepare(long item) const {
    Item it;
    it.item = item; it.buf = 0; it.bulk = 0; it.ci = 0;
    UpCall c;
    if (calls == nullptr) {
      c = one;
    } else {
      int lo = 0, hi = ncalls;                       // calls[lo].item0 <= item < calls[hi].item0
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ldg(&calls[mid].item0) <= item) lo = mid; else hi = mid;
      }
      c.in = ldg(&calls[lo].in); c.out = const_cast<float*>(ldg(const_cast<const float* const*>(&calls[lo].out)));
      c.mixed = const_cast<float*>(ldg(const_cast<const float* const*>(&calls[lo].mixed)));
      c.L = ldg(&calls[lo].L); c.item0 = ldg(&calls[lo].item0);
      it.ci = lo;
    }
    const long local = item - (long)c.item0;
    const long ipr = (long)((c.L + kPos - 1) / kPos);
    const int row0 = STEREO ? 0 : (int)(local / ipr);
    it.n0 = (int64_t)(local % ipr) * kPos;
    it.L = c.L;
    it.src = pcm16 == 2 ? reinterpret_cast<const float*>(reinterpret_cast<const int16_t*>(c.in) + row0)
           : pcm16     ? reinterpret_cast<const float*>(reinterpret_cast<const int16_t*>(c.in) + (int64_t)row0 * c.L)
                       : c.in + (int64_t)row0 * c.L;
    it.dst = c.out + (int64_t)row0 * 2 * c.L;
    it.mix = c.mixed;
    return it;
  }

  // item cur.item + 1 of a stereo launch without a search: the next 1024 positions of the same call, else the
  // first item of the next call that has samples
  AFE_HD Item next(const Item& cur, long item) const {
    Item it = cur;
    it.item = item;
    if (cur.n0 + kPos < cur.L) { it.n0 = cur.n0 + kPos; return it; }
    int ci = cur.ci + 1;
    while (ci + 1 < ncalls && ldg(&calls[ci].L) <= 0) ++ci;
    it.ci = ci; it.n0 = 0;
    it.src = ldg(&calls[ci].in);
    it.dst = const_cast<float*>(ldg(const_cast<const float* const*>(&calls[ci].out)));
    it.mix = const_cast<float*>(ldg(const_cast<const float* const*>(&calls[ci].mixed)));
    it.L = ldg(&calls[ci].L);
    return it;
  }
  // interior tiles of float input: rows x[n0-8 .. n0+1032) are two aligned contiguous copies
  AFE_HD bool stage_src2(const Item& it, const float* (&src)[2], int (&nf)[2], int (&off)[2]) const {
    const float* p0 = it.src + (it.n0 - 8);
    src[0] = p0; src[1] = p0 + it.L;
    nf[0] = nf[1] = kTile; off[0] = 0; off[1] = kTile;
    return STEREO && pcm16 == 0 && it.n0 >= 8 && it.n0 - 8 + kTile <= it.L &&
           ((reinterpret_cast<uintptr_t>(p0) | reinterpret_cast<uintptr_t>(p0 + it.L)) & 15u) == 0;
  }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    float* xs = reinterpret_cast<float*>(smem + (size_t)it.buf * kStageBytes);
    const int64_t n0 = it.n0, L = it.L;
    if (ph == 0 && pcm16 == 2) {
      // frames as they lie in a wav file: sample (g, r) at src[g * rows + r]; it.src already points at row0
      constexpr float kScale = 1.0f / 32768.0f;
      const int16_t* src = reinterpret_cast<const int16_t*>(it.src);
      if (STEREO) {
        const int16_t* p0 = src + 2 * (n0 - 8);
        if (n0 >= 8 && n0 - 8 + kTile <= L && (reinterpret_cast<uintptr_t>(p0) & 15u) == 0) {
          // one 128-bit word = 4 frames; low half of each 32-bit lane is channel 0
          const int4* s4 = reinterpret_cast<const int4*>(p0);
          float4* a4 = reinterpret_cast<float4*>(xs);
          float4* c4 = reinterpret_cast<float4*>(xs + kTile);
          for (int i = tid; i < kTile / 4; i += kThreads) {
            const int4 v = s4[i];
            a4[i] = float4{(float)(int16_t)v.x * kScale, (float)(int16_t)v.y * kScale,
                           (float)(int16_t)v.z * kScale, (float)(int16_t)v.w * kScale};
            c4[i] = float4{(float)(v.x >> 16) * kScale, (float)(v.y >> 16) * kScale,
                           (float)(v.z >> 16) * kScale, (float)(v.w >> 16) * kScale};
          }
        } else {
          for (int i = tid; i < kTile; i += kThreads) {
            const int64_t g = n0 - 8 + i;
            const bool in = (g >= 0 && g < L);
            xs[i] = in ? (float)src[2 * g] * kScale : 0.f;
            xs[kTile + i] = in ? (float)src[2 * g + 1] * kScale : 0.f;
          }
        }
      } else {
        for (int i = tid; i < kTile; i += kThreads) {
          const int64_t g = n0 - 8 + i;
          xs[i] = (g >= 0 && g < L) ? (float)src[g * rows] * kScale : 0.f;
        }
      }
      return;
    }
    if (ph == 0 && pcm16) {
      constexpr float kScale = 1.0f / 32768.0f;            // exact: the same value as x / 32768
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        const int16_t* src = reinterpret_cast<const int16_t*>(it.src) + (int64_t)r * L;
        const int16_t* p0 = src + (n0 - 8);
        if (n0 >= 8 && n0 - 8 + kTile <= L && (reinterpret_cast<uintptr_t>(p0) & 7u) == 0) {
          const short4* s4 = reinterpret_cast<const short4*>(p0);
          float4* d4 = reinterpret_cast<float4*>(xs + r * kTile);
          for (int i = tid; i < kTile / 4; i += kThreads) {
            const short4 v = s4[i];
            d4[i] = float4{(float)v.x * kScale, (float)v.y * kScale, (float)v.z * kScale, (float)v.w * kScale};
          }
        } else {
          for (int i = tid; i < kTile; i += kThreads) {
            const int64_t g = n0 - 8 + i;
            xs[r * kTile + i] = (g >= 0 && g < L) ? (float)src[g] * kScale : 0.f;
          }
        }
      }
      return;
    }
    if (ph == 0) {
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        const float* src = it.src + (int64_t)r * L;
        const float* p0 = src + (n0 - 8);
        if (n0 >= 8 && n0 - 8 + kTile <= L && (reinterpret_cast<uintptr_t>(p0) & 15u) == 0) {
          const float4* s4 = reinterpret_cast<const float4*>(p0);
          float4* d4 = reinterpret_cast<float4*>(xs + r * kTile);
          for (int i = tid; i < kTile / 4; i += kThreads) d4[i] = s4[i];
        } else {
          for (int i = tid; i < kTile; i += kThreads) {
            const int64_t g = n0 - 8 + i;
            xs[r * kTile + i] = (g >= 0 && g < L) ? src[g] : 0.f;
          }
        }
      }
      return;
    }
    // ph == 1: pair q = tid + 256 j covers positions n0 + 2q, n0 + 2q + 1 -> outputs 4q .. 4q+3 of
    // the item.  xp[n + k] = x[n + k - 7] = xs[(n - n0) + k + 1]; the window xs[2q .. 2q+17]
    // is read as nine 64-bit words.
    const bool vec_out = ((reinterpret_cast<uintptr_t>(it.dst) & 15u) == 0) && ((L & 1) == 0 || !STEREO) &&
                         (!STEREO || (reinterpret_cast<uintptr_t>(it.mix) & 15u) == 0);
#pragma unroll
    for (int j = 0; j < kPos / 2 / kThreads; ++j) {
      const int q = tid + kThreads * j;
      const int64_t n = n0 + 2 * q;
      if (n >= L) break;
      float y[kRows][4];
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        float w[18];
        const float2* w2 = reinterpret_cast<const float2*>(xs + r * kTile + 2 * q);
#pragma unroll
        for (int i = 0; i < 9; ++i) { const float2 v = w2[i]; w[2 * i] = v.x; w[2 * i + 1] = v.y; }
#pragma unroll
        for (int p = 0; p < 2; ++p)
#pragma unroll
          for (int ph2 = 0; ph2 < 2; ++ph2) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < 15; ++k) acc = fma_f(h[ph2][k], w[p + k + 1], acc);
            y[r][2 * p + ph2] = acc;
          }
      }
      const bool full = (n + 1 < L);
#pragma unroll
      for (int r = 0; r < kRows; ++r) {
        float* dst = it.dst + (int64_t)r * 2 * L + 2 * n;
        if (full && vec_out) {
          *reinterpret_cast<float4*>(dst) = float4{y[r][0], y[r][1], y[r][2], y[r][3]};
        } else {
          dst[0] = y[r][0]; dst[1] = y[r][1];
          if (full) { dst[2] = y[r][2]; dst[3] = y[r][3]; }
        }
      }
      if (STEREO) {
        float* dst = it.mix + 2 * n;
        float m[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) m[i] = (y[0][i] + y[kRows - 1][i]) * 0.5f;
        if (full && vec_out) {
          *reinterpret_cast<float4*>(dst) = float4{m[0], m[1], m[2], m[3]};
        } else {
          dst[0] = m[0]; dst[1] = m[1];
          if (full) { dst[2] = m[2]; dst[3] = m[3]; }
        }
      }
    }
  }
};

// ------------------------------------------------------------------ generic L/M polyphase
// One item = (row, group of 32 phases, 16 periods).  Thread (nl = tid/32, jj = tid%32)
// computes periods n0 + nl and n0 + nl + 8 for phase j0 + jj.
struct PolyphaseKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 1;
  static constexpr int kPhases = 2;
  static constexpr int kPeriods = 16;

  const float* in;        // [rows][L]
  float* out;             // [rows][out_len]
  const float* tapsT;     // [K][N]
  int64_t L, out_len;
  int rows, O, N, K, width;
  int jgroups;            // ceil(N / 32)
  long ntiles;            // ceil(periods_total / 16)

  AFE_HD int xs_len() const { return (kPeriods - 1) * O + K; }
  AFE_HD size_t smem_bytes() const { return sizeof(float) * ((size_t)xs_len() + 32 * (size_t)K); }
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return (long)rows * jgroups * ntiles; }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    const long item = it.item;
    float* xs = reinterpret_cast<float*>(smem);
    float* ts = xs + xs_len();
    const long tile = item % ntiles;
    const int jg = (int)((item / ntiles) % jgroups);
    const int row = (int)(item / (ntiles * jgroups));
    const int64_t n0 = tile * kPeriods;
    const int j0 = jg * 32;
    if (ph == 0) {
      const float* src = in + (int64_t)row * L;
      const int64_t g0 = n0 * O - width;
      const int len = xs_len();
      for (int i = tid; i < len; i += kThreads) {
        const int64_t g = g0 + i;
        xs[i] = (g >= 0 && g < L) ? src[g] : 0.f;
      }
      for (int i = tid; i < 32 * K; i += kThreads) {
        const int k = i >> 5, jj = i & 31;
        ts[i] = (j0 + jj < N) ? ldg(tapsT + (int64_t)k * N + j0 + jj) : 0.f;
      }
      return;
    }
    const int nl = tid >> 5, jj = tid & 31;
    const float* xa = xs + nl * O;
    const float* xb = xs + (nl + 8) * O;
    float a = 0.f, b = 0.f;
    for (int k = 0; k < K; ++k) {
      const float w = ts[k * 32 + jj];
      a = fma_f(w, xa[k], a);
      b = fma_f(w, xb[k], b);
    }
    if (j0 + jj >= N) return;
    float* dst = out + (int64_t)row * out_len;
    const int64_t ia = (n0 + nl) * N + j0 + jj, ib = (n0 + nl + 8) * N + j0 + jj;
    if (ia < out_len) dst[ia] = a;
    if (ib < out_len) dst[ib] = b;
  }
};

}  // namespace afe
