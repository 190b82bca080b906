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
// O = 1, N = 2, K = 15 (width 7).  One item = 1024 input positions of one row group;
// 256 threads x 4 positions.  STEREO: both rows + mixed = (a + c) * 0.5 in one pass.
template <bool STEREO> struct Upsample2xKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 4;
  static constexpr int kPhases = 2;
  static constexpr int kPos = 1024;                 // input positions per item
  static constexpr int kTile = kPos + 16;           // staged samples per row: [n0-8, n0+1032)
  static constexpr int kRows = STEREO ? 2 : 1;
  static constexpr size_t kSmemBytes = sizeof(float) * kTile * kRows;

  const float* in;        // [rows][L]
  float* out;             // [rows][2L]
  float* mixed;           // [2L] (STEREO only)
  int64_t L;
  int rows;               // total rows (STEREO: 2)
  long items_per_row;     // ceil(L / 1024)
  float h[2][15];

  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return items_per_row * (STEREO ? 1 : rows); }

  AFE_HD void phase(int ph, const Item& it, unsigned char* smem, int tid) const {
    const long item = it.item;
    float* xs = reinterpret_cast<float*>(smem);
    const int row0 = STEREO ? 0 : (int)(item / items_per_row);
    const int64_t n0 = (item % items_per_row) * kPos;
    if (ph == 0) {
      for (int r = 0; r < kRows; ++r) {
        const float* src = in + (int64_t)(row0 + r) * L;
        for (int i = tid; i < kTile; i += kThreads) {
          const int64_t g = n0 - 8 + i;
          xs[r * kTile + i] = (g >= 0 && g < L) ? src[g] : 0.f;
        }
      }
      return;
    }
    // ph == 1: positions n0 + 4*tid .. +3; xp[n + k] = x[n + k - 7] = xs[(n - n0) + k + 1]
    float y[kRows][8];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      float x[20];
#pragma unroll
      for (int i = 0; i < 20; ++i) x[i] = xs[r * kTile + 4 * tid + i];
#pragma unroll
      for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          float acc = 0.f;
#pragma unroll
          for (int k = 0; k < 15; ++k) acc = fma_f(h[j][k], x[p + k + 1], acc);
          y[r][2 * p + j] = acc;
        }
    }
    const int64_t n = n0 + 4 * tid;
    if (n >= L) return;
    const int nout = (int)tmin<int64_t>(8, 2 * (L - n));
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
      float* dst = out + (int64_t)(row0 + r) * 2 * L + 2 * n;
#pragma unroll
      for (int i = 0; i < 8; ++i) if (i < nout) dst[i] = y[r][i];
    }
    if (STEREO) {
      float* dst = mixed + 2 * n;
#pragma unroll
      for (int i = 0; i < 8; ++i) if (i < nout) dst[i] = (y[0][i] + y[kRows - 1][i]) * 0.5f;
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
