// afe_host.cc -- host-only integer / float64 contracts of the C ABI (no device work).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "afe_ctx.h"

using namespace afe;

extern "C" {

int64_t afe_resampled_length(int64_t length, int orig_hz, int new_hz) {
  if (orig_hz <= 0 || new_hz <= 0 || length < 0) return -1;
  if (orig_hz == new_hz) return length;
  return resampled_length(length, orig_hz, new_hz);
}

int64_t afe_fbank_num_frames(int64_t length) {
  return length < kFrame ? 0 : 1 + (length - kFrame) / kHop;
}

int64_t afe_fbank_padded_frames(const int64_t* h_lengths, int batch) {
  int64_t tmax = 0;
  for (int b = 0; b < batch; ++b) tmax = std::max(tmax, afe_fbank_num_frames(h_lengths[b]));
  return tmax + (tmax & 1);
}

// services/audio_processor.py:115-158: start = 0; loop { end = min(start + chunk, total);
// emit; start += step; if end >= total break } while start < total.
int afe_chunk_table(int64_t total, int64_t chunk, int64_t step, int64_t* h_starts, int64_t* h_ends,
                    int cap, int* n) {
  if (!n) return fail(AFE_ERR_INVALID, "afe_chunk_table: n is NULL");
  if (chunk <= 0 || step <= 0) return fail(AFE_ERR_INVALID, "afe_chunk_table: chunk and step must be positive");
  int count = 0;
  int64_t start = 0;
  while (start < total) {
    const int64_t end = std::min(start + chunk, total);
    if (count < cap && h_starts && h_ends) { h_starts[count] = start; h_ends[count] = end; }
    ++count;
    start += step;
    if (end >= total) break;
  }
  *n = count;
  return AFE_OK;
}

// services/speech_segment.py:385-418, float64 throughout, int() truncates toward zero.
int afe_segment_table(int64_t num_samples, int sample_rate, const double* iv, int n_iv, double padding,
                      double max_len, double overlap, int64_t* s0, int64_t* s1, double* t0, double* t1,
                      int cap, int* n) {
  if (!n || sample_rate <= 0) return fail(AFE_ERR_INVALID, "afe_segment_table: bad arguments");
  const double total = (double)num_samples / (double)sample_rate;
  int count = 0;
  for (int i = 0; i < n_iv; ++i) {
    const double ps = std::max(0.0, iv[2 * i] - padding);
    const double pe = std::min(total, iv[2 * i + 1] + padding);
    double cur = ps;
    while (cur < pe) {
      const double we = std::min(pe, cur + max_len);
      if (count < cap) {
        if (s0) s0[count] = (int64_t)(cur * sample_rate);
        if (s1) s1[count] = (int64_t)(we * sample_rate);
        if (t0) t0[count] = cur;
        if (t1) t1[count] = we;
      }
      ++count;
      if (we >= pe) break;
      cur = std::max(cur + max_len - overlap, cur + 0.1);
    }
  }
  *n = count;
  return AFE_OK;
}

int afe_resample_taps(int orig_hz, int new_hz, float* h_taps, size_t cap, int* width, int* in_stride,
                      int* phases, int* taps_per_phase) {
  if (orig_hz <= 0 || new_hz <= 0) return fail(AFE_ERR_INVALID, "afe_resample_taps: rates must be positive");
  ResampleTaps r = make_resample_taps(orig_hz, new_hz);
  if (width) *width = r.width;
  if (in_stride) *in_stride = r.O;
  if (phases) *phases = r.N;
  if (taps_per_phase) *taps_per_phase = r.K;
  if (h_taps) {
    if (cap < r.taps.size()) return fail(AFE_ERR_INVALID, "afe_resample_taps: buffer too small");
    std::memcpy(h_taps, r.taps.data(), sizeof(float) * r.taps.size());
  }
  return AFE_OK;
}

}  // extern "C"
