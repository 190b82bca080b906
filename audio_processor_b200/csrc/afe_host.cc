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

// RIFF/WAVE header walk.  Layout: "RIFF" <u32 size> "WAVE" { <4-byte id> <u32 size> <size bytes> [pad to even] }*.
// "fmt ": u16 tag, u16 channels, u32 rate, u32 byte rate, u16 block align, u16 bits
//         [, u16 cbSize, u16 valid bits, u32 channel mask, 16-byte sub-format GUID whose first u16 is the tag].
namespace {
inline uint32_t rd_u16(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
inline uint32_t rd_u32(const unsigned char* p) { return rd_u16(p) | (rd_u16(p + 2) << 16); }
}  // namespace

int afe_wav_parse(const void* bytes, size_t nbytes, int64_t total_bytes, afe_wav_info* out) {
  if (!bytes || !out) return fail(AFE_ERR_INVALID, "afe_wav_parse: NULL argument");
  if (total_bytes < (int64_t)nbytes) total_bytes = (int64_t)nbytes;
  const unsigned char* b = static_cast<const unsigned char*>(bytes);
  if (nbytes < 12 || std::memcmp(b, "RIFF", 4) != 0 || std::memcmp(b + 8, "WAVE", 4) != 0)
    return fail(AFE_ERR_INVALID, "afe_wav_parse: not a RIFF/WAVE file");
  std::memset(out, 0, sizeof(*out));
  bool have_fmt = false;
  uint32_t tag = 0;
  size_t pos = 12;
  while (true) {
    if (pos + 8 > nbytes) return fail(AFE_ERR_INVALID, have_fmt ? "afe_wav_parse: no data chunk within the bytes given"
                                                                 : "afe_wav_parse: no fmt chunk within the bytes given");
    const unsigned char* c = b + pos;
    const uint32_t size = rd_u32(c + 4);
    if (std::memcmp(c, "fmt ", 4) == 0) {
      if (size < 16 || pos + 8 + 16 > nbytes) return fail(AFE_ERR_INVALID, "afe_wav_parse: truncated fmt chunk");
      const unsigned char* f = c + 8;
      tag = rd_u16(f);
      out->channels = (int32_t)rd_u16(f + 2);
      out->sample_rate = (int32_t)rd_u32(f + 4);
      out->block_align = (int32_t)rd_u16(f + 12);
      out->bits_per_sample = (int32_t)rd_u16(f + 14);
      if (tag == 0xFFFEu) {                                    // WAVE_FORMAT_EXTENSIBLE
        if (size < 40 || pos + 8 + 40 > nbytes) return fail(AFE_ERR_INVALID, "afe_wav_parse: truncated extensible fmt chunk");
        tag = rd_u16(f + 24);
      }
      have_fmt = true;
    } else if (std::memcmp(c, "data", 4) == 0) {
      if (!have_fmt) return fail(AFE_ERR_INVALID, "afe_wav_parse: data chunk before fmt chunk");
      const int bits = out->bits_per_sample;
      int fmt = 0;
      if (tag == 1u) fmt = bits == 8 ? AFE_WAV_U8 : bits == 16 ? AFE_WAV_S16 : bits == 24 ? AFE_WAV_S24 : bits == 32 ? AFE_WAV_S32 : 0;
      else if (tag == 3u) fmt = bits == 32 ? AFE_WAV_F32 : bits == 64 ? AFE_WAV_F64 : 0;
      if (fmt == 0) return fail(AFE_ERR_UNSUPPORTED, "afe_wav_parse: unsupported format tag / bit depth");
      if (out->channels <= 0 || out->sample_rate <= 0) return fail(AFE_ERR_INVALID, "afe_wav_parse: bad channel count or rate");
      const int64_t frame = (int64_t)out->channels * (bits / 8);
      if (out->block_align != frame) return fail(AFE_ERR_INVALID, "afe_wav_parse: block align does not match channels * bytes");
      out->format = fmt;
      out->data_offset = (int64_t)pos + 8;
      const int64_t rest = total_bytes - out->data_offset;
      int64_t n = (int64_t)size;
      if (size == 0u || size == 0xFFFFFFFFu || n > rest) n = rest;   // streamed header, or a truncated file
      out->num_frames = n / frame;
      out->data_bytes = out->num_frames * frame;
      return AFE_OK;
    }
    pos += 8 + (size_t)size + (size & 1u);
  }
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
