// afe_resample.cu -- C ABI entry points for resampling and the stereo mix.
#include <cstring>

#include "afe_ctx.h"
#include "afe_resample_kernels.h"

using namespace afe;

namespace afe {
// mixed = (row0 + row1) * 0.5 when no resampling is needed (sample_rate == target)
struct MixKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 8;
  static constexpr int kPhases = 1;
  static constexpr int kItem = 4096;
  const float* in; float* mixed; int64_t L;
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return (long)((L + kItem - 1) / kItem); }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const long item = it.item;
    for (int v = tid; v < kItem; v += kThreads) {
      const int64_t i = item * kItem + v;
      if (i < L) mixed[i] = (in[i] + in[L + i]) * 0.5f;
    }
  }
};
// float32 = int16 / 32768 (what torchaudio.load(normalize=True) returns for 16-bit PCM)
struct Pcm16Kernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 8;
  static constexpr int kPhases = 1;
  static constexpr int kItem = 4096;
  const int16_t* in; float* out; int64_t n;
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return (long)((n + kItem - 1) / kItem); }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    for (int v = tid; v < kItem; v += kThreads) {
      const int64_t i = it.item * kItem + v;
      if (i < n) out[i] = (float)in[i] * (1.0f / 32768.0f);
    }
  }
};
// torchaudio.load(normalize=True) of an interleaved data chunk -> planar float32 [channels][frames].
// Byte-wise loads: a data chunk starts wherever the header ends, so nothing is assumed aligned.
struct WavDecodeKernel {
  static constexpr int kThreads = 256;
  static constexpr int kMinBlocks = 8;
  static constexpr int kPhases = 1;
  static constexpr int kItem = 2048;                 // frames per item
  const unsigned char* in; float* out; int64_t frames; int channels; int format;
  AFE_PLAIN_ITEM
  AFE_HD long num_items() const { return (long)((frames + kItem - 1) / kItem); }
  AFE_HD static float sample(const unsigned char* p, int format) {
    switch (format) {
      case AFE_WAV_U8:  return (float)((int)p[0] - 128) * (1.0f / 128.0f);
      case AFE_WAV_S16: return (float)(int16_t)((unsigned)p[0] | ((unsigned)p[1] << 8)) * (1.0f / 32768.0f);
      case AFE_WAV_S24: {   // sign-extend through the top of an int32, as a decoder that widens to s32 does
        const int32_t v = (int32_t)(((uint32_t)p[0] << 8) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 24));
        return (float)v * (1.0f / 2147483648.0f);
      }
      case AFE_WAV_S32: {
        const int32_t v = (int32_t)((uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24));
        return (float)v * (1.0f / 2147483648.0f);
      }
      case AFE_WAV_F32: {
        const uint32_t u = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
        float f; memcpy(&f, &u, 4); return f;
      }
      default: {
        uint64_t u = 0;
        for (int i = 7; i >= 0; --i) u = (u << 8) | p[i];
        double d; memcpy(&d, &u, 8); return (float)d;
      }
    }
  }
  AFE_HD void phase(int, const Item& it, unsigned char*, int tid) const {
    const int bps = format == AFE_WAV_U8 ? 1 : format == AFE_WAV_S16 ? 2 : format == AFE_WAV_S24 ? 3 : format == AFE_WAV_F64 ? 8 : 4;
    const int64_t f0 = it.item * kItem;
    const int nf = (int)tmin<int64_t>(kItem, frames - f0);
    // consecutive threads take consecutive samples of the interleaved run (coalesced reads);
    // the planar writes are `channels` coalesced runs
    const int total = nf * channels;
    const unsigned char* base = in + f0 * channels * bps;
    for (int i = tid; i < total; i += kThreads) {
      const int f = i / channels, c = i - f * channels;
      out[(int64_t)c * frames + f0 + f] = sample(base + (int64_t)i * bps, format);
    }
  }
};
}  // namespace afe

namespace {
// one launch of the 1:2 kernel over `ncalls` calls (device descriptors) or the inline call
int launch_up2x(afe_ctx* ctx, cudaStream_t s, const ResampleTaps& t, const UpCall* d_calls, const UpCall& one,
                int ncalls, int rows, bool stereo_mix, long nitems = -1, int pcm16 = 0) {
  if (nitems < 0) nitems = (long)((one.L + 1023) / 1024) * (stereo_mix ? 1 : rows);
  if (stereo_mix) {
    Upsample2xKernel<true> k;
    k.calls = d_calls; k.one = one; k.ncalls = ncalls; k.rows = 2; k.nitems = nitems; k.pcm16 = pcm16;
    for (int j = 0; j < 2; ++j) for (int i = 0; i < 15; ++i) k.h[j][i] = t.taps[j * 15 + i];
    return afe_launch(k, Upsample2xKernel<true>::kSmemBytes, ctx->num_sms * Upsample2xKernel<true>::kMinBlocks, s);
  }
  Upsample2xKernel<false> k;
  k.calls = d_calls; k.one = one; k.ncalls = ncalls; k.rows = rows; k.nitems = nitems; k.pcm16 = pcm16;
  for (int j = 0; j < 2; ++j) for (int i = 0; i < 15; ++i) k.h[j][i] = t.taps[j * 15 + i];
  return afe_launch(k, Upsample2xKernel<false>::kSmemBytes, ctx->num_sms * Upsample2xKernel<false>::kMinBlocks, s);
}
}  // namespace

extern "C" {

int afe_mix_stereo(afe_ctx* ctx, void* stream, const float* d_in, int64_t length, float* d_mixed) {
  AFE_NVTX("afe_mix_stereo");
  if (!ctx || !d_in || !d_mixed || length < 0) return fail(AFE_ERR_INVALID, "afe_mix_stereo: bad arguments");
  MixKernel k{d_in, d_mixed, length};
  ProfScope ps(ctx, AFE_K_MIX, (cudaStream_t)stream);
  if (afe_launch(k, 0, ctx->num_sms * 8, (cudaStream_t)stream) != 0) return fail(AFE_ERR_CUDA, "afe_mix_stereo: launch failed");
  return AFE_OK;
}

int afe_resample(afe_ctx* ctx, void* stream, const float* d_in, int rows, int64_t length, int orig_hz,
                 int new_hz, float* d_out, float* d_mixed, int64_t out_length) {
  AFE_NVTX("afe_resample");
  if (!ctx || !d_in || !d_out || rows <= 0 || length < 0 || orig_hz <= 0 || new_hz <= 0)
    return fail(AFE_ERR_INVALID, "afe_resample: bad arguments");
  if (orig_hz == new_hz) return fail(AFE_ERR_INVALID, "afe_resample: equal rates need no resampling");
  if (out_length != resampled_length(length, orig_hz, new_hz))
    return fail(AFE_ERR_INVALID, "afe_resample: out_length must equal ceil(new*L/orig)");
  if (d_mixed && rows != 2) return fail(AFE_ERR_INVALID, "afe_resample: mixed output needs exactly 2 rows");
  if (length == 0) return AFE_OK;
  const DeviceTaps* dt = ctx->get_taps(orig_hz, new_hz);
  if (!dt) return AFE_ERR_CUDA;
  const ResampleTaps& t = dt->host;
  cudaStream_t s = (cudaStream_t)stream;
  int rc;
  {
  ProfScope ps(ctx, AFE_K_RESAMPLE, s);
  if (t.O == 1 && t.N == 2 && t.K == 15) {
    UpCall one{d_in, d_out, d_mixed, length, 0};
    rc = launch_up2x(ctx, s, t, nullptr, one, 1, rows, d_mixed != nullptr);
  } else {
    PolyphaseKernel k;
    k.in = d_in; k.out = d_out; k.tapsT = dt->d_tapsT; k.L = length; k.out_len = out_length;
    k.rows = rows; k.O = t.O; k.N = t.N; k.K = t.K; k.width = t.width;
    k.jgroups = (t.N + 31) / 32;
    const int64_t periods = (out_length + t.N - 1) / t.N;
    k.ntiles = (long)((periods + PolyphaseKernel::kPeriods - 1) / PolyphaseKernel::kPeriods);
    const size_t smem = k.smem_bytes();
    if (smem > 200 * 1024) return fail(AFE_ERR_UNSUPPORTED, "afe_resample: tap tile exceeds shared memory for this rate pair");
    rc = afe_launch(k, smem, ctx->num_sms * 2, s);
  }
  }
  if (!(t.O == 1 && t.N == 2 && t.K == 15)) {
    if (rc == 0 && d_mixed) {
      // generic path: mix the two resampled rows afterwards
      MixKernel m{d_out, d_mixed, out_length};
      ProfScope pm(ctx, AFE_K_MIX, s);
      rc = afe_launch(m, 0, ctx->num_sms * 8, s);
    }
  }
  if (rc != 0) return fail(AFE_ERR_CUDA, std::string("afe_resample: launch failed: ") + dev_last_error_string());
  return AFE_OK;
}

size_t afe_resample_batch_workspace_bytes(int ncalls) {
  return (sizeof(UpCall) * (size_t)(ncalls > 0 ? ncalls : 1) + 255) / 256 * 256;
}

int afe_pcm16_to_float(afe_ctx* ctx, void* stream, const int16_t* d_in, int64_t count, float* d_out) {
  AFE_NVTX("afe_pcm16_to_float");
  if (!ctx || count < 0 || (count > 0 && (!d_in || !d_out))) return fail(AFE_ERR_INVALID, "afe_pcm16_to_float: bad arguments");
  if (count == 0) return AFE_OK;
  Pcm16Kernel k{d_in, d_out, count};
  ProfScope ps(ctx, AFE_K_MIX, (cudaStream_t)stream);
  if (afe_launch(k, 0, ctx->num_sms * 8, (cudaStream_t)stream) != 0) return fail(AFE_ERR_CUDA, "afe_pcm16_to_float: launch failed");
  return AFE_OK;
}

static int resample_batch_impl(afe_ctx* ctx, void* stream, const void* const* h_in, int pcm16, const int64_t* h_lengths,
                               int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                               float* const* h_mixed, void* d_workspace, size_t workspace_bytes);

int afe_resample_batch(afe_ctx* ctx, void* stream, const float* const* h_in, const int64_t* h_lengths, int ncalls,
                       int rows, int orig_hz, int new_hz, float* const* h_out, float* const* h_mixed,
                       void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_resample_batch");
  return resample_batch_impl(ctx, stream, reinterpret_cast<const void* const*>(h_in), 0, h_lengths, ncalls, rows,
                             orig_hz, new_hz, h_out, h_mixed, d_workspace, workspace_bytes);
}

int afe_resample_batch_pcm16(afe_ctx* ctx, void* stream, const int16_t* const* h_in, const int64_t* h_lengths,
                             int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                             float* const* h_mixed, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_resample_batch_pcm16");
  return resample_batch_impl(ctx, stream, reinterpret_cast<const void* const*>(h_in), 1, h_lengths, ncalls, rows,
                             orig_hz, new_hz, h_out, h_mixed, d_workspace, workspace_bytes);
}

int afe_resample_batch_wav16(afe_ctx* ctx, void* stream, const int16_t* const* h_in, const int64_t* h_lengths,
                             int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                             float* const* h_mixed, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_resample_batch_wav16");
  return resample_batch_impl(ctx, stream, reinterpret_cast<const void* const*>(h_in), 2, h_lengths, ncalls, rows,
                             orig_hz, new_hz, h_out, h_mixed, d_workspace, workspace_bytes);
}

int afe_wav_decode(afe_ctx* ctx, void* stream, const void* d_frames, int format, int channels, int64_t num_frames,
                   float* d_out) {
  AFE_NVTX("afe_wav_decode");
  if (!ctx || channels <= 0 || num_frames < 0 || format < AFE_WAV_U8 || format > AFE_WAV_F64 ||
      (num_frames > 0 && (!d_frames || !d_out)))
    return fail(AFE_ERR_INVALID, "afe_wav_decode: bad arguments");
  if (num_frames == 0) return AFE_OK;
  WavDecodeKernel k{static_cast<const unsigned char*>(d_frames), d_out, num_frames, channels, format};
  ProfScope ps(ctx, AFE_K_MIX, (cudaStream_t)stream);
  if (afe_launch(k, 0, ctx->num_sms * 8, (cudaStream_t)stream) != 0) return fail(AFE_ERR_CUDA, "afe_wav_decode: launch failed");
  return AFE_OK;
}

static int resample_batch_impl(afe_ctx* ctx, void* stream, const void* const* h_in, int pcm16, const int64_t* h_lengths,
                               int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                               float* const* h_mixed, void* d_workspace, size_t workspace_bytes) {
  if (!ctx || ncalls < 0 || rows <= 0 || orig_hz <= 0 || new_hz <= 0)
    return fail(AFE_ERR_INVALID, "afe_resample_batch: bad arguments");
  if (ncalls == 0) return AFE_OK;
  if (!h_in || !h_lengths || !h_out) return fail(AFE_ERR_INVALID, "afe_resample_batch: NULL argument");
  if (orig_hz == new_hz) return fail(AFE_ERR_INVALID, "afe_resample_batch: equal rates need no resampling");
  if (h_mixed && rows != 2) return fail(AFE_ERR_INVALID, "afe_resample_batch: mixed output needs exactly 2 rows");
  const DeviceTaps* dt = ctx->get_taps(orig_hz, new_hz);
  if (!dt) return AFE_ERR_CUDA;
  const ResampleTaps& t = dt->host;
  if (pcm16 && !(t.O == 1 && t.N == 2 && t.K == 15))
    return fail(AFE_ERR_UNSUPPORTED, "afe_resample_batch_pcm16: only the 1:2 rate pair stages PCM directly; "
                                     "convert with afe_pcm16_to_float first");
  if (!pcm16 && (!(t.O == 1 && t.N == 2 && t.K == 15) || ncalls == 1)) {
    // other rate pairs: one launch per call through the generic path
    for (int c = 0; c < ncalls; ++c) {
      const int rc = afe_resample(ctx, stream, static_cast<const float*>(h_in[c]), rows, h_lengths[c], orig_hz, new_hz, h_out[c],
                                  h_mixed ? h_mixed[c] : nullptr, resampled_length(h_lengths[c], orig_hz, new_hz));
      if (rc != AFE_OK) return rc;
    }
    return AFE_OK;
  }
  const size_t need = afe_resample_batch_workspace_bytes(ncalls);
  if (!d_workspace || workspace_bytes < need) return fail(AFE_ERR_WORKSPACE, "afe_resample_batch: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;
  StagingBlock* blk = ctx->staging_acquire(need);
  if (!blk) return AFE_ERR_CUDA;
  UpCall* hc = static_cast<UpCall*>(blk->host);
  const bool mix = h_mixed != nullptr;
  long items = 0;
  bool bad = false;
  for (int c = 0; c < ncalls; ++c) {
    if (h_lengths[c] < 0 || (h_lengths[c] > 0 && (!h_in[c] || !h_out[c] || (mix && !h_mixed[c])))) bad = true;
    hc[c] = UpCall{static_cast<const float*>(h_in[c]), h_out[c], mix ? h_mixed[c] : nullptr, h_lengths[c], (int64_t)items};
    items += (long)((h_lengths[c] + 1023) / 1024) * (mix ? 1 : rows);
  }
  if (bad) { ctx->staging_submit(blk, s); return fail(AFE_ERR_INVALID, "afe_resample_batch: NULL buffer or negative length"); }
  const int up = dev_upload_async(d_workspace, hc, sizeof(UpCall) * (size_t)ncalls, s);
  ctx->staging_submit(blk, s);
  if (up != 0) return AFE_ERR_CUDA;
  int rc;
  {
    ProfScope ps(ctx, AFE_K_RESAMPLE, s);
    rc = launch_up2x(ctx, s, t, static_cast<const UpCall*>(d_workspace), UpCall{}, ncalls, rows, mix, items, pcm16);
  }
  if (rc != 0) return fail(AFE_ERR_CUDA, std::string("afe_resample_batch: launch failed: ") + dev_last_error_string());
  return AFE_OK;
}

}  // extern "C"
