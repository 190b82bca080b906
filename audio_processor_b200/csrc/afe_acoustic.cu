// afe_acoustic.cu -- C ABI entry point of the batched acoustic features.
#include <algorithm>

#include "afe_acoustic_kernels.h"
#include "afe_ctx.h"

using namespace afe;

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct AcLayout {
  int B = 0;
  long ntiles = 0;
  int64_t rows = 0;
  size_t off_wave = 0, off_len = 0, off_f1 = 0, off_f2 = 0, off_tprefix = 0, off_rprefix = 0, off_oidx = 0, desc_bytes = 0;
  size_t off_meldb = 0, off_tile = 0, off_mfcc = 0, off_bound = 0, off_gmax = 0, total = 0;
};

inline int64_t frames2_of(int64_t L) { return L < kFrame ? 0 : 1 + L / kHop; }
inline int64_t frames1_of(int64_t L) { return L < kFrame ? 0 : (L - kFrame) / kHop + 1; }

AcLayout ac_layout(const int64_t* lengths, int B) {
  AcLayout l;
  l.B = B;
  for (int b = 0; b < B; ++b) {
    const int64_t F2 = frames2_of(lengths[b]);
    l.rows += F2;
    l.ntiles += (long)((F2 + kTileFrames - 1) / kTileFrames);
  }
  size_t o = 0;
  l.off_wave = o;    o = align_up(o + sizeof(void*) * (size_t)B, 16);
  l.off_len = o;     o = align_up(o + sizeof(int64_t) * (size_t)B, 16);
  l.off_tprefix = o; o = align_up(o + sizeof(int64_t) * ((size_t)B + 1), 16);
  l.off_rprefix = o; o = align_up(o + sizeof(int64_t) * ((size_t)B + 1), 16);
  l.off_f1 = o;      o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.off_f2 = o;      o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.off_oidx = o;    o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.desc_bytes = o;
  l.off_tile = align_up(o, 256);
  l.off_mfcc = align_up(l.off_tile + sizeof(double) * kTileRec * (size_t)l.ntiles, 256);
  l.off_bound = align_up(l.off_mfcc + sizeof(double) * 2 * 14 * (size_t)l.ntiles, 256);
  l.off_gmax = align_up(l.off_bound + sizeof(float) * 2 * kBoundary * (size_t)l.ntiles, 256);
  l.off_meldb = align_up(l.off_gmax + sizeof(float) * (size_t)B, 256);
  l.total = align_up(l.off_meldb + sizeof(float) * kAcMel * (size_t)l.rows, 256);
  return l;
}

}  // namespace

extern "C" {

size_t afe_acoustic_workspace_bytes(const int64_t* h_lengths, int batch) {
  if (!h_lengths || batch <= 0) return 256;
  return ac_layout(h_lengths, batch).total;
}

int afe_acoustic_batch(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths, int batch,
                       int sample_rate, float* d_out, int32_t* d_status, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_acoustic_batch");
  return afe_acoustic_batch_placed(ctx, stream, h_waves, h_lengths, batch, sample_rate, nullptr, d_out, d_status,
                                   d_workspace, workspace_bytes);
}

int afe_acoustic_batch_placed(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                              int batch, int sample_rate, const int32_t* h_out_index, float* d_out,
                              int32_t* d_status, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_acoustic_batch_placed");
  return afe_acoustic_batch_scaled(ctx, stream, h_waves, h_lengths, batch, sample_rate, h_out_index, d_out, d_status,
                                   nullptr, nullptr, 0, nullptr, d_workspace, workspace_bytes);
}

int afe_acoustic_batch_scaled(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                              int batch, int sample_rate, const int32_t* h_out_index, float* d_out,
                              int32_t* d_status, const double* d_mean, const double* d_scale, int n_features,
                              double* d_scaled, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_acoustic_batch_scaled");
  if (!ctx || batch < 0) return fail(AFE_ERR_INVALID, "afe_acoustic_batch: bad arguments");
  if (d_scaled && (!d_mean || !d_scale || n_features <= 0))
    return fail(AFE_ERR_INVALID, "afe_acoustic_batch_scaled: scaler mean / scale / n_features missing");
  if (batch == 0) return AFE_OK;
  if (sample_rate != 16000)
    return fail(AFE_ERR_UNSUPPORTED, "afe_acoustic_batch: only sample_rate 16000 (frame 400 / hop 160) is built");
  if (!h_waves || !h_lengths || !d_out || !d_status || !d_workspace)
    return fail(AFE_ERR_INVALID, "afe_acoustic_batch: NULL argument");
  const AcLayout l = ac_layout(h_lengths, batch);
  if (workspace_bytes < l.total) return fail(AFE_ERR_WORKSPACE, "afe_acoustic_batch: workspace too small");
  cudaStream_t s = (cudaStream_t)stream;

  StagingBlock* blk = ctx->staging_acquire(l.desc_bytes);
  if (!blk) return AFE_ERR_CUDA;
  unsigned char* h = static_cast<unsigned char*>(blk->host);
  const float** hw = reinterpret_cast<const float**>(h + l.off_wave);
  int64_t* hl = reinterpret_cast<int64_t*>(h + l.off_len);
  int64_t* htp = reinterpret_cast<int64_t*>(h + l.off_tprefix);
  int64_t* hrp = reinterpret_cast<int64_t*>(h + l.off_rprefix);
  int32_t* h1 = reinterpret_cast<int32_t*>(h + l.off_f1);
  int32_t* h2 = reinterpret_cast<int32_t*>(h + l.off_f2);
  int32_t* hoi = reinterpret_cast<int32_t*>(h + l.off_oidx);
  int64_t tiles = 0, rows = 0;
  bool bad = false;
  for (int b = 0; b < batch; ++b) {
    const int64_t F1 = frames1_of(h_lengths[b]), F2 = frames2_of(h_lengths[b]);
    if (F2 > 0 && !h_waves[b]) bad = true;
    hw[b] = h_waves[b]; hl[b] = h_lengths[b]; h1[b] = (int32_t)F1; h2[b] = (int32_t)F2;
    htp[b] = tiles; hrp[b] = rows;
    hoi[b] = h_out_index ? h_out_index[b] : b;
    if (hoi[b] < 0) bad = true;
    tiles += (F2 + kTileFrames - 1) / kTileFrames;
    rows += F2;
  }
  htp[batch] = tiles; hrp[batch] = rows;
  if (bad) { ctx->staging_submit(blk, s); return fail(AFE_ERR_INVALID, "afe_acoustic_batch: NULL waveform or negative output row"); }
  unsigned char* ws = static_cast<unsigned char*>(d_workspace);
  const int up = dev_upload_async(ws, h, l.desc_bytes, s);
  ctx->staging_submit(blk, s);
  if (up != 0) return AFE_ERR_CUDA;

  AcDesc d;
  d.wave = reinterpret_cast<const float* const*>(ws + l.off_wave);
  d.len = reinterpret_cast<const int64_t*>(ws + l.off_len);
  d.frames1 = reinterpret_cast<const int32_t*>(ws + l.off_f1);
  d.frames2 = reinterpret_cast<const int32_t*>(ws + l.off_f2);
  d.tile_prefix = reinterpret_cast<const int64_t*>(ws + l.off_tprefix);
  d.row_prefix = reinterpret_cast<const int64_t*>(ws + l.off_rprefix);
  d.out_index = reinterpret_cast<const int32_t*>(ws + l.off_oidx);
  d.B = batch; d.ntiles = l.ntiles;
  double* tile_rec = reinterpret_cast<double*>(ws + l.off_tile);
  double* mfcc_rec = reinterpret_cast<double*>(ws + l.off_mfcc);
  float* boundary = reinterpret_cast<float*>(ws + l.off_bound);
  float* gmax = reinterpret_cast<float*>(ws + l.off_gmax);
  float* meldb = reinterpret_cast<float*>(ws + l.off_meldb);

  AcousticTileKernel tk;
  tk.d = d; tk.hann = ctx->d_hann32; tk.tw512 = ctx->d_tw32; tk.mel_w = ctx->htk_slots.w; tk.mel_start = ctx->htk_slots.start;
  tk.meldb = meldb; tk.tile_rec = tile_rec; tk.boundary = boundary;
  int rc;
  { ProfScope ps(ctx, AFE_K_AC_TILE, s); rc = afe_launch(tk, AcSmem::kBytes, ctx->num_sms, s); }
  if (rc == 0) {
    AcousticMaxKernel mk{d, tile_rec, gmax};
    ProfScope ps(ctx, AFE_K_AC_MAX, s);
    rc = afe_launch(mk, 0, 1 << 20, s);
  }
  if (rc == 0) {
    AcousticMfccKernel ck;
    ck.d = d; ck.meldb = meldb; ck.gmax = gmax; ck.boundary = boundary; ck.mfcc_rec = mfcc_rec;
    ProfScope ps(ctx, AFE_K_AC_MFCC, s);
    rc = afe_launch(ck, AcousticMfccKernel::kSmemBytes, ctx->num_sms * 8, s);
  }
  if (rc == 0) {
    AcousticFinalKernel fk{d, tile_rec, mfcc_rec, boundary, d_out, d_status, d_mean, d_scale, d_scaled, n_features};
    ProfScope ps(ctx, AFE_K_AC_FINAL, s);
    rc = afe_launch(fk, 0, 1 << 20, s);
  }
  if (rc != 0) return fail(AFE_ERR_CUDA, std::string("afe_acoustic_batch: launch failed: ") + dev_last_error_string());
  return AFE_OK;
}

}  // extern "C"
