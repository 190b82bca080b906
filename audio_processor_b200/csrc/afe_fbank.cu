// afe_fbank.cu -- C ABI entry point of the batched fbank featuriser.
#include <algorithm>
#include <string>
#include <vector>

#include "afe_ctx.h"
#include "afe_fbank_kernels.h"
#include "afe_fbank_pipe.h"

using namespace afe;

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// AFE_FBANK_PIPE=1 selects the three-group tile pipeline (afe_fbank_pipe.h, 16-frame tiles) instead of the
// phase-serial kernel (32-frame tiles); read once per process.  Both give bit-identical features.
#ifndef AFE_FBANK_PIPE_DEFAULT
#define AFE_FBANK_PIPE_DEFAULT 0
#endif
bool use_pipe() {
  static const bool on = [] { const char* e = std::getenv("AFE_FBANK_PIPE"); return e ? e[0] == '1' : (AFE_FBANK_PIPE_DEFAULT != 0); }();
  return on;
}
int fbank_tile() { return use_pipe() ? kPipeTile : kFbTile; }

// workspace layout (device): [descriptors][partials: ntiles*80*2 f64][stats: B*80*2 f32]
struct FbankLayout {
  int B = 0;
  long ntiles = 0;
  size_t off_wave = 0, off_len = 0, off_frames = 0, off_prefix = 0, off_row = 0, off_tp = 0, off_whole = 0, off_sel = 0,
         desc_bytes = 0;
  size_t off_partials = 0, off_stats = 0, total = 0;
};

FbankLayout fbank_layout(const int64_t* lengths, int B) {
  FbankLayout l;
  l.B = B;
  for (int b = 0; b < B; ++b) {
    const int64_t T = afe_fbank_num_frames(lengths[b]);
    l.ntiles += (long)((T + fbank_tile() - 1) / fbank_tile());
  }
  size_t o = 0;
  l.off_wave = o;   o = align_up(o + sizeof(void*) * (size_t)B, 16);
  l.off_len = o;    o = align_up(o + sizeof(int64_t) * (size_t)B, 16);
  l.off_prefix = o; o = align_up(o + sizeof(int64_t) * ((size_t)B + 1), 16);
  l.off_row = o;    o = align_up(o + sizeof(int64_t) * (size_t)B, 16);
  l.off_frames = o; o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.off_tp = o;     o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.off_whole = o;  o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.off_sel = o;    o = align_up(o + sizeof(int32_t) * (size_t)B, 16);
  l.desc_bytes = o;
  l.off_partials = align_up(o, 256);
  l.off_stats = align_up(l.off_partials + sizeof(double) * 2 * kMel * (size_t)l.ntiles, 256);
  l.total = align_up(l.off_stats + sizeof(float) * 2 * kMel * (size_t)B, 256);
  return l;
}

template <typename T>
int run_fbank(afe_ctx* ctx, cudaStream_t s, const BatchDesc& d, float* feat, double* partials,
              const T* window, const T* tw, int32_t* mask, float padding_value) {
  ProfScope ps(ctx, AFE_K_FBANK, s);
  if (use_pipe()) {
    FbankPipeKernel<T> k;
    k.d = d; k.out = feat; k.partials = partials; k.window = window; k.tw512 = tw;
    k.mask = mask; k.padding_value = padding_value;
    k.log_floor = ctx->log_floor;
    k.mel_w = ctx->kaldi_slots.w; k.mel_start = ctx->kaldi_slots.start;
    return afe_launch_fbank_pipe(k, ctx->num_sms, s);
  }
  FbankKernel<T> k;
  k.d = d; k.out = feat; k.partials = partials; k.window = window; k.tw512 = tw;
  k.mask = mask; k.padding_value = padding_value;
  k.log_floor = ctx->log_floor;
  k.mel_w = ctx->kaldi_slots.w; k.mel_start = ctx->kaldi_slots.start;
  return afe_launch(k, FbankKernel<T>::S::kBytes, ctx->num_sms * FbankKernel<T>::kMinBlocks, s);
}

}  // namespace

extern "C" {

size_t afe_fbank_workspace_bytes(const int64_t* h_lengths, int batch) {
  if (!h_lengths || batch <= 0) return 256;
  return fbank_layout(h_lengths, batch).total;
}

int afe_fbank_layout(const int64_t* h_lengths, int batch, int group_size, int64_t* h_group_tp, int64_t* total_rows) {
  if (batch < 0 || (batch > 0 && !h_lengths)) return fail(AFE_ERR_INVALID, "afe_fbank_layout: bad arguments");
  const int G = group_size > 0 ? group_size : (batch > 0 ? batch : 1);
  int64_t rows = 0;
  for (int g0 = 0, g = 0; g0 < batch; g0 += G, ++g) {
    const int n = std::min(G, batch - g0);
    const int64_t tp = afe_fbank_padded_frames(h_lengths + g0, n);
    if (h_group_tp) h_group_tp[g] = tp;
    rows += tp * n;
  }
  if (total_rows) *total_rows = rows;
  return AFE_OK;
}

}  // extern "C"

namespace {
// shared body: utterance b lands at frame row rows_of[b] of d_features, padded to tp_of[b] frames
int fbank_run(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths, int batch,
              const int64_t* rows_of, const int64_t* tp_of, float padding_value, int precision, int flags,
              float* d_features, int32_t* d_mask, void* d_workspace, size_t workspace_bytes, const char* who,
              unsigned char* h_desc = nullptr /* caller-owned host block for the descriptors instead of the pinned pool */) {
  const bool raw_only = (flags & AFE_FBANK_RAW) != 0;
  const FbankLayout l = fbank_layout(h_lengths, batch);
  if (workspace_bytes < l.total) return fail(AFE_ERR_WORKSPACE, std::string(who) + ": workspace too small");
  cudaStream_t s = (cudaStream_t)stream;

  // descriptors: pinned staging -> head of the workspace
  StagingBlock* blk = h_desc ? nullptr : ctx->staging_acquire(l.desc_bytes);
  if (!h_desc && !blk) return AFE_ERR_CUDA;
  unsigned char* h = h_desc ? h_desc : static_cast<unsigned char*>(blk->host);
  const float** hw = reinterpret_cast<const float**>(h + l.off_wave);
  int64_t* hl = reinterpret_cast<int64_t*>(h + l.off_len);
  int64_t* hp = reinterpret_cast<int64_t*>(h + l.off_prefix);
  int64_t* hr = reinterpret_cast<int64_t*>(h + l.off_row);
  int32_t* hf = reinterpret_cast<int32_t*>(h + l.off_frames);
  int32_t* ht = reinterpret_cast<int32_t*>(h + l.off_tp);
  int32_t* hwh = reinterpret_cast<int32_t*>(h + l.off_whole);
  int32_t* hsel = reinterpret_cast<int32_t*>(h + l.off_sel);
  // The fbank launch gives block i the tiles [i*per, (i+1)*per).  An utterance whose tiles all fall
  // into one block is normalised by that block right after its last tile (its raw rows are still
  // in L2); the others -- at most one per block boundary -- and those without frames go to the
  // stats / norm kernels.
  const long grid = afe_async_grid(l.ntiles, use_pipe() ? ctx->num_sms : ctx->num_sms * FbankKernel<double>::kMinBlocks);
  const long per = grid > 0 ? (l.ntiles + grid - 1) / grid : 1;
  int nsel = 0;
  int64_t acc = 0, rows = 0, tp_max = 0;
  bool bad = false;
  for (int b = 0; b < batch; ++b) {
    const int64_t T = afe_fbank_num_frames(h_lengths[b]);
    const int64_t tp = tp_of[b];
    if (T > 0 && !h_waves[b]) bad = true;
    if (tp < T || (tp & 1) || rows_of[b] < 0 || (rows_of[b] & 1)) bad = true;
    hw[b] = h_waves[b]; hl[b] = h_lengths[b]; hf[b] = (int32_t)T; hp[b] = acc; hr[b] = rows_of[b]; ht[b] = (int32_t)tp;
    const int64_t nt = (T + fbank_tile() - 1) / fbank_tile();
    const bool whole = !raw_only && nt > 0 && (acc / per) == ((acc + nt - 1) / per);
    hwh[b] = whole ? 1 : 0;
    if (!whole) { hsel[nsel++] = b; tp_max = std::max(tp_max, tp); }
    acc += nt;
    rows += tp;
  }
  hp[batch] = acc;
  if (bad || (rows > 0 && (!d_features || (!raw_only && !d_mask)))) {
    if (blk) ctx->staging_submit(blk, s);
    return fail(AFE_ERR_INVALID, std::string(who) + ": NULL waveform / output or inconsistent placement");
  }
  unsigned char* ws = static_cast<unsigned char*>(d_workspace);
  const int up = dev_upload_async(ws, h, l.desc_bytes, s);
  if (blk) ctx->staging_submit(blk, s);
  if (up != 0) return AFE_ERR_CUDA;

  BatchDesc d;
  d.wave = reinterpret_cast<const float* const*>(ws + l.off_wave);
  d.len = reinterpret_cast<const int64_t*>(ws + l.off_len);
  d.frames = reinterpret_cast<const int32_t*>(ws + l.off_frames);
  d.tile_prefix = reinterpret_cast<const int64_t*>(ws + l.off_prefix);
  d.out_row = reinterpret_cast<const int64_t*>(ws + l.off_row);
  d.tp = reinterpret_cast<const int32_t*>(ws + l.off_tp);
  d.whole = reinterpret_cast<const int32_t*>(ws + l.off_whole);
  d.sel = reinterpret_cast<const int32_t*>(ws + l.off_sel);
  d.nsel = nsel;
  d.B = batch; d.ntiles = l.ntiles;
  double* partials = reinterpret_cast<double*>(ws + l.off_partials);
  float* stats = reinterpret_cast<float*>(ws + l.off_stats);

  int rc = (precision == AFE_PRECISION_F64)
               ? run_fbank<double>(ctx, s, d, d_features, partials, ctx->d_povey64, ctx->d_tw64, d_mask, padding_value)
               : run_fbank<float>(ctx, s, d, d_features, partials, ctx->d_povey32, ctx->d_tw32, d_mask, padding_value);
  if (rc == 0 && !raw_only && rows > 0 && nsel > 0) {
    FbankStatsKernel sk{d, partials, stats};
    { ProfScope ps(ctx, AFE_K_FBANK_STATS, s); rc = afe_launch(sk, 0, 1 << 20, s); }
    if (rc == 0) {
      FbankNormKernel nk;
      nk.d = d; nk.feat = d_features; nk.mask = d_mask; nk.stats = stats;
      nk.padding_value = padding_value;
      nk.items_per_utt = (long)((tp_max * kMel + FbankNormKernel::kItemFloats - 1) / FbankNormKernel::kItemFloats);
      ProfScope ps(ctx, AFE_K_FBANK_NORM, s);
      rc = afe_launch(nk, 0, ctx->num_sms * 8, s);
    }
  }
  if (rc != 0) return fail(AFE_ERR_CUDA, std::string(who) + ": launch failed: " + dev_last_error_string());
  return AFE_OK;
}
}  // namespace

extern "C" {

int afe_fbank_batch(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths, int batch,
                    int group_size, float padding_value, int precision, int flags, float* d_features,
                    int32_t* d_mask, void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_fbank_batch");
  if (!ctx || batch < 0) return fail(AFE_ERR_INVALID, "afe_fbank_batch: bad arguments");
  if (batch == 0) return AFE_OK;
  if (!h_waves || !h_lengths || !d_workspace) return fail(AFE_ERR_INVALID, "afe_fbank_batch: NULL argument");
  if (precision != AFE_PRECISION_F64 && precision != AFE_PRECISION_F32)
    return fail(AFE_ERR_INVALID, "afe_fbank_batch: unknown precision");
  const int G = group_size > 0 ? group_size : batch;
  std::vector<int64_t> rows_of((size_t)batch), tp_of((size_t)batch);
  int64_t rows = 0;
  for (int g0 = 0; g0 < batch; g0 += G) {
    const int n = std::min(G, batch - g0);
    const int64_t tp = afe_fbank_padded_frames(h_lengths + g0, n);
    for (int b = g0; b < g0 + n; ++b) { rows_of[(size_t)b] = rows; tp_of[(size_t)b] = tp; rows += tp; }
  }
  return fbank_run(ctx, stream, h_waves, h_lengths, batch, rows_of.data(), tp_of.data(), padding_value, precision,
                   flags, d_features, d_mask, d_workspace, workspace_bytes, "afe_fbank_batch");
}

int afe_fbank_batch_placed(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                           int batch, const int64_t* h_out_rows, const int64_t* h_padded_frames,
                           float padding_value, int precision, int flags, float* d_features, int32_t* d_mask,
                           void* d_workspace, size_t workspace_bytes) {
  AFE_NVTX("afe_fbank_batch_placed");
  if (!ctx || batch < 0) return fail(AFE_ERR_INVALID, "afe_fbank_batch_placed: bad arguments");
  if (batch == 0) return AFE_OK;
  if (!h_waves || !h_lengths || !h_out_rows || !h_padded_frames || !d_workspace)
    return fail(AFE_ERR_INVALID, "afe_fbank_batch_placed: NULL argument");
  if (precision != AFE_PRECISION_F64 && precision != AFE_PRECISION_F32)
    return fail(AFE_ERR_INVALID, "afe_fbank_batch_placed: unknown precision");
  return fbank_run(ctx, stream, h_waves, h_lengths, batch, h_out_rows, h_padded_frames, padding_value, precision,
                   flags, d_features, d_mask, d_workspace, workspace_bytes, "afe_fbank_batch_placed");
}

}  // extern "C"

// ---- afe_fbank_batch_static: the same featuriser for callers that must not allocate ----------------------
// (no counterpart in the reference, see include/afe.h)
// The workspace (descriptors | tile partials | utterance statistics) is a file-scope device array that every
// GPU loading the library holds until exit.  Its name is only a host-side stand-in in host code: the device
// address is taken once per device with cudaGetSymbolAddress at set-up (afe_create, that device current) and
// kept in the context.  Device code never names the array; nothing in it has to start at zero (every byte a
// kernel reads was written earlier in the same call), and calls that overlap on the GPU race on it.
namespace {
constexpr size_t kStaticScratchBytes = (size_t)AFE_FBANK_STATIC_SCRATCH_BYTES;
#if AFE_CUDA
static __device__ __align__(256) unsigned char g_fbank_static_scratch[kStaticScratchBytes];
#else
alignas(256) static unsigned char g_fbank_static_scratch[kStaticScratchBytes];
#endif
}  // namespace

namespace afe {
int fbank_static_setup(afe_ctx* ctx) {
#if AFE_CUDA
  void* p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&p, g_fbank_static_scratch);
  if (e != cudaSuccess || !p) { note_error("cudaGetSymbolAddress(fbank static scratch)", cudaGetErrorString(e)); return -1; }
  ctx->fbank_static_scratch = p;
#else
  ctx->fbank_static_scratch = g_fbank_static_scratch;
#endif
  return 0;
}
}  // namespace afe

extern "C" {

int afe_fbank_batch_static(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths, int batch,
                           int group_size, float padding_value, int precision, int flags, float* d_features,
                           int32_t* d_mask) {
  AFE_NVTX("afe_fbank_batch_static");
  if (!ctx || batch < 0) return fail(AFE_ERR_INVALID, "afe_fbank_batch_static: bad arguments");
  if (batch == 0) return AFE_OK;
  if (!h_waves || !h_lengths) return fail(AFE_ERR_INVALID, "afe_fbank_batch_static: NULL argument");
  if (precision != AFE_PRECISION_F64 && precision != AFE_PRECISION_F32)
    return fail(AFE_ERR_INVALID, "afe_fbank_batch_static: unknown precision");
  if (batch > AFE_FBANK_STATIC_MAX_BATCH)
    return fail(AFE_ERR_WORKSPACE, "afe_fbank_batch_static: batch of " + std::to_string(batch) + " waveforms exceeds the maximum of " +
                                       std::to_string(AFE_FBANK_STATIC_MAX_BATCH) + "; use afe_fbank_batch with a caller workspace");
  const FbankLayout l = fbank_layout(h_lengths, batch);
  if (l.total > kStaticScratchBytes)
    return fail(AFE_ERR_WORKSPACE, "afe_fbank_batch_static: the batch needs " + std::to_string(l.total) + " bytes of scratch, the static array holds " +
                                       std::to_string(kStaticScratchBytes) + " (about " + std::to_string(kStaticScratchBytes / (2 * kMel * sizeof(double)) * fbank_tile() / 100) +
                                       " s of audio per call); use afe_fbank_batch with a caller workspace");
  if (!ctx->fbank_static_scratch) return fail(AFE_ERR_CUDA, "afe_fbank_batch_static: the static scratch was not set up for this device");
  // descriptors and placement on the stack: no pinned pool, no heap block whose size depends on the call
  alignas(16) unsigned char h_desc[16 * 1024];
  static_assert(sizeof(h_desc) >= (size_t)AFE_FBANK_STATIC_MAX_BATCH * 56 + 16 * 9, "descriptor block of the largest batch");
  if (l.desc_bytes > sizeof(h_desc)) return fail(AFE_ERR_WORKSPACE, "afe_fbank_batch_static: descriptor block too large");
  int64_t rows_of[AFE_FBANK_STATIC_MAX_BATCH], tp_of[AFE_FBANK_STATIC_MAX_BATCH];
  const int G = group_size > 0 ? group_size : batch;
  int64_t rows = 0;
  for (int g0 = 0; g0 < batch; g0 += G) {
    const int n = std::min(G, batch - g0);
    const int64_t tp = afe_fbank_padded_frames(h_lengths + g0, n);
    for (int b = g0; b < g0 + n; ++b) { rows_of[b] = rows; tp_of[b] = tp; rows += tp; }
  }
  return fbank_run(ctx, stream, h_waves, h_lengths, batch, rows_of, tp_of, padding_value, precision, flags, d_features,
                   d_mask, ctx->fbank_static_scratch, kStaticScratchBytes, "afe_fbank_batch_static", h_desc);
}

}  // extern "C"
