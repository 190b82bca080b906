/* afe.h -- C ABI of the B200 audio front end (libafe.so).
 *
 * Drop-in boundary for the waveform hot path of jaratM/audio_processor.  The reference is
 * pure Python with no FFI of its own; each entry point below names
 * the reference call site whose arithmetic it replaces.  Conventions:
 *   - every function returns 0 on success or a negative afe_status; afe_last_error() gives
 *     the message of the calling thread's last failure; no C++ exception crosses the ABI;
 *   - the caller allocates every input, output and workspace buffer; "d_" parameters are
 *     device pointers (cudaMalloc / torch CUDA tensors), "h_" parameters are host pointers;
 *   - work is enqueued on the caller's stream (a cudaStream_t passed as void*), nothing
 *     synchronises; the library keeps no mutable global state and is re-entrant;
 *   - an afe_ctx owns only immutable constant tables for one device.
 */
#ifndef AFE_H_
#define AFE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct afe_ctx afe_ctx;

typedef enum afe_status {
  AFE_OK = 0,
  AFE_ERR_INVALID = -1,      /* bad argument (also: sampling_rate != 16000 -> ValueError upstream) */
  AFE_ERR_CUDA = -2,         /* CUDA runtime failure, message in afe_last_error */
  AFE_ERR_WORKSPACE = -3,    /* workspace too small */
  AFE_ERR_UNSUPPORTED = -4   /* e.g. a rate pair whose tap tile does not fit shared memory */
} afe_status;

#define AFE_FBANK_FRAME 400
#define AFE_FBANK_HOP 160
#define AFE_FBANK_MELS 80
#define AFE_ACOUSTIC_FEATURES 38

/* precision of the linear path (pre-emphasis .. FFT) of afe_fbank_batch */
#define AFE_PRECISION_F64 0   /* default: matches the reference's float64 path to <= 1e-3 */
#define AFE_PRECISION_F32 1   /* opt-in fast path, ~3e-3 worst case in empty bands */
/* flags of afe_fbank_batch */
#define AFE_FBANK_RAW 1       /* stop after the log-mel: no normalisation, padding or mask */

/* ---- library / context ------------------------------------------------------------- */
const char* afe_version(void);
/* 1 = CUDA sm_100a build (the product); 2 = host emulation of the kernels (tests only) */
int afe_backend(void);
int afe_create(int device, afe_ctx** out);
void afe_destroy(afe_ctx* ctx);
const char* afe_last_error(void);
int afe_num_sms(const afe_ctx* ctx);

/* ---- host-side integer / float64 contracts (no device work) ------------------------- */
/* ceil(new*L/orig): torchaudio _apply_sinc_resample_kernel, functional.py:1425 */
int64_t afe_resampled_length(int64_t length, int orig_hz, int new_hz);
/* 1 + (L - 400) / 160, 0 if L < 400: transformers audio_utils.spectrogram :778 */
int64_t afe_fbank_num_frames(int64_t length);
/* longest frame count rounded up to a multiple of 2: SequenceFeatureExtractor.pad,
 * feature_extraction_seamless_m4t.py:267-275 */
int64_t afe_fbank_padded_frames(const int64_t* h_lengths, int batch);
/* AudioProcessor.split_audio chunk loop, services/audio_processor.py:115-158.
 * Writes up to cap (start, end) pairs; *n receives the number of chunks. */
int afe_chunk_table(int64_t total_samples, int64_t chunk_samples, int64_t step_samples,
                    int64_t* h_starts, int64_t* h_ends, int cap, int* n);
/* SpeechSegment._extract_segments, services/speech_segment.py:385-418 (float64 arithmetic,
 * int() truncation).  intervals = n_intervals (start, end) pairs in seconds. */
int afe_segment_table(int64_t num_samples, int sample_rate, const double* h_intervals, int n_intervals,
                      double padding, double max_len_sec, double overlap_sec,
                      int64_t* h_start_sample, int64_t* h_end_sample, double* h_start_sec, double* h_end_sec,
                      int cap, int* n);
/* the tap bank of torchaudio.transforms.Resample(orig, new).kernel, functional.py:1341-1400.
 * h_taps may be NULL to query sizes; layout [phases][taps]. */
int afe_resample_taps(int orig_hz, int new_hz, float* h_taps, size_t cap_floats,
                      int* width, int* in_stride, int* phases, int* taps_per_phase);

/* ---- device kernels ----------------------------------------------------------------- */
/* torchaudio.transforms.Resample.forward on a [rows, length] float32 tensor
 * (AudioProcessor.get_resampler, services/audio_processor.py:99-102).  If d_mixed is not
 * NULL and rows == 2 it also writes mean(dim=0) = (row0 + row1)/2 (split_audio :123-131). */
int afe_resample(afe_ctx* ctx, void* stream, const float* d_in, int rows, int64_t length,
                 int orig_hz, int new_hz, float* d_out, float* d_mixed, int64_t out_length);

/* The same for a batch of calls in ONE launch (AudioProcessor.process_batch iterates
 * split_audio over the files of a batch, services/audio_processor.py:324-347).  h_in[c] /
 * h_out[c] / h_mixed[c] are DEVICE pointers of call c ([rows, h_lengths[c]] in,
 * [rows, ceil(new*L/orig)] out); h_mixed may be NULL.  d_workspace holds the per-call
 * descriptors (afe_resample_batch_workspace_bytes). */
size_t afe_resample_batch_workspace_bytes(int ncalls);
int afe_resample_batch(afe_ctx* ctx, void* stream, const float* const* h_in, const int64_t* h_lengths,
                       int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                       float* const* h_mixed, void* d_workspace, size_t workspace_bytes);

/* The same reading 16-bit PCM: h_in[c] points to int16 samples [rows, h_lengths[c]] on the
 * device and x / 32768 -- torchaudio.load's normalisation (AudioProcessor.load_audio,
 * services/audio_processor.py:72-93) -- is applied while the tile is staged, so a call
 * crosses PCIe at 2 bytes per sample.  Only the 1:2 rate pair (8 k -> 16 k). */
int afe_resample_batch_pcm16(afe_ctx* ctx, void* stream, const int16_t* const* h_in, const int64_t* h_lengths,
                             int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                             float* const* h_mixed, void* d_workspace, size_t workspace_bytes);
/* float32 = int16 / 32768 for `count` samples (other rate pairs, or no resampling at all) */
int afe_pcm16_to_float(afe_ctx* ctx, void* stream, const int16_t* d_in, int64_t count, float* d_out);

/* ---- audio decode + staging ------------------------------------------ */
/* Sample encodings of a RIFF/WAVE data chunk (little endian, frames interleaved by channel). */
typedef enum afe_wav_format {
  AFE_WAV_U8 = 1, AFE_WAV_S16 = 2, AFE_WAV_S24 = 3, AFE_WAV_S32 = 4, AFE_WAV_F32 = 5, AFE_WAV_F64 = 6
} afe_wav_format;

typedef struct afe_wav_info {
  int32_t format;           /* afe_wav_format */
  int32_t channels;
  int32_t sample_rate;
  int32_t bits_per_sample;
  int32_t block_align;      /* bytes per frame */
  int32_t reserved;
  int64_t data_offset;      /* byte offset of the first frame */
  int64_t data_bytes;       /* bytes of frames present (clamped to the file, whole frames only) */
  int64_t num_frames;
} afe_wav_info;

/* Header walk of a RIFF/WAVE file (host only, integers only): what torchaudio.info reports to
 * AudioFileScanner._validate_file (services/pipeline.py:84-108) and what torchaudio.load needs
 * to find the samples (AudioProcessor.load_audio, services/audio_processor.py:72-93).  `bytes`
 * holds the first `nbytes` of a file that is `total_bytes` long (total_bytes >= nbytes; a
 * header-only read passes the file size).  PCM 8/16/24/32-bit, IEEE float 32/64,
 * WAVE_FORMAT_EXTENSIBLE; chunks before "data" (LIST, fact, ...) are skipped; a data size of
 * 0 or 0xFFFFFFFF (ffmpeg writing to a pipe -- the reference's .ogg path) means "to the end".
 * AFE_ERR_INVALID: not RIFF/WAVE, or the header does not end within nbytes;
 * AFE_ERR_UNSUPPORTED: a compressed format tag. */
int afe_wav_parse(const void* bytes, size_t nbytes, int64_t total_bytes, afe_wav_info* out);

/* torchaudio.load's (waveform [channels, frames] float32 in [-1, 1), normalize=True) from the
 * interleaved frames of a data chunk that already lie in device memory: u8 -> (x-128)/128,
 * s16 -> x/32768, s24 -> x/2^23, s32 -> float(x)/2^31, f32 -> x, f64 -> float(x).
 * d_frames needs no alignment beyond 1 byte. */
int afe_wav_decode(afe_ctx* ctx, void* stream, const void* d_frames, int format, int channels,
                   int64_t num_frames, float* d_out /*[channels][num_frames]*/);

/* afe_resample_batch_pcm16 reading the frames as they lie in a 16-bit wav file: h_in[c] is a
 * DEVICE pointer to h_lengths[c] interleaved frames of `rows` int16 samples (L0 R0 L1 R1 ...).
 * De-interleaving, x / 32768, the 1:2 resampling and the stereo mean happen in one pass: the
 * 8 kHz float32 waveform of load_audio is never materialised.  Only the 1:2 rate pair. */
int afe_resample_batch_wav16(afe_ctx* ctx, void* stream, const int16_t* const* h_in, const int64_t* h_lengths,
                             int ncalls, int rows, int orig_hz, int new_hz, float* const* h_out,
                             float* const* h_mixed, void* d_workspace, size_t workspace_bytes);

/* mixed = (row0 + row1) * 0.5 for waveforms that need no resampling (split_audio :123-131) */
int afe_mix_stereo(afe_ctx* ctx, void* stream, const float* d_in, int64_t length, float* d_mixed);

/* processor(list_of_waveforms, sampling_rate=16000, return_tensors="pt", padding=True):
 * SeamlessM4TFeatureExtractor.__call__ (services/audio_processor.py:253-261,
 * services/speech_segment.py:226-231).  h_waves[b] is a DEVICE pointer to waveform b
 * (h_lengths[b] float32 samples).
 *
 * group_size > 0 featurises several processor() calls in one launch: consecutive groups of
 * group_size waveforms (the reference's batches of `batch_size` chunks,
 * audio_processor.py:176-181) are each padded to their OWN longest member.  group_size <= 0
 * means one group.  Group g occupies rows_g = n_g * Tp_g consecutive frame rows; d_features
 * is (total_rows/2, 160) float32 and d_mask (total_rows/2) int32, i.e. the groups'
 * (n_g, Tp_g/2, 160) and (n_g, Tp_g/2) tensors laid end to end.  afe_fbank_layout reports
 * Tp_g per group (h_group_padded_frames may be NULL) and total_rows. */
int afe_fbank_layout(const int64_t* h_lengths, int batch, int group_size,
                     int64_t* h_group_padded_frames, int64_t* total_rows);
size_t afe_fbank_workspace_bytes(const int64_t* h_lengths, int batch);
int afe_fbank_batch(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                    int batch, int group_size, float padding_value, int precision, int flags,
                    float* d_features, int32_t* d_mask, void* d_workspace, size_t workspace_bytes);

/* afe_fbank_batch with explicit placement: utterance b is written at frame row
 * h_out_rows[b] (even) of d_features / mask row h_out_rows[b]/2 and padded to
 * h_padded_frames[b] frames (even, >= its own frame count).  Lets a caller featurise the
 * members of the reference's length-sorted batches in any order and in several launches
 * (e.g. per file, overlapped with the upload of the next file) into ONE output laid out as
 * afe_fbank_layout describes.  Workspace: afe_fbank_workspace_bytes(h_lengths, batch). */
int afe_fbank_batch_placed(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                           int batch, const int64_t* h_out_rows, const int64_t* h_padded_frames,
                           float padding_value, int precision, int flags, float* d_features, int32_t* d_mask,
                           void* d_workspace, size_t workspace_bytes);

/* afe_fbank_batch for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as afe_fbank_batch minus the workspace: the scratch (descriptors,
 * tile partials, utterance statistics) is a static device array of AFE_FBANK_STATIC_SCRATCH_BYTES inside the
 * library (every GPU that loads the library holds it until exit), its device address taken once per device in
 * afe_create.  The call performs no cudaMalloc / cudaFree / pinned or event allocation (keep afe_profile_enable
 * off: profiling creates events).  Limits: batch <= AFE_FBANK_STATIC_MAX_BATCH and
 * afe_fbank_workspace_bytes(h_lengths, batch) <= AFE_FBANK_STATIC_SCRATCH_BYTES (about 80 full 25 s chunks);
 * beyond either it returns AFE_ERR_WORKSPACE with a message naming the limit.  All calls on a device share the one
 * array: calls that overlap on the GPU (two streams, two threads) race on it -- serialise them on one stream. */
#define AFE_FBANK_STATIC_SCRATCH_BYTES (8u << 20)
#define AFE_FBANK_STATIC_MAX_BATCH 256
int afe_fbank_batch_static(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                           int batch, int group_size, float padding_value, int precision, int flags,
                           float* d_features, int32_t* d_mask);

/* AcousticSentimentAnalyzer._extract_acoustic_features for a batch of mono 16 kHz
 * waveforms (services/sentiment_analysis.py:807-894).  d_out is (batch, 38) float32 in
 * the reference's key order, with slot 11 ("tempo") holding the spectral-flux energy
 * flag (> 0 -> tempo = 60 / ((F-2) * hop / sr), else 0; the host finishes it in float64);
 * d_status[b] = 0 ok, 1 = "Audio too short" (reference returns None). */
size_t afe_acoustic_workspace_bytes(const int64_t* h_lengths, int batch);
int afe_acoustic_batch(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                       int batch, int sample_rate, float* d_out, int32_t* d_status,
                       void* d_workspace, size_t workspace_bytes);

/* the same, writing waveform b to row h_out_index[b] of d_out / d_status (NULL: row b) */
int afe_acoustic_batch_placed(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                              int batch, int sample_rate, const int32_t* h_out_index, float* d_out,
                              int32_t* d_status, void* d_workspace, size_t workspace_bytes);

/* The same with the StandardScaler step of AcousticSentimentAnalyzer.analyze_batch_sentiment
 * (services/sentiment_analysis.py:736-752) as an epilogue of the last kernel: besides d_out,
 * row b of d_scaled (float64, [rows][n_features]) receives (x - d_mean) / d_scale of the feature
 * LIST the reference builds for scaler.transform -- the 38 values in key order with tempo
 * finished (float64, as Python computes it), all zeros for a waveform whose extraction failed
 * (status 1 -> _get_zero_features), padded with zeros / truncated to n_features
 * (scaler.n_features_in_).  Same float64 operations as numpy's (X - mean_) / scale_. */
int afe_acoustic_batch_scaled(afe_ctx* ctx, void* stream, const float* const* h_waves, const int64_t* h_lengths,
                              int batch, int sample_rate, const int32_t* h_out_index, float* d_out,
                              int32_t* d_status, const double* d_mean, const double* d_scale, int n_features,
                              double* d_scaled, void* d_workspace, size_t workspace_bytes);

/* ---- per-kernel device timing (CUDA events on the caller's stream) --------------------
 * While enabled every kernel launch of this context is bracketed by cudaEvents.
 * afe_profile_read synchronises those events and returns, per kernel id, the summed device
 * time in milliseconds and the number of launches since the last reset. */
enum {
  AFE_K_RESAMPLE = 0, AFE_K_MIX = 1, AFE_K_FBANK = 2, AFE_K_FBANK_STATS = 3, AFE_K_FBANK_NORM = 4,
  AFE_K_AC_TILE = 5, AFE_K_AC_MAX = 6, AFE_K_AC_MFCC = 7, AFE_K_AC_FINAL = 8, AFE_K_COUNT = 9
};
int afe_profile_enable(afe_ctx* ctx, int enabled);
int afe_profile_reset(afe_ctx* ctx);
int afe_profile_read(afe_ctx* ctx, double* h_ms /*[AFE_K_COUNT]*/, int64_t* h_launches /*[AFE_K_COUNT]*/);
const char* afe_kernel_name(int kernel_id);

#ifdef __cplusplus
}
#endif
#endif /* AFE_H_ */
