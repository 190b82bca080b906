// afe_tables.h -- constant tables of the front end, built on the host in float64.
//
// Sources restated (never copied): torchaudio functional.py:1341-1400 (sinc taps),
// transformers audio_utils.py:453-544/593-602 (Kaldi mel bank, Povey window),
// torchaudio functional.py:518-588/636-667 (HTK mel bank, DCT-II), torch.hann_window.
#pragma once
#include <cstdint>
#include <vector>

namespace afe {

constexpr int kFrame = 400;       // 25 ms @ 16 kHz
constexpr int kHop = 160;         // 10 ms
constexpr int kNfft = 512;
constexpr int kBins = 257;
constexpr int kMel = 80;          // fbank mel bins
constexpr int kAcMel = 128;       // MFCC mel bins
constexpr int kMfcc = 13;
constexpr int kAcFeat = 38;
constexpr float kMelFloor = 1.192092955078125e-07f;

struct ResampleTaps {
  int orig = 0, target = 0;       // requested rates
  int O = 1, N = 1;               // gcd-reduced input stride / number of phases
  int width = 0, K = 0;           // K = 2*width + O taps per phase
  std::vector<float> taps;        // [N][K]   (reference layout)
  std::vector<float> tapsT;       // [K][N]   (phase-fastest, for coalesced loads)
};
ResampleTaps make_resample_taps(int orig_hz, int new_hz);
int64_t resampled_length(int64_t len, int orig_hz, int new_hz);

// sparse triangular filter bank: filter m covers bins [start[m], start[m]+count[m]) with
// weights w[offset[m] + i]
struct SparseBank {
  int filters = 0;
  std::vector<int32_t> start, count, offset;
  std::vector<float> w;
  std::vector<double> dense;      // [bins][filters] float64 (tests)
  // interval form: the F+2 edge points cut the axis into F+1 intervals; a bin in interval i
  // feeds the UP slope of filter i (i < F) and the DOWN slope of filter i-1 (i >= 1), so
  //   filter m = U[m] + D[m+1],  U[i] = sum_{k in i} up[k] P[k],  D[i] = sum_{k in i} dn[k] P[k].
  std::vector<float> up, dn;              // [bins]
  std::vector<int32_t> iv_start, iv_count;   // [F+1] bins of interval i (contiguous, may be empty)
  std::vector<int32_t> role_iv;           // [17] intervals [role_iv[r], role_iv[r+1]) of role r, balanced by bins
};
SparseBank make_kaldi_mel_bank();        // 80 filters, 20..8000 Hz, mel-space triangles
SparseBank make_htk_mel_bank();          // 128 filters, 0..8000 Hz, Hz-space triangles (float32 like torchaudio)

// Slot form for the device (afe_mel_slots.h): filter (role r, slot s) = 16 s + (s even ? r : 15 - r); slot s is
// padded to slot_len[s] bins (zero weights; the first bin moves down where the padding would pass bin 256).
// w[16][kMelSlotRow] rows of weights, slots back to back; start[16][kMelSlotStarts] first bin of every slot.
// ok = false when a filter is wider than its slot (the kernels' slot lengths are compile-time constants).
constexpr int kMelSlotRow = 40;          // floats per role (both banks: 4+4+6+10+16 = 40, 2+2+2+3+4+6+8+11 = 38)
constexpr int kMelSlotStarts = 8;
constexpr int kKaldiSlotLen[5] = {4, 4, 6, 10, 16};
constexpr int kHtkSlotLen[8] = {2, 2, 2, 3, 4, 6, 8, 11};
struct MelSlots {
  std::vector<float> w;
  std::vector<int32_t> start;
  bool ok = false;
};
MelSlots make_mel_slots(const SparseBank& bank, const int* slot_len, int nslots);

std::vector<double> make_povey_window();           // 400, symmetric hann^0.85
std::vector<double> make_hann_periodic();          // 400
std::vector<double> make_twiddle512();             // [16][16][2]: e^{-2 pi i n2 k1 / 512}
std::vector<float> make_dct_matrix();              // [13][128] ortho DCT-II rows

}  // namespace afe
