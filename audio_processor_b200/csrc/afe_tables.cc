// afe_tables.cc -- see afe_tables.h.  Plain host C++ (built by nvcc and by g++).
#include "afe_tables.h"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace afe {

static const double kPi = 3.14159265358979323846264338327950288;

int64_t resampled_length(int64_t len, int orig_hz, int new_hz) {
  const int g = std::gcd(orig_hz, new_hz);
  const int64_t O = orig_hz / g, N = new_hz / g;
  // torch.ceil(torch.as_tensor(new * L / orig)) -- Python true division in float64
  return (int64_t)std::ceil((double)(N * len) / (double)O);
}

ResampleTaps make_resample_taps(int orig_hz, int new_hz) {
  ResampleTaps r;
  r.orig = orig_hz; r.target = new_hz;
  const int g = std::gcd(orig_hz, new_hz);
  r.O = orig_hz / g; r.N = new_hz / g;
  const double lowpass_width = 6.0, rolloff = 0.99;
  const double base = std::min(r.O, r.N) * rolloff;
  r.width = (int)std::ceil(lowpass_width * r.O / base);
  r.K = 2 * r.width + r.O;
  r.taps.assign((size_t)r.N * r.K, 0.f);
  r.tapsT.assign((size_t)r.N * r.K, 0.f);
  const double scale = base / r.O;
  for (int j = 0; j < r.N; ++j) {
    // the reference forms -j/N in float32 (int tensor / int) before promoting to float64
    const double phase = (double)(float)((double)(-j) / (double)r.N);
    for (int k = 0; k < r.K; ++k) {
      const double idx = (double)(k - r.width) / (double)r.O;
      double t = (phase + idx) * base;
      t = std::min(std::max(t, -lowpass_width), lowpass_width);
      const double cw = std::cos(t * kPi / lowpass_width / 2.0);
      const double window = cw * cw;
      t *= kPi;
      const double sinc = (t == 0.0) ? 1.0 : std::sin(t) / t;
      const float v = (float)(sinc * (window * scale));
      r.taps[(size_t)j * r.K + k] = v;
      r.tapsT[(size_t)k * r.N + j] = v;
    }
  }
  return r;
}

static void sparsify(const std::vector<double>& dense, int bins, int filters, SparseBank& b) {
  b.filters = filters;
  b.dense = dense;
  b.start.assign(filters, 0); b.count.assign(filters, 0); b.offset.assign(filters, 0);
  b.w.clear();
  for (int m = 0; m < filters; ++m) {
    int lo = -1, hi = -1;
    for (int k = 0; k < bins; ++k)
      if (dense[(size_t)k * filters + m] != 0.0) { if (lo < 0) lo = k; hi = k; }
    b.offset[m] = (int32_t)b.w.size();
    if (lo < 0) { b.start[m] = 0; b.count[m] = 0; continue; }     // all-zero filter (MFCC low bands)
    b.start[m] = lo; b.count[m] = hi - lo + 1;
    for (int k = lo; k <= hi; ++k) b.w.push_back((float)dense[(size_t)k * filters + m]);
  }
}

// interval form from the edge points `pts` (F+2, ascending) and the bin coordinates `x`
static void intervalise(const std::vector<double>& pts, const std::vector<double>& x, SparseBank& b) {
  const int F = b.filters, bins = (int)x.size();
  b.up.assign(bins, 0.f); b.dn.assign(bins, 0.f);
  b.iv_start.assign(F + 1, 0); b.iv_count.assign(F + 1, 0);
  std::vector<int> iv(bins, 0);
  for (int k = 0; k < bins; ++k) {
    int i = 0;
    while (i < F && x[k] >= pts[i + 1]) ++i;          // pts[i] <= x < pts[i+1]  (clamped to 0..F)
    iv[k] = i;
    if (i < F) b.up[k] = (float)b.dense[(size_t)k * F + i];
    if (i >= 1) b.dn[k] = (float)b.dense[(size_t)k * F + i - 1];
  }
  for (int i = 0, k = 0; i <= F; ++i) {                 // x ascending -> intervals are contiguous bin runs
    b.iv_start[i] = k;
    while (k < bins && iv[k] == i) ++k;
    b.iv_count[i] = k - b.iv_start[i];
  }
  // 16 contiguous interval ranges with balanced cost (1 unit per bin + 1 per interval)
  b.role_iv.assign(17, F + 1);
  b.role_iv[0] = 0;
  const double total = bins + (F + 1);
  double acc = 0.0;
  int r = 1;
  for (int i = 0; i <= F && r < 16; ++i) {
    acc += b.iv_count[i] + 1;
    if (acc >= total * r / 16.0) b.role_iv[r++] = i + 1;
  }
  for (; r < 16; ++r) b.role_iv[r] = F + 1;
}

SparseBank make_kaldi_mel_bank() {
  auto mel = [](double f) { return 1127.0 * std::log(1.0 + f / 700.0); };
  const int F = kMel;
  const double lo = mel(20.0), hi = mel(8000.0);
  std::vector<double> pts(F + 2);
  const double step = (hi - lo) / (double)(F + 1);          // np.linspace: arange * step + start
  for (int i = 0; i < F + 2; ++i) pts[i] = (double)i * step + lo;
  pts[F + 1] = hi;
  std::vector<double> dense((size_t)kBins * F, 0.0);
  const double bin_hz = 16000.0 / ((kBins - 1) * 2.0);
  for (int k = 0; k < kBins; ++k) {
    const double mk = mel(bin_hz * k);
    for (int m = 0; m < F; ++m) {
      const double up = (mk - pts[m]) / (pts[m + 1] - pts[m]);
      const double down = (pts[m + 2] - mk) / (pts[m + 2] - pts[m + 1]);
      dense[(size_t)k * F + m] = std::max(0.0, std::min(up, down));
    }
  }
  SparseBank b; sparsify(dense, kBins, F, b);
  std::vector<double> x(kBins);
  for (int k = 0; k < kBins; ++k) x[k] = mel(bin_hz * k);
  intervalise(pts, x, b);
  return b;
}

SparseBank make_htk_mel_bank() {
  const int F = kAcMel;
  const double m_min = 2595.0 * std::log10(1.0 + 0.0 / 700.0);
  const double m_max = 2595.0 * std::log10(1.0 + 8000.0 / 700.0);
  std::vector<double> fp(F + 2);
  for (int i = 0; i < F + 2; ++i) {
    const double m = m_min + (m_max - m_min) * (double)i / (double)(F + 1);
    fp[i] = 700.0 * (std::pow(10.0, m / 2595.0) - 1.0);
  }
  std::vector<double> dense((size_t)kBins * F, 0.0);
  for (int k = 0; k < kBins; ++k) {
    const double f = 8000.0 * (double)k / (double)(kBins - 1);
    for (int m = 0; m < F; ++m) {
      const double down = -(fp[m] - f) / (fp[m + 1] - fp[m]);
      const double up = (fp[m + 2] - f) / (fp[m + 2] - fp[m + 1]);
      dense[(size_t)k * F + m] = std::max(0.0, std::min(down, up));
    }
  }
  SparseBank b; sparsify(dense, kBins, F, b);
  std::vector<double> x(kBins);
  for (int k = 0; k < kBins; ++k) x[k] = 8000.0 * (double)k / (double)(kBins - 1);
  intervalise(fp, x, b);
  return b;
}

std::vector<double> make_povey_window() {
  std::vector<double> w(kFrame);
  for (int n = 0; n < kFrame; ++n) {
    const double h = 0.5 - 0.5 * std::cos(2.0 * kPi * n / (kFrame - 1));
    w[n] = std::pow(h, 0.85);
  }
  return w;
}

std::vector<double> make_hann_periodic() {
  std::vector<double> w(kFrame);
  for (int n = 0; n < kFrame; ++n) w[n] = 0.5 - 0.5 * std::cos(2.0 * kPi * n / kFrame);
  return w;
}

std::vector<double> make_twiddle512() {
  std::vector<double> t(16 * 16 * 2);
  for (int n2 = 0; n2 < 16; ++n2)
    for (int k1 = 0; k1 < 16; ++k1) {
      const double a = -2.0 * kPi * (double)(n2 * k1) / 512.0;
      t[(n2 * 16 + k1) * 2 + 0] = std::cos(a);
      t[(n2 * 16 + k1) * 2 + 1] = std::sin(a);
    }
  return t;
}

std::vector<float> make_dct_matrix() {
  std::vector<float> d((size_t)kMfcc * kAcMel);
  for (int k = 0; k < kMfcc; ++k)
    for (int n = 0; n < kAcMel; ++n) {
      double v = std::cos(kPi / kAcMel * (n + 0.5) * k);
      if (k == 0) v *= 1.0 / std::sqrt(2.0);
      v *= std::sqrt(2.0 / kAcMel);
      d[(size_t)k * kAcMel + n] = (float)v;
    }
  return d;
}


MelSlots make_mel_slots(const SparseBank& b, const int* slot_len, int nslots) {
  MelSlots out;
  out.w.assign((size_t)16 * kMelSlotRow, 0.f);
  out.start.assign((size_t)16 * kMelSlotStarts, 0);
  out.ok = (b.filters == 16 * nslots) && nslots <= kMelSlotStarts;
  if (!out.ok) return out;
  for (int r = 0; r < 16; ++r) {
    int off = 0;
    for (int s = 0; s < nslots; ++s) {
      const int L = slot_len[s];
      const int m = 16 * s + ((s & 1) ? 15 - r : r);
      const int count = b.count[(size_t)m];
      if (count > L || off + L > kMelSlotRow) { out.ok = false; return out; }
      int first = b.start[(size_t)m];
      if (first + L > kBins) first = kBins - L;               // keep the padded reads inside the spectrum
      const int shift = b.start[(size_t)m] - first;
      for (int i = 0; i < count; ++i) out.w[(size_t)r * kMelSlotRow + off + shift + i] = b.w[(size_t)b.offset[(size_t)m] + i];
      out.start[(size_t)r * kMelSlotStarts + s] = first;
      off += L;
    }
  }
  return out;
}

}  // namespace afe
