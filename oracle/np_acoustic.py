"""TEST INFRASTRUCTURE -- numpy (float64) restatement of the 38 acoustic features.

Reference: ``AcousticSentimentAnalyzer._extract_acoustic_features``
(services/sentiment_analysis.py:807-894) and ``_get_zero_features`` (:896-913).  Its
arithmetic is ``torch.stft`` (periodic Hann 400 centred in a 512 buffer, center=True,
reflect pad) and ``torchaudio.transforms.MFCC`` (transforms/_transforms.py:634-718 ->
MelSpectrogram :515-632, melscale_fbanks functional.py:518-588, amplitude_to_DB
:356-404 with top_db 80, create_dct :636-667), all in float32; pinned torch/torchaudio
2.6.0, installed 2.11.0.  This restatement works in float64.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np

FRAME = 400
HOP = 160
NFFT = 512
NBINS = 257
NMELS = 128
NMFCC = 13

FEATURE_KEYS = (
    ["rms_mean", "rms_std", "rms_range", "zcr_mean", "zcr_std",
     "spectral_centroid_mean", "spectral_centroid_std",
     "spectral_bandwidth_mean", "spectral_bandwidth_std",
     "spectral_rolloff_mean", "spectral_rolloff_std", "tempo"]
    + [f"mfcc_{i}_{s}" for i in range(NMFCC) for s in ("mean", "std")]
)


def zero_features() -> "OrderedDict[str, float]":
    """sentiment_analysis.py:896-913 -- 38 zeros in the contract's key order."""
    return OrderedDict((k, 0.0) for k in FEATURE_KEYS)


def hann_periodic() -> np.ndarray:
    n = np.arange(FRAME, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / FRAME)


def htk_mel_bank(sample_rate: int = 16000) -> np.ndarray:
    """(257, 128) triangles in Hz between 130 HTK-mel-spaced points, no norm
    (functional.py:518-588, _create_triangular_filterbank :490-515)."""
    all_freqs = np.linspace(0.0, sample_rate // 2, NBINS)
    m_min = 2595.0 * math.log10(1.0 + 0.0 / 700.0)
    m_max = 2595.0 * math.log10(1.0 + (sample_rate / 2.0) / 700.0)
    m_pts = np.linspace(m_min, m_max, NMELS + 2)
    f_pts = 700.0 * (10.0 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    return np.maximum(0.0, np.minimum(down, up))


def dct_matrix() -> np.ndarray:
    """(128, 13) ortho DCT-II (functional.py:636-667)."""
    n = np.arange(NMELS, dtype=np.float64)
    k = np.arange(NMFCC, dtype=np.float64)[:, None]
    dct = np.cos(math.pi / NMELS * (n + 0.5) * k)
    dct[0] *= 1.0 / math.sqrt(2.0)
    dct *= math.sqrt(2.0 / NMELS)
    return dct.T


_HANN = hann_periodic()
_FB = htk_mel_bank()
_DCT = dct_matrix()


def _std(v: np.ndarray) -> float:
    """torch.std default: unbiased; n == 1 -> NaN."""
    n = v.shape[-1]
    if n < 2:
        return float("nan")
    return float(np.sqrt(np.sum((v - v.mean()) ** 2) / (n - 1)))


def stft_mag(y: np.ndarray) -> np.ndarray:
    """|STFT| (257, 1 + L//160): reflect-pad 256, frames of 512 at hop 160, the 400-tap
    periodic Hann centred at offset 56 (torch.stft, sentiment_analysis.py:835-840)."""
    yp = np.pad(y, (NFFT // 2, NFFT // 2), mode="reflect")
    F2 = 1 + (yp.size - NFFT) // HOP
    idx = HOP * np.arange(F2)[:, None] + np.arange(NFFT)[None, :]
    win = np.zeros(NFFT)
    off = (NFFT - FRAME) // 2
    win[off:off + FRAME] = _HANN
    return np.abs(np.fft.rfft(yp[idx] * win[None, :], axis=1)).T


def acoustic_features(waveform: np.ndarray, sample_rate: int = 16000):
    """The 38-feature ordered dict, or ``None`` when the reference raises (L < 400)."""
    y = np.asarray(waveform, dtype=np.float32).reshape(-1).astype(np.float64)
    sr = sample_rate
    fl, hop = int(0.025 * sr), int(0.010 * sr)
    if (fl, hop, sr) != (FRAME, HOP, 16000):
        raise ValueError("oracle restates the 16 kHz configuration only")
    if y.size < fl:
        return None                                            # 'Audio too short' (:820-821)
    F1 = (y.size - fl) // hop + 1
    frames = y[hop * np.arange(F1)[:, None] + np.arange(fl)[None, :]]
    rms = np.sqrt(np.mean(frames ** 2, axis=1))
    sg = np.sign(frames)
    zc = ((sg[:, :-1] * sg[:, 1:]) < 0).sum(axis=1).astype(np.float64) / fl

    mag = stft_mag(y)                                          # (257, F2)
    freqs = np.linspace(0.0, sr / 2, NBINS)
    energy = mag.sum(axis=0) + 1e-8
    centroid = (mag * freqs[:, None]).sum(axis=0) / energy
    bandwidth = np.sqrt((mag * (freqs[:, None] - centroid[None, :]) ** 2).sum(axis=0) / energy)
    cum = np.cumsum(mag, axis=0)
    thr = 0.85 * (cum[-1] + 1e-8)
    roll_idx = np.argmax(cum >= thr[None, :], axis=0)          # 0 when no bin qualifies
    roll = freqs[roll_idx]

    mel = (mag ** 2).T @ _FB                                   # (F2, 128)
    db = 10.0 * np.log10(np.maximum(mel, 1e-10))
    db = np.maximum(db, db.max() - 80.0)
    mfcc = (db @ _DCT).T                                       # (13, F2)

    flux = np.maximum(mag[:, 1:] - mag[:, :-1], 0.0).sum(axis=0)
    Tf = flux.size
    # full cross-correlation, index 0 (lag -(Tf-1)) zeroed -- NOT lag 0 (:871-872)
    if Tf <= 8192:
        corr = np.correlate(flux, flux, mode="full")
        corr[0] = 0.0
        lag = int(np.argmax(corr))
    else:
        # O(T^2) is impractical for long-form input; by Cauchy-Schwarz the zero-lag term
        # (index Tf-1) is the strict maximum unless flux is identically zero.
        lag = Tf - 1 if float(np.sum(flux * flux)) > 0.0 else 0
    period = lag * hop / sr if lag > 0 else 0.0
    tempo = 60.0 / period if period > 0 else 0.0

    out = OrderedDict()
    out["rms_mean"] = float(rms.mean()); out["rms_std"] = _std(rms)
    out["rms_range"] = float(rms.max() - rms.min())
    out["zcr_mean"] = float(zc.mean()); out["zcr_std"] = _std(zc)
    out["spectral_centroid_mean"] = float(centroid.mean()); out["spectral_centroid_std"] = _std(centroid)
    out["spectral_bandwidth_mean"] = float(bandwidth.mean()); out["spectral_bandwidth_std"] = _std(bandwidth)
    out["spectral_rolloff_mean"] = float(roll.mean()); out["spectral_rolloff_std"] = _std(roll)
    out["tempo"] = tempo
    for i in range(NMFCC):
        out[f"mfcc_{i}_mean"] = float(mfcc[i].mean())
        out[f"mfcc_{i}_std"] = _std(mfcc[i])
    return out
