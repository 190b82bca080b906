"""TEST INFRASTRUCTURE -- numpy (float64) restatement of the reference's fbank featuriser.

Reference call sites: ``services/audio_processor.py:253-261`` and
``services/speech_segment.py:226-231,264`` -- ``processor(arrays, sampling_rate=16000,
return_tensors="pt", padding=True)``.  The arithmetic is transformers'
``SeamlessM4TFeatureExtractor`` (models/seamless_m4t/feature_extraction_seamless_m4t.py
:60-87 constants, :112-139 fbank, :243-302 normalise/pad/stack/mask) on top of
``audio_utils.spectrogram`` (:624-832), ``mel_filter_bank`` (:453-544) and
``window_function`` (:560-620); padding is ``SequenceFeatureExtractor.pad``
(feature_extraction_sequence_utils.py:196-291).  Pinned version 4.48.0, installed 5.5.0.

Written from the behaviour spec in SURVEY.md Appendix A.2, vectorised over frames.
"""
from __future__ import annotations

import numpy as np

FRAME = 400
HOP = 160
NFFT = 512
NBINS = 257
NMEL = 80
MEL_FLOOR = 1.192092955078125e-07
PREEMPH = 0.97
SAMPLE_RATE = 16000


def povey_window() -> np.ndarray:
    """Symmetric Hann(400)^0.85 in float64 (audio_utils.py:593-602, periodic=False)."""
    n = np.arange(FRAME, dtype=np.float64)
    hann = 0.5 - 0.5 * np.cos(2.0 * np.pi * n / (FRAME - 1))
    return np.power(hann, 0.85)


def kaldi_mel_bank() -> np.ndarray:
    """(257, 80) float64 triangles, Kaldi mel scale, triangularised in mel space,
    20..8000 Hz, no area normalisation (audio_utils.py:453-544 as configured at
    feature_extraction_seamless_m4t.py:73-82)."""
    def mel(f):
        return 1127.0 * np.log(1.0 + f / 700.0)
    pts = np.linspace(mel(20.0), mel(SAMPLE_RATE / 2.0), NMEL + 2)
    bin_mel = mel((SAMPLE_RATE / ((NBINS - 1) * 2.0)) * np.arange(NBINS, dtype=np.float64))
    left, centre, right = pts[:-2], pts[1:-1], pts[2:]
    up = (bin_mel[:, None] - left[None, :]) / (centre - left)[None, :]
    down = (right[None, :] - bin_mel[:, None]) / (right - centre)[None, :]
    return np.maximum(0.0, np.minimum(up, down))


_WINDOW = povey_window()
_MEL = kaldi_mel_bank()


def num_frames(length: int) -> int:
    """``1 + floor((L - 400) / 160)`` clamped at 0 (audio_utils.py:778)."""
    return max(0, 1 + (int(length) - FRAME) // HOP) if length >= FRAME else 0


def raw_log_mel(waveform: np.ndarray) -> np.ndarray:
    """Un-normalised log-mel ``(T, 80)`` float32 of one mono float32 waveform."""
    x = np.asarray(waveform)
    if x.ndim == 2:                       # stereo arrays: channel 0 only (:121-122)
        x = x[0]
    x = np.squeeze(x).astype(np.float32) * np.float32(32768.0)   # float32 scale (:124)
    x = x.astype(np.float64)
    T = num_frames(x.size)
    if T <= 0:
        return np.zeros((NMEL, 0), dtype=np.float32).T
    idx = HOP * np.arange(T)[:, None] + np.arange(FRAME)[None, :]
    fr = x[idx]
    fr = fr - fr.mean(axis=1, keepdims=True)                     # remove_dc_offset
    pe = np.empty_like(fr)
    pe[:, 1:] = fr[:, 1:] - PREEMPH * fr[:, :-1]                 # pre-update neighbours
    pe[:, 0] = fr[:, 0] * (1.0 - PREEMPH)
    pe *= _WINDOW[None, :]
    spec = np.fft.rfft(pe, n=NFFT, axis=1).astype(np.complex64)  # stored as complex64 (:781)
    power = np.abs(spec, dtype=np.float64) ** 2                  # (:808)
    melspec = np.maximum(MEL_FLOOR, _MEL.T @ power.T)            # (80, T) like the reference
    # the reference returns ``spectrogram(...).T``: a (T, 80) VIEW of a C-ordered (80, T)
    # float32 array.  The layout matters: numpy's float32 mean/var over axis 0 then runs
    # pairwise along contiguous memory, and for zero-variance bins (digital silence) the
    # normalised value is that rounding noise divided by sqrt(1e-7).
    return np.log(melspec).astype(np.float32).T


def normalise(feat: np.ndarray) -> np.ndarray:
    """Per-mel-bin ``(x - mean) / sqrt(var_ddof1 + 1e-7)`` over the utterance's own frames,
    in numpy float32 like the reference (:257-262); T==1 gives NaN (0/0 variance)."""
    with np.errstate(invalid="ignore", divide="ignore"):
        mean = np.expand_dims(feat.mean(0), 0)
        var = np.expand_dims(feat.var(0, ddof=1), 0)
        return (feat - mean) / np.sqrt(var + 1e-7)


def padded_frames(lengths) -> int:
    """Padded frame count of a batch: longest T rounded UP to a multiple of 2."""
    tmax = max([num_frames(int(n)) for n in lengths] + [0])
    return tmax + (tmax % 2)


def featurise(waveforms, padding_value: float = 0.0, sampling_rate: int = 16000):
    """The whole ``processor(...)`` call for a list of mono waveforms.

    Returns ``(input_features float32 (B, Tp/2, 160), attention_mask int32 (B, Tp/2))``."""
    if sampling_rate != SAMPLE_RATE:
        raise ValueError("sampling_rate must be 16000")       # (:215-221)
    if isinstance(waveforms, np.ndarray) and waveforms.ndim == 1:
        waveforms = [waveforms]
    feats = [normalise(raw_log_mel(w)) for w in waveforms]
    B = len(feats)
    Tp = max([f.shape[0] for f in feats] + [0])
    Tp += Tp % 2
    out = np.full((B, Tp, NMEL), padding_value, dtype=np.float32)
    mask = np.zeros((B, Tp), dtype=np.int32)
    for b, f in enumerate(feats):
        out[b, : f.shape[0]] = f
        mask[b, : f.shape[0]] = 1
    out = out.reshape(B, Tp // 2, 2 * NMEL)
    mask = mask[:, 1::2]                                       # the odd frame's bit (:292-293)
    return out, mask
