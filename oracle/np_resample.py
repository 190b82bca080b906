"""TEST INFRASTRUCTURE -- numpy restatement of the reference's resampler.

Reference call sites: ``services/audio_processor.py:99-102`` (``get_resampler`` ->
``torchaudio.transforms.Resample(orig, new)``) and ``:108-112`` (applied to the whole
``[C, L]`` waveform).  Algorithm: torchaudio 2.x ``_get_sinc_resample_kernel`` /
``_apply_sinc_resample_kernel`` (torchaudio/functional/functional.py:1305-1432,
resampling_method="sinc_interp_hann", lowpass_filter_width=6, rolloff=0.99).

Measured in the build container: torchaudio's conv1d accumulates each output as an
in-order chain ``acc = fma(h[j,k], xp[n*O+k], acc)`` for k = 0..K-1 starting from 0, in
float32.  ``apply_taps`` restates exactly that chain and is BIT-EXACT against torchaudio
for 8 k->16 k and 44.1 k->16 k (tests/test_oracle_golden.py), so the CUDA kernel's bar
for this stage is bit-exactness, not a tolerance.
"""
from __future__ import annotations

import math

import numpy as np


def sinc_taps(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """Polyphase Hann-windowed sinc bank: float64 maths, float32 result ``(N, K)``.

    Follows functional.py:1341-1400.  Returns ``(taps, width, O, N)`` with
    ``O = orig/gcd`` (input stride), ``N = new/gcd`` (phases), ``K = 2*width + O``."""
    g = math.gcd(int(orig_freq), int(new_freq))
    O = int(orig_freq) // g
    N = int(new_freq) // g
    base = min(O, N) * rolloff
    width = math.ceil(lowpass_filter_width * O / base)
    idx = np.arange(-width, width + O, dtype=np.float64)[None, :] / O          # (1, K)
    # torch.arange(0, -N, -1) is int64 and `/ N` yields float32 (default dtype) before the
    # sum with the float64 grid promotes it (functional.py:1373): the phase is float32-rounded.
    phase = (np.arange(0, -N, -1, dtype=np.float64) / N).astype(np.float32).astype(np.float64)[:, None]
    t = (phase + idx) * base
    t = np.clip(t, -lowpass_filter_width, lowpass_filter_width)
    window = np.cos(t * math.pi / lowpass_filter_width / 2.0) ** 2
    t = t * math.pi
    scale = base / O
    with np.errstate(invalid="ignore", divide="ignore"):
        kern = np.where(t == 0.0, 1.0, np.sin(t) / t)
    kern = kern * (window * scale)
    return kern.astype(np.float32), width, O, N


def resampled_length(length: int, orig_freq: int, new_freq: int) -> int:
    """``ceil(new * L / orig)`` on the gcd-reduced rates (functional.py:1425)."""
    g = math.gcd(int(orig_freq), int(new_freq))
    O = int(orig_freq) // g
    N = int(new_freq) // g
    return int(math.ceil(N * length / O))


def _fma32(a: np.ndarray, b: np.ndarray, c: np.ndarray) -> np.ndarray:
    # float32 fma: the product of two float32 is exact in float64 (48 <= 53 bits)
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def apply_taps(x: np.ndarray, taps: np.ndarray, width: int, O: int, N: int) -> np.ndarray:
    """``y[n*N + j] = sum_k h[j,k] * xp[n*O + k]`` with ``xp = pad(x, (width, width+O))``
    (functional.py:1416-1428), accumulated as the in-order float32 FMA chain."""
    x = np.asarray(x, dtype=np.float32)
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    L = x2.shape[1]
    K = taps.shape[1]
    xp = np.pad(x2, ((0, 0), (width, width + O)))
    nper = (xp.shape[1] - K) // O + 1
    starts = np.arange(nper) * O
    out = np.zeros((x2.shape[0], nper, N), dtype=np.float32)
    for j in range(N):
        acc = np.zeros((x2.shape[0], nper), dtype=np.float32)
        for k in range(K):
            acc = _fma32(np.broadcast_to(taps[j, k], acc.shape), xp[:, starts + k], acc)
        out[:, :, j] = acc
    tgt = int(math.ceil(N * L / O))
    y = out.reshape(x2.shape[0], -1)[:, :tgt]
    return y.reshape(lead + (tgt,))


def resample(x: np.ndarray, orig_freq: int, new_freq: int) -> np.ndarray:
    """``torchaudio.transforms.Resample(orig, new)(x)`` restated (identity if equal)."""
    if int(orig_freq) == int(new_freq):
        return np.asarray(x, dtype=np.float32)
    taps, width, O, N = sinc_taps(orig_freq, new_freq)
    return apply_taps(x, taps, width, O, N)


def split_channels(y: np.ndarray):
    """services/audio_processor.py:123-131: stereo -> agent=ch0, client=ch1,
    mixed=mean over channels (float32 ``(a + c) / 2``); mono -> all three alias."""
    y = np.asarray(y, dtype=np.float32)
    if y.shape[0] == 2:
        agent = y[0:1]
        client = y[1:2]
        mixed = ((y[0] + y[1]) * np.float32(0.5))[None, :]
        return agent, client, mixed
    return y, y, y
