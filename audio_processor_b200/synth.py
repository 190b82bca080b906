"""Seeded synthetic waveforms for the five BASELINE.json configs.

Every generator is deterministic in (seed, length, sample_rate) and runs on the CPU with a
``torch.Generator`` so the CPU oracle and the CUDA path see bit-identical float32 samples.

Two value distributions:

* ``speech_like`` -- sum of 5 sinusoids (100-3400 Hz, random phase), each amplitude-modulated
  at 2-6 Hz, plus Gaussian noise (sigma 0.01), peak-normalised to 0.5, with ~20 % of the
  signal hard-zeroed in 0.3-2 s gaps (exercises mel_floor / amin / silent-frame rolloff).
* ``white`` -- U(-0.5, 0.5), the distribution the survey-time CPU probe used.
"""
from __future__ import annotations

import math
from typing import List, Tuple

import torch

__all__ = [
    "speech_like", "white", "make_call", "cfg1_call", "cfg2_calls", "cfg3_segments",
    "cfg5_longform", "CFG3_EDGE_LENGTHS",
]

CFG3_EDGE_LENGTHS = (8000, 8001, 16000, 479999, 480000)


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def speech_like(n: int, sample_rate: int, seed: int) -> torch.Tensor:
    """One mono speech-like waveform, float32, shape ``(n,)``."""
    g = _gen(seed)
    p = torch.rand(5, 4, generator=g, dtype=torch.float64)
    t = torch.arange(n, dtype=torch.float64) / float(sample_rate)
    x = torch.zeros(n, dtype=torch.float64)
    for i in range(5):
        f = 100.0 + 3300.0 * float(p[i, 0])
        ph = 2.0 * math.pi * float(p[i, 1])
        fm = 2.0 + 4.0 * float(p[i, 2])
        pm = 2.0 * math.pi * float(p[i, 3])
        x += torch.sin(2.0 * math.pi * f * t + ph) * (0.5 + 0.5 * torch.sin(2.0 * math.pi * fm * t + pm))
    x += 0.01 * torch.randn(n, generator=g, dtype=torch.float64)
    peak = float(x.abs().max())
    if peak > 0.0:
        x *= 0.5 / peak
    # hard-zero ~20 % of the signal in 0.3-2 s gaps
    target = int(0.2 * n)
    zeroed = 0
    guard = 0
    while zeroed < target and guard < 10000:
        guard += 1
        u = torch.rand(2, generator=g, dtype=torch.float64)
        glen = int((0.3 + 1.7 * float(u[0])) * sample_rate)
        glen = max(1, min(glen, n // 5))      # short signals keep most of their content
        start = int(float(u[1]) * max(1, n - glen))
        x[start:start + glen] = 0.0
        zeroed += glen
    return x.to(torch.float32)


def white(n: int, seed: int) -> torch.Tensor:
    """U(-0.5, 0.5) float32 noise, shape ``(n,)``."""
    g = _gen(seed)
    return torch.rand(n, generator=g, dtype=torch.float32) - 0.5


def make_call(seconds: float, sample_rate: int, seed: int, channels: int = 2,
              kind: str = "speech_like") -> torch.Tensor:
    """A synthetic call ``(channels, n)``; agent (ch0) and client (ch1) are independent."""
    n = int(round(seconds * sample_rate))
    chans = []
    for c in range(channels):
        s = seed * 7919 + c
        chans.append(speech_like(n, sample_rate, s) if kind == "speech_like" else white(n, s))
    return torch.stack(chans, 0).contiguous()


def cfg1_call(kind: str = "speech_like") -> Tuple[torch.Tensor, int]:
    """cfg1: one 60 s stereo 8 kHz call, seed 0."""
    return make_call(60.0, 8000, 0, 2, kind), 8000


def cfg2_calls(n_calls: int = 64, seconds: float = 300.0, kind: str = "speech_like",
               distinct: int = 0) -> Tuple[List[torch.Tensor], int]:
    """cfg2: ``n_calls`` stereo 8 kHz 5-minute calls, seeds 100.. .

    ``distinct`` > 0 generates only that many distinct calls and cycles them (generation of
    64 x 2 x 2.4 M samples takes tens of seconds of host time; the kernels do not care)."""
    m = n_calls if distinct <= 0 else min(distinct, n_calls)
    pool = [make_call(seconds, 8000, 100 + i, 2, kind) for i in range(m)]
    return [pool[i % m] for i in range(n_calls)], 8000


def cfg3_segments(n_segments: int = 4096, seed: int = 7, kind: str = "speech_like",
                  sample_rate: int = 16000) -> Tuple[List[torch.Tensor], int]:
    """cfg3: VAD-style mono segments, 0.5-30 s, plus the forced edge lengths."""
    g = _gen(seed)
    u = torch.rand(n_segments, generator=g, dtype=torch.float64)
    lengths = [int(sample_rate * (0.5 + 29.5 * float(v))) for v in u]
    lengths = list(CFG3_EDGE_LENGTHS) + lengths[: max(0, n_segments - len(CFG3_EDGE_LENGTHS))]
    segs = []
    for i, n in enumerate(lengths):
        s = 1000 + i
        w = speech_like(n, sample_rate, s) if kind == "speech_like" else white(n, s)
        segs.append(w.unsqueeze(0))
    return segs, sample_rate


def cfg5_longform(seconds: float = 3600.0, kind: str = "speech_like") -> Tuple[torch.Tensor, int]:
    """cfg5: one mono 44.1 kHz long-form waveform, seed 5."""
    return make_call(seconds, 44100, 5, 1, kind), 44100


def wav_image(call: torch.Tensor, sample_rate: int) -> torch.Tensor:
    """The bytes of a canonical 16-bit PCM wav file holding ``call`` ((C, L) float32 in [-1, 1)): a 44-byte
    RIFF header followed by the interleaved frames, as a uint8 tensor."""
    import struct
    pcm = (call * 32768.0).round().clamp(-32768, 32767).to(torch.int16).t().contiguous()     # (L, C)
    channels, nbytes = int(call.shape[0]), int(pcm.numel()) * 2
    head = (b"RIFF" + struct.pack("<I", 36 + nbytes) + b"WAVEfmt " +
            struct.pack("<IHHIIHH", 16, 1, channels, sample_rate, sample_rate * channels * 2, channels * 2, 16) +
            b"data" + struct.pack("<I", nbytes))
    out = torch.empty((44 + nbytes,), dtype=torch.uint8)
    out[:44] = torch.frombuffer(bytearray(head), dtype=torch.uint8)
    out[44:] = pcm.view(-1).view(torch.uint8)
    return out
