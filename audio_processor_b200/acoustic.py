"""Drop-in for ``AcousticSentimentAnalyzer._extract_acoustic_features`` and
``_get_zero_features`` (services/sentiment_analysis.py:807-913): the 38 acoustic scalars
that feed the reference's SVM / RandomForest sentiment models, in the reference's key order.

``extract_acoustic_features_batch`` is the batched form the reference's per-waveform loop
(:736-748) collapses into: ONE launch sequence and ONE device->host copy for the whole
batch instead of 38 ``.item()`` syncs per waveform.
"""
from __future__ import annotations

import sys
from collections import OrderedDict
from typing import List, Optional, Sequence

import torch

from . import runtime

FEATURE_KEYS = (
    ["rms_mean", "rms_std", "rms_range", "zcr_mean", "zcr_std",
     "spectral_centroid_mean", "spectral_centroid_std",
     "spectral_bandwidth_mean", "spectral_bandwidth_std",
     "spectral_rolloff_mean", "spectral_rolloff_std", "tempo"]
    + [f"mfcc_{i}_{s}" for i in range(13) for s in ("mean", "std")]
)
TEMPO_SLOT = 11


def zero_features() -> "OrderedDict[str, float]":
    return OrderedDict((k, 0.0) for k in FEATURE_KEYS)


def _tempo(flux_energy: float, num_samples: int, sample_rate: int) -> float:
    """The reference's bug-compatible tempo (:869-875): ``autocorr[0] = 0`` clears lag
    -(Tf-1), so arg-max is the zero-lag peak at index Tf-1 = F2-2 unless the flux is all zero."""
    hop_length = int(0.010 * sample_rate)
    f2 = 1 + num_samples // hop_length
    max_lag = (f2 - 2) if flux_energy > 0.0 else 0
    period = max_lag * hop_length / sample_rate if max_lag > 0 else 0.0
    return 60.0 / period if period > 0 else 0.0


def extract_acoustic_features_batch(waveforms: Sequence, sample_rate: int = 16000, device=None
                                    ) -> List[Optional["OrderedDict[str, float]"]]:
    """One ordered dict of 38 Python floats per waveform, ``None`` where the reference
    returns ``None`` (fewer than 400 samples)."""
    if not waveforms:
        return []
    fe = runtime.get_frontend(device)
    rows, status = fe.acoustic(list(waveforms), sample_rate)
    rows = rows.cpu()
    status = status.cpu().tolist()
    values = rows.tolist()
    out: List[Optional[OrderedDict]] = []
    for w, st, vals in zip(waveforms, status, values):
        if st != 0:
            out.append(None)
            continue
        n = int(w.numel() if isinstance(w, torch.Tensor) else w.size)
        vals[TEMPO_SLOT] = _tempo(vals[TEMPO_SLOT], n, sample_rate)
        out.append(OrderedDict(zip(FEATURE_KEYS, vals)))
    return out


class AcousticFeatureExtractor:
    """Mix-in / stand-alone object exposing the reference's two method names, so it can be
    bound onto an ``AcousticSentimentAnalyzer``."""

    def __init__(self, device=None):
        self.device = device

    def _extract_acoustic_features(self, waveform: torch.Tensor, sample_rate: int):
        try:
            res = extract_acoustic_features_batch([waveform], sample_rate, self.device)[0]
            if res is None:
                raise ValueError("Audio too short")
            return res
        except Exception as exc:                      # the reference swallows everything (:890-894)
            print(f"[Feature Extraction Error]: {exc}", file=sys.stderr)
            return None

    def _get_zero_features(self):
        return zero_features()

    def extract_batch(self, waveforms: Sequence, sample_rate: int):
        """Feature lists in the reference's order with zero rows for failures (:736-748)."""
        feats = extract_acoustic_features_batch(list(waveforms), sample_rate, self.device)
        return [list((f if f is not None else zero_features()).values()) for f in feats]
