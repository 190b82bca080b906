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


_NO_AUDIO = torch.zeros(0, dtype=torch.float32)


def _as_reference_mono(w):
    """What ``waveform.squeeze(0)`` followed by ``unfold(0, 400, 160)`` accepts in the reference (:813-823): a 1-D
    signal or a ``(1, L)`` tensor.  Anything else -- ``(2, L)`` stereo, ``(1, 1, L)`` -- makes ``unfold`` raise there,
    which the reference turns into ``None`` / the zero feature vector; an empty signal gives the same result here."""
    shape = tuple(w.shape)
    if len(shape) == 1:
        return w
    if len(shape) == 2 and shape[0] == 1:
        return w[0]
    return _NO_AUDIO


def extract_acoustic_features_batch(waveforms: Sequence, sample_rate: int = 16000, device=None
                                    ) -> List[Optional["OrderedDict[str, float]"]]:
    """One ordered dict of 38 Python floats per waveform, ``None`` where the reference
    returns ``None`` (fewer than 400 samples)."""
    if not waveforms:
        return []
    fe = runtime.get_frontend(device)
    waveforms = [_as_reference_mono(w) for w in waveforms]
    rows, status = fe.acoustic(waveforms, sample_rate)
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


_EMPTY_RESULT = {"prediction": "", "confidence": 0.0, "probabilities": []}


class AcousticSentimentAnalyzer(AcousticFeatureExtractor):
    """The acoustic classifier glue of the reference's class of the same name
    (``analyze_batch_sentiment``, services/sentiment_analysis.py:696-805): all valid waveforms of a batch are
    featurised in one launch sequence, the ``StandardScaler`` step runs as an epilogue of the last kernel
    (float64, the operations of ``scaler.transform``), ONE (B, n_features) matrix comes back to the host and
    goes straight into the sklearn model's ``predict`` / ``predict_proba`` (SVM / RandomForest stay sklearn).

    The fitted models are handed in (the reference loads them with joblib): ``agent_acoustic_model`` /
    ``agent_acoustic_scaler`` / ``agent_acoustic_id2label`` and the ``client_*`` twins."""

    def __init__(self, device=None, agent_acoustic_model=None, agent_acoustic_scaler=None, agent_acoustic_id2label=None,
                 client_acoustic_model=None, client_acoustic_scaler=None, client_acoustic_id2label=None, logger=None):
        super().__init__(device)
        self.agent_acoustic_model, self.agent_acoustic_scaler = agent_acoustic_model, agent_acoustic_scaler
        self.client_acoustic_model, self.client_acoustic_scaler = client_acoustic_model, client_acoustic_scaler
        self.agent_acoustic_id2label, self.client_acoustic_id2label = agent_acoustic_id2label, client_acoustic_id2label
        self.models_available = all(m is not None for m in (agent_acoustic_model, agent_acoustic_scaler,
                                                            client_acoustic_model, client_acoustic_scaler))
        import logging
        self.logger = logger or logging.getLogger(__name__)
        self._scalers = {}

    @classmethod
    def from_reference(cls, analyzer, device=None) -> "AcousticSentimentAnalyzer":
        """Take the fitted models of a constructed reference analyzer."""
        names = ("agent_acoustic_model", "agent_acoustic_scaler", "agent_acoustic_id2label",
                 "client_acoustic_model", "client_acoustic_scaler", "client_acoustic_id2label")
        return cls(device, **{n: getattr(analyzer, n, None) for n in names}, logger=getattr(analyzer, "logger", None))

    def _scaler_vectors(self, scaler):
        """(mean, scale) float64 device vectors of ``scaler.n_features_in_`` entries (``with_mean`` /
        ``with_std`` = False become 0 / 1), cached per scaler object."""
        import numpy as np
        key = id(scaler)
        hit = self._scalers.get(key)
        if hit is None or hit[0] is not scaler:
            n = int(scaler.n_features_in_)
            mean = getattr(scaler, "mean_", None) if getattr(scaler, "with_mean", True) else None
            scale = getattr(scaler, "scale_", None) if getattr(scaler, "with_std", True) else None
            mean = np.zeros(n) if mean is None else np.asarray(mean, np.float64)
            scale = np.ones(n) if scale is None else np.asarray(scale, np.float64)
            fe = runtime.get_frontend(self.device)
            hit = (scaler, torch.from_numpy(mean.copy()).to(fe.device), torch.from_numpy(scale.copy()).to(fe.device))
            self._scalers[key] = hit
        return hit[1], hit[2]

    def _scaled_features(self, waveforms: Sequence[torch.Tensor], sample_rate: int, scaler):
        """``acoustic_scaler.transform(feature_lists)`` of :736-752 as a float64 numpy matrix."""
        fe = runtime.get_frontend(self.device)
        dev_waves, lens = fe._to_device_1d([_as_reference_mono(w) for w in waveforms])
        _, _, scaled = fe.acoustic_ptrs([t.data_ptr() if t.numel() else 0 for t in dev_waves], lens, sample_rate,
                                        scaler=self._scaler_vectors(scaler))
        return scaled.cpu().numpy()

    def analyze_batch_sentiment(self, waveforms: List[torch.Tensor], sample_rate: int, speaker: str):
        import numpy as np
        empty = lambda: dict(_EMPTY_RESULT, probabilities=[])          # noqa: E731
        try:
            if not waveforms:
                return []
            if not getattr(self, "models_available", False):
                return [empty() for _ in waveforms]
            valid = [i for i, w in enumerate(waveforms) if w is not None and w.numel() > 0]
            if not valid:
                return [empty() for _ in waveforms]
            if speaker == "client":
                model, scaler = self.client_acoustic_model, self.client_acoustic_scaler
            else:
                model, scaler = self.agent_acoustic_model, self.agent_acoustic_scaler
            if model is None or scaler is None:
                return [empty() for _ in waveforms]
            scaled = self._scaled_features([waveforms[i] for i in valid], sample_rate, scaler)
            ok = ~(np.all(scaled == 0, axis=1) | np.any(np.isnan(scaled), axis=1))
            predictions = model.predict(scaled)
            probabilities = model.predict_proba(scaled)
            results = [empty() for _ in waveforms]
            for j, i in enumerate(valid):
                if ok[j]:
                    results[i] = {"prediction": predictions[j], "confidence": float(probabilities[j].max()),
                                  "probabilities": probabilities[j].tolist()}
            return results
        except Exception as e:
            self.logger.error(f"Batch acoustic sentiment analysis error: {e}")
            return [empty() for _ in waveforms]

