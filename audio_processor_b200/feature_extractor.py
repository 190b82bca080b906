"""Drop-in for the reference's ``processor(...)`` featurise call.

The reference builds ``Wav2Vec2BertProcessor.from_pretrained(...)`` and calls
``processor(arrays, sampling_rate=16000, return_tensors="pt", padding=True)``
(services/audio_processor.py:61-69, :253-261; services/speech_segment.py:226-231, :264),
which forwards to ``SeamlessM4TFeatureExtractor.__call__``
(transformers feature_extraction_seamless_m4t.py:141-302).  ``FbankFeatureExtractor``
keeps that call signature and result mapping; the arithmetic runs in ``libafe.so``.

Output placement: tensors are returned on the CUDA device (the
reference moves them there itself right after, ``.to(self.device)`` is then a no-op); pass
``return_tensors="np"`` for host numpy arrays.
"""
from __future__ import annotations

from collections import UserDict
from typing import Any, List, Optional, Sequence, Union

import numpy as np
import torch

from . import _native, runtime


class FeatureBatch(UserDict):
    """Minimal ``BatchFeature``: mapping + attribute access + ``.to(device)``."""

    def __getattr__(self, item):
        try:
            return self.data[item]
        except KeyError as e:
            raise AttributeError(item) from e

    def to(self, *args, **kwargs) -> "FeatureBatch":
        self.data = {k: (v.to(*args, **kwargs) if isinstance(v, torch.Tensor) else v) for k, v in self.data.items()}
        return self


def _is_batched(raw_speech) -> bool:
    if isinstance(raw_speech, np.ndarray):
        return raw_speech.ndim > 1
    if isinstance(raw_speech, torch.Tensor):
        return raw_speech.dim() > 1
    if isinstance(raw_speech, (list, tuple)):
        return len(raw_speech) > 0 and isinstance(raw_speech[0], (np.ndarray, torch.Tensor, list, tuple))
    return False


class FbankFeatureExtractor:
    """Same constructor defaults and ``__call__`` keywords as ``SeamlessM4TFeatureExtractor``."""

    model_input_names = ["input_features", "attention_mask"]

    def __init__(self, feature_size: int = 80, sampling_rate: int = 16000, num_mel_bins: int = 80,
                 padding_value: float = 0.0, stride: int = 2, device=None, precision: str = "f64", **kwargs):
        if (feature_size, sampling_rate, num_mel_bins, stride) != (80, 16000, 80, 2):
            raise ValueError("libafe implements the w2v-BERT / SeamlessM4T front end: 80 mel bins, 16 kHz, stride 2")
        self.feature_size = feature_size
        self.sampling_rate = sampling_rate
        self.num_mel_bins = num_mel_bins
        self.padding_value = float(padding_value)
        self.stride = stride
        self.return_attention_mask = True
        self.padding_side = "right"
        self.precision = _native.PRECISION_F64 if precision == "f64" else _native.PRECISION_F32
        self._device = device

    @classmethod
    def from_reference(cls, hf_feature_extractor, **kw) -> "FbankFeatureExtractor":
        """Mirror the checkpoint-specific knobs (``padding_value``) of a loaded HF extractor."""
        return cls(padding_value=float(getattr(hf_feature_extractor, "padding_value", 0.0)), **kw)

    def _frontend(self) -> runtime.Frontend:
        return runtime.get_frontend(self._device)

    def __call__(self, raw_speech, padding: Union[bool, str] = True, pad_to_multiple_of: Optional[int] = 2,
                 max_length: Optional[int] = None, truncation: bool = False,
                 return_tensors: Optional[str] = None, sampling_rate: Optional[int] = None,
                 return_attention_mask: Optional[bool] = None, do_normalize_per_mel_bins: Optional[bool] = True,
                 **kwargs) -> FeatureBatch:
        if sampling_rate is not None and sampling_rate != self.sampling_rate:
            raise ValueError(
                f"The model corresponding to this feature extractor: {self.__class__.__name__} was trained using a "
                f"sampling rate of {self.sampling_rate}. Please make sure that the provided `raw_speech` input was "
                f"sampled with {self.sampling_rate} and not {sampling_rate}.")
        if padding not in (True, "longest") or pad_to_multiple_of != 2 or max_length is not None or truncation:
            raise NotImplementedError("only padding=True / pad_to_multiple_of=2 (the reference's call) is built")
        if not do_normalize_per_mel_bins:
            raise NotImplementedError("do_normalize_per_mel_bins=False is not used by the reference")
        waves: Sequence[Any] = list(raw_speech) if _is_batched(raw_speech) else [raw_speech]
        waves = [np.asarray(w, dtype=np.float32) if isinstance(w, (list, tuple)) else w for w in waves]
        feat, mask = self._frontend().fbank(waves, padding_value=self.padding_value, precision=self.precision)
        want_mask = self.return_attention_mask if return_attention_mask is None else return_attention_mask
        out = {"input_features": feat}
        if want_mask:
            out["attention_mask"] = mask
        if return_tensors in ("np", "numpy"):
            out = {k: v.cpu().numpy() for k, v in out.items()}
        elif return_tensors not in (None, "pt"):
            raise ValueError("return_tensors must be 'pt', 'np' or None")
        return FeatureBatch(out)


class FbankProcessor:
    """``Wav2Vec2BertProcessor`` stand-in: audio goes to ``FbankFeatureExtractor``, everything
    tokenizer-side (``batch_decode`` / ``decode``) is forwarded untouched to the HF tokenizer
    or processor the reference loaded (out of scope here)."""

    def __init__(self, feature_extractor: Optional[FbankFeatureExtractor] = None, tokenizer=None):
        self.feature_extractor = feature_extractor or FbankFeatureExtractor()
        self.tokenizer = tokenizer

    @classmethod
    def wrap(cls, hf_processor, **kw) -> "FbankProcessor":
        """``processor = FbankProcessor.wrap(Wav2Vec2BertProcessor.from_pretrained(path))``"""
        fe = FbankFeatureExtractor.from_reference(getattr(hf_processor, "feature_extractor", hf_processor), **kw)
        return cls(fe, getattr(hf_processor, "tokenizer", hf_processor))

    def __call__(self, audio=None, text=None, **kwargs):
        if audio is None and "raw_speech" in kwargs:
            audio = kwargs.pop("raw_speech")
        if text is not None:
            if self.tokenizer is None:
                raise ValueError("no tokenizer attached")
            return self.tokenizer(text, **kwargs)
        if audio is None:
            raise ValueError("You need to specify either an `audio` or `text` input to process.")
        return self.feature_extractor(audio, **kwargs)

    def batch_decode(self, *args, **kwargs):
        if self.tokenizer is None:
            raise ValueError("no tokenizer attached")
        return self.tokenizer.batch_decode(*args, **kwargs)

    def decode(self, *args, **kwargs):
        if self.tokenizer is None:
            raise ValueError("no tokenizer attached")
        return self.tokenizer.decode(*args, **kwargs)
