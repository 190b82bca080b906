"""Drop-in for the hot-path methods of the reference's ``SpeechSegment`` /
``SpeechBatchTranscriber`` (services/speech_segment.py): ``_merge_segments`` (:366-383),
``_extract_segments`` (:385-418), the sort / batch / featurise part of
``transcribe_segments_batched`` (:190-290, same name, signature and error contract) and the
``transcribe_mono`` glue (:420-427).

The VAD model (pyannote) and the CTC model are out of scope: ``vad_fn``, ``transcription_model`` (+ the
tokenizer behind ``processor.batch_decode``) and ``decode_fn`` are injection points.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch

import logging

from . import _native, runtime
from .audio_processor import remove_special_characters
from .feature_extractor import FbankFeatureExtractor, FbankProcessor

logger = logging.getLogger(__name__)


def merge_segments(timeline: Sequence[Any], gap_threshold: float) -> List[Tuple[float, float]]:
    """Fuse consecutive intervals whose gap is <= ``gap_threshold``.  Items may be objects
    with ``.start`` / ``.end`` (pyannote segments) or ``(start, end)`` pairs."""
    def se(t):
        return (t.start, t.end) if hasattr(t, "start") else (t[0], t[1])
    if not timeline:
        return []
    merged = []
    cs, ce = se(timeline[0])
    for turn in timeline[1:]:
        s, e = se(turn)
        if s - ce <= gap_threshold:
            ce = e
        else:
            merged.append((cs, ce))
            cs, ce = s, e
    merged.append((cs, ce))
    return merged


def segment_table(num_samples: int, sample_rate: int, segments: Sequence[Tuple[float, float]], padding: float,
                  max_len_sec: float, overlap: float):
    """``(start_sample, end_sample, start_sec, end_sec)`` rows from the C ABI (float64, int() truncation)."""
    lib = _native.load()
    n_iv = len(segments)
    iv = (C.c_double * max(1, 2 * n_iv))(*[float(v) for p in segments for v in p])
    n = C.c_int(0)
    args = (int(num_samples), int(sample_rate), iv, n_iv, float(padding), float(max_len_sec), float(overlap))
    lib.check(lib.afe_segment_table(*args, None, None, None, None, 0, C.byref(n)))
    cap = max(1, n.value)
    s0, s1 = (C.c_int64 * cap)(), (C.c_int64 * cap)()
    t0, t1 = (C.c_double * cap)(), (C.c_double * cap)()
    lib.check(lib.afe_segment_table(*args, s0, s1, t0, t1, cap, C.byref(n)))
    return [(int(s0[i]), int(s1[i]), float(t0[i]), float(t1[i])) for i in range(n.value)]


class SpeechSegment:
    def __init__(self, config: dict, vad_fn: Optional[Callable] = None, processor=None, frontend=None):
        self.config = config
        gpu_index = int(config.get("gpu_index", 0))
        self.device = torch.device(f"cuda:{gpu_index}")
        self._frontend = frontend
        self.vad_fn = vad_fn                      # (waveform [1, L], sample_rate) -> [(start, end), ...]
        self.processor = processor or FbankProcessor(FbankFeatureExtractor(device=self.device))
        self.transcription_model = None           # callable(input_features) -> object with .logits (Wav2Vec2BertForCTC)

    @property
    def frontend(self) -> "runtime.Frontend":
        if self._frontend is None:
            self._frontend = runtime.get_frontend(self.device)
        return self._frontend

    def _merge_segments(self, timeline, gap_threshold: float):
        return merge_segments(timeline, gap_threshold)

    def _extract_segments(self, waveform: torch.Tensor, sample_rate: int, segments, speaker_label: str,
                          padding: float) -> List[Dict[str, Any]]:
        max_len_sec = float(self.config.get("vad_max_segment_sec", 20.0))
        overlap = float(self.config.get("vad_window_overlap_sec", 0.5))
        rows = segment_table(waveform.shape[1], sample_rate, list(segments), padding, max_len_sec, overlap)
        return [{"segment_waveform": waveform[:, a:b], "start": st, "end": en, "speaker": speaker_label}
                for a, b, st, en in rows]

    def get_speech_segments(self, waveform: torch.Tensor, sample_rate: int, speaker_label: str):
        if self.vad_fn is None:
            raise RuntimeError("no VAD configured: pass vad_fn (the pyannote pipeline is out of scope)")
        timeline = self.vad_fn(waveform, sample_rate)
        merged = self._merge_segments(list(timeline), float(self.config.get("vad_gap_threshold", 0.8)))
        return self._extract_segments(waveform, sample_rate, merged, speaker_label,
                                      float(self.config.get("vad_padding", 0.5)))

    def featurise_segments(self, segments: Sequence[Dict[str, Any]], sample_rate: int):
        """The featurise half of ``transcribe_segments_batched``: stable sort by length, slices
        of ``batch_size``, one processor() call per slice (one launch for all slices).
        Returns ``[(original_indices, input_features, attention_mask), ...]``."""
        if sample_rate != 16000:
            raise ValueError(f"processor was trained with sampling rate 16000, not {sample_rate}")
        if not segments:
            return []
        items = [(idx, seg["segment_waveform"]) for idx, seg in enumerate(segments)]
        items.sort(key=lambda x: x[1].shape[-1])
        batch_size = int(self.config.get("batch_size", self.config.get("chunk_batch_size", 8)))
        fx = self.processor.feature_extractor
        waves = [w.reshape(-1) if w.dim() > 1 and w.shape[0] == 1 else w for _, w in items]
        groups = self.frontend.fbank(waves, padding_value=fx.padding_value, precision=fx.precision,
                                     group_size=batch_size)
        out = []
        for gi, (feat, mask) in enumerate(groups):
            idxs = [i for i, _ in items[gi * batch_size:(gi + 1) * batch_size]]
            out.append((idxs, feat, mask))
        return out


    # -------------------------------------------------------------- reference :190-290
    def transcribe_segments_batched(self, segments: List[Dict[str, Any]], sample_rate: int) -> List[Dict[str, Any]]:
        """Same signature, ordering, result dicts and error contract as the reference method
        (services/speech_segment.py:190-290): stable sort of the segments by length, slices of
        ``batch_size`` through ``processor`` -> ``transcription_model`` -> argmax -> ``batch_decode`` ->
        ``remove_special_characters`` (+ optional ``converter``); an exception mentioning "out of memory"
        halves the batch size and retries the same slice (:253-260); any other sends the slice through
        the one-segment-at-a-time path (:262-283), where a failing segment gets ""; afterwards the cursor
        moves on by the batch size then in force.  Results keep the input order as ``{**segment, "text": ...}``.

        Batch assembly stays on the device: with the CUDA processor the features of ALL remaining slices
        come from one fbank launch (``group_size = batch_size``: every slice padded to its own longest
        segment, as one ``processor(...)`` call per slice would) and are handed to the model as views.
        The model (``self.transcription_model``) and tokenizer are injected, like ``AudioProcessor.model``."""
        if not segments:
            return []
        order = sorted(range(len(segments)), key=lambda k: int(segments[k]["segment_waveform"].shape[-1]))
        waves = [self._mono(segments[k]["segment_waveform"]) for k in order]
        model, processor = self.transcription_model, self.processor
        batch_size = int(self.config.get("batch_size", self.config.get("chunk_batch_size", 8)))
        converter = getattr(self, "converter", None)
        texts: Dict[int, str] = {}
        ahead = None                      # (first position, batch size, [input_features per slice])
        pos = 0
        while pos < len(order):
            members = order[pos:pos + batch_size]
            try:
                if ahead is None or ahead[1] != batch_size or not (0 <= (pos - ahead[0]) // batch_size < len(ahead[2])):
                    ahead = (pos, batch_size, self._features_ahead(waves[pos:], batch_size, sample_rate))
                features = ahead[2][(pos - ahead[0]) // batch_size]
                if features is None:      # a foreign processor: one call per slice, as the reference
                    inputs = processor(waves[pos:pos + batch_size], sampling_rate=sample_rate, return_tensors="pt", padding=True)
                    features = inputs["input_features"].to(self.device)
                decoded = processor.batch_decode(self._predict(model, features), skip_special_tokens=True)
                for k, text in zip(members, decoded):
                    cleaned = remove_special_characters(text or "")
                    if converter:
                        try:
                            cleaned = converter.convert_text(cleaned)
                        except Exception:
                            pass
                    texts[k] = cleaned.strip()
            except Exception as e:
                logger.error(f"segment slice at position {pos} failed: {e}")
                if "out of memory" in str(e).lower() and batch_size > 1:
                    torch.cuda.empty_cache()
                    batch_size = max(1, batch_size // 2)
                    ahead = None
                    logger.warning(f"out of memory: segment batch size lowered to {batch_size}, same slice again")
                    continue
                for k, wave in zip(members, waves[pos:pos + batch_size]):
                    try:
                        inputs = processor(wave, sampling_rate=sample_rate, return_tensors="pt", padding=True)
                        ids = self._predict(model, inputs["input_features"].to(self.device))
                        texts[k] = remove_special_characters(processor.batch_decode(ids, skip_special_tokens=True)[0] or "").strip()
                    except Exception:
                        texts[k] = ""
            pos += batch_size
        return [{**seg, "text": texts.get(k, "")} for k, seg in enumerate(segments)]

    @staticmethod
    def _mono(w: torch.Tensor) -> torch.Tensor:
        return w.reshape(-1) if w.dim() > 1 and w.shape[0] == 1 else w

    def _features_ahead(self, waves: Sequence[torch.Tensor], batch_size: int, sample_rate: int):
        """``input_features`` of consecutive slices of ``batch_size`` waveforms from one launch, or ``None`` per slice
        when the processor is not the CUDA one."""
        n_slices = (len(waves) + batch_size - 1) // batch_size
        fx = getattr(self.processor, "feature_extractor", None)
        if not isinstance(fx, FbankFeatureExtractor):
            return [None] * n_slices
        if sample_rate != fx.sampling_rate:
            raise ValueError(f"processor was trained with sampling rate {fx.sampling_rate}, not {sample_rate}")
        groups = self.frontend.fbank(list(waves), padding_value=fx.padding_value, precision=fx.precision, group_size=batch_size)
        return [feat for feat, _ in groups]

    def _predict(self, model, features: torch.Tensor) -> torch.Tensor:
        with torch.no_grad():
            if torch.cuda.is_available():
                with torch.autocast("cuda"):
                    logits = model(features).logits
            else:
                logits = model(features).logits
        return torch.argmax(logits, dim=-1)


class SpeechBatchTranscriber:
    def __init__(self, config: dict, vad_fn=None, decode_fn: Optional[Callable] = None, model=None, processor=None,
                 frontend=None):
        """``model`` + ``processor`` (with a tokenizer): the reference's path through
        ``transcribe_segments_batched``.  ``decode_fn`` ((input_features) -> list[str]) is the light-weight
        injection point used when only the featurise half is wanted."""
        self.segmenter = SpeechSegment(config, vad_fn=vad_fn, processor=processor, frontend=frontend)
        self.segmenter.transcription_model = model
        self.decode_fn = decode_fn

    def transcribe_mono(self, waveform: torch.Tensor, sample_rate: int, speaker_label: str = "unknown"):
        """reference :420-427: ``get_speech_segments`` -> ``transcribe_segments_batched``."""
        segments = self.segmenter.get_speech_segments(waveform, sample_rate, speaker_label)
        if self.segmenter.transcription_model is not None:
            return self.segmenter.transcribe_segments_batched(segments, sample_rate)
        texts: Dict[int, str] = {}
        for idxs, feat, _mask in self.segmenter.featurise_segments(segments, sample_rate):
            decoded = self.decode_fn(feat) if self.decode_fn is not None else [""] * len(idxs)
            for i, t in zip(idxs, decoded):
                texts[i] = t
        return [{**seg, "text": texts.get(i, "")} for i, seg in enumerate(segments)]
