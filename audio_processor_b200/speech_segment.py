"""Drop-in for the hot-path methods of the reference's ``SpeechSegment`` /
``SpeechBatchTranscriber`` (services/speech_segment.py): ``_merge_segments`` (:366-383),
``_extract_segments`` (:385-418), the sort / batch / featurise part of
``transcribe_segments_batched`` (:199-232) and the ``transcribe_mono`` glue (:420-427).

The VAD model (pyannote) and the CTC model are out of scope: ``vad_fn`` and ``decode_fn``
are injection points that default to "not configured".
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch

from . import _native, runtime
from .feature_extractor import FbankFeatureExtractor, FbankProcessor


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
    def __init__(self, config: dict, vad_fn: Optional[Callable] = None, processor=None):
        self.config = config
        gpu_index = int(config.get("gpu_index", 0))
        self.device = torch.device(f"cuda:{gpu_index}")
        self.frontend = runtime.get_frontend(self.device)
        self.vad_fn = vad_fn                      # (waveform [1, L], sample_rate) -> [(start, end), ...]
        self.processor = processor or FbankProcessor(FbankFeatureExtractor(device=self.device))
        self.transcription_model = None

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


class SpeechBatchTranscriber:
    def __init__(self, config: dict, vad_fn=None, decode_fn: Optional[Callable] = None):
        self.segmenter = SpeechSegment(config, vad_fn=vad_fn)
        self.decode_fn = decode_fn                # (input_features) -> list[str]; the CTC model, out of scope

    def transcribe_mono(self, waveform: torch.Tensor, sample_rate: int, speaker_label: str = "unknown"):
        segments = self.segmenter.get_speech_segments(waveform, sample_rate, speaker_label)
        texts: Dict[int, str] = {}
        for idxs, feat, _mask in self.segmenter.featurise_segments(segments, sample_rate):
            decoded = self.decode_fn(feat) if self.decode_fn is not None else [""] * len(idxs)
            for i, t in zip(idxs, decoded):
                texts[i] = t
        return [{**seg, "text": texts.get(i, "")} for i, seg in enumerate(segments)]
