"""TEST INFRASTRUCTURE -- container-only bridge to the UNMODIFIED reference.

Imports the reference's own hot-path modules from ``/root/reference`` (read-only mount,
absent on the GPU box) after stubbing the third-party modules that are not installed
here and are not on the hot path (pyannote, rapidfuzz, boto3, psycopg2, minio, openpyxl).
Used only by ``oracle/make_golden.py`` to generate ``tests/golden/*.npz`` and by the
optional ``tests/test_oracle_vs_reference.py`` (skipped when the mount is missing).

Nothing in the product package imports this file.
"""
from __future__ import annotations

import os
import sys
from unittest.mock import MagicMock

REFERENCE_ROOT = os.environ.get("AFE_REFERENCE_ROOT", "/root/reference")

_STUBS = ("pyannote", "pyannote.audio", "rapidfuzz", "boto3", "psycopg2", "psycopg2.extras",
          "minio", "minio.error", "openpyxl", "tenacity", "botocore", "botocore.exceptions",
          "botocore.config")

DEFAULT_CONFIG = {
    # the hot-path keys of config.yaml:7,21-24,56-65
    "chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000,
    "batch_size": 16, "io_workers": 2, "max_retries": 1, "gpu_index": 0,
    "vad_max_segment_sec": 25.0, "vad_window_overlap_sec": 0.5,
    "vad_gap_threshold": 0.8, "vad_padding": 0.5,
}


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "services"))


def _install_stubs() -> None:
    for name in _STUBS:
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = MagicMock()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


_cache = {}


def audio_processor(config=None):
    """The reference's ``AudioProcessor`` (services/audio_processor.py:17) on the CPU."""
    _install_stubs()
    key = ("ap", repr(sorted((config or DEFAULT_CONFIG).items())))
    if key not in _cache:
        import logging
        logging.disable(logging.CRITICAL)
        from services.audio_processor import AudioProcessor  # type: ignore
        _cache[key] = AudioProcessor(dict(config or DEFAULT_CONFIG))
        logging.disable(logging.NOTSET)
    return _cache[key]


def speech_segment(config=None):
    """The reference's ``SpeechSegment`` (services/speech_segment.py:160)."""
    return audio_processor(config).transcriber.segmenter


def processor():
    """What ``Wav2Vec2BertProcessor.__call__`` forwards audio to: the class-default
    ``SeamlessM4TFeatureExtractor`` (the fine-tuned checkpoint is not shipped; its
    fbank parameters are the class defaults)."""
    if "proc" not in _cache:
        from transformers import SeamlessM4TFeatureExtractor
        _cache["proc"] = SeamlessM4TFeatureExtractor()
    return _cache["proc"]


def acoustic_analyzer():
    """The reference's ``AcousticSentimentAnalyzer`` without its joblib models: only
    ``_extract_acoustic_features`` / ``_get_zero_features`` are exercised."""
    _install_stubs()
    if "ac" not in _cache:
        import logging
        import torch
        from services.sentiment_analysis import AcousticSentimentAnalyzer  # type: ignore
        obj = AcousticSentimentAnalyzer.__new__(AcousticSentimentAnalyzer)
        obj.device = torch.device("cpu")
        obj.logger = logging.getLogger("reference.acoustic")
        _cache["ac"] = obj
    return _cache["ac"]
