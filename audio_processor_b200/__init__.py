"""audio_processor_b200 -- B200-native waveform front end for jaratM/audio_processor.

Resample + channel split + chunking, Wav2Vec2-BERT / SeamlessM4T fbank features and the
38 acoustic sentiment features, computed by hand-written sm_100a CUDA kernels in
``libafe.so`` (C ABI: ``include/afe.h``) behind the reference's own Python call signatures.
There is no CPU fallback: importing is cheap, but the first use raises ``AfeLibraryError``
unless the CUDA library and a CUDA device are present.
"""
from ._native import AfeError, AfeLibraryError  # noqa: F401

__all__ = ["AfeError", "AfeLibraryError", "AudioProcessor", "Resampler", "FbankFeatureExtractor", "FbankProcessor",
           "SpeechSegment", "SpeechBatchTranscriber", "AcousticFeatureExtractor", "extract_acoustic_features_batch",
           "CallFrontend", "get_frontend"]


def __getattr__(name):       # lazy: keep `import audio_processor_b200` free of torch / CUDA work
    if name in ("AudioProcessor", "Resampler"):
        from . import audio_processor as m
    elif name in ("FbankFeatureExtractor", "FbankProcessor"):
        from . import feature_extractor as m
    elif name in ("SpeechSegment", "SpeechBatchTranscriber"):
        from . import speech_segment as m
    elif name in ("AcousticFeatureExtractor", "extract_acoustic_features_batch"):
        from . import acoustic as m
    elif name == "CallFrontend":
        from . import frontend as m
    elif name == "get_frontend":
        from . import runtime as m
    else:
        raise AttributeError(name)
    return getattr(m, name)
