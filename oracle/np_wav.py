"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the wav side of the reference's audio loading
(SURVEY.md section 8f-1).  Nothing in ``audio_processor_b200`` may import this module.

Follows ``AudioProcessor.load_audio`` (services/audio_processor.py:72-93: ``torchaudio.load(path)``,
default ``normalize=True``, ``channels_first=True``) and ``AudioFileScanner._validate_file``
(services/pipeline.py:84-108: ``torchaudio.info`` -> ``num_frames`` / ``num_channels``).

The arithmetic is in torchaudio (pinned 2.6.0), not under /root/reference.  Its published
behaviour for wav data: the result is float32 ``(channels, frames)``; integer PCM of b bits is
divided by 2**(b-1) (8-bit wav data is unsigned with a bias of 128); 24-bit data is widened to
32 bits first; float data is passed through (float64 is cast).

PARITY UNPINNED AGAINST TORCHAUDIO: the installed torchaudio 2.11 needs TorchCodec (absent) for
``load`` and has no ``info``.  The restatement is pinned instead against two independent wav
readers that ARE in this container -- Python's stdlib ``wave`` (header fields, raw frames) and
``scipy.io.wavfile`` (sample values of every encoding) -- in tests/test_wav.py.
"""
import struct

import numpy as np

U8, S16, S24, S32, F32, F64 = 1, 2, 3, 4, 5, 6
_BYTES = {U8: 1, S16: 2, S24: 3, S32: 4, F32: 4, F64: 8}


def parse(data: bytes, total_bytes=None):
    """dict(format, channels, sample_rate, bits_per_sample, data_offset, data_bytes, num_frames)."""
    total = len(data) if total_bytes is None else max(int(total_bytes), len(data))
    if len(data) < 12 or data[:4] != b"RIFF" or data[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file")
    pos, fmt = 12, None
    while True:
        if pos + 8 > len(data):
            raise ValueError("header does not end within the bytes given")
        cid, size = data[pos:pos + 4], struct.unpack_from("<I", data, pos + 4)[0]
        if cid == b"fmt ":
            if size < 16 or pos + 24 > len(data):
                raise ValueError("truncated fmt chunk")
            tag, ch, rate, _, align, bits = struct.unpack_from("<HHIIHH", data, pos + 8)
            if tag == 0xFFFE:
                if size < 40 or pos + 48 > len(data):
                    raise ValueError("truncated extensible fmt chunk")
                tag = struct.unpack_from("<H", data, pos + 8 + 24)[0]
            fmt = (tag, ch, rate, align, bits)
        elif cid == b"data":
            if fmt is None:
                raise ValueError("data chunk before fmt chunk")
            tag, ch, rate, align, bits = fmt
            code = {(1, 8): U8, (1, 16): S16, (1, 24): S24, (1, 32): S32, (3, 32): F32, (3, 64): F64}.get((tag, bits))
            if code is None:
                raise NotImplementedError("unsupported format tag / bit depth")
            frame = ch * (bits // 8)
            if ch <= 0 or rate <= 0 or align != frame:
                raise ValueError("inconsistent fmt chunk")
            off = pos + 8
            rest = total - off
            n = rest if (size in (0, 0xFFFFFFFF) or size > rest) else size
            frames = n // frame
            return dict(format=code, channels=ch, sample_rate=rate, bits_per_sample=bits, data_offset=off,
                        data_bytes=frames * frame, num_frames=frames)
        pos += 8 + size + (size & 1)


def decode(frames: bytes, fmt: int, channels: int) -> np.ndarray:
    """Interleaved data-chunk bytes -> float32 (channels, frames), torchaudio.load(normalize=True)."""
    b = _BYTES[fmt]
    n = len(frames) // (b * channels)
    raw = np.frombuffer(frames, np.uint8, n * channels * b)
    if fmt == U8:
        x = (raw.astype(np.int32) - 128).astype(np.float32) / np.float32(128.0)
    elif fmt == S16:
        x = raw.view("<i2").astype(np.float32) / np.float32(32768.0)
    elif fmt == S24:
        t = raw.reshape(-1, 3).astype(np.uint32)
        v = ((t[:, 0] << 8) | (t[:, 1] << 16) | (t[:, 2] << 24)).astype(np.uint32).view(np.int32)   # widened to s32
        x = v.astype(np.float32) / np.float32(2147483648.0)
    elif fmt == S32:
        x = raw.view("<i4").astype(np.float32) / np.float32(2147483648.0)
    elif fmt == F32:
        x = raw.view("<f4").astype(np.float32)
    else:
        x = raw.view("<f8").astype(np.float32)
    return np.ascontiguousarray(x.reshape(n, channels).T)


def load(data: bytes):
    """``torchaudio.load`` on the bytes of a wav file: (waveform float32 (C, L), sample_rate)."""
    h = parse(data)
    return decode(data[h["data_offset"]:h["data_offset"] + h["data_bytes"]], h["format"], h["channels"]), h["sample_rate"]


def validate(data: bytes) -> bool:
    """AudioFileScanner._validate_file's decision for a file with these bytes."""
    try:
        h = parse(data)
    except Exception:
        return False
    return h["num_frames"] != 0 and h["channels"] != 1
