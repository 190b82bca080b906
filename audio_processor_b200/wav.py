"""Audio decode + staging: the wav side of ``AudioProcessor.load_audio``
(services/audio_processor.py:72-93, ``torchaudio.load``) and of
``AudioFileScanner._validate_file`` (services/pipeline.py:84-108, ``torchaudio.info``).

The header walk is ``afe_wav_parse`` (host, integers only); the samples cross PCIe as they lie
in the file -- interleaved 16-bit frames, 2 bytes per sample -- and are de-interleaved,
normalised (``x / 32768``) and, for 8 kHz calls, resampled in ONE kernel
(``afe_resample_batch_wav16``).  ``load_audio`` itself (the reference signature: a float32
``(channels, frames)`` waveform + sample rate) decodes on the device with ``afe_wav_decode``.

Other containers stay with the reference: ``.ogg`` is piped through ffmpeg into wav bytes
exactly as the reference does and those bytes take the same path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from pathlib import Path
from typing import Optional, Tuple, Union

import torch

from . import _native, runtime

_ENCODING = {_native.WAV_U8: "PCM_U", _native.WAV_S16: "PCM_S", _native.WAV_S24: "PCM_S", _native.WAV_S32: "PCM_S",
             _native.WAV_F32: "PCM_F", _native.WAV_F64: "PCM_F"}
_DTYPE = {_native.WAV_U8: torch.uint8, _native.WAV_S16: torch.int16, _native.WAV_S32: torch.int32,
          _native.WAV_F32: torch.float32, _native.WAV_F64: torch.float64}


@dataclass(frozen=True)
class WavHeader:
    """The fields of ``torchaudio.info``'s AudioMetaData plus where the frames are."""
    sample_rate: int
    num_frames: int
    num_channels: int
    bits_per_sample: int
    encoding: str
    format: int            # afe_wav_format
    data_offset: int
    data_bytes: int


def parse_wav(buf, total_bytes: Optional[int] = None) -> WavHeader:
    """Header of a RIFF/WAVE file from its first bytes (``bytes`` / ``bytearray`` / uint8 tensor).
    ``total_bytes``: size of the whole file when ``buf`` is only a prefix."""
    lib = _native.load()
    if isinstance(buf, torch.Tensor):
        ptr, n, keep = C.c_void_p(buf.data_ptr()), int(buf.numel() * buf.element_size()), buf
    elif isinstance(buf, bytes):
        keep = C.c_char_p(buf)                                   # no copy: points into the bytes object
        ptr, n = C.cast(keep, C.c_void_p), len(buf)
    else:
        keep = (C.c_char * len(buf)).from_buffer(buf)            # bytearray / writable memoryview
        ptr, n = C.cast(keep, C.c_void_p), len(buf)
    info = _native.WavInfo()
    code = lib.afe_wav_parse(ptr, n, int(total_bytes if total_bytes is not None else n), C.byref(info))
    del keep
    if code != _native.AFE_OK:
        msg = lib.afe_last_error()
        raise ValueError(msg.decode() if msg else f"afe_wav_parse failed ({code})")
    return WavHeader(int(info.sample_rate), int(info.num_frames), int(info.channels), int(info.bits_per_sample),
                     _ENCODING[int(info.format)], int(info.format), int(info.data_offset), int(info.data_bytes))


def info(path: Union[str, Path]) -> WavHeader:
    """``torchaudio.info(path)`` for wav files: reads the header only (4 KB, growing while the data
    chunk hides behind large LIST / bext chunks)."""
    size = os.path.getsize(path)
    want = 4096
    with open(path, "rb") as f:
        while True:
            f.seek(0)
            head = f.read(min(want, size))
            try:
                return parse_wav(head, size)
            except ValueError as e:
                if "within the bytes given" not in str(e) and "truncated" not in str(e):
                    raise
                if len(head) >= size:
                    raise
                want *= 16


def validate_file(path: Union[str, Path]) -> bool:
    """``AudioFileScanner._validate_file`` (services/pipeline.py:84-108): exists, not empty, and
    NOT mono (the reference rejects ``num_channels == 1``: a call needs both speakers)."""
    try:
        p = Path(path)
        if not p.exists():
            return False
        meta = info(p)
        if meta.num_frames == 0:
            return False
        if meta.num_channels == 1:
            return False
        return True
    except Exception:
        return False


def read_pinned(path: Union[str, Path]) -> torch.Tensor:
    """The whole file in page-locked memory (one read, no intermediate bytes object)."""
    size = os.path.getsize(path)
    buf = torch.empty((max(size, 1),), dtype=torch.uint8, pin_memory=torch.cuda.is_available())
    with open(path, "rb") as f:
        got = f.readinto(memoryview(buf.numpy())[:size]) if size else 0
    if got != size:
        raise IOError(f"{path}: short read ({got} of {size} bytes)")
    return buf[:size]


def _as_tensor(data) -> torch.Tensor:
    if isinstance(data, torch.Tensor):
        return data
    t = torch.frombuffer(bytearray(data), dtype=torch.uint8)
    return t.pin_memory() if torch.cuda.is_available() else t


def load_pcm16(source, header: Optional[WavHeader] = None) -> Tuple[torch.Tensor, int]:
    """``(pcm, sample_rate)`` for a 16-bit wav file (path, bytes or uint8 tensor): ``pcm`` is an int16
    ``(channels, frames)`` VIEW of the interleaved frames where they lie in the (pinned) file buffer
    -- nothing is copied or converted on the host.  ``CallFrontend.process_calls`` and
    ``Frontend.resample_batch`` take it as a call; ``Frontend.to_float`` gives the float waveform."""
    buf = read_pinned(source) if isinstance(source, (str, Path)) else _as_tensor(source)
    h = header or parse_wav(buf)
    if h.format != _native.WAV_S16:
        raise ValueError(f"load_pcm16: {h.bits_per_sample}-bit {h.encoding} data; use load_audio")
    frames = buf[h.data_offset:h.data_offset + h.data_bytes].view(torch.int16).view(h.num_frames, h.num_channels)
    return frames.t(), h.sample_rate


def load_audio(source, device=None) -> Tuple[torch.Tensor, int]:
    """``torchaudio.load(path)`` for wav data: ``(waveform float32 (channels, frames) on the device,
    sample_rate)``.  The file's bytes are copied to the device as they are and decoded there."""
    fe = runtime.get_frontend(device)
    buf = read_pinned(source) if isinstance(source, (str, Path)) else _as_tensor(source)
    h = parse_wav(buf)
    data = buf[h.data_offset:h.data_offset + h.data_bytes].to(fe.device, non_blocking=True)
    return fe.wav_decode(data, h.format, h.num_channels, h.num_frames), h.sample_rate


def ffmpeg_wav_bytes(path: Union[str, Path]) -> bytes:
    """The reference's .ogg route (services/audio_processor.py:76-85): ffmpeg -> 16-bit wav on a pipe
    (whose RIFF sizes are placeholders; ``afe_wav_parse`` reads such data chunks to the end)."""
    command = ["ffmpeg", "-i", str(path), "-f", "wav", "-acodec", "pcm_s16le", "-"]
    proc = subprocess.Popen(command, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
    wav_bytes, _ = proc.communicate()
    return wav_bytes
