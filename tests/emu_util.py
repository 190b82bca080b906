"""TEST ONLY: drive the C ABI of the host-emulated kernels with numpy buffers.

The emulation library (tests/_emu/libafe_emu.so) is the SAME kernel source compiled as
plain C++; "device pointers" are numpy buffers.  It checks kernel index arithmetic in the
GPU-less container; the product package never loads it."""
import ctypes as C
import functools

import numpy as np

from audio_processor_b200 import _native
from audio_processor_b200.csrc import build


@functools.lru_cache(maxsize=1)
def emu():
    lib = _native.Library(build.build_emu())
    assert lib.backend == 2
    return lib, lib.create(0)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def resample(x, orig, new, want_mixed=False):
    lib, ctx = emu()
    x = np.ascontiguousarray(x, np.float32)
    rows, L = x.shape
    Lo = lib.afe_resampled_length(L, orig, new)
    out = np.zeros((rows, Lo), np.float32)
    mixed = np.zeros((1, Lo), np.float32) if want_mixed else None
    lib.check(lib.afe_resample(ctx, None, _p(x), rows, L, orig, new, _p(out), _p(mixed) if want_mixed else None, Lo))
    return (out, mixed) if want_mixed else out


def wav_parse(data: bytes, total_bytes=None):
    lib, _ = emu()
    info = _native.WavInfo()
    buf = C.create_string_buffer(data, len(data))
    code = lib.afe_wav_parse(C.cast(buf, C.c_void_p), len(data), len(data) if total_bytes is None else total_bytes,
                             C.byref(info))
    if code != 0:
        return code, lib.afe_last_error().decode()
    return 0, {k: int(getattr(info, k)) for k, _ in _native.WavInfo._fields_ if k != "reserved"}


def wav_decode(frames: bytes, fmt: int, channels: int, misalign: int = 0):
    """afe_wav_decode on the bytes of a data chunk placed `misalign` bytes off an aligned address."""
    lib, ctx = emu()
    bps = {1: 1, 2: 2, 3: 3, 4: 4, 5: 4, 6: 8}[fmt]
    n = len(frames) // (bps * channels)
    buf = np.zeros(len(frames) + misalign + 16, np.uint8)
    buf[misalign:misalign + len(frames)] = np.frombuffer(frames, np.uint8)
    out = np.full((channels, n), np.float32(-9.0), np.float32)
    lib.check(lib.afe_wav_decode(ctx, None, C.c_void_p(buf.ctypes.data + misalign), fmt, channels, n, _p(out)))
    return out


def resample_batch(calls, orig, new, want_mixed=False, pcm16=False, wav=False):
    """``wav``: the calls are int16 (rows, L) arrays handed over as interleaved frames [L][rows]."""
    lib, ctx = emu()
    n, rows = len(calls), calls[0].shape[0]
    lens = [c.shape[1] for c in calls]
    if wav:
        calls = [np.ascontiguousarray(np.asarray(c, np.int16).T) for c in calls]
    else:
        calls = [np.ascontiguousarray(c, np.int16 if pcm16 else np.float32) for c in calls]
    outs = [np.full((rows, lib.afe_resampled_length(L, orig, new)), np.float32(-9.0), np.float32) for L in lens]
    mixes = [np.full((1, o.shape[1]), np.float32(-9.0), np.float32) for o in outs] if want_mixed else None
    ws = np.zeros(lib.afe_resample_batch_workspace_bytes(n), np.uint8)
    fn = lib.afe_resample_batch_wav16 if wav else lib.afe_resample_batch_pcm16 if pcm16 else lib.afe_resample_batch
    lib.check(fn(ctx, None, _native.ptr_array([c.ctypes.data for c in calls]),
                                     _native.i64_array(lens), n, rows, orig, new,
                                     _native.ptr_array([o.ctypes.data for o in outs]),
                                     _native.ptr_array([m.ctypes.data for m in mixes]) if want_mixed else None,
                                     _p(ws), ws.size))
    return (outs, mixes) if want_mixed else outs


def fbank_grouped(waves, group_size, padding_value=0.0, precision=0):
    """Returns a list of (features, mask) per group."""
    lib, ctx = emu()
    waves = [np.ascontiguousarray(w, np.float32).reshape(-1) for w in waves]
    B = len(waves)
    lens = _native.i64_array([w.size for w in waves])
    ptrs = _native.ptr_array([w.ctypes.data for w in waves])
    ng = (B + group_size - 1) // group_size
    tps = (C.c_int64 * ng)()
    rows = C.c_int64()
    lib.check(lib.afe_fbank_layout(lens, B, group_size, tps, C.byref(rows)))
    ws = np.zeros(lib.afe_fbank_workspace_bytes(lens, B), np.uint8)
    feat = np.full((rows.value // 2, 160), np.float32(-777.0), np.float32)
    mask = np.full((rows.value // 2,), -7, np.int32)
    lib.check(lib.afe_fbank_batch(ctx, None, ptrs, lens, B, group_size, padding_value, precision, 0,
                                  _p(feat), _p(mask), _p(ws), ws.size))
    out, r = [], 0
    for g in range(ng):
        n = min(group_size, B - g * group_size)
        h = tps[g] // 2
        out.append((feat[r:r + n * h].reshape(n, h, 160), mask[r:r + n * h].reshape(n, h)))
        r += n * h
    return out


def fbank(waves, padding_value=0.0, precision=0, raw=False):
    lib, ctx = emu()
    waves = [np.ascontiguousarray(w, np.float32).reshape(-1) for w in waves]
    B = len(waves)
    lens = _native.i64_array([w.size for w in waves])
    ptrs = _native.ptr_array([w.ctypes.data for w in waves])
    Tp = lib.afe_fbank_padded_frames(lens, B)
    ws = np.zeros(lib.afe_fbank_workspace_bytes(lens, B), np.uint8)
    feat = np.full((B, Tp // 2, 160), np.float32(-777.0), np.float32)
    mask = np.full((B, Tp // 2), -7, np.int32)
    lib.check(lib.afe_fbank_batch(ctx, None, ptrs, lens, B, 0, padding_value, precision, 1 if raw else 0,
                                  _p(feat), _p(mask), _p(ws), ws.size))
    return feat, mask


def acoustic(waves, sample_rate=16000):
    lib, ctx = emu()
    waves = [np.ascontiguousarray(w, np.float32).reshape(-1) for w in waves]
    B = len(waves)
    lens = _native.i64_array([w.size for w in waves])
    ptrs = _native.ptr_array([w.ctypes.data for w in waves])
    ws = np.zeros(lib.afe_acoustic_workspace_bytes(lens, B), np.uint8)
    out = np.zeros((B, 38), np.float32)
    status = np.full((B,), -1, np.int32)
    lib.check(lib.afe_acoustic_batch(ctx, None, ptrs, lens, B, sample_rate, _p(out), _p(status), _p(ws), ws.size))
    return out, status


def acoustic_scaled(waves, mean, scale, sample_rate=16000):
    """afe_acoustic_batch_scaled: (rows (B, 38) f32, status, scaled (B, n_features) f64)."""
    lib, ctx = emu()
    waves = [np.ascontiguousarray(w, np.float32).reshape(-1) for w in waves]
    B = len(waves)
    lens = _native.i64_array([w.size for w in waves])
    ptrs = _native.ptr_array([w.ctypes.data for w in waves])
    ws = np.zeros(lib.afe_acoustic_workspace_bytes(lens, B), np.uint8)
    out = np.zeros((B, 38), np.float32)
    status = np.full((B,), -1, np.int32)
    mean = np.ascontiguousarray(mean, np.float64); scale = np.ascontiguousarray(scale, np.float64)
    scaled = np.full((B, mean.size), np.nan, np.float64)
    lib.check(lib.afe_acoustic_batch_scaled(ctx, None, ptrs, lens, B, sample_rate, None, _p(out), _p(status), _p(mean),
                                            _p(scale), mean.size, _p(scaled), _p(ws), ws.size))
    return out, status, scaled

