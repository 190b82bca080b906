"""Device runtime: one ``Frontend`` per CUDA device, a thin typed layer over the C ABI.

PyTorch is plumbing here (device memory, the caching allocator, the current stream); all
arithmetic happens in ``libafe.so``.  Every method enqueues on ``torch.cuda.current_stream``
and never synchronises unless it has to hand Python scalars back.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _native

Wave = Union[np.ndarray, torch.Tensor]

_frontends: Dict[int, "Frontend"] = {}
_frontends_lock = threading.Lock()


def get_frontend(device: Union[int, str, torch.device, None] = None) -> "Frontend":
    """The per-device singleton.  Raises if CUDA or the native library is unavailable."""
    if not torch.cuda.is_available():
        raise _native.AfeLibraryError("audio_processor_b200 needs a CUDA device (no CPU fallback)")
    if device is None:
        idx = torch.cuda.current_device()
    else:
        dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
        idx = dev.index if dev.index is not None else torch.cuda.current_device()
    with _frontends_lock:
        fe = _frontends.get(idx)
        if fe is None:
            fe = Frontend(idx)
            _frontends[idx] = fe
        return fe


class Frontend:
    def __init__(self, device_index: int):
        self.lib = _native.load()
        self.device = torch.device("cuda", device_index)
        self.ctx = self.lib.create(device_index)
        self.num_sms = int(self.lib.afe_num_sms(self.ctx))
        self._tls = threading.local()

    # ------------------------------------------------------------------ helpers
    def _stream(self) -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _workspace(self, nbytes: int) -> torch.Tensor:
        ws = getattr(self._tls, "ws", None)
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
            self._tls.ws = ws
        return ws

    def _workspace2(self, nbytes: int) -> torch.Tensor:
        ws = getattr(self._tls, "ws2", None)        # the acoustic workspace (holds the mel-dB rows)
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
            self._tls.ws2 = ws
        return ws

    def _workspace3(self, nbytes: int) -> torch.Tensor:
        ws = getattr(self._tls, "ws3", None)        # resample batch descriptors
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=self.device)
            self._tls.ws3 = ws
        return ws

    def _to_device_1d(self, waves: Sequence[Wave]) -> Tuple[List[torch.Tensor], List[int]]:
        """Mono float32 device views of every waveform; host inputs share ONE pinned staging
        buffer and ONE H2D copy (offsets rounded to 4 floats so rows stay 16-byte aligned)."""
        out: List[Optional[torch.Tensor]] = [None] * len(waves)
        host_idx, host_arr = [], []
        for i, w in enumerate(waves):
            if isinstance(w, torch.Tensor):
                t = w.detach()
                if t.dim() == 2:
                    t = t[0]                       # stereo arrays: channel 0 only (reference :121-122)
                t = t.reshape(-1)
                if t.is_cuda:
                    out[i] = t.to(device=self.device, dtype=torch.float32).contiguous()
                    continue
                w = t.to(torch.float32).contiguous().numpy()
            a = np.asarray(w)
            if a.ndim == 2:
                a = a[0]
            host_idx.append(i)
            host_arr.append(np.ascontiguousarray(a.reshape(-1), dtype=np.float32))
        if host_idx:
            offs, total = [], 0
            for a in host_arr:
                offs.append(total)
                total += (a.size + 3) // 4 * 4
            pinned = torch.empty(max(total, 1), dtype=torch.float32, pin_memory=True)
            pv = pinned.numpy()
            for a, o in zip(host_arr, offs):
                pv[o:o + a.size] = a
            dev = pinned.to(self.device, non_blocking=True)
            self._tls.last_pinned = pinned          # keep alive until the next call on this thread
            for i, a, o in zip(host_idx, host_arr, offs):
                out[i] = dev[o:o + a.size]
        lens = [int(t.numel()) for t in out]
        return out, lens  # type: ignore[return-value]

    # ------------------------------------------------------------------ resample
    def resampled_length(self, length: int, orig: int, new: int) -> int:
        return int(self.lib.afe_resampled_length(int(length), int(orig), int(new)))

    def resample(self, x: torch.Tensor, orig: int, new: int, want_mixed: bool = False):
        """``torchaudio.transforms.Resample(orig, new)(x)`` on a (..., L) float32 CUDA tensor."""
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        lead, L = x.shape[:-1], int(x.shape[-1])
        rows = int(np.prod(lead)) if lead else 1
        Lo = self.resampled_length(L, orig, new)
        out = torch.empty((rows, Lo), dtype=torch.float32, device=self.device)
        mixed = torch.empty((1, Lo), dtype=torch.float32, device=self.device) if want_mixed else None
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_resample(
                self.ctx, self._stream(), C.c_void_p(x.data_ptr()), rows, L, int(orig), int(new),
                C.c_void_p(out.data_ptr()), C.c_void_p(mixed.data_ptr()) if want_mixed else None, Lo))
        out = out.reshape(tuple(lead) + (Lo,))
        return (out, mixed) if want_mixed else out

    def resample_batch(self, calls: Sequence[torch.Tensor], orig: int, new: int, want_mixed: bool = False):
        """``Resample(orig, new)`` over a batch of (rows, L_c) float32 CUDA tensors in ONE launch
        (the per-file loop of ``AudioProcessor.process_batch``).  Returns a list of outputs (and a
        list of (1, L') mixed tensors with ``want_mixed``); all outputs are views of one
        allocation, every view 16-byte aligned."""
        n = len(calls)
        if n == 0:
            return ([], []) if want_mixed else []
        # 16-bit PCM (planar, or interleaved frames as in a wav file): x / 32768 while staging
        kinds = {self.pcm_kind(c) for c in calls}
        if kinds != {"f32"} and (len(kinds) > 1 or self.resampled_length(1000, orig, new) != 2000):
            calls = [self.to_float(c) for c in calls]; kinds = {"f32"}
        kind = next(iter(kinds))
        pcm16 = kind != "f32"
        if kind == "wav":                                       # keep the (1, C) strides of the frames
            xs = [c.to(device=self.device) for c in calls]
        else:
            xs = [c.to(device=self.device, dtype=torch.int16 if pcm16 else torch.float32).contiguous() for c in calls]
        rows = int(xs[0].shape[0])
        if any(x.dim() != 2 or int(x.shape[0]) != rows for x in xs):
            raise ValueError("resample_batch: every call must be (rows, L) with the same number of rows")
        lens = [int(x.shape[1]) for x in xs]
        louts = [self.resampled_length(L, orig, new) for L in lens]
        per = rows + (1 if want_mixed else 0)
        offs, total = [], 0
        for lo in louts:
            offs.append(total)
            total += per * ((lo + 3) // 4 * 4)
        buf = torch.empty((max(total, 1),), dtype=torch.float32, device=self.device)
        outs, mixes, p_out, p_mix = [], [], [], []
        for o, lo in zip(offs, louts):
            pad = (lo + 3) // 4 * 4                 # the block is padded, rows stay contiguous
            y = buf[o:o + rows * lo].view(rows, lo)
            outs.append(y); p_out.append(y.data_ptr())
            if want_mixed:
                m = buf[o + rows * pad:o + rows * pad + lo].view(1, lo)
                mixes.append(m); p_mix.append(m.data_ptr())
        ws = self._workspace3(int(self.lib.afe_resample_batch_workspace_bytes(n)))
        fn = {"f32": self.lib.afe_resample_batch, "pcm": self.lib.afe_resample_batch_pcm16,
              "wav": self.lib.afe_resample_batch_wav16}[kind]
        with torch.cuda.device(self.device):
            self.lib.check(fn(
                self.ctx, self._stream(), _native.ptr_array([x.data_ptr() for x in xs]), _native.i64_array(lens), n,
                rows, int(orig), int(new), _native.ptr_array(p_out), _native.ptr_array(p_mix) if want_mixed else None,
                C.c_void_p(ws.data_ptr()), ws.numel()))
        self._tls.keep_rs = xs                      # inputs stay alive until the next call on this thread
        return (outs, mixes) if want_mixed else outs

    @staticmethod
    def pcm_kind(x: torch.Tensor) -> str:
        """"f32": a float waveform; "pcm": int16 planar (C, L); "wav": an int16 (C, L) VIEW of the
        interleaved frames of a wav data chunk (strides (1, C)), as ``wav.load_pcm16`` returns."""
        if x.dtype != torch.int16:
            return "f32"
        if x.dim() == 2 and int(x.shape[0]) > 1 and tuple(x.stride()) == (1, int(x.shape[0])):
            return "wav"
        return "pcm"

    def to_float(self, x: torch.Tensor) -> torch.Tensor:
        """What ``torchaudio.load`` would have returned for this call: float32 (C, L) on the device."""
        kind = self.pcm_kind(x)
        if kind == "pcm":
            return self.pcm16_to_float(x)
        if kind == "wav":
            return self.wav_decode(x.to(self.device).t(), _native.WAV_S16, int(x.shape[0]), int(x.shape[1]))
        return x.to(device=self.device, dtype=torch.float32).contiguous()

    def wav_decode(self, frames: torch.Tensor, fmt: int, channels: int, num_frames: int) -> torch.Tensor:
        """Interleaved frames of a wav data chunk (any dtype, contiguous, on the device) ->
        ``torchaudio.load``'s (channels, num_frames) float32 waveform."""
        frames = frames.to(self.device)
        if not frames.is_contiguous():
            raise ValueError("wav_decode: frames must be the contiguous bytes of the data chunk")
        out = torch.empty((channels, num_frames), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_wav_decode(self.ctx, self._stream(), C.c_void_p(frames.data_ptr()), int(fmt),
                                                   int(channels), int(num_frames), C.c_void_p(out.data_ptr())))
        self._tls.keep_wav = frames
        return out

    def pcm16_to_float(self, x: torch.Tensor) -> torch.Tensor:
        """int16 PCM -> float32 in [-1, 1): ``x / 32768`` (torchaudio.load's normalisation)."""
        x = x.to(device=self.device, dtype=torch.int16).contiguous()
        out = torch.empty(x.shape, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_pcm16_to_float(self.ctx, self._stream(), C.c_void_p(x.data_ptr()),
                                                       int(x.numel()), C.c_void_p(out.data_ptr())))
        return out

    def mix_stereo(self, x: torch.Tensor) -> torch.Tensor:
        x = x.to(device=self.device, dtype=torch.float32).contiguous()
        L = int(x.shape[-1])
        mixed = torch.empty((1, L), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_mix_stereo(self.ctx, self._stream(), C.c_void_p(x.data_ptr()), L,
                                                   C.c_void_p(mixed.data_ptr())))
        return mixed

    # ------------------------------------------------------------------ fbank
    def fbank(self, waves: Sequence[Wave], padding_value: float = 0.0, precision: int = _native.PRECISION_F64,
              raw: bool = False, group_size: int = 0):
        """One ``processor(...)`` call: (input_features (B, Tp/2, 160) float32, attention_mask
        (B, Tp/2) int32) on the device.  With ``group_size`` > 0 several processor() calls are
        featurised in one launch and a list of (features, mask) per group is returned."""
        dev_waves, lens = self._to_device_1d(waves)
        out = self.fbank_ptrs([t.data_ptr() if t.numel() else 0 for t in dev_waves], lens, padding_value,
                              precision, raw, group_size)
        self._tls.keep = dev_waves                  # stream-ordered lifetime of the staged inputs
        return out

    def fbank_ptrs(self, ptrs: Sequence[int], lens: Sequence[int], padding_value: float = 0.0,
                   precision: int = _native.PRECISION_F64, raw: bool = False, group_size: int = 0):
        """Same as ``fbank`` for waveforms already on this device, given as raw device addresses
        (float32, contiguous) and sample counts -- no per-waveform tensor objects."""
        B = len(ptrs)
        G = int(group_size) if group_size and group_size > 0 else 0
        if B == 0:
            empty = (torch.empty((0, 0, 160), dtype=torch.float32, device=self.device),
                     torch.empty((0, 0), dtype=torch.int32, device=self.device))
            return [] if G else empty
        lens_c = _native.i64_array(lens)
        ng = (B + G - 1) // G if G else 1
        tps = (C.c_int64 * ng)()
        rows = C.c_int64()
        self.lib.check(self.lib.afe_fbank_layout(lens_c, B, G, tps, C.byref(rows)))
        half = rows.value // 2
        feat = torch.empty((half, 160), dtype=torch.float32, device=self.device)
        mask = torch.empty((half,), dtype=torch.int32, device=self.device)
        ws = self._workspace(int(self.lib.afe_fbank_workspace_bytes(lens_c, B)))
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_fbank_batch(
                self.ctx, self._stream(), _native.ptr_array(ptrs), lens_c, B, G, float(padding_value), int(precision),
                _native.FBANK_RAW if raw else 0, C.c_void_p(feat.data_ptr()), C.c_void_p(mask.data_ptr()),
                C.c_void_p(ws.data_ptr()), ws.numel()))
        out, r = [], 0
        for g in range(ng):
            n = min(G, B - g * G) if G else B
            h = int(tps[g]) // 2
            out.append((feat[r:r + n * h].view(n, h, 160), mask[r:r + n * h].view(n, h)))
            r += n * h
        return out if G else out[0]

    def fbank_layout(self, lens: Sequence[int], group_size: int = 0):
        """(padded frames per group, total frame rows) of the grouped output (afe_fbank_layout)."""
        B = len(lens)
        G = int(group_size) if group_size and group_size > 0 else 0
        ng = ((B + G - 1) // G if G else 1) if B else 0
        tps = (C.c_int64 * max(ng, 1))()
        rows = C.c_int64()
        self.lib.check(self.lib.afe_fbank_layout(_native.i64_array(lens), B, G, tps, C.byref(rows)))
        return [int(tps[g]) for g in range(ng)], int(rows.value)

    def fbank_placed(self, ptrs: Sequence[int], lens: Sequence[int], out_rows: Sequence[int],
                     padded_frames: Sequence[int], feat: torch.Tensor, mask: torch.Tensor,
                     padding_value: float = 0.0, precision: int = _native.PRECISION_F64) -> None:
        """Featurise device waveforms into caller-placed rows of ``feat`` (rows/2, 160) and
        ``mask`` (rows/2,): utterance b starts at frame row ``out_rows[b]`` and is padded to
        ``padded_frames[b]`` frames (afe_fbank_batch_placed)."""
        B = len(ptrs)
        if B == 0:
            return
        lens_c = _native.i64_array(lens)
        ws = self._workspace(int(self.lib.afe_fbank_workspace_bytes(lens_c, B)))
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_fbank_batch_placed(
                self.ctx, self._stream(), _native.ptr_array(ptrs), lens_c, B, _native.i64_array(out_rows),
                _native.i64_array(padded_frames), float(padding_value), int(precision), 0,
                C.c_void_p(feat.data_ptr()), C.c_void_p(mask.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel()))

    # ------------------------------------------------------------------ acoustic
    def acoustic(self, waves: Sequence[Wave], sample_rate: int = 16000) -> Tuple[torch.Tensor, torch.Tensor]:
        """((B, 38) float32 feature rows, (B,) int32 status) on the device; see afe.h."""
        dev_waves, lens = self._to_device_1d(waves)
        out = self.acoustic_ptrs([t.data_ptr() if t.numel() else 0 for t in dev_waves], lens, sample_rate)
        self._tls.keep = dev_waves
        return out

    def acoustic_ptrs(self, ptrs: Sequence[int], lens: Sequence[int], sample_rate: int = 16000,
                      out: Optional[torch.Tensor] = None, status: Optional[torch.Tensor] = None,
                      out_index: Optional[Sequence[int]] = None, scaler=None, scaled: Optional[torch.Tensor] = None):
        """With ``out`` / ``status`` / ``out_index`` waveform b is written to row ``out_index[b]``
        of the caller's (N, 38) / (N,) tensors (afe_acoustic_batch_placed).  ``scaler`` = (mean, scale)
        float64 device vectors of n_features entries: also returns / fills the (N, n_features) float64
        matrix ``scaler.transform`` would produce (afe_acoustic_batch_scaled)."""
        B = len(ptrs)
        if scaler is not None:
            mean, scale = scaler
            nf = int(mean.numel())
            if scaled is None:
                scaled = torch.empty((B if out is None else out.shape[0], nf), dtype=torch.float64, device=self.device)
            if out is None:
                out = torch.empty((B, 38), dtype=torch.float32, device=self.device)
                status = torch.empty((B,), dtype=torch.int32, device=self.device)
            if B:
                lens_c = _native.i64_array(lens)
                idx_c = (C.c_int32 * B)(*[int(i) for i in out_index]) if out_index is not None else None
                ws = self._workspace2(int(self.lib.afe_acoustic_workspace_bytes(lens_c, B)))
                with torch.cuda.device(self.device):
                    self.lib.check(self.lib.afe_acoustic_batch_scaled(
                        self.ctx, self._stream(), _native.ptr_array(ptrs), lens_c, B, int(sample_rate), idx_c,
                        C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()), C.c_void_p(mean.data_ptr()),
                        C.c_void_p(scale.data_ptr()), nf, C.c_void_p(scaled.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel()))
            return out, status, scaled
        if out is None:
            out = torch.empty((B, 38), dtype=torch.float32, device=self.device)
            status = torch.empty((B,), dtype=torch.int32, device=self.device)
        if B == 0:
            return out, status
        lens_c = _native.i64_array(lens)
        idx_c = (C.c_int32 * B)(*[int(i) for i in out_index]) if out_index is not None else None
        ws = self._workspace2(int(self.lib.afe_acoustic_workspace_bytes(lens_c, B)))
        with torch.cuda.device(self.device):
            self.lib.check(self.lib.afe_acoustic_batch_placed(
                self.ctx, self._stream(), _native.ptr_array(ptrs), lens_c, B, int(sample_rate), idx_c,
                C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel()))
        return out, status

    # ------------------------------------------------------------------ per-kernel device timing
    def profile(self, enabled: bool) -> None:
        self.lib.check(self.lib.afe_profile_enable(self.ctx, 1 if enabled else 0))
        self.lib.check(self.lib.afe_profile_reset(self.ctx))

    def profile_read(self):
        """{kernel name: (summed device ms, launches)} since the last ``profile()`` call."""
        ms = (C.c_double * _native.K_COUNT)()
        n = (C.c_int64 * _native.K_COUNT)()
        self.lib.check(self.lib.afe_profile_read(self.ctx, ms, n))
        return {self.lib.afe_kernel_name(i).decode(): (float(ms[i]), int(n[i])) for i in range(_native.K_COUNT)}
