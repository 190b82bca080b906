"""Whole front end for a batch of calls: the path ``AudioProcessor.process_batch`` ->
``split_audio`` -> ``transcribe_batch`` featurise (x3 streams) -> acoustic features for
agent / client chunks (services/audio_processor.py:324-367, sentiment_analysis.py:98-125)
collapses into, with a handful of launches per batch instead of per-chunk Python loops.

``process_calls`` is the public entry the benchmark times end to end (host waveforms in,
device features + host acoustic rows out).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _native, runtime
from .acoustic import TEMPO_SLOT, _tempo
from .audio_processor import chunk_table


@dataclass
class FrontendResult:
    n_calls: int = 0
    call_seconds: float = 0.0
    chunks: List[Tuple[int, int, int, int]] = field(default_factory=list)    # (call, chunk_idx, start, end)
    order: List[int] = field(default_factory=list)                            # length-sorted chunk order
    # per stream ("stereo", "agent", "client"): list over batches of (input_features, attention_mask)
    features: Dict[str, List[Tuple[torch.Tensor, torch.Tensor]]] = field(default_factory=dict)
    acoustic_rows: Optional[torch.Tensor] = None      # (2 * n_chunks, 38): agent chunks then client chunks
    acoustic_status: Optional[torch.Tensor] = None
    waveforms: List[Tuple[torch.Tensor, torch.Tensor]] = field(default_factory=list)   # (stereo16k (2,L'), mixed (1,L'))


class CallFrontend:
    def __init__(self, device=None, chunk_duration_sec=25, overlap_sec=1, target_sample_rate=16000,
                 batch_size=16, padding_value=0.0, precision=_native.PRECISION_F64, acoustic=True):
        self.fe = runtime.get_frontend(device)
        self.device = self.fe.device
        self.chunk_duration_sec = chunk_duration_sec
        self.overlap_sec = overlap_sec
        self.target_sample_rate = target_sample_rate
        self.batch_size = int(batch_size)
        self.padding_value = float(padding_value)
        self.precision = precision
        self.acoustic = acoustic
        self._tables = {}
        self._copy_stream = None

    def process_calls(self, calls: Sequence[torch.Tensor], sample_rate: int, acoustic_to_host: bool = True,
                      sub_batch: Optional[int] = None) -> FrontendResult:
        """``calls``: (C, L) float32 waveforms, or int16 PCM as stored in a 16-bit wav file
        (``torchaudio.load(..., normalize=False)``; converted on the device exactly like
        ``torchaudio.load``'s ``x / 32768``, so a call crosses PCIe at 2 bytes per sample);
        host -- ideally pinned -- or device tensors.

        Every integer of the result (chunk table, length-sorted order, the padded length of
        each batch of ``batch_size`` chunks, the output row of every chunk) depends on the call
        lengths only, so the whole output layout is planned on the host first.  The calls are
        then processed in sub-batches of ``sub_batch`` calls: while sub-batch k is resampled,
        featurised (all three streams in one launch, each chunk written to its final rows)
        and its acoustic rows computed, sub-batch k+1 is uploaded on a second stream."""
        fe, sr_t, dev = self.fe, self.target_sample_rate, self.device
        res = FrontendResult(n_calls=len(calls))
        chunk_samples = int(self.chunk_duration_sec * sr_t)
        step = chunk_samples - int(self.overlap_sec * sr_t)
        # ---- plan (host, integers only)
        louts, first_chunk, lens = [], [], []
        for ci, call in enumerate(calls):
            L = int(call.shape[1])
            res.call_seconds += L / float(sample_rate)
            Lr = fe.resampled_length(L, sample_rate, sr_t) if sample_rate != sr_t else L
            louts.append(Lr)
            first_chunk.append(len(res.chunks))
            for idx, (s, e) in enumerate(self._chunk_table(Lr, chunk_samples, step)):
                res.chunks.append((ci, idx, s, e)); lens.append(e - s)
        first_chunk.append(len(res.chunks))
        n = len(res.chunks)
        if n == 0:
            return res
        G = self.batch_size
        res.order = sorted(range(n), key=lens.__getitem__)       # transcribe_batch: stable sort by length
        slens = [lens[i] for i in res.order]
        tps, total_rows = fe.fbank_layout(slens, G)
        row_of, tp_of, pos = [0] * n, [0] * n, 0
        for g, tp in enumerate(tps):
            for j in range(g * G, min(n, (g + 1) * G)):
                row_of[res.order[j]] = pos; tp_of[res.order[j]] = tp; pos += tp
        half = total_rows // 2
        feat_all = torch.empty((3 * half, 160), dtype=torch.float32, device=dev)
        mask_all = torch.empty((3 * half,), dtype=torch.int32, device=dev)
        if self.acoustic:
            rows = torch.empty((2 * n, 38), dtype=torch.float32, device=dev)
            status = torch.empty((2 * n,), dtype=torch.int32, device=dev)
        # ---- pipeline over sub-batches of calls
        any_host = any(not c.is_cuda for c in calls)
        sb = int(sub_batch) if sub_batch else (8 if any_host else len(calls))
        main = torch.cuda.current_stream(dev)
        if any_host and self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(dev)
        # Device copies of host calls live in ONE buffer allocated on the compute stream (so the
        # caching allocator never has to cudaMalloc / defer a free on the side stream); the
        # copy stream first waits for the compute stream to reach this point.
        xdev: List[Optional[torch.Tensor]] = [None] * len(calls)
        if any_host:
            offs, total = [], 0                                   # in bytes, 16-byte aligned
            for c in calls:
                offs.append(total)
                if not c.is_cuda:
                    total += (c.numel() * c.element_size() + 15) // 16 * 16
            xin = torch.empty((max(total, 16),), dtype=torch.uint8, device=dev)
            for i, c in enumerate(calls):
                if not c.is_cuda:
                    if c.dtype not in (torch.float32, torch.int16):
                        raise TypeError("process_calls: calls must be float32 waveforms or int16 PCM")
                    xdev[i] = xin[offs[i]:offs[i] + c.numel() * c.element_size()].view(c.dtype).view(c.shape)
            self._copy_stream.wait_stream(main)

        def upload(k0):
            sub = range(k0, min(k0 + sb, len(calls)))
            if not any_host:
                return [calls[i].to(dev) for i in sub], None
            xs = []
            with torch.cuda.stream(self._copy_stream):
                for i in sub:
                    if xdev[i] is None:
                        xs.append(calls[i].to(dev))
                    else:
                        xdev[i].copy_(calls[i], non_blocking=True)
                        xs.append(xdev[i])
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            return xs, ev

        starts = list(range(0, len(calls), sb))
        pending = upload(starts[0])
        res.waveforms = [None] * len(calls)                       # type: ignore[list-item]
        for si, k0 in enumerate(starts):
            xs, ev = pending
            if si + 1 < len(starts):
                pending = upload(starts[si + 1])                  # enqueue the next upload before this compute
            if ev is not None:
                main.wait_event(ev)
            # resample / mix: stereo and mono calls of the sub-batch, one launch each
            for want_rows in (2, 1):
                ids = [k0 + j for j, x in enumerate(xs) if int(x.shape[0]) == want_rows]
                if not ids:
                    continue
                grp = [xs[i - k0] for i in ids]
                if sample_rate != sr_t:
                    if want_rows == 2:
                        ys, ms = fe.resample_batch(grp, sample_rate, sr_t, want_mixed=True)
                    else:
                        ys = fe.resample_batch(grp, sample_rate, sr_t); ms = ys
                else:
                    ys = [fe.pcm16_to_float(x) if x.dtype == torch.int16 else x.contiguous() for x in grp]
                    ms = [fe.mix_stereo(y) for y in ys] if want_rows == 2 else ys
                for i, y, m in zip(ids, ys, ms):
                    res.waveforms[i] = (y, m)
            bad = [k0 + j for j, x in enumerate(xs) if int(x.shape[0]) not in (1, 2)]
            if bad:
                raise ValueError(f"process_calls: call {bad[0]} has {int(xs[bad[0] - k0].shape[0])} channels")
            c0, c1 = first_chunk[k0], first_chunk[min(k0 + sb, len(calls))]
            if c1 == c0:
                continue
            mixed_p, agent_p, client_p = [], [], []
            for ci, _, s, _ in res.chunks[c0:c1]:
                y, m = res.waveforms[ci]
                Lr = int(y.shape[1])
                pa = y.data_ptr()
                pc = pa + 4 * Lr if y.shape[0] == 2 else pa
                mixed_p.append(m.data_ptr() + 4 * s); agent_p.append(pa + 4 * s); client_p.append(pc + 4 * s)
            sub_lens = lens[c0:c1]
            sub_rows = row_of[c0:c1]
            fe.fbank_placed(mixed_p + agent_p + client_p, sub_lens * 3,
                            sub_rows + [r + total_rows for r in sub_rows] + [r + 2 * total_rows for r in sub_rows],
                            tp_of[c0:c1] * 3, feat_all, mask_all, padding_value=self.padding_value,
                            precision=self.precision)
            if self.acoustic:
                fe.acoustic_ptrs(agent_p + client_p, sub_lens * 2, sr_t, out=rows, status=status,
                                 out_index=list(range(c0, c1)) + list(range(n + c0, n + c1)))
        for si, key in enumerate(("stereo", "agent", "client")):
            views, r = [], si * half
            for g, tp in enumerate(tps):
                ng, h = min(G, n - g * G), tp // 2
                views.append((feat_all[r:r + ng * h].view(ng, h, 160), mask_all[r:r + ng * h].view(ng, h)))
                r += ng * h
            res.features[key] = views
        if self.acoustic:
            if acoustic_to_host:
                rows = rows.cpu(); status = status.cpu()
            res.acoustic_rows, res.acoustic_status = rows, status
        return res

    def _chunk_table(self, total: int, chunk: int, step: int):
        key = (total, chunk, step)
        tab = self._tables.get(key)
        if tab is None:
            tab = self._tables[key] = chunk_table(total, chunk, step)
        return tab

    def acoustic_dicts(self, res: FrontendResult):
        """Finish the host side of the acoustic rows: (agent dicts, client dicts) per chunk."""
        from .acoustic import FEATURE_KEYS
        from collections import OrderedDict
        rows = res.acoustic_rows.cpu().tolist()
        status = res.acoustic_status.cpu().tolist()
        n = len(res.chunks)
        out = []
        for j, (vals, st) in enumerate(zip(rows, status)):
            _, _, s, e = res.chunks[j % n]
            if st != 0:
                out.append(None); continue
            vals[TEMPO_SLOT] = _tempo(vals[TEMPO_SLOT], e - s, self.target_sample_rate)
            out.append(OrderedDict(zip(FEATURE_KEYS, vals)))
        return out[:n], out[n:]
