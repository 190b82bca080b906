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


@dataclass
class _Plan:
    """Everything about a batch of calls that depends on their lengths only."""
    call_seconds: float
    chunks: List[Tuple[int, int, int, int]]
    lens: List[int]
    first_chunk: List[int]
    order: List[int]
    tps: List[int]
    total_rows: int
    row_of: List[int]
    tp_of: List[int]


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
        self._plans: Dict[tuple, _Plan] = {}
        self._copy_stream = None
        self._side_stream = None
        self.trace = None        # debug: set to a dict to get CUDA-event timelines of the uploads and sub-batches (tools/e2e_timeline.py)

    def process_calls(self, calls: Sequence[torch.Tensor], sample_rate: int, acoustic_to_host: bool = True,
                      sub_batch: Optional[int] = None) -> FrontendResult:
        """``calls``: (C, L) float32 waveforms, or int16 PCM as stored in a 16-bit wav file
        (``torchaudio.load(..., normalize=False)``; converted on the device exactly like
        ``torchaudio.load``'s ``x / 32768``, so a call crosses PCIe at 2 bytes per sample);
        host -- ideally pinned -- or device tensors.

        Every integer of the result (chunk table, length-sorted order, the padded length of
        each batch of ``batch_size`` chunks, the output row of every chunk) depends on the call
        lengths only, so the whole output layout is planned on the host first.  The calls are
        then processed in sub-batches (``sub_batch`` calls each, or adaptively -- ``_next_group``):
        every upload is enqueued on a second stream at once; while sub-batch k is resampled,
        featurised (all three streams in one launch, each chunk written to its final rows)
        and its acoustic rows computed, the later calls keep crossing PCIe."""
        fe, sr_t, dev = self.fe, self.target_sample_rate, self.device
        res = FrontendResult(n_calls=len(calls))
        # ---- uploads first: the copy of sub-batch 0 crosses PCIe while the host plans the layout
        any_host = any(not c.is_cuda for c in calls)
        groups = self._schedule(len(calls), sub_batch, any_host)       # [(first call, end call)] per sub-batch
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
                    flat = xin[offs[i]:offs[i] + c.numel() * c.element_size()].view(c.dtype)
                    # wav frames keep their interleaved layout: a (C, L) view with strides (1, C)
                    xdev[i] = flat.view(c.shape[1], c.shape[0]).t() if fe.pcm_kind(c) == "wav" else flat.view(c.shape)
            self._copy_stream.wait_stream(main)

        # every host call is enqueued on the copy stream right away, one event per call; device calls
        # (and host calls of an unsupported layout) take the ordinary .to(dev) path
        xs_all: List[torch.Tensor] = [None] * len(calls)           # type: ignore[list-item]
        evs: List[Optional[torch.cuda.Event]] = [None] * len(calls)
        if any_host:
            with torch.cuda.stream(self._copy_stream):
                for i, c in enumerate(calls):
                    if xdev[i] is None:
                        xs_all[i] = c.to(dev)
                    else:
                        if fe.pcm_kind(c) == "wav":
                            xdev[i].t().copy_(c.t(), non_blocking=True)      # both sides contiguous (L, C)
                        else:
                            xdev[i].copy_(c, non_blocking=True)
                        xs_all[i] = xdev[i]
                        evs[i] = torch.cuda.Event(enable_timing=self.trace is not None)
                        evs[i].record(self._copy_stream)
        else:
            xs_all = [c.to(dev) for c in calls]
        # ---- plan (host, integers only; cached by the call lengths it depends on)
        plan = self._plan(tuple(int(c.shape[1]) for c in calls), int(sample_rate))
        res.call_seconds = plan.call_seconds
        res.chunks, res.order = plan.chunks, plan.order
        lens, first_chunk, tps, total_rows = plan.lens, plan.first_chunk, plan.tps, plan.total_rows
        row_of, tp_of = plan.row_of, plan.tp_of
        n = len(res.chunks)
        if n == 0:
            return res
        G = self.batch_size
        half = total_rows // 2
        feat_all = torch.empty((3 * half, 160), dtype=torch.float32, device=dev)
        mask_all = torch.empty((3 * half,), dtype=torch.int32, device=dev)
        if self.acoustic:
            rows = torch.empty((2 * n, 38), dtype=torch.float32, device=dev)
            status = torch.empty((2 * n,), dtype=torch.int32, device=dev)
        res.waveforms = [None] * len(calls)                       # type: ignore[list-item]
        done_prev: Optional[torch.cuda.Event] = None
        if self.trace is not None:
            t_start = torch.cuda.Event(enable_timing=True); t_start.record(main)
            self.trace.update(start=t_start, uploads=evs, groups=[])
        k0 = 0
        while k0 < len(calls):
            k1 = self._next_group(k0, len(calls), evs, groups)
            if self.trace is not None:
                g0 = torch.cuda.Event(enable_timing=True); g0.record(main)
            xs = xs_all[k0:k1]
            last_ev = next((evs[i] for i in range(k1 - 1, k0 - 1, -1) if evs[i] is not None), None)
            if last_ev is not None:
                main.wait_event(last_ev)                          # copies complete in order on the copy stream
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
                    ys = [fe.to_float(x) for x in grp]
                    ms = [fe.mix_stereo(y) for y in ys] if want_rows == 2 else ys
                for i, y, m in zip(ids, ys, ms):
                    res.waveforms[i] = (y, m)
            bad = [k0 + j for j, x in enumerate(xs) if int(x.shape[0]) not in (1, 2)]
            if bad:
                raise ValueError(f"process_calls: call {bad[0]} has {int(xs[bad[0] - k0].shape[0])} channels")
            c0, c1 = first_chunk[k0], first_chunk[k1]
            if c1 == c0:
                k0 = k1
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
            # fbank and the acoustic features both read the resampled waveforms and nothing of each other: the
            # acoustic kernels go to a second stream, so their blocks fill the SMs the fbank kernel's last blocks
            # leave idle (each tile kernel occupies a whole SM) and the small kernels of the two chains overlap
            if self.acoustic and self.kTwoStreams:
                if self._side_stream is None:
                    self._side_stream = torch.cuda.Stream(dev)
                side = self._side_stream
                side.wait_stream(main)                            # the resampled waveforms are ready
            fe.fbank_placed(mixed_p + agent_p + client_p, sub_lens * 3,
                            sub_rows + [r + total_rows for r in sub_rows] + [r + 2 * total_rows for r in sub_rows],
                            tp_of[c0:c1] * 3, feat_all, mask_all, padding_value=self.padding_value,
                            precision=self.precision)
            if self.acoustic:
                oi = list(range(c0, c1)) + list(range(n + c0, n + c1))
                if self.kTwoStreams:
                    with torch.cuda.stream(side):
                        fe.acoustic_ptrs(agent_p + client_p, sub_lens * 2, sr_t, out=rows, status=status, out_index=oi)
                    main.wait_stream(side)
                else:
                    fe.acoustic_ptrs(agent_p + client_p, sub_lens * 2, sr_t, out=rows, status=status, out_index=oi)
            if self.trace is not None:
                g1 = torch.cuda.Event(enable_timing=True); g1.record(main)
                self.trace["groups"].append((k0, k1, g0, g1))
            if groups is None and k1 < len(calls):
                # adaptive grouping: at most two groups in flight, so the next group is formed from
                # what has ARRIVED by the time the GPU is about to need it
                done = torch.cuda.Event(); done.record(main)
                if done_prev is not None:
                    done_prev.synchronize()
                done_prev = done
            k0 = k1
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

    @staticmethod
    def _schedule(n_calls: int, sub_batch: Optional[int], any_host: bool) -> Optional[List[Tuple[int, int]]]:
        """Fixed sub-batches of calls, or None for the adaptive grouping of host calls (_next_group)."""
        if n_calls <= 0:
            return []
        if sub_batch:
            sb = int(sub_batch)
            return [(k, min(k + sb, n_calls)) for k in range(0, n_calls, sb)]
        if not any_host:
            return [(0, n_calls)]
        return None

    kFirstGroup, kMinGroup, kMaxGroup = 2, 4, 16
    kTwoStreams = True       # acoustic kernels on a second stream beside the fbank kernels

    def _next_group(self, k0: int, n_calls: int, evs, groups) -> int:
        """End of the sub-batch that starts at call k0.  Adaptive mode: wait for ``kMinGroup`` calls to
        have crossed PCIe, then take every further call that has already arrived, up to ``kMaxGroup``.
        When PCIe is the bound the groups stay small (the step ends ``upload(all) + compute(last
        group)``); when the kernels are, uploads run ahead and the groups grow to launches that fill
        the GPU and normalise most utterances inside the fbank kernel."""
        if groups is not None:
            return next(k1 for (g0, k1) in groups if g0 == k0)
        k1 = min(k0 + (self.kMinGroup if k0 > 0 else self.kFirstGroup), n_calls)   # nothing to overlap the first upload with
        ev = next((evs[i] for i in range(k1 - 1, k0 - 1, -1) if evs[i] is not None), None)
        if ev is not None:
            ev.synchronize()
        while k1 < n_calls and k1 - k0 < self.kMaxGroup and (evs[k1] is None or evs[k1].query()):
            k1 += 1
        return k1

    def _plan(self, lengths: Tuple[int, ...], sample_rate: int) -> _Plan:
        """Chunk table of every call (split_audio, audio_processor.py:115-160), the stable length sort of
        transcribe_batch (:170-181), the padded length of each batch of ``batch_size`` chunks and the
        output row of every chunk.  Cached: batches of equally long calls repeat."""
        # every attribute the layout depends on is part of the key: they are public and may be changed between calls
        key = (lengths, sample_rate, self.chunk_duration_sec, self.overlap_sec, self.target_sample_rate, self.batch_size)
        plan = self._plans.get(key)
        if plan is not None:
            return plan
        fe, sr_t = self.fe, self.target_sample_rate
        chunk_samples = int(self.chunk_duration_sec * sr_t)
        step = chunk_samples - int(self.overlap_sec * sr_t)
        chunks, lens, first_chunk, seconds = [], [], [], 0.0
        for ci, L in enumerate(lengths):
            seconds += L / float(sample_rate)
            Lr = fe.resampled_length(L, sample_rate, sr_t) if sample_rate != sr_t else L
            first_chunk.append(len(chunks))
            for idx, (s, e) in enumerate(self._chunk_table(Lr, chunk_samples, step)):
                chunks.append((ci, idx, s, e)); lens.append(e - s)
        first_chunk.append(len(chunks))
        n, G = len(chunks), self.batch_size
        order = sorted(range(n), key=lens.__getitem__)           # transcribe_batch: stable sort by length
        tps, total_rows = fe.fbank_layout([lens[i] for i in order], G) if n else ([], 0)
        row_of, tp_of, pos = [0] * n, [0] * n, 0
        for g, tp in enumerate(tps):
            for j in range(g * G, min(n, (g + 1) * G)):
                row_of[order[j]] = pos; tp_of[order[j]] = tp; pos += tp
        plan = _Plan(seconds, chunks, lens, first_chunk, order, list(tps), int(total_rows), row_of, tp_of)
        if len(self._plans) >= 64:
            self._plans.pop(next(iter(self._plans)))
        self._plans[key] = plan
        return plan

    def clear_plans(self) -> None:
        """Forget the cached layouts (bench.py: the end-to-end steps pay the planning, as ragged batches would)."""
        self._plans.clear()
        self._tables.clear()

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
