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

    def process_calls(self, calls: Sequence[torch.Tensor], sample_rate: int, acoustic_to_host: bool = True
                      ) -> FrontendResult:
        """``calls``: (C, L) float32 tensors (host -- ideally pinned -- or device)."""
        fe, sr_t = self.fe, self.target_sample_rate
        res = FrontendResult(n_calls=len(calls))
        chunk_samples = int(self.chunk_duration_sec * sr_t)
        step = chunk_samples - int(self.overlap_sec * sr_t)
        mixed_p, agent_p, client_p, lens = [], [], [], []
        for ci, call in enumerate(calls):
            res.call_seconds += call.shape[1] / float(sample_rate)
            x = call.to(self.device, non_blocking=True)
            if sample_rate != sr_t:
                if x.shape[0] == 2:
                    y, mixed = fe.resample(x, sample_rate, sr_t, want_mixed=True)
                else:
                    y = fe.resample(x, sample_rate, sr_t); mixed = y
            else:
                y = x.contiguous(); mixed = fe.mix_stereo(y) if y.shape[0] == 2 else y
            res.waveforms.append((y, mixed))
            Lr = int(y.shape[1])
            pa = y.data_ptr()
            pc = pa + 4 * Lr if y.shape[0] == 2 else pa
            pm = mixed.data_ptr()
            for idx, (s, e) in enumerate(self._chunk_table(Lr, chunk_samples, step)):
                res.chunks.append((ci, idx, s, e))
                mixed_p.append(pm + 4 * s); agent_p.append(pa + 4 * s); client_p.append(pc + 4 * s); lens.append(e - s)
        n = len(res.chunks)
        if n == 0:
            return res
        # transcribe_batch: stable sort by mixed length, slices of batch_size
        res.order = sorted(range(n), key=lens.__getitem__)
        slens = [lens[i] for i in res.order]
        for key, lst in (("stereo", mixed_p), ("agent", agent_p), ("client", client_p)):
            res.features[key] = fe.fbank_ptrs([lst[i] for i in res.order], slens, padding_value=self.padding_value,
                                              precision=self.precision, group_size=self.batch_size)
        if self.acoustic:
            rows, status = fe.acoustic_ptrs(agent_p + client_p, lens + lens, sr_t)
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
