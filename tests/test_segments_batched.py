"""``SpeechSegment.transcribe_segments_batched`` / ``SpeechBatchTranscriber.transcribe_mono``
(services/speech_segment.py:190-290, 366-427): control flow against the UNMODIFIED reference methods driven
with the same stub processor / model (container only -- /root/reference is not on the GPU box), its documented
contract everywhere, and the product path on a B200."""
import numpy as np
import pytest
import torch

from oracle import reference_calls as rc

CFG = {"batch_size": 4, "gpu_index": 0, "vad_max_segment_sec": 25.0, "vad_window_overlap_sec": 0.5,
       "vad_gap_threshold": 0.8, "vad_padding": 0.5}
LENGTHS = [4000, 900, 2500, 400, 16000, 1200, 1201, 8000, 700, 3300, 5000]


class _Logits:
    def __init__(self, logits):
        self.logits = logits


class _StubProcessor:
    """HF-processor shaped; raises for the waveform lengths in ``poison`` (a broken segment)."""
    def __init__(self, poison=()):
        self.poison = set(poison)
        self.calls = []

    def __call__(self, arrays, sampling_rate=16000, return_tensors="pt", padding=True):
        single = not isinstance(arrays, (list, tuple))
        arrays = [arrays] if single else list(arrays)
        arrays = [torch.as_tensor(np.asarray(a.cpu() if torch.is_tensor(a) else a)).reshape(-1) for a in arrays]
        self.calls.append([a.numel() for a in arrays])
        if any(a.numel() in self.poison for a in arrays):
            raise ValueError("poisoned segment")
        T = max(1 + (a.numel() - 400) // 160 for a in arrays)
        feat = torch.zeros((len(arrays), T, 8))
        for b, a in enumerate(arrays):
            t = 1 + (a.numel() - 400) // 160
            feat[b, :t] = a[:t * 160].reshape(t, 160)[:, :8]
        return {"input_features": feat}

    def batch_decode(self, ids, skip_special_tokens=True):
        return ["".join("Ab,.?c!-"[int(v)] for v in row[:12]) + " " for row in ids]


def _segments(lengths, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [{"segment_waveform": torch.rand((1, n), generator=g) - 0.5, "start": 0.25 * i, "end": 0.25 * i + n / 16000,
             "speaker": "agent"} for i, n in enumerate(lengths)]


def _make_model(log, oom_calls=()):
    def model(features):
        log.append(int(features.shape[0]))
        if len(log) in oom_calls:
            raise RuntimeError("CUDA out of memory. Tried to allocate 2.00 GiB")
        return _Logits(features[..., :7].float())
    return model


def _mine(processor, model):
    from audio_processor_b200.speech_segment import SpeechSegment
    seg = SpeechSegment(dict(CFG), processor=processor, frontend=object())
    seg.device = torch.device("cpu")
    seg.transcription_model = model
    return seg


def test_contract_order_cleaning_and_fallback(monkeypatch):
    monkeypatch.setattr(torch.cuda, "is_available", lambda: False)
    log = []
    proc = _StubProcessor(poison={2500})
    out = _mine(proc, _make_model(log)).transcribe_segments_batched(_segments(LENGTHS), 16000)
    assert [o["start"] for o in out] == [0.25 * i for i in range(len(LENGTHS))]           # input order kept
    assert all(set(o) == {"segment_waveform", "start", "end", "speaker", "text"} for o in out)
    order = sorted(range(len(LENGTHS)), key=LENGTHS.__getitem__)
    # slices of 4 in length order; the slice holding the poisoned segment is redone one by one
    assert proc.calls[0] == [LENGTHS[k] for k in order[0:4]]
    bad_slice = [LENGTHS[k] for k in order[4:8]]
    assert proc.calls[1] == bad_slice and proc.calls[2:6] == [[n] for n in bad_slice]
    assert out[2]["text"] == "" and all(o["text"] for i, o in enumerate(out) if i != 2)
    assert all(o["text"] == o["text"].strip().lower() and not set(",.?!-") & set(o["text"]) for o in out)
    assert _mine(proc, _make_model([])).transcribe_segments_batched([], 16000) == []


@pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")
@pytest.mark.parametrize("poison,oom", [((), ()), ((2500,), ()), ((), (1, 2)), ((900, 16000), (3,))])
def test_matches_reference_method(monkeypatch, poison, oom):
    import contextlib
    ref = rc.speech_segment()
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)        # the reference only halves "on CUDA"
    monkeypatch.setattr(torch.cuda, "empty_cache", lambda: None)
    monkeypatch.setattr(torch.cuda.amp, "autocast", lambda *a, **k: contextlib.nullcontext())
    monkeypatch.setattr(torch, "autocast", lambda *a, **k: contextlib.nullcontext())
    log_ref, log_mine = [], []
    p_ref, p_mine = _StubProcessor(poison), _StubProcessor(poison)
    saved = (ref.processor, ref.transcription_model, ref.device, dict(ref.config), getattr(ref, "converter", None))
    ref.processor, ref.transcription_model, ref.device = p_ref, _make_model(log_ref, oom), torch.device("cpu")
    ref.config = {**ref.config, "batch_size": CFG["batch_size"]}
    ref.converter = None
    try:
        a = ref.transcribe_segments_batched(_segments(LENGTHS, seed=5), 16000)
    finally:
        ref.processor, ref.transcription_model, ref.device, ref.config, ref.converter = saved
    b = _mine(p_mine, _make_model(log_mine, oom)).transcribe_segments_batched(_segments(LENGTHS, seed=5), 16000)
    assert p_ref.calls == p_mine.calls and log_ref == log_mine
    assert [(x["start"], x["end"], x["speaker"], x["text"]) for x in a] == [(x["start"], x["end"], x["speaker"], x["text"]) for x in b]
    assert any(x["text"] for x in b)


@pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")
def test_transcribe_mono_glue_matches_reference(monkeypatch):
    """``_merge_segments`` -> ``_extract_segments`` -> sort / batch order, reference :366-427, with a stub VAD."""
    monkeypatch.setattr(torch.cuda, "is_available", lambda: False)
    from audio_processor_b200.speech_segment import SpeechBatchTranscriber
    timeline = [(0.2, 1.1), (1.5, 2.0), (4.0, 31.5), (31.9, 32.4), (40.0, 40.3)]
    wave = torch.rand((1, 16000 * 45), generator=torch.Generator().manual_seed(3)) - 0.5
    ref = rc.speech_segment()
    merged = ref._merge_segments([type("T", (), {"start": s, "end": e})() for s, e in timeline], CFG["vad_gap_threshold"])
    saved = dict(ref.config)
    ref.config = {**ref.config, **CFG}
    try:
        ref_segments = ref._extract_segments(wave, 16000, merged, "client", CFG["vad_padding"])
    finally:
        ref.config = saved
    log = []
    proc = _StubProcessor()
    tr = SpeechBatchTranscriber(dict(CFG), vad_fn=lambda w, sr: timeline, model=_make_model(log), processor=proc, frontend=object())
    tr.segmenter.device = torch.device("cpu")
    out = tr.transcribe_mono(wave, 16000, "client")
    assert [(o["start"], o["end"], o["speaker"]) for o in out] == [(r["start"], r["end"], r["speaker"]) for r in ref_segments]
    assert all(torch.equal(o["segment_waveform"], r["segment_waveform"]) for o, r in zip(out, ref_segments))
    lens = sorted(int(r["segment_waveform"].shape[-1]) for r in ref_segments)
    assert [n for call in proc.calls for n in call] == lens and all(len(c) <= 4 for c in proc.calls)
    assert all(isinstance(o["text"], str) and o["text"] for o in out)


@pytest.mark.gpu
def test_gpu_transcribe_mono_product_path():
    """The CUDA processor inside ``transcribe_mono``: features of all slices from one launch equal one
    ``processor(...)`` call per slice; segment boundaries equal the oracle's table; texts land on their segments."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from oracle import np_fbank, np_tables
    from audio_processor_b200 import synth
    from audio_processor_b200.speech_segment import SpeechBatchTranscriber
    timeline = [(0.2, 1.1), (1.5, 2.0), (4.0, 31.5), (31.9, 32.4), (40.0, 40.3)]
    call = synth.make_call(45.0, 16000, 21, 1)
    seen = []

    def model(features):
        seen.append(features.clone())
        return _Logits(features[..., :7].float())
    tr = SpeechBatchTranscriber(dict(CFG), vad_fn=lambda w, sr: timeline, model=model)
    tr.segmenter.processor.batch_decode = _StubProcessor().batch_decode
    wave = call.to(tr.segmenter.device)
    out = tr.transcribe_mono(wave, 16000, "agent")
    merged = np_tables.merge_intervals(timeline, CFG["vad_gap_threshold"])
    rows = np_tables.segment_table(wave.shape[1], 16000, merged, CFG["vad_padding"], CFG["vad_max_segment_sec"],
                                   CFG["vad_window_overlap_sec"])
    assert [(o["start"], o["end"]) for o in out] == [(r[2], r[3]) for r in rows]
    assert [int(o["segment_waveform"].shape[1]) for o in out] == [r[1] - r[0] for r in rows]
    order = sorted(range(len(out)), key=lambda k: out[k]["segment_waveform"].shape[1])
    host = call.numpy()[0]
    for g, feats in enumerate(seen):
        members = order[4 * g:4 * g + 4]
        ref_f, _ = np_fbank.featurise([host[rows[k][0]:rows[k][1]] for k in members])
        got = feats.cpu().numpy()
        assert got.shape == ref_f.shape
        ok = ~np.isnan(ref_f)
        assert np.max(np.abs(got[ok] - ref_f[ok])) <= 1e-3
    assert len(seen) == (len(out) + 3) // 4 and all(o["text"] for o in out)
