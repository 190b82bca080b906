"""Batch assembly for the CTC model: ``transcribe_batch`` bucketing + OOM halving,
``_process_single_batch``, ``_run_inference`` (services/audio_processor.py:162-309).

CPU: the control flow against the UNMODIFIED reference methods driven with the same stub model (container
only -- /root/reference is not on the GPU box) and against its documented contract (everywhere).
GPU: features produced ahead in one launch per stream == one ``processor(...)`` call per slice."""
import types

import numpy as np
import pytest
import torch

from oracle import reference_calls as rc

CFG = {"chunk_duration_sec": 25, "overlap_sec": 1, "target_sample_rate": 16000, "batch_size": 8, "gpu_index": 0}


def _chunks(lengths, seed=0):
    g = torch.Generator().manual_seed(seed)
    out = []
    for i, n in enumerate(lengths):
        w = torch.rand((3, n), generator=g) - 0.5
        out.append({"file_name": f"f{i}", "stereo_waveform": w[0:1], "agent_waveform": w[1:2], "client_waveform": w[2:3],
                    "chunk_idx": i, "start_time": 0.0, "end_time": n / 16000})
    return out


class _Logits:
    def __init__(self, logits):
        self.logits = logits


def _stub_model(input_features):
    # deterministic "CTC head": 7 classes from the first 7 feature columns
    return _Logits(input_features[..., :7].float())


class _StubProcessor:
    """HF-processor shaped: callable -> {"input_features"}, and batch_decode(ids) -> strings."""
    def __call__(self, arrays, sampling_rate=16000, return_tensors="pt", padding=True):
        assert sampling_rate == 16000
        arrays = [torch.as_tensor(np.asarray(a)) for a in arrays]
        T = max(1 + (a.numel() - 400) // 160 for a in arrays)
        feat = torch.zeros((len(arrays), T, 8))
        for b, a in enumerate(arrays):
            t = 1 + (a.numel() - 400) // 160
            feat[b, :t] = a[:t * 160].reshape(t, 160)[:, :8]
        return {"input_features": feat}

    def batch_decode(self, ids):
        return ["".join("Ab,.?c!-"[int(v)] for v in row[:12]) for row in ids]


def _mine(processor=None, model=_stub_model, cfg=CFG):
    from audio_processor_b200.audio_processor import AudioProcessor
    return AudioProcessor(dict(cfg), processor=processor or _StubProcessor(), model=model, frontend=object())


LENGTHS = [4000, 900, 2500, 400, 16000, 1200, 1201, 8000, 700, 3300, 5000, 450, 9100, 640, 2222, 12000, 560, 7777, 1000]


# ------------------------------------------------------------------ contract, no reference needed
def test_transcribe_batch_buckets_and_results():
    ap = _mine()
    seen = []
    orig = ap._process_single_batch
    ap._process_single_batch = lambda batch, feats=None: (seen.append([c["chunk_idx"] for c in batch]), orig(batch, feats))[1]
    chunks = _chunks(LENGTHS)
    res = ap.transcribe_batch(chunks)
    order = sorted(range(len(LENGTHS)), key=LENGTHS.__getitem__)
    assert [c["chunk_idx"] for c in res] == order                      # stable sort by stereo length
    assert seen == [order[0:8], order[8:16], order[16:19]]
    assert all(c["error"] == "" and isinstance(c["transcription_chunk"], str) for c in res)
    assert all(set("ABCDEFGHIJKLMNOPQRSTUVWXYZ,.?!-") & set(c["agent_transcription"]) == set() for c in res)   # cleaned
    assert ap.transcribe_batch([]) == []
    ap2 = _mine(cfg={**CFG, "enable_length_bucketing": False})
    assert [c["chunk_idx"] for c in ap2.transcribe_batch(_chunks(LENGTHS))] == list(range(len(LENGTHS)))


def test_transcribe_batch_oom_halving_and_error_marking():
    ap = _mine()
    seen = []

    def flaky(batch, feats=None):
        seen.append(len(batch))
        if len(batch) > 2 and len(seen) < 4:
            raise RuntimeError("CUDA out of memory. Tried to allocate 20.00 GiB")
        if any(c["chunk_idx"] == 4 for c in batch):
            raise ValueError("boom")
        for c in batch:
            c.update({"transcription_chunk": "x", "agent_transcription": "y", "client_transcription": "z", "error": ""})
        return batch
    ap._process_single_batch = flaky
    res = ap.transcribe_batch(_chunks(LENGTHS))
    assert seen[:3] == [8, 4, 2] and set(seen[3:]) <= {2, 1}             # 8 -> 4 -> 2, then stays at 2
    assert len(res) == len(LENGTHS)
    bad = [c for c in res if c["error"]]
    # chunk 4 is the longest: after the sort it sits alone in the last slice of 2
    assert [c["chunk_idx"] for c in bad] == [4] and bad[0]["error"] == "boom" and bad[0]["transcription_chunk"] == ""


def test_run_inference_swallows_and_reports():
    def broken(input_features):
        raise RuntimeError("CUDA out of memory in the model")
    ap = _mine(model=broken)
    res = ap.transcribe_batch(_chunks(LENGTHS[:5]))
    # _run_inference turns ANY exception into empty results (reference :305-309), so no halving happens
    assert [c["error"] for c in res] == ["CUDA out of memory in the model"] * 5
    assert all(c["agent_transcription"] == "" for c in res)


# ------------------------------------------------------------------ against the reference's own methods
@pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")
def test_control_flow_matches_reference(monkeypatch):
    ref = rc.audio_processor({**rc.DEFAULT_CONFIG, "batch_size": 8})
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)       # the reference only halves "on CUDA"
    monkeypatch.setattr(torch.cuda, "empty_cache", lambda: None)
    logs = {"ref": [], "mine": []}

    def make(tag):
        calls = []

        def fn(batch, feats=None):
            calls.append(1)
            logs[tag].append([c["chunk_idx"] for c in batch])
            if len(batch) > 4 and len(calls) in (1, 2, 5):
                raise RuntimeError("CUDA error: out of memory")
            if len(calls) == 4:
                raise KeyError("bad batch")
            for c in batch:
                c.update({"transcription_chunk": f"t{c['chunk_idx']}", "agent_transcription": "a", "client_transcription": "c", "error": ""})
            return batch
        return fn
    monkeypatch.setattr(ref, "_process_single_batch", make("ref"))
    monkeypatch.setattr(ref, "_cleanup_memory", lambda: None)
    ap = _mine()
    ap._process_single_batch = make("mine")
    r_ref = ref.transcribe_batch(_chunks(LENGTHS))
    r_mine = ap.transcribe_batch(_chunks(LENGTHS))
    assert logs["ref"] == logs["mine"]
    keys = ("chunk_idx", "transcription_chunk", "agent_transcription", "client_transcription", "error")
    assert [[c[k] for k in keys] for c in r_ref] == [[c[k] for k in keys] for c in r_mine]


@pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")
def test_single_batch_and_inference_match_reference():
    ref = rc.audio_processor({**rc.DEFAULT_CONFIG, "batch_size": 8})
    proc = _StubProcessor()
    saved = (getattr(ref, "processor", None), getattr(ref, "model", None), ref.device)
    ref.processor, ref.model, ref.device = proc, _stub_model, torch.device("cpu")
    try:
        a = ref._process_single_batch(_chunks(LENGTHS[:6], seed=3))
    finally:
        ref.processor, ref.model, ref.device = saved
    b = _mine(processor=proc)._process_single_batch(_chunks(LENGTHS[:6], seed=3))
    keys = ("transcription_chunk", "agent_transcription", "client_transcription", "error")
    assert [[c[k] for k in keys] for c in a] == [[c[k] for k in keys] for c in b]
    assert any(c["transcription_chunk"] for c in b)


# ------------------------------------------------------------------ product path (B200)
@pytest.mark.gpu
def test_gpu_features_ahead_equal_per_slice_processor_calls():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from audio_processor_b200 import AudioProcessor, synth
    ap = AudioProcessor({**CFG, "batch_size": 4})
    decode = types.SimpleNamespace(batch_decode=_StubProcessor().batch_decode)
    ap.processor.batch_decode = decode.batch_decode
    ap.model = _stub_model
    calls, sr = synth.cfg2_calls(2, 130.0, distinct=2)
    chunks = []
    for i, c in enumerate(calls):
        chunks += ap.split_audio(c, sr, f"call{i}.wav")[0]
    assert len(chunks) == 12
    # features produced ahead (one launch per stream for all slices) ...
    ahead = ap._featurise_ahead(sorted(chunks, key=lambda c: c["stereo_waveform"].shape[1]), 4)
    # ... equal one processor(...) call per slice, which is what the reference does
    ordered = sorted(chunks, key=lambda c: c["stereo_waveform"].shape[1])
    for g in range(3):
        per = ap.featurise_batch(ordered[4 * g:4 * g + 4])
        for k in ("stereo", "agent", "client"):
            assert torch.equal(ahead[g][k], per[k])
    res = ap.transcribe_batch([dict(c) for c in chunks])
    res2 = []
    for g in range(3):
        res2 += ap._process_single_batch([dict(c) for c in ordered[4 * g:4 * g + 4]])
    keys = ("chunk_idx", "file_name", "transcription_chunk", "agent_transcription", "client_transcription", "error")
    assert [[c[k] for k in keys] for c in res] == [[c[k] for k in keys] for c in res2]
    assert all(c["error"] == "" for c in res) and len({c["transcription_chunk"] for c in res}) > 1
