"""Container-only: the oracle and the CPU-baseline port against the UNMODIFIED reference
imported from /root/reference (skipped on the GPU box where the mount does not exist)."""
import numpy as np
import pytest
import torch

from oracle import libref, np_acoustic, np_fbank, np_resample, reference_calls as rc

pytestmark = pytest.mark.skipif(not rc.available(), reason="/root/reference not mounted")


def test_split_audio_and_processor_match_reference():
    from audio_processor_b200 import synth
    call = synth.make_call(27.0, 8000, 5, 2)
    ap = rc.audio_processor()
    ref_chunks, ref_a, ref_c = ap.split_audio(call.clone(), 8000, "x.wav")
    chunks, a, c = libref.split_audio(call.clone(), 8000, "x.wav")
    assert torch.equal(a, ref_a) and torch.equal(c, ref_c) and len(chunks) == len(ref_chunks) == 2
    for x, y in zip(chunks, ref_chunks):
        assert list(x.keys()) == list(y.keys())
        assert (x["chunk_idx"], x["start_time"], x["end_time"]) == (y["chunk_idx"], y["start_time"], y["end_time"])
        assert torch.equal(x["stereo_waveform"], y["stereo_waveform"])
    # oracle resample is bit-exact with the reference's own split_audio
    assert np.array_equal(np_resample.resample(call.numpy(), 8000, 16000)[0:1], ref_a.numpy())
    # oracle fbank == reference processor on the reference's chunk
    w = ref_chunks[1]["agent_waveform"].squeeze().numpy()
    r = rc.processor()([w], sampling_rate=16000, return_tensors="pt", padding=True)
    f, m = np_fbank.featurise([w])
    assert np.array_equal(m, r["attention_mask"].numpy())
    assert np.max(np.abs(f - r["input_features"].numpy())) <= 5e-5


def test_acoustic_port_and_oracle_match_reference():
    from audio_processor_b200 import synth
    w = synth.speech_like(40000, 16000, 21)
    ref = rc.acoustic_analyzer()._extract_acoustic_features(w[None, :], 16000)
    port = libref.acoustic_features(w[None, :], 16000)
    assert list(port.keys()) == list(ref.keys())
    assert np.allclose(list(port.values()), list(ref.values()), rtol=1e-6, atol=1e-6)
    orc = np_acoustic.acoustic_features(w.numpy())
    a, b = np.asarray(list(orc.values())), np.asarray(list(ref.values()))
    assert np.all(np.abs(a - b) <= np.maximum(1e-3, 1e-4 * np.abs(b)))
